// Device group: 1/2/4/8 GPUs behind ONE call of ONE process -- the reference's placement switch re-targeted.
//
// The reference picks NUMA placement when the program starts: shl__init(nthreads, partitioned)
// (apps/output_cpp/src/common_main.h:95-99), scripts/run.sh:86-92 (-r replicate, -p partition).  Here the same switch
// selects how many B200s the unchanged drivers use (GM_CUDA_DEVICES=1|2|4|8, host/shl.cc): the CSR is replicated on every
// device (Shoal's read-only arrays), the rows are partitioned into destination ranges (its indexed arrays), and every
// generated procedure runs as one host thread per device over the sharded entry points of this library:
//   * one worker thread per device, alive as long as the group (each keeps its device current, its communicator rank,
//     its graph replica and its retired state block);
//   * the communicator is NCCL with several ranks in one process; the peers' state blocks are reached with plain
//     device pointers after cudaDeviceEnablePeerAccess (no CUDA IPC handles to export, open and close);
//   * copy-in: every device copies its 1/W slice of each CSR array over its own PCIe link, the slices are allgathered
//     over NVLink (gmc_graph_upload_sharded); copy-out: every device copies its 1/W slice of the result into the
//     caller's ONE host array.
#include <condition_variable>
#include <functional>
#include <mutex>
#include <string>
#include <thread>
#include <vector>
#include "gmc_common.cuh"
#include "gmc_comm.h"

using namespace gmc;

struct gmc_group {
    int W = 0;
    std::vector<int> dev;
    std::vector<std::thread> th;
    std::vector<gmc_comm*> comm;
    std::vector<gmc_graph*> graph;
    std::vector<int> rc;
    std::vector<std::string> err;
    // job hand-off
    std::mutex mu;
    std::condition_variable cv_job, cv_done;
    std::function<int(int)> job;
    unsigned long long job_no = 0;
    int pending = 0;
    bool stop = false;
    int32_t N = 0, M = 0;
    bool has_reverse = false;
};

namespace {

void worker(gmc_group* G, int r) {
    int rc = gmc_init(G->dev[r]);
    unsigned long long seen = 0;
    for (;;) {
        std::function<int(int)> job;
        {
            std::unique_lock<std::mutex> lk(G->mu);
            G->cv_job.wait(lk, [&] { return G->stop || G->job_no != seen; });
            if (G->stop) return;
            seen = G->job_no;
            job = G->job;
        }
        int jrc = rc != GMC_OK ? rc : job(r);
        std::string e = jrc != GMC_OK ? std::string(gmc_last_error()) : std::string();
        {
            std::lock_guard<std::mutex> lk(G->mu);
            G->rc[r] = jrc;
            G->err[r] = e;
            if (--G->pending == 0) G->cv_done.notify_all();
        }
    }
}

// runs fn(rank) on every worker (its device current), returns the first failure (error text copied to the caller's thread)
int run_all(gmc_group* G, const std::function<int(int)>& fn) {
    {
        std::unique_lock<std::mutex> lk(G->mu);
        G->job = fn;
        G->pending = G->W;
        G->job_no++;
    }
    G->cv_job.notify_all();
    {
        std::unique_lock<std::mutex> lk(G->mu);
        G->cv_done.wait(lk, [&] { return G->pending == 0; });
    }
    for (int r = 0; r < G->W; r++)
        if (G->rc[r] != GMC_OK) { set_error("device group, rank %d (device %d): %s", r, G->dev[r], G->err[r].c_str()); return G->rc[r]; }
    return GMC_OK;
}

// [lo, hi) = rank r's slice of n elements
void slice_of(int64_t n, int W, int r, int64_t* lo, int64_t* hi) {
    const int64_t per = (n + W - 1) / W;
    *lo = std::min<int64_t>(n, per * r);
    *hi = std::min<int64_t>(n, per * (r + 1));
}

}  // namespace

extern "C" {

int gmc_group_create(gmc_group** out, int ndev, const int* devices) {
    GMC_REQUIRE(out && ndev >= 1 && ndev <= 8, GMC_ERR_ARG, "gmc_group_create: 1..8 devices (one NVSwitch domain), got %d", ndev);
    int have = 0;
    GMC_CUDA(cudaGetDeviceCount(&have));
    GMC_REQUIRE(have >= ndev, GMC_ERR_ARG, "gmc_group_create: %d devices asked for, %d visible", ndev, have);
    gmc_group* G = new gmc_group();
    G->W = ndev;
    G->dev.resize(ndev); G->comm.assign(ndev, nullptr); G->graph.assign(ndev, nullptr); G->rc.assign(ndev, GMC_OK); G->err.resize(ndev);
    for (int r = 0; r < ndev; r++) G->dev[r] = devices ? devices[r] : r;
    for (int r = 0; r < ndev; r++) G->th.emplace_back(worker, G, r);
    unsigned char id[GMC_COMM_ID_BYTES] = {0};
    int rc = GMC_OK;
    if (ndev > 1) rc = gmc_comm_unique_id(id);
    if (rc == GMC_OK)
        rc = run_all(G, [&](int r) -> int {
            // every peer reachable with plain loads / stores (NVLink): enabled once per pair and direction
            for (int q = 0; q < G->W; q++) {
                if (q == r) continue;
                int can = 0;
                GMC_CUDA(cudaDeviceCanAccessPeer(&can, G->dev[r], G->dev[q]));
                GMC_REQUIRE(can, GMC_ERR_STATE, "device %d cannot access device %d", G->dev[r], G->dev[q]);
                cudaError_t e = cudaDeviceEnablePeerAccess(G->dev[q], 0);
                if (e == cudaErrorPeerAccessAlreadyEnabled) cudaGetLastError();
                else GMC_CUDA(e);
            }
            GMC_TRY(gmc_comm_create(&G->comm[r], G->W, r, id));          // ncclCommInitRank from W threads of one process
            G->comm[r]->local = true;
            G->comm[r]->device = G->dev[r];
            return (int)GMC_OK;
        });
    if (rc != GMC_OK) { gmc_group_destroy(G); return rc; }
    *out = G;
    return GMC_OK;
}

int gmc_group_graph_free(gmc_group* G) {
    if (!G) return GMC_OK;
    return run_all(G, [&](int r) -> int {
        if (G->graph[r]) {
            if (G->comm[r]) gmc_comm_barrier(G->comm[r], nullptr);     // nobody frees a state block a peer may still store into
            gmc_graph_destroy(G->graph[r]);
            G->graph[r] = nullptr;
        }
        return (int)GMC_OK;
    });
}

int gmc_group_destroy(gmc_group* G) {
    if (!G) return GMC_OK;
    gmc_group_graph_free(G);
    run_all(G, [&](int r) -> int {
        if (G->comm[r]) { gmc_comm_destroy(G->comm[r]); G->comm[r] = nullptr; }
        return (int)GMC_OK;
    });
    {
        std::lock_guard<std::mutex> lk(G->mu);
        G->stop = true;
    }
    G->cv_job.notify_all();
    for (auto& t : G->th) t.join();
    delete G;
    return GMC_OK;
}

int gmc_group_size(const gmc_group* G, int* ndev) {
    GMC_REQUIRE(G && ndev, GMC_ERR_ARG, "gmc_group_size: NULL");
    *ndev = G->W;
    return GMC_OK;
}

int gmc_group_graph_upload(gmc_group* G, int32_t n, int32_t m, const int32_t* begin, const int32_t* node_idx, const int32_t* r_begin,
                           const int32_t* r_node_idx) {
    GMC_REQUIRE(G && begin && n >= 0 && m >= 0, GMC_ERR_ARG, "gmc_group_graph_upload: bad args");
    GMC_TRY(gmc_group_graph_free(G));
    G->N = n; G->M = m; G->has_reverse = r_begin != nullptr;
    return run_all(G, [&](int r) -> int {
        GMC_TRY(gmc_graph_create(&G->graph[r], n, m));
        return gmc_graph_upload_sharded(G->graph[r], G->comm[r], begin, node_idx, r_begin, r_node_idx, nullptr);
    });
}

int gmc_group_pagerank(gmc_group* G, double e, double d, int32_t max_iter, double* rank_host, gmc_pagerank_info* info) {
    GMC_REQUIRE(G && G->graph[0] && (rank_host || G->N == 0), GMC_ERR_ARG, "gmc_group_pagerank: no graph / NULL");
    if (G->N == 0) { if (info) memset(info, 0, sizeof(*info)); return GMC_OK; }
    std::vector<gmc_pagerank_info> infos((size_t)G->W);
    GMC_TRY(run_all(G, [&](int r) -> int {
        double* out = nullptr;
        GMC_TRY(gmc_malloc((void**)&out, sizeof(double) * (size_t)G->N));
        int rc = gmc_pagerank_sharded_dev(G->graph[r], G->W > 1 ? G->comm[r] : nullptr, e, d, max_iter, out, &infos[r], nullptr);
        if (rc == GMC_OK) {        // the result is complete on every device: each copies its slice into the caller's array
            int64_t lo, hi;
            slice_of(G->N, G->W, r, &lo, &hi);
            if (hi > lo) rc = gmc_d2h(rank_host + lo, out + lo, sizeof(double) * (size_t)(hi - lo), nullptr);
        }
        cudaFree(out);
        return rc;
    }));
    if (info) {
        *info = infos[0];
        for (int r = 1; r < G->W; r++) { info->launches += infos[r].launches; if (infos[r].kernel_ms > info->kernel_ms) info->kernel_ms = infos[r].kernel_ms; }
    }
    return GMC_OK;
}

int gmc_group_hop_dist(gmc_group* G, int32_t root, int32_t* dist_host, gmc_bfs_info* info) {
    GMC_REQUIRE(G && G->graph[0] && (dist_host || G->N == 0), GMC_ERR_ARG, "gmc_group_hop_dist: no graph / NULL");
    if (G->N == 0) { if (info) memset(info, 0, sizeof(*info)); return GMC_OK; }
    std::vector<gmc_bfs_info> infos((size_t)G->W);
    GMC_TRY(run_all(G, [&](int r) -> int {
        int32_t* out = nullptr;
        GMC_TRY(gmc_malloc((void**)&out, sizeof(int32_t) * (size_t)G->N));
        int rc = gmc_hop_dist_sharded_dev(G->graph[r], G->W > 1 ? G->comm[r] : nullptr, root, out, &infos[r], nullptr);
        if (rc == GMC_OK) {
            int64_t lo, hi;
            slice_of(G->N, G->W, r, &lo, &hi);
            if (hi > lo) rc = gmc_d2h(dist_host + lo, out + lo, sizeof(int32_t) * (size_t)(hi - lo), nullptr);
        }
        cudaFree(out);
        return rc;
    }));
    if (info) {
        *info = infos[0];
        for (int r = 1; r < G->W; r++) { info->launches += infos[r].launches; if (infos[r].kernel_ms > info->kernel_ms) info->kernel_ms = infos[r].kernel_ms; }
    }
    return GMC_OK;
}

int gmc_group_exchange_path(gmc_group* G, int* path) {
    GMC_REQUIRE(G && G->graph[0] && path, GMC_ERR_ARG, "gmc_group_exchange_path: no graph / NULL");
    return gmc_pagerank_exchange_path(G->graph[0], path);
}

}  // extern "C"
