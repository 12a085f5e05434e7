// gmc_comm: one NCCL communicator per process (one process per GPU), used by the sharded PageRank (rank
// allgather per iteration) and the sharded BFS (frontier bitmap exchange per level) -- SURVEY.md 8e.
//
// NCCL is bound at run time (dlopen) so that libgm_cuda.so has no link-time dependency on one particular
// libnccl: inside a Python process the copy torch already loaded is reused, the reference's C++ drivers
// pick up the system libnccl.so.2.  Only the types come from <nccl.h>.
//
// Also here: gmc_shard_cuts, the host-side balanced partition of a prefix-sum array into `world`
// contiguous ranges (used for destination-range ownership; pure host code, tested on CPU).
#include "gmc_common.cuh"
#include "gmc_comm.h"
#include <dlfcn.h>
#include <stdlib.h>
#include <mutex>
#include <vector>

using namespace gmc;

namespace {

struct NcclApi {
    void* handle = nullptr;
    ncclResult_t (*GetUniqueId)(ncclUniqueId*) = nullptr;
    ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int) = nullptr;
    ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
    const char* (*GetErrorString)(ncclResult_t) = nullptr;
    ncclResult_t (*AllGather)(const void*, void*, size_t, ncclDataType_t, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*AllReduce)(const void*, void*, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*Broadcast)(const void*, void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*Send)(const void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*Recv)(void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*GroupStart)() = nullptr;
    ncclResult_t (*GroupEnd)() = nullptr;
    ncclResult_t (*GetVersion)(int*) = nullptr;
};

NcclApi g_api;
std::mutex g_api_mu;

int load_nccl() {
    std::lock_guard<std::mutex> lk(g_api_mu);
    if (g_api.handle) return GMC_OK;
    void* h = nullptr;
    if (const char* p = getenv("GMC_NCCL_LIB")) h = dlopen(p, RTLD_NOW | RTLD_GLOBAL);
    if (!h) h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_NOLOAD);     // already in the process (torch's copy)
    if (!h) h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
    if (!h) h = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
    GMC_REQUIRE(h, GMC_ERR_STATE, "gmc_comm: cannot load libnccl.so.2 (%s); set GMC_NCCL_LIB", dlerror());
#define BIND(field, sym)                                                               \
    do {                                                                               \
        *(void**)(&g_api.field) = dlsym(h, sym);                                       \
        GMC_REQUIRE(g_api.field, GMC_ERR_STATE, "gmc_comm: %s missing in libnccl", sym); \
    } while (0)
    BIND(GetUniqueId, "ncclGetUniqueId");
    BIND(CommInitRank, "ncclCommInitRank");
    BIND(CommDestroy, "ncclCommDestroy");
    BIND(GetErrorString, "ncclGetErrorString");
    BIND(AllGather, "ncclAllGather");
    BIND(AllReduce, "ncclAllReduce");
    BIND(Broadcast, "ncclBroadcast");
    BIND(Send, "ncclSend");
    BIND(Recv, "ncclRecv");
    BIND(GroupStart, "ncclGroupStart");
    BIND(GroupEnd, "ncclGroupEnd");
    BIND(GetVersion, "ncclGetVersion");
#undef BIND
    g_api.handle = h;
    return GMC_OK;
}

}  // namespace

#define GMC_NCCL(call)                                                                              \
    do {                                                                                            \
        ncclResult_t _r = (call);                                                                   \
        if (_r != ncclSuccess) {                                                                    \
            gmc::set_error("%s:%d: %s -> %s", __FILE__, __LINE__, #call, g_api.GetErrorString(_r)); \
            return GMC_ERR_CUDA;                                                                    \
        }                                                                                           \
    } while (0)

// ---- internal API used by the kernels' host code (gmc_comm.h) ----------------------------------------
int gmc_comm_allgather(gmc_comm* c, const void* send, void* recv, size_t bytes_per_rank, cudaStream_t st) {
    if (!c || c->world == 1) {
        if (send != recv) GMC_CUDA(cudaMemcpyAsync(recv, send, bytes_per_rank, cudaMemcpyDeviceToDevice, st));
        return GMC_OK;
    }
    GMC_NCCL(g_api.AllGather(send, recv, bytes_per_rank, ncclUint8, c->nccl, st));
    return GMC_OK;
}

int gmc_comm_allreduce_sum_f64(gmc_comm* c, const double* send, double* recv, size_t count, cudaStream_t st) {
    if (!c || c->world == 1) {
        if (send != recv) GMC_CUDA(cudaMemcpyAsync(recv, send, 8 * count, cudaMemcpyDeviceToDevice, st));
        return GMC_OK;
    }
    GMC_NCCL(g_api.AllReduce(send, recv, count, ncclFloat64, ncclSum, c->nccl, st));
    return GMC_OK;
}

int gmc_comm_allreduce_sum_i64(gmc_comm* c, const long long* send, long long* recv, size_t count, cudaStream_t st) {
    if (!c || c->world == 1) {
        if (send != recv) GMC_CUDA(cudaMemcpyAsync(recv, send, 8 * count, cudaMemcpyDeviceToDevice, st));
        return GMC_OK;
    }
    GMC_NCCL(g_api.AllReduce(send, recv, count, ncclInt64, ncclSum, c->nccl, st));
    return GMC_OK;
}

// slice r of `send` (bytes_per_rank each) goes to rank r, which stores the slice it gets from rank q at recv + q * bytes_per_rank;
// the own slice is not moved (recv's own slot is left untouched).  One grouped launch of 2 * (world - 1) point-to-point transfers.
int gmc_comm_alltoall(gmc_comm* c, const void* send, void* recv, size_t bytes_per_rank, cudaStream_t st) {
    if (!c || c->world == 1) return GMC_OK;
    GMC_NCCL(g_api.GroupStart());
    for (int r = 0; r < c->world; r++) {
        if (r == c->rank) continue;
        ncclResult_t res = g_api.Send((const char*)send + (size_t)r * bytes_per_rank, bytes_per_rank, ncclUint8, r, c->nccl, st);
        if (res == ncclSuccess) res = g_api.Recv((char*)recv + (size_t)r * bytes_per_rank, bytes_per_rank, ncclUint8, r, c->nccl, st);
        if (res != ncclSuccess) {
            g_api.GroupEnd();
            set_error("ncclSend/ncclRecv: %s", g_api.GetErrorString(res));
            return GMC_ERR_CUDA;
        }
    }
    GMC_NCCL(g_api.GroupEnd());
    return GMC_OK;
}

// every rank r owns `nseg` segments of one replicated buffer; after the call every rank holds all of them.
// seg_off/seg_len: [world][nseg] in elements of `elem_bytes`.  One grouped launch of world*nseg broadcasts.
int gmc_comm_allgather_segments(gmc_comm* c, void* buf, size_t elem_bytes, const int64_t* seg_off, const int64_t* seg_len,
                                int nseg, cudaStream_t st) {
    if (!c || c->world == 1) return GMC_OK;
    GMC_NCCL(g_api.GroupStart());
    for (int r = 0; r < c->world; r++)
        for (int s = 0; s < nseg; s++) {
            const int64_t len = seg_len[r * nseg + s];
            if (len <= 0) continue;
            char* p = (char*)buf + (size_t)seg_off[r * nseg + s] * elem_bytes;
            ncclResult_t res = g_api.Broadcast(p, p, (size_t)len * elem_bytes, ncclUint8, r, c->nccl, st);
            if (res != ncclSuccess) {
                g_api.GroupEnd();
                set_error("ncclBroadcast: %s", g_api.GetErrorString(res));
                return GMC_ERR_CUDA;
            }
        }
    GMC_NCCL(g_api.GroupEnd());
    return GMC_OK;
}

// Peer mapping of one device allocation per rank (cudaMalloc'ed, `local_base` = its base address): the CUDA IPC
// handles travel through the communicator itself (in-place allgather of 64-byte handles), every rank opens the
// other ranks' handles.  peer_bases[r] = rank r's block as seen from this process (own rank: local_base).
// Returns GMC_ERR_STATE when IPC is not available (callers fall back to NCCL for the data path).
int gmc_comm_ipc_open_all(gmc_comm* c, void* local_base, void** peer_bases, cudaStream_t st) {
    GMC_REQUIRE(c && c->world > 1 && local_base && peer_bases, GMC_ERR_ARG, "gmc_comm_ipc_open_all: bad args");
    static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle size");
    if (getenv("GMC_NO_P2P")) { set_error("peer mapping disabled by GMC_NO_P2P"); return GMC_ERR_STATE; }
    if (c->local) {
        // all ranks in one process (gmc_group, peer access already enabled): the pointers themselves are exchanged
        unsigned long long mine_p = (unsigned long long)(uintptr_t)local_base, *stage_p = nullptr;
        std::vector<unsigned long long> all_p((size_t)c->world);
        GMC_TRY(gmc_malloc((void**)&stage_p, 8 * (size_t)c->world));
        int rc2 = GMC_OK;
        if (cudaMemcpyAsync(stage_p + c->rank, &mine_p, 8, cudaMemcpyHostToDevice, st) != cudaSuccess) rc2 = GMC_ERR_CUDA;
        if (rc2 == GMC_OK) rc2 = gmc_comm_allgather(c, stage_p + c->rank, stage_p, 8, st);
        if (rc2 == GMC_OK && (cudaMemcpyAsync(all_p.data(), stage_p, 8 * (size_t)c->world, cudaMemcpyDeviceToHost, st) != cudaSuccess ||
                              cudaStreamSynchronize(st) != cudaSuccess)) rc2 = GMC_ERR_CUDA;
        cudaFree(stage_p);
        if (rc2 != GMC_OK) { set_error("gmc_comm_ipc_open_all: pointer exchange failed"); return rc2; }
        for (int r = 0; r < c->world; r++) peer_bases[r] = (void*)(uintptr_t)all_p[(size_t)r];
        return GMC_OK;
    }
    // No rank may leave before the collectives below (the others would hang in them): a local failure is carried as ok = 0
    // through the handle exchange and decided by all ranks together in the agreement allreduce
    int ok = 1;
    cudaIpcMemHandle_t mine;
    memset(&mine, 0, sizeof(mine));
    if (cudaIpcGetMemHandle(&mine, local_base) != cudaSuccess) {
        set_error("cudaIpcGetMemHandle: %s", cudaGetErrorString(cudaGetLastError()));
        ok = 0;
    }
    std::vector<cudaIpcMemHandle_t> all((size_t)c->world);
    char* stage = nullptr;
    GMC_TRY(gmc_malloc((void**)&stage, 64 * (size_t)c->world));
    int rc = GMC_OK;
    if (cudaMemcpyAsync(stage + 64 * (size_t)c->rank, &mine, 64, cudaMemcpyHostToDevice, st) != cudaSuccess) rc = GMC_ERR_CUDA;
    if (rc == GMC_OK) rc = gmc_comm_allgather(c, stage + 64 * (size_t)c->rank, stage, 64, st);
    if (rc == GMC_OK && (cudaMemcpyAsync(all.data(), stage, 64 * (size_t)c->world, cudaMemcpyDeviceToHost, st) != cudaSuccess ||
                         cudaStreamSynchronize(st) != cudaSuccess)) rc = GMC_ERR_CUDA;
    cudaFree(stage);
    if (rc != GMC_OK) { set_error("gmc_comm_ipc_open_all: handle exchange failed"); return rc; }
    for (int r = 0; r < c->world; r++) {
        peer_bases[r] = nullptr;
        if (r == c->rank) { peer_bases[r] = local_base; continue; }
        if (ok && cudaIpcOpenMemHandle(&peer_bases[r], all[r], cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) {
            set_error("cudaIpcOpenMemHandle(rank %d): %s", r, cudaGetErrorString(cudaGetLastError()));
            peer_bases[r] = nullptr;
            ok = 0;
        }
    }
    // every rank must take the same path: agree on success
    long long *flag = nullptr, h = ok;
    GMC_TRY(gmc_malloc((void**)&flag, 16));
    cudaMemcpyAsync(flag, &h, 8, cudaMemcpyHostToDevice, st);
    rc = gmc_comm_allreduce_sum_i64(c, flag, flag + 1, 1, st);
    if (rc == GMC_OK && (cudaMemcpyAsync(&h, flag + 1, 8, cudaMemcpyDeviceToHost, st) != cudaSuccess || cudaStreamSynchronize(st) != cudaSuccess)) rc = GMC_ERR_CUDA;
    cudaFree(flag);
    if (rc == GMC_OK && h != c->world) rc = GMC_ERR_STATE;
    if (rc != GMC_OK) {
        for (int r = 0; r < c->world; r++)
            if (r != c->rank && peer_bases[r]) { cudaIpcCloseMemHandle(peer_bases[r]); peer_bases[r] = nullptr; }
    }
    return rc;
}

int gmc_comm_stream_barrier(gmc_comm* c, cudaStream_t st) {
    if (!c || c->world == 1) return GMC_OK;
    if (!c->scratch) GMC_TRY(gmc_malloc((void**)&c->scratch, 64));
    GMC_NCCL(g_api.AllReduce(c->scratch, c->scratch, 1, ncclFloat64, ncclSum, c->nccl, st));
    return GMC_OK;
}

void gmc_comm_ipc_close_all(int world, int rank, void** peer_bases, bool local) {
    if (local) { for (int r = 0; r < world; r++) if (r != rank) peer_bases[r] = nullptr; return; }      // plain pointers: nothing to unmap
    for (int r = 0; r < world; r++)
        if (r != rank && peer_bases[r]) { cudaIpcCloseMemHandle(peer_bases[r]); peer_bases[r] = nullptr; }
}

extern "C" {

int gmc_comm_unique_id(void* id128) {
    GMC_REQUIRE(id128, GMC_ERR_ARG, "gmc_comm_unique_id: NULL");
    GMC_TRY(load_nccl());
    ncclUniqueId id;
    GMC_NCCL(g_api.GetUniqueId(&id));
    memcpy(id128, id.internal, GMC_COMM_ID_BYTES);
    return GMC_OK;
}

int gmc_comm_create(gmc_comm** out, int world, int rank, const void* id128) {
    GMC_REQUIRE(out && world >= 1 && rank >= 0 && rank < world, GMC_ERR_ARG, "gmc_comm_create: bad world/rank %d/%d", rank, world);
    gmc_comm* c = new gmc_comm();
    c->world = world;
    c->rank = rank;
    if (world > 1) {
        GMC_REQUIRE(id128, GMC_ERR_ARG, "gmc_comm_create: NULL id");
        int rc = load_nccl();
        if (rc != GMC_OK) { delete c; return rc; }
        ncclUniqueId id;
        memcpy(id.internal, id128, GMC_COMM_ID_BYTES);
        ncclResult_t r = g_api.CommInitRank(&c->nccl, world, id, rank);
        if (r != ncclSuccess) {
            set_error("ncclCommInitRank(rank %d of %d): %s", rank, world, g_api.GetErrorString(r));
            delete c;
            return GMC_ERR_CUDA;
        }
    }
    *out = c;
    return GMC_OK;
}

extern "C" void gmc_pr_release_cached_state(void);     // gmc_pagerank.cu: retired peer-mapped state block

int gmc_comm_destroy(gmc_comm* c) {
    if (!c) return GMC_OK;
    if (c->world > 1) gmc_pr_release_cached_state();
    if (c->nccl) g_api.CommDestroy(c->nccl);
    delete c;
    return GMC_OK;
}

int gmc_comm_info(const gmc_comm* c, int* world, int* rank, int* nccl_version) {
    GMC_REQUIRE(c, GMC_ERR_ARG, "gmc_comm_info: NULL");
    if (world) *world = c->world;
    if (rank) *rank = c->rank;
    if (nccl_version) { *nccl_version = 0; if (g_api.GetVersion) g_api.GetVersion(nccl_version); }
    return GMC_OK;
}

int gmc_comm_barrier(gmc_comm* c, void* stream) {
    GMC_REQUIRE(c, GMC_ERR_ARG, "gmc_comm_barrier: NULL");
    cudaStream_t st = as_stream(stream);
    if (c->world > 1) {
        if (!c->scratch) GMC_TRY(gmc_malloc((void**)&c->scratch, 64));
        GMC_NCCL(g_api.AllReduce(c->scratch, c->scratch, 1, ncclFloat64, ncclSum, c->nccl, st));
    }
    GMC_CUDA(cudaStreamSynchronize(st));
    return GMC_OK;
}

// Balanced contiguous partition: prefix[0..n] non-decreasing (prefix[0] = 0); cuts[r] = first index i with
// prefix[i] >= r * prefix[n] / world, rounded down to a multiple of `align`; cuts[0] = 0, cuts[world] = n.
// Rank r owns [cuts[r], cuts[r+1]).  Pure host code.
int gmc_shard_cuts(const int32_t* prefix, int32_t n, int world, int32_t align, int32_t* cuts) {
    GMC_REQUIRE(prefix && cuts && n >= 0 && world >= 1 && align >= 1, GMC_ERR_ARG, "gmc_shard_cuts: bad args");
    const int64_t total = prefix[n];
    cuts[0] = 0;
    for (int r = 1; r < world; r++) {
        const int64_t target = (total * r + world - 1) / world;
        int32_t lo = 0, hi = n;      // first i with prefix[i] >= target
        while (lo < hi) {
            const int32_t mid = lo + ((hi - lo) >> 1);
            if (prefix[mid] >= target) hi = mid; else lo = mid + 1;
        }
        int32_t cut = (lo / align) * align;
        if (cut < cuts[r - 1]) cut = cuts[r - 1];
        cuts[r] = cut;
    }
    cuts[world] = n;
    return GMC_OK;
}

}  // extern "C"
