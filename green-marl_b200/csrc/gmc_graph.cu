// Device-resident CSR mirror of gm_graph: upload/download, semi-sort (K5), reverse-edge build (K4).
// Reference behaviour followed: apps/output_cpp/gm_graph/src/gm_graph.cc:205-304 (make_reverse_edges),
// :380-423,468-503 (semi sort).  Sorting uses CUB's radix / segmented sort from the CUDA toolkit
// (library code, untimed graph-build step); histogram, scan and scatter/gather kernels are ours.
#include "gmc_common.cuh"
#include "gmc_comm.h"
#include <cub/device/device_radix_sort.cuh>
#include <cub/device/device_segmented_sort.cuh>
#include <algorithm>
#include <vector>

using namespace gmc;

void gmc_pr_plan_free(gmc_graph* g);   // gmc_pagerank.cu
void gmc_bfs_ws_free(gmc_graph* g);    // gmc_bfs.cu

void gmc_graph_invalidate_plans(gmc_graph* g) {
    g->tc_rows_checked = false;
    gmc_pr_plan_free(g);
    gmc_bfs_ws_free(g);
}

namespace {

__global__ void iota_i32(int32_t* __restrict__ p, int32_t n) {
    int32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) p[i] = i;
}

// in-degree histogram: the fetch-and-add of gm_graph.cc:238-245 (the returned slot is not needed:
// order inside a row comes from the stable sort)
__global__ void count_in_degree(const int32_t* __restrict__ node_idx, int32_t m, int32_t* __restrict__ deg) {
    int32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    int32_t stride = gridDim.x * blockDim.x;
    for (; i < m; i += stride) atomicAdd(&deg[ld_stream(node_idx + i)], 1);
}

// source vertex of forward edge e = the row r with begin[r] <= e < begin[r+1] (last such r: rows may be empty)
__device__ __forceinline__ int32_t row_of_edge(const int32_t* __restrict__ begin, int32_t n, int32_t e) {
    int32_t lo = 0, hi = n;  // invariant: begin[lo] <= e < begin[hi]
    while (hi - lo > 1) {
        int32_t mid = lo + ((hi - lo) >> 1);
        if (begin[mid] <= e) lo = mid; else hi = mid;
    }
    return lo;
}

__global__ void gather_sources(const int32_t* __restrict__ begin, int32_t n, const int32_t* __restrict__ edge_sorted,
                               int32_t m, int32_t* __restrict__ r_node_idx) {
    int32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < m) r_node_idx[i] = row_of_edge(begin, n, edge_sorted[i]);
}

int bits_for(int32_t n) {
    int b = 1;
    while (b < 31 && (1 << b) < n) b++;
    return b;
}

}  // namespace

extern "C" {

int gmc_graph_create(gmc_graph** out, int32_t n, int32_t m) {
    GMC_REQUIRE(out, GMC_ERR_ARG, "gmc_graph_create: NULL out");
    GMC_REQUIRE(n >= 0 && m >= 0, GMC_ERR_ARG, "gmc_graph_create: negative size");
    gmc_graph* g = new gmc_graph();
    g->N = n;
    g->M = m;
    int rc = gmc_malloc((void**)&g->begin, sizeof(int32_t) * ((size_t)n + 1));
    if (rc == GMC_OK) rc = gmc_malloc((void**)&g->node_idx, sizeof(int32_t) * (size_t)(m ? m : 1));
    if (rc != GMC_OK) { gmc_graph_destroy(g); return rc; }
    *out = g;
    return GMC_OK;
}

int gmc_graph_destroy(gmc_graph* g) {
    if (!g) return GMC_OK;
    gmc_graph_invalidate_plans(g);
    cudaFree(g->begin); cudaFree(g->node_idx); cudaFree(g->r_begin); cudaFree(g->r_node_idx);
    cudaFree(g->e_rev2idx); cudaFree(g->e_idx2idx);
    delete g;
    return GMC_OK;
}

int gmc_graph_num_nodes(const gmc_graph* g, int32_t* n) { GMC_REQUIRE(g && n, GMC_ERR_ARG, "NULL"); *n = g->N; return GMC_OK; }
int gmc_graph_num_edges(const gmc_graph* g, int32_t* m) { GMC_REQUIRE(g && m, GMC_ERR_ARG, "NULL"); *m = g->M; return GMC_OK; }
int gmc_graph_has_reverse(const gmc_graph* g, int* yes) { GMC_REQUIRE(g && yes, GMC_ERR_ARG, "NULL"); *yes = g->has_reverse; return GMC_OK; }

int gmc_graph_upload(gmc_graph* g, const int32_t* begin, const int32_t* node_idx, void* stream) {
    GMC_REQUIRE(g && begin, GMC_ERR_ARG, "gmc_graph_upload: NULL");
    gmc_graph_invalidate_plans(g);
    cudaStream_t st = as_stream(stream);
    GMC_CUDA(cudaMemcpyAsync(g->begin, begin, sizeof(int32_t) * ((size_t)g->N + 1), cudaMemcpyHostToDevice, st));
    // node_idx == NULL: a pull-only graph (what the generated pagerank() copies in: begin, r_begin, r_node_idx);
    // everything that follows out-edges (BFS, semi-sort, reverse build) then fails with GMC_ERR_STATE
    g->has_forward = node_idx != nullptr || g->M == 0;
    if (g->M && node_idx) GMC_CUDA(cudaMemcpyAsync(g->node_idx, node_idx, sizeof(int32_t) * (size_t)g->M, cudaMemcpyHostToDevice, st));
    GMC_CUDA(cudaStreamSynchronize(st));
    g->semi_sorted = false;
    g->has_reverse = false;
    return GMC_OK;
}

int gmc_graph_upload_reverse(gmc_graph* g, const int32_t* r_begin, const int32_t* r_node_idx, void* stream) {
    GMC_REQUIRE(g && r_begin && (r_node_idx || g->M == 0), GMC_ERR_ARG, "gmc_graph_upload_reverse: NULL");
    gmc_graph_invalidate_plans(g);
    cudaStream_t st = as_stream(stream);
    if (!g->r_begin) GMC_TRY(gmc_malloc((void**)&g->r_begin, sizeof(int32_t) * ((size_t)g->N + 1)));
    if (!g->r_node_idx) GMC_TRY(gmc_malloc((void**)&g->r_node_idx, sizeof(int32_t) * (size_t)(g->M ? g->M : 1)));
    GMC_CUDA(cudaMemcpyAsync(g->r_begin, r_begin, sizeof(int32_t) * ((size_t)g->N + 1), cudaMemcpyHostToDevice, st));
    if (g->M) GMC_CUDA(cudaMemcpyAsync(g->r_node_idx, r_node_idx, sizeof(int32_t) * (size_t)g->M, cudaMemcpyHostToDevice, st));
    GMC_CUDA(cudaStreamSynchronize(st));
    g->has_reverse = true;
    return GMC_OK;
}

// K5.  Segmented stable sort of every adjacency row by neighbour id, e_idx2idx carried along
// (gm_graph.cc:468-503).  Stable => ties keep ascending original edge index.
int gmc_graph_semi_sort(gmc_graph* g, int want_idx2idx, void* stream) {
    GMC_REQUIRE(g, GMC_ERR_ARG, "gmc_graph_semi_sort: NULL");
    GMC_REQUIRE(g->has_forward, GMC_ERR_STATE, "gmc_graph_semi_sort: node_idx was not uploaded");
    if (g->semi_sorted && (!want_idx2idx || g->e_idx2idx)) return GMC_OK;
    cudaStream_t st = as_stream(stream);
    gmc_graph_invalidate_plans(g);
    const int32_t n = g->N, m = g->M;
    if (m == 0 || n == 0) { g->semi_sorted = true; return GMC_OK; }
    int32_t *keys_out = nullptr, *vals_in = nullptr, *vals_out = nullptr;
    void* tmp = nullptr;
    size_t tmp_bytes = 0;
    GMC_TRY(gmc_malloc((void**)&keys_out, sizeof(int32_t) * (size_t)m));
    int rc = GMC_OK;
    if (want_idx2idx) {
        rc = gmc_malloc((void**)&vals_in, sizeof(int32_t) * (size_t)m);
        if (rc == GMC_OK) rc = gmc_malloc((void**)&vals_out, sizeof(int32_t) * (size_t)m);
        if (rc != GMC_OK) { cudaFree(keys_out); cudaFree(vals_in); return rc; }
        iota_i32<<<div_up(m, 256), 256, 0, st>>>(vals_in, m);
        cub::DeviceSegmentedSort::StableSortPairs(nullptr, tmp_bytes, g->node_idx, keys_out, vals_in, vals_out, (int64_t)m, (int64_t)n, g->begin, g->begin + 1, st);
        rc = gmc_malloc(&tmp, tmp_bytes);
        if (rc == GMC_OK) {
            cudaError_t e = cub::DeviceSegmentedSort::StableSortPairs(tmp, tmp_bytes, g->node_idx, keys_out, vals_in, vals_out, (int64_t)m, (int64_t)n, g->begin, g->begin + 1, st);
            if (e != cudaSuccess) { set_error("segmented sort: %s", cudaGetErrorString(e)); rc = GMC_ERR_CUDA; }
        }
    } else {
        cub::DeviceSegmentedSort::StableSortKeys(nullptr, tmp_bytes, g->node_idx, keys_out, (int64_t)m, (int64_t)n, g->begin, g->begin + 1, st);
        rc = gmc_malloc(&tmp, tmp_bytes);
        if (rc == GMC_OK) {
            cudaError_t e = cub::DeviceSegmentedSort::StableSortKeys(tmp, tmp_bytes, g->node_idx, keys_out, (int64_t)m, (int64_t)n, g->begin, g->begin + 1, st);
            if (e != cudaSuccess) { set_error("segmented sort: %s", cudaGetErrorString(e)); rc = GMC_ERR_CUDA; }
        }
    }
    if (rc == GMC_OK && cudaStreamSynchronize(st) != cudaSuccess) { set_error("semi_sort: %s", cudaGetErrorString(cudaGetLastError())); rc = GMC_ERR_CUDA; }
    cudaFree(tmp);
    cudaFree(vals_in);
    if (rc != GMC_OK) { cudaFree(keys_out); cudaFree(vals_out); return rc; }
    cudaFree(g->node_idx);
    g->node_idx = keys_out;
    if (want_idx2idx) { cudaFree(g->e_idx2idx); g->e_idx2idx = vals_out; }
    g->semi_sorted = true;
    // an existing reverse CSR stays valid as a graph, but its e_rev2idx refers to the old edge order
    if (g->has_reverse && g->e_rev2idx) { cudaFree(g->e_rev2idx); g->e_rev2idx = nullptr; }
    return GMC_OK;
}

// K4.  in-degree histogram -> exclusive scan -> stable radix sort of edge ids by destination -> gather
// sources.  Forward edges are ordered by source, so a STABLE sort by destination leaves every reverse
// row ascending by source id with ties in ascending edge index: the state the reference reaches after
// make_reverse_edges + do_semi_sort_reverse (gm_graph.cc:233-293), minus its nondeterministic tie order.
int gmc_graph_make_reverse(gmc_graph* g, int want_rev2idx, void* stream) {
    GMC_REQUIRE(g, GMC_ERR_ARG, "gmc_graph_make_reverse: NULL");
    GMC_REQUIRE(g->has_forward || g->has_reverse, GMC_ERR_STATE, "gmc_graph_make_reverse: node_idx was not uploaded");
    if (g->has_reverse && (!want_rev2idx || g->e_rev2idx)) return GMC_OK;
    cudaStream_t st = as_stream(stream);
    gmc_graph_invalidate_plans(g);
    const int32_t n = g->N, m = g->M;
    if (!g->r_begin) GMC_TRY(gmc_malloc((void**)&g->r_begin, sizeof(int32_t) * ((size_t)n + 1)));
    if (!g->r_node_idx) GMC_TRY(gmc_malloc((void**)&g->r_node_idx, sizeof(int32_t) * (size_t)(m ? m : 1)));
    GMC_CUDA(cudaMemsetAsync(g->r_begin, 0, sizeof(int32_t) * ((size_t)n + 1), st));
    if (m == 0 || n == 0) { GMC_CUDA(cudaStreamSynchronize(st)); g->has_reverse = true; return GMC_OK; }
    count_in_degree<<<sm_count() * 8, 256, 0, st>>>(g->node_idx, m, g->r_begin);
    GMC_CUDA(cudaGetLastError());
    GMC_TRY(exclusive_scan_i32(g->r_begin, g->r_begin, (size_t)n + 1, nullptr, st));

    int32_t *keys_out = nullptr, *vals_in = nullptr, *vals_out = nullptr;
    void* tmp = nullptr;
    size_t tmp_bytes = 0;
    int rc = gmc_malloc((void**)&keys_out, sizeof(int32_t) * (size_t)m);
    if (rc == GMC_OK) rc = gmc_malloc((void**)&vals_in, sizeof(int32_t) * (size_t)m);
    if (rc == GMC_OK) rc = gmc_malloc((void**)&vals_out, sizeof(int32_t) * (size_t)m);
    if (rc == GMC_OK) {
        iota_i32<<<div_up(m, 256), 256, 0, st>>>(vals_in, m);
        const int end_bit = bits_for(n);
        cub::DeviceRadixSort::SortPairs(nullptr, tmp_bytes, g->node_idx, keys_out, vals_in, vals_out, (int64_t)m, 0, end_bit, st);
        rc = gmc_malloc(&tmp, tmp_bytes);
        if (rc == GMC_OK) {
            cudaError_t e = cub::DeviceRadixSort::SortPairs(tmp, tmp_bytes, g->node_idx, keys_out, vals_in, vals_out, (int64_t)m, 0, end_bit, st);
            if (e != cudaSuccess) { set_error("radix sort: %s", cudaGetErrorString(e)); rc = GMC_ERR_CUDA; }
        }
    }
    if (rc == GMC_OK) {
        gather_sources<<<div_up(m, 256), 256, 0, st>>>(g->begin, n, vals_out, m, g->r_node_idx);
        if (cudaStreamSynchronize(st) != cudaSuccess) { set_error("make_reverse: %s", cudaGetErrorString(cudaGetLastError())); rc = GMC_ERR_CUDA; }
    }
    cudaFree(tmp); cudaFree(keys_out); cudaFree(vals_in);
    if (rc != GMC_OK) { cudaFree(vals_out); return rc; }
    cudaFree(g->e_rev2idx);
    g->e_rev2idx = nullptr;
    if (want_rev2idx) g->e_rev2idx = vals_out; else cudaFree(vals_out);
    g->has_reverse = true;
    return GMC_OK;
}

// Sharded copy-in (multi-GPU e2e path): every rank copies only its 1/world slice of each CSR array from its host
// memory over PCIe, the slices are then allgathered between the GPUs over NVLink, so the job moves the graph
// host->device once instead of `world` times.  Host arrays are the full arrays (only the rank's slice is read).
int gmc_graph_upload_sharded(gmc_graph* g, gmc_comm* comm, const int32_t* begin, const int32_t* node_idx, const int32_t* r_begin,
                             const int32_t* r_node_idx, void* stream) {
    GMC_REQUIRE(g && begin, GMC_ERR_ARG, "gmc_graph_upload_sharded: NULL");
    if (!comm || comm->world == 1) {
        GMC_TRY(gmc_graph_upload(g, begin, node_idx, stream));
        return r_begin ? gmc_graph_upload_reverse(g, r_begin, r_node_idx, stream) : GMC_OK;
    }
    gmc_graph_invalidate_plans(g);
    cudaStream_t st = as_stream(stream);
    const int W = comm->world, R = comm->rank;
    if (r_begin) {
        if (!g->r_begin) GMC_TRY(gmc_malloc((void**)&g->r_begin, sizeof(int32_t) * ((size_t)g->N + 1)));
        if (!g->r_node_idx) GMC_TRY(gmc_malloc((void**)&g->r_node_idx, sizeof(int32_t) * (size_t)(g->M ? g->M : 1)));
    }
    struct Arr { int32_t* dev; const int32_t* host; int64_t len; };
    const Arr arrs[4] = {{g->begin, begin, (int64_t)g->N + 1}, {g->node_idx, node_idx, (int64_t)g->M},
                         {g->r_begin, r_begin, (int64_t)g->N + 1}, {g->r_node_idx, r_node_idx, (int64_t)g->M}};
    std::vector<int64_t> off((size_t)W), len((size_t)W);
    g->has_forward = node_idx != nullptr || g->M == 0;
    for (int a = 0; a < (r_begin ? 4 : 2); a++) {
        if (!arrs[a].host) continue;      // node_idx not given: pull-only graph
        const int64_t L = arrs[a].len, per = (L + W - 1) / W;
        for (int r = 0; r < W; r++) { off[r] = std::min<int64_t>(L, r * per); len[r] = std::min<int64_t>(L, (r + 1) * per) - off[r]; }
        if (len[R] > 0)
            GMC_CUDA(cudaMemcpyAsync(arrs[a].dev + off[R], arrs[a].host + off[R], sizeof(int32_t) * (size_t)len[R], cudaMemcpyHostToDevice, st));
        GMC_TRY(gmc_comm_allgather_segments(comm, arrs[a].dev, sizeof(int32_t), off.data(), len.data(), 1, st));
    }
    GMC_CUDA(cudaStreamSynchronize(st));
    g->semi_sorted = false;
    g->has_reverse = r_begin != nullptr;
    return GMC_OK;
}

static int array_of(const gmc_graph* g, int which, const int32_t** p, size_t* len) {
    switch (which) {
        case GMC_ARR_BEGIN: *p = g->begin; *len = (size_t)g->N + 1; break;
        case GMC_ARR_NODE_IDX: *p = g->has_forward ? g->node_idx : nullptr; *len = (size_t)g->M; break;
        case GMC_ARR_R_BEGIN: *p = g->has_reverse ? g->r_begin : nullptr; *len = (size_t)g->N + 1; break;
        case GMC_ARR_R_NODE_IDX: *p = g->has_reverse ? g->r_node_idx : nullptr; *len = (size_t)g->M; break;
        case GMC_ARR_E_REV2IDX: *p = g->e_rev2idx; *len = (size_t)g->M; break;
        case GMC_ARR_E_IDX2IDX: *p = g->e_idx2idx; *len = (size_t)g->M; break;
        default: set_error("unknown array id %d", which); return GMC_ERR_ARG;
    }
    if (!*p) { set_error("array %d has not been built", which); return GMC_ERR_STATE; }
    return GMC_OK;
}

int gmc_graph_download(const gmc_graph* g, int which, int32_t* host_dst, size_t count, void* stream) {
    GMC_REQUIRE(g && (host_dst || count == 0), GMC_ERR_ARG, "gmc_graph_download: NULL");
    const int32_t* p; size_t len;
    GMC_TRY(array_of(g, which, &p, &len));
    GMC_REQUIRE(count == len, GMC_ERR_ARG, "gmc_graph_download: count %zu != array length %zu", count, len);
    return gmc_d2h(host_dst, p, len * sizeof(int32_t), stream);
}

int gmc_graph_device_ptr(const gmc_graph* g, int which, const int32_t** dev_ptr) {
    GMC_REQUIRE(g && dev_ptr, GMC_ERR_ARG, "gmc_graph_device_ptr: NULL");
    size_t len;
    return array_of(g, which, dev_ptr, &len);
}

}  // extern "C"
