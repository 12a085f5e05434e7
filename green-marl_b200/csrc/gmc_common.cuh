// Shared plumbing for the libgm_cuda kernels (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdarg.h>
#include <string.h>
#include "gm_cuda.h"

namespace gmc {

void set_error(const char* fmt, ...);

#define GMC_CUDA(call)                                                                         \
    do {                                                                                       \
        cudaError_t _e = (call);                                                               \
        if (_e != cudaSuccess) {                                                               \
            gmc::set_error("%s:%d: %s -> %s", __FILE__, __LINE__, #call, cudaGetErrorString(_e)); \
            return GMC_ERR_CUDA;                                                               \
        }                                                                                      \
    } while (0)

#define GMC_REQUIRE(cond, status, ...)      \
    do {                                    \
        if (!(cond)) {                      \
            gmc::set_error(__VA_ARGS__);    \
            return (status);                \
        }                                   \
    } while (0)

#define GMC_TRY(expr)                 \
    do {                              \
        int _s = (expr);              \
        if (_s != GMC_OK) return _s;  \
    } while (0)

static inline cudaStream_t as_stream(void* s) { return (cudaStream_t)s; }

int sm_count();

template <typename T>
static inline T div_up(T a, T b) { return (a + b - 1) / b; }

// --- device helpers -----------------------------------------------------------------------
// streaming (read-once) loads: keep the index stream from displacing the gathered vector in L2
__device__ __forceinline__ int32_t ld_stream(const int32_t* p) { return __ldcs(p); }
__device__ __forceinline__ int4 ld_stream(const int4* p) { return __ldcs(p); }

__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_down_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ int warp_sum(int v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_down_sync(0xffffffffu, v, o);
    return v;
}

// exclusive prefix sum of an int32 device array (in place allowed), returns total in *total_dev (optional)
int exclusive_scan_i32(const int32_t* in, int32_t* out, size_t n, int32_t* total_dev, cudaStream_t st);

}  // namespace gmc

extern "C" int gmc_malloc_ipc(void** p, size_t bytes);   // internal: plain cudaMalloc (IPC-exportable); gmc_malloc uses the memory pool

struct gmc_pending_upload {
    static constexpr int MAX_CHUNKS = 8;
    cudaEvent_t small = nullptr;
    int nchunks = 0;
    cudaEvent_t chunk[MAX_CHUNKS] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
    int32_t end[MAX_CHUNKS] = {0, 0, 0, 0, 0, 0, 0, 0};
};

// The device-resident CSR mirror behind gmc_graph.
struct gmc_pr_plan;
struct gmc_bfs_ws;
struct gmc_graph {
    int32_t N = 0, M = 0;
    int32_t* begin = nullptr;       // [N+1]
    int32_t* node_idx = nullptr;    // [M]
    int32_t* r_begin = nullptr;     // [N+1]
    int32_t* r_node_idx = nullptr;  // [M]
    int32_t* e_rev2idx = nullptr;   // [M] optional
    int32_t* e_idx2idx = nullptr;   // [M] optional
    bool semi_sorted = false;
    bool tc_rows_checked = false;   // triangle_counting has verified that the rows of both CSRs are ascending
    bool has_reverse = false;
    bool has_forward = true;        // false: node_idx was not copied in (pull-only graph: PageRank needs begin + the reverse CSR)
    // copy-in still in flight (gmc_pagerank_host_csr): begin/r_begin are complete once `small` has fired, r_node_idx[0, end[k])
    // once chunk[k] has; the PageRank plan builder waits on them on its own stream, chunk by chunk (no host sync)
    struct gmc_pending_upload* pending = nullptr;
    int32_t pr_iters_hint = 0;      // max_iter of the call the next PageRank plan is built for (0: unknown / build the full plan)
    // lazily built per-algorithm state (invalidated when the CSR changes)
    gmc_pr_plan* pr = nullptr;      // PageRank plan: relabelled entry array, tasks, state buffers (gmc_pagerank.cu)
    gmc_bfs_ws* bfs = nullptr;
};

void gmc_graph_invalidate_plans(gmc_graph* g);
