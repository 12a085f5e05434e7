// Triangle counting: the generated triangle_counting() (reference apps/src/triangle_counting.gm:1-13) on the device.
//
//   Foreach(v: G.Nodes) Foreach(u: v.Nbrs)(u > v) Foreach(w: v.Nbrs)(w > u) If (w.HasEdgeTo(u)) T += 1;
//
// The C++ back end emits it as three nested loops over the out-edge slots of v (schedule(dynamic,128) over v, a private
// counter per thread combined with ATOMIC_ADD) with G.has_edge_to(w, u) -- a binary search of u in the semi-sorted row
// of w (gm_graph.cc:60-66,589-633) -- innermost: d(v)^2/2 searches per vertex.  Counted here is the same multiset
// (edge SLOTS i, j of row v with idx[i] > v, idx[j] > idx[i] and an edge idx[j] -> idx[i]; parallel edges of row v count
// per slot, the membership test is boolean), but per slot i the inner loop becomes one sorted-list intersection:
//     A = slots of row v with value > u            (a suffix of the semi-sorted row, multiplicity kept)
//     B = sources w > u of the in-edges of u       (a suffix of the reverse row of u, ascending by source)
//     count_i = #{ a in A : value(a) in B }
// evaluated by binary-searching the shorter side in the longer one.  Work item = one edge slot; a warp takes 32
// consecutive slots, lanes finish the small intersections alone and the warp goes through the large ones together.
// The count is an integer sum, so the result is exact and independent of the order of evaluation.
#include <cuda_runtime.h>
#include <stdint.h>
#include <mutex>
#include "gmc_common.cuh"
#include "gmc_comm.h"

using namespace gmc;

namespace {

constexpr int TC_THREADS = 256;
constexpr int TC_TILE = 2048;          // edge slots per row hint
constexpr int TC_SOLO = 12;            // intersections whose shorter side is at most this long are done by one lane
constexpr int TC_HEAVY = 2048;         // intersections whose shorter side is longer than this are cut into segments ...
constexpr int TC_SEG = 1024;           // ... of this many elements, handed to the second kernel (one warp per segment)
constexpr unsigned FULLM = 0xffffffffu;

struct TcSeg { int32_t a_lo, a_hi, b_lo, b_hi, k0, k1; };

__global__ void tc_row_hints(const int32_t* __restrict__ begin, int32_t n, int32_t m, int32_t n_hints, int32_t* __restrict__ hint) {
    int32_t b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= n_hints) return;
    const int64_t e64 = (int64_t)b * TC_TILE;
    if (e64 >= m) { hint[b] = n; return; }
    const int32_t e = (int32_t)e64;
    int32_t lo = 0, hi = n;            // begin[lo] <= e < begin[hi]
    while (hi - lo > 1) {
        const int32_t mid = lo + ((hi - lo) >> 1);
        if (begin[mid] <= e) lo = mid; else hi = mid;
    }
    hint[b] = lo;
}

// first position in [lo, hi) whose value is > x (values ascending)
__device__ __forceinline__ int32_t upper_bound(const int32_t* __restrict__ a, int32_t lo, int32_t hi, int32_t x) {
    while (lo < hi) {
        const int32_t mid = lo + ((hi - lo) >> 1);
        if (__ldg(a + mid) <= x) lo = mid + 1; else hi = mid;
    }
    return lo;
}
__device__ __forceinline__ int32_t lower_bound(const int32_t* __restrict__ a, int32_t lo, int32_t hi, int32_t x) {
    while (lo < hi) {
        const int32_t mid = lo + ((hi - lo) >> 1);
        if (__ldg(a + mid) < x) lo = mid + 1; else hi = mid;
    }
    return lo;
}
__device__ __forceinline__ bool contains(const int32_t* __restrict__ a, int32_t lo, int32_t hi, int32_t x) {
    const int32_t p = lower_bound(a, lo, hi, x);
    return p < hi && __ldg(a + p) == x;
}

// contribution of positions [k0, k1) (stride `step`) of the shorter side of one intersection
__device__ __forceinline__ unsigned long long intersect_part(const int32_t* __restrict__ idx, const int32_t* __restrict__ ridx,
                                                             int32_t a_lo, int32_t a_hi, int32_t b_lo, int32_t b_hi, int32_t k0, int32_t step,
                                                             int32_t k1 = INT32_MAX) {
    unsigned long long c = 0;
    if (a_hi - a_lo <= b_hi - b_lo) {                 // walk the slots of row v, look each value up among the in-neighbours of u
        const int32_t a_end = (int32_t)min((int64_t)a_hi, (int64_t)a_lo + k1);
        for (int32_t a = a_lo + k0; a < a_end; a += step) c += contains(ridx, b_lo, b_hi, __ldg(idx + a)) ? 1u : 0u;
    } else {                                          // walk the distinct in-neighbours of u, count their slots in row v
        const int32_t b_end = (int32_t)min((int64_t)b_hi, (int64_t)b_lo + k1);
        for (int32_t b = b_lo + k0; b < b_end; b += step) {
            const int32_t w = __ldg(ridx + b);
            if (b > b_lo && __ldg(ridx + b - 1) == w) continue;          // parallel edges w -> u: the test is boolean
            const int32_t p = lower_bound(idx, a_lo, a_hi, w);
            if (p < a_hi && __ldg(idx + p) == w) c += (unsigned long long)(upper_bound(idx, p, a_hi, w) - p);
        }
    }
    return c;
}

__global__ void __launch_bounds__(TC_THREADS) tc_count(const int32_t* __restrict__ begin, const int32_t* __restrict__ idx,
                                                        const int32_t* __restrict__ rbegin, const int32_t* __restrict__ ridx,
                                                        const int32_t* __restrict__ hint, int32_t n, int32_t e_lo, int32_t e_hi,
                                                        unsigned int* __restrict__ next_chunk, unsigned long long* __restrict__ total,
                                                        TcSeg* __restrict__ segs, unsigned int* __restrict__ n_segs, unsigned int seg_cap) {
    const int lane = threadIdx.x & 31;
    unsigned long long cnt = 0;
    const int64_t n_chunks = ((int64_t)e_hi - e_lo + 31) >> 5;
    for (;;) {
        unsigned int ch = 0;
        if (lane == 0) ch = atomicAdd(next_chunk, 1u);
        ch = __shfl_sync(FULLM, ch, 0);
        if ((int64_t)ch >= n_chunks) break;
        const int64_t e64 = (int64_t)e_lo + ((int64_t)ch << 5) + lane;
        int32_t a_lo = 0, a_hi = 0, b_lo = 0, b_hi = 0;
        if (e64 < e_hi) {
            const int32_t e = (int32_t)e64;
            int32_t lo = __ldg(hint + e / TC_TILE), hi = min(__ldg(hint + e / TC_TILE + 1) + 1, n);    // begin[lo] <= e < begin[hi]
            while (hi - lo > 1) {
                const int32_t mid = lo + ((hi - lo) >> 1);
                if (__ldg(begin + mid) <= e) lo = mid; else hi = mid;
            }
            const int32_t v = lo, u = __ldg(idx + e);
            if (u > v) {
                a_hi = __ldg(begin + v + 1);
                a_lo = upper_bound(idx, e + 1, a_hi, u);                        // slots of row v with value > u
                b_hi = __ldg(rbegin + u + 1);
                if (a_lo < a_hi) b_lo = upper_bound(ridx, __ldg(rbegin + u), b_hi, u); else b_hi = 0;   // in-neighbours of u that are > u
            }
        }
        int32_t work = min(a_hi - a_lo, b_hi - b_lo);
        // heavy intersections (hub x hub: up to ~1e5 elements on the shorter side, tens of ms for one warp) are cut into
        // fixed segments of the shorter side and left to tc_heavy; if the segment list is full they stay here
        if (work > TC_HEAVY) {
            const unsigned ns = (unsigned)((work + TC_SEG - 1) / TC_SEG);
            const unsigned at = atomicAdd(n_segs, ns);
            if (at + ns <= seg_cap) {
                for (unsigned k = 0; k < ns; k++) segs[at + k] = TcSeg{a_lo, a_hi, b_lo, b_hi, (int32_t)(k * TC_SEG), (int32_t)min((int64_t)work, (int64_t)(k + 1) * TC_SEG)};
                work = 0;
            }      // list full: the counter stays beyond the capacity, so every later request fails too; unwritten slots are zero (no work)
        }
        if (work > 0 && work <= TC_SOLO) cnt += intersect_part(idx, ridx, a_lo, a_hi, b_lo, b_hi, 0, 1);
        unsigned big = __ballot_sync(FULLM, work > TC_SOLO);
        while (big) {
            const int src = __ffs(big) - 1;
            big &= big - 1;
            const int32_t A0 = __shfl_sync(FULLM, a_lo, src), A1 = __shfl_sync(FULLM, a_hi, src);
            const int32_t B0 = __shfl_sync(FULLM, b_lo, src), B1 = __shfl_sync(FULLM, b_hi, src);
            cnt += intersect_part(idx, ridx, A0, A1, B0, B1, lane, 32);
        }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) cnt += __shfl_xor_sync(FULLM, cnt, o);
    if (lane == 0 && cnt) atomicAdd(total, cnt);
}

// second pass: one warp per segment of a heavy intersection
__global__ void __launch_bounds__(TC_THREADS) tc_heavy(const int32_t* __restrict__ idx, const int32_t* __restrict__ ridx, const TcSeg* __restrict__ segs,
                                                        const unsigned int* __restrict__ n_segs, unsigned int seg_cap,
                                                        unsigned int* __restrict__ next_seg, unsigned long long* __restrict__ total) {
    const int lane = threadIdx.x & 31;
    const unsigned int n = min(*n_segs, seg_cap);
    unsigned long long cnt = 0;
    for (;;) {
        unsigned int sgi = 0;
        if (lane == 0) sgi = atomicAdd(next_seg, 1u);
        sgi = __shfl_sync(FULLM, sgi, 0);
        if (sgi >= n) break;
        const TcSeg sg = segs[sgi];
        cnt += intersect_part(idx, ridx, sg.a_lo, sg.a_hi, sg.b_lo, sg.b_hi, sg.k0 + lane, 32, sg.k1);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) cnt += __shfl_xor_sync(FULLM, cnt, o);
    if (lane == 0 && cnt) atomicAdd(total, cnt);
}

}  // namespace

__global__ void k_rows_ascending(const int32_t* __restrict__ begin, const int32_t* __restrict__ idx, int32_t n, int32_t* __restrict__ bad) {
    const int32_t v = blockIdx.x * blockDim.x + threadIdx.x;
    if (v >= n) return;
    for (int32_t e = begin[v] + 1; e < begin[v + 1]; e++)
        if (idx[e] < idx[e - 1]) { *bad = 1; return; }
}

extern "C" {

// edge slots [e_lo, e_hi) of the forward CSR: the count is a plain sum over slots, so a multi-GPU job gives every rank a
// slice of the slots and adds the partial counts (no exchange step on the path)
int gmc_triangle_counting_range(gmc_graph* g, int64_t e_lo, int64_t e_hi, int64_t* count, float* kernel_ms, void* stream) {
    GMC_REQUIRE(g && count, GMC_ERR_ARG, "gmc_triangle_counting: NULL");
    GMC_REQUIRE(g->has_forward, GMC_ERR_STATE, "triangle_counting: the forward CSR (node_idx) was not uploaded for this graph");
    // rows must be ascending (semi-sorted), the assumption has_edge_to's binary search makes in the reference too (gm_graph.cc:592)
    GMC_REQUIRE(g->has_reverse, GMC_ERR_STATE, "triangle_counting needs the reverse CSR: call gmc_graph_make_reverse first");
    GMC_REQUIRE(e_lo >= 0 && e_lo <= e_hi && e_hi <= g->M, GMC_ERR_ARG, "gmc_triangle_counting: slot range [%lld, %lld) outside [0, %d]",
                (long long)e_lo, (long long)e_hi, g->M);
    *count = 0;
    if (kernel_ms) *kernel_ms = 0.f;
    if (e_hi == e_lo) return GMC_OK;
    cudaStream_t st = as_stream(stream);
    // the intersections assume ascending rows in both CSRs; a graph filled through prepare_external_creation need not have them.
    // Checked once per graph (one pass over the edges), not assumed: unsorted rows would silently give a wrong count.
    if (!g->tc_rows_checked) {
        int32_t* bad = nullptr;
        GMC_TRY(gmc_malloc((void**)&bad, 4));
        cudaMemsetAsync(bad, 0, 4, st);
        k_rows_ascending<<<div_up(g->N, 256), 256, 0, st>>>(g->begin, g->node_idx, g->N, bad);
        k_rows_ascending<<<div_up(g->N, 256), 256, 0, st>>>(g->r_begin, g->r_node_idx, g->N, bad);
        int32_t hbad = 0;
        cudaMemcpyAsync(&hbad, bad, 4, cudaMemcpyDeviceToHost, st);
        cudaStreamSynchronize(st);
        cudaFree(bad);
        GMC_REQUIRE(!hbad, GMC_ERR_STATE, "triangle_counting: rows are not ascending (semi-sort the graph and rebuild the reverse CSR first)");
        g->tc_rows_checked = true;
    }
    const int32_t n = g->N, m = g->M;
    const int32_t n_hints = div_up(m, TC_TILE) + 1;
    int32_t* hints = nullptr;
    unsigned long long* total = nullptr;      // [0] = count, [1] (as 2 x u32) = chunk counter, segment count, [2] = segment cursor
    TcSeg* segs = nullptr;
    const unsigned seg_cap = 1u << 22;        // 4 M segments of 1024 elements (96 MB); beyond that heavy items are done in place
    GMC_TRY(gmc_malloc((void**)&hints, 4 * (size_t)n_hints));
    int rc = gmc_malloc((void**)&total, 32);
    if (rc == GMC_OK) rc = gmc_malloc((void**)&segs, sizeof(TcSeg) * (size_t)seg_cap);
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    if (rc == GMC_OK && (cudaEventCreate(&ev0) != cudaSuccess || cudaEventCreate(&ev1) != cudaSuccess)) { set_error("triangle_counting: event"); rc = GMC_ERR_CUDA; }
    unsigned long long h = 0;
    if (rc == GMC_OK) {
        cudaMemsetAsync(total, 0, 32, st);
        cudaMemsetAsync(segs, 0, sizeof(TcSeg) * (size_t)seg_cap, st);
        cudaEventRecord(ev0, st);
        tc_row_hints<<<div_up(n_hints, 256), 256, 0, st>>>(g->begin, n, m, n_hints, hints);
        tc_count<<<sm_count() * 8, TC_THREADS, 0, st>>>(g->begin, g->node_idx, g->r_begin, g->r_node_idx, hints, n, (int32_t)e_lo, (int32_t)e_hi,
                                                         reinterpret_cast<unsigned int*>(total + 1), total, segs,
                                                         reinterpret_cast<unsigned int*>(total + 1) + 1, seg_cap);
        tc_heavy<<<sm_count() * 8, TC_THREADS, 0, st>>>(g->node_idx, g->r_node_idx, segs, reinterpret_cast<unsigned int*>(total + 1) + 1, seg_cap,
                                                         reinterpret_cast<unsigned int*>(total + 2), total);
        cudaEventRecord(ev1, st);
        if (cudaMemcpyAsync(&h, total, 8, cudaMemcpyDeviceToHost, st) != cudaSuccess || cudaStreamSynchronize(st) != cudaSuccess) {
            set_error("triangle_counting: %s", cudaGetErrorString(cudaGetLastError())); rc = GMC_ERR_CUDA;
        }
    }
    if (rc == GMC_OK) {
        *count = (int64_t)h;
        if (kernel_ms) cudaEventElapsedTime(kernel_ms, ev0, ev1);
    }
    if (ev0) cudaEventDestroy(ev0);
    if (ev1) cudaEventDestroy(ev1);
    cudaFree(hints); cudaFree(total); cudaFree(segs);
    return rc;
}

// one process per GPU: rank r counts the slots [r*M/world, (r+1)*M/world) of the (replicated) graph; the partial counts
// are added with one 8-byte allreduce -- the only exchange the procedure needs.  Every rank returns the total.
int gmc_triangle_counting_sharded(gmc_graph* g, gmc_comm* comm, int64_t* count, float* kernel_ms, void* stream) {
    GMC_REQUIRE(g && count, GMC_ERR_ARG, "gmc_triangle_counting_sharded: NULL");
    const int world = comm ? comm->world : 1, rank = comm ? comm->rank : 0;
    const int64_t lo = (int64_t)g->M * rank / world, hi = (int64_t)g->M * (rank + 1) / world;
    int64_t part = 0;
    GMC_TRY(gmc_triangle_counting_range(g, lo, hi, &part, kernel_ms, stream));
    if (world == 1) { *count = part; return GMC_OK; }
    cudaStream_t st = as_stream(stream);
    long long* buf = nullptr;
    GMC_TRY(gmc_malloc((void**)&buf, 16));
    long long h = (long long)part;
    int rc = GMC_OK;
    if (cudaMemcpyAsync(buf, &h, 8, cudaMemcpyHostToDevice, st) != cudaSuccess) rc = GMC_ERR_CUDA;
    if (rc == GMC_OK) rc = gmc_comm_allreduce_sum_i64(comm, buf, buf + 1, 1, st);
    if (rc == GMC_OK && (cudaMemcpyAsync(&h, buf + 1, 8, cudaMemcpyDeviceToHost, st) != cudaSuccess || cudaStreamSynchronize(st) != cudaSuccess)) rc = GMC_ERR_CUDA;
    cudaFree(buf);
    if (rc != GMC_OK) { set_error("gmc_triangle_counting_sharded: the partial counts could not be added"); return rc; }
    *count = (int64_t)h;
    return GMC_OK;
}

int gmc_triangle_counting(gmc_graph* g, int64_t* count, float* kernel_ms) {
    GMC_REQUIRE(g, GMC_ERR_ARG, "gmc_triangle_counting: NULL");
    return gmc_triangle_counting_range(g, 0, g->M, count, kernel_ms, nullptr);
}

}  // extern "C"

// ---- the same count for callers that must not allocate (no counterpart in the reference)
// -- code under CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc.  The scratch of the three
// kernels (row hints, counters, segment list of the heavy intersections, the "rows not ascending" flag) lives in ONE file-scope
// device array of a fixed size, so every GPU that loads the library holds TC_NA_BYTES (about 0.9 MB) until exit, and graphs of
// more than GMC_TC_NOALLOC_MAX_EDGES edge slots are refused.  All calls on a device share the array and it keeps the last call's
// contents: every call clears the counters, the flag and the segment list first, and two calls that overlap on the GPU (two
// streams, two host threads) RACE -- callers serialise them.
constexpr int64_t TC_NA_MAX_EDGES = GMC_TC_NOALLOC_MAX_EDGES;
constexpr unsigned TC_NA_SEG_CAP = 1u << 15;                   // 32 K segments (768 KB); heavy items beyond that are done in place
constexpr size_t TC_NA_HDR = 256;                              // total[4] (u64) at 0, flag (i32) at 64
constexpr size_t TC_NA_HINTS = ((size_t)(TC_NA_MAX_EDGES / TC_TILE + 2) * 4 + 255) / 256 * 256;
constexpr size_t TC_NA_SEGS = sizeof(TcSeg) * (size_t)TC_NA_SEG_CAP;
constexpr size_t TC_NA_BYTES = TC_NA_HDR + TC_NA_HINTS + TC_NA_SEGS;
static __device__ __align__(256) unsigned char g_tc_na_scratch[TC_NA_BYTES];

namespace {

// count_dev = the count, or -1 when a row of either CSR was not ascending (the allocating entry point returns GMC_ERR_STATE there)
__global__ void tc_na_publish(const unsigned long long* __restrict__ total, const int32_t* __restrict__ bad, int64_t* __restrict__ count_dev) {
    *count_dev = *bad ? (int64_t)-1 : (int64_t)*total;
}

// In host code g_tc_na_scratch is only a stand-in for the symbol, not a device address: the address is asked for once per
// device, with that device current, and cached.
constexpr int TC_NA_MAX_DEVICES = 64;
unsigned char* g_tc_na_ptr[TC_NA_MAX_DEVICES];
std::mutex g_tc_na_mu;

int tc_na_scratch(unsigned char** out) {
    int dev = -1;
    GMC_CUDA(cudaGetDevice(&dev));
    GMC_REQUIRE(dev >= 0 && dev < TC_NA_MAX_DEVICES, GMC_ERR_STATE, "gmc_triangle_counting_noalloc: device ordinal %d out of range", dev);
    std::lock_guard<std::mutex> lk(g_tc_na_mu);
    if (!g_tc_na_ptr[dev]) {
        void* p = nullptr;
        const cudaError_t e = cudaGetSymbolAddress(&p, g_tc_na_scratch);
        if (e != cudaSuccess || !p) {
            set_error("gmc_triangle_counting_noalloc: cudaGetSymbolAddress: %s", cudaGetErrorString(e));
            return GMC_ERR_CUDA;
        }
        g_tc_na_ptr[dev] = (unsigned char*)p;
    }
    *out = g_tc_na_ptr[dev];
    return GMC_OK;
}

}  // namespace

extern "C" int gmc_triangle_counting_noalloc(gmc_graph* g, int64_t* count_dev, void* stream) {
    GMC_REQUIRE(g && count_dev, GMC_ERR_ARG, "gmc_triangle_counting_noalloc: NULL");
    GMC_REQUIRE(g->has_forward, GMC_ERR_STATE, "triangle_counting: the forward CSR (node_idx) was not uploaded for this graph");
    GMC_REQUIRE(g->has_reverse, GMC_ERR_STATE, "triangle_counting needs the reverse CSR: call gmc_graph_make_reverse first");
    GMC_REQUIRE((int64_t)g->M <= TC_NA_MAX_EDGES, GMC_ERR_ARG,
                "gmc_triangle_counting_noalloc: %d edge slots, the fixed scratch holds at most %lld (use gmc_triangle_counting)", g->M,
                (long long)TC_NA_MAX_EDGES);
    unsigned char* base = nullptr;
    GMC_TRY(tc_na_scratch(&base));
    cudaStream_t st = as_stream(stream);
    unsigned long long* total = reinterpret_cast<unsigned long long*>(base);
    int32_t* bad = reinterpret_cast<int32_t*>(base + 64);
    int32_t* hints = reinterpret_cast<int32_t*>(base + TC_NA_HDR);
    TcSeg* segs = reinterpret_cast<TcSeg*>(base + TC_NA_HDR + TC_NA_HINTS);
    const int32_t n = g->N, m = g->M;
    const int32_t n_hints = div_up(m, TC_TILE) + 1;
    // the array holds whatever the last call left: counters, flag and segment list start at zero
    GMC_CUDA(cudaMemsetAsync(base, 0, TC_NA_HDR, st));
    GMC_CUDA(cudaMemsetAsync(segs, 0, TC_NA_SEGS, st));
    if (m > 0) {
        // no host-side state, no synchronisation: the ascending-rows check runs on every call and travels with the result
        k_rows_ascending<<<div_up(n, 256), 256, 0, st>>>(g->begin, g->node_idx, n, bad);
        k_rows_ascending<<<div_up(n, 256), 256, 0, st>>>(g->r_begin, g->r_node_idx, n, bad);
        tc_row_hints<<<div_up(n_hints, 256), 256, 0, st>>>(g->begin, n, m, n_hints, hints);
        tc_count<<<sm_count() * 8, TC_THREADS, 0, st>>>(g->begin, g->node_idx, g->r_begin, g->r_node_idx, hints, n, 0, m,
                                                         reinterpret_cast<unsigned int*>(total + 1), total, segs,
                                                         reinterpret_cast<unsigned int*>(total + 1) + 1, TC_NA_SEG_CAP);
        tc_heavy<<<sm_count() * 8, TC_THREADS, 0, st>>>(g->node_idx, g->r_node_idx, segs, reinterpret_cast<unsigned int*>(total + 1) + 1,
                                                         TC_NA_SEG_CAP, reinterpret_cast<unsigned int*>(total + 2), total);
    }
    tc_na_publish<<<1, 1, 0, st>>>(total, bad, count_dev);
    GMC_CUDA(cudaGetLastError());
    return GMC_OK;
}
