// Betweenness centrality: the generated comp_BC() (reference apps/src/bc.gm:4-29) with its two traversal callbacks as
// device kernels.
//
//   Foreach(s: Seeds.Items) {
//     G.sigma = 0; s.sigma = 1;
//     InBFS(v: G.Nodes From s)      { v.sigma = Sum(w: v.UpNbrs)   { w.sigma }; }
//     InReverse                     { v.delta = Sum(w: v.DownNbrs) { v.sigma / w.sigma * (1 + w.delta) }; v.BC += v.delta; }
//   }
//
// gm_comp emits one gm_bfs_template<short, ..., save_child> subclass per InBFS/InReverse pair
// (src/backend_cpp/gm_cpp_gen_bfs.cc:88-221): visit_fw(v) walks the in-edges of v and keeps the sources one level up
// (gm_cpp_gen_foreach.cc:163-175), visit_rv(v) walks the out-edge slots of v that save_child marked as down edges
// (gm_bfs_template.h:496-524: an edge into level + 1, marked by the top-down expansion only), levels in ascending order forward (do_bfs_forward :69-271) and
// descending order backward (do_bfs_reverse :273-312).  Here: the traversal is the device BFS (gmc_bfs.cu), the visited
// vertices are bucketed by level once, then two launches per level and direction.  A warp owns a vertex: it loads 32
// edge slots at a time (coalesced ids, gathered levels and values) and every lane adds the matching terms in slot order
// with __fadd_rn / __fdiv_rn / __fmul_rn -- the float sums are the reference's, term by term, whatever the degree; rows of
// 1024 slots or more are handed to a second kernel (a CTA per row, matching terms compacted into shared memory in slot
// order, one in-order add chain).
//
// Kept reference behaviour: the InBFS body has no (v != s) filter, so the root's sigma = 1 is overwritten by its empty
// UpNbrs sum; with the shipped bc.gm every sigma is 0 and delta is NaN wherever a vertex has a BFS child.
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdlib.h>
#include <vector>
#include "gmc_common.cuh"

using namespace gmc;

// gmc_bfs.cu: forward traversal from `root`, int16 levels (-2 = not reached) left in the graph's BFS workspace
int gmc_bfs_levels_resident(gmc_graph* g, int32_t root, bool save_child, const int16_t** level_dev, const int32_t** level_count,
                            const int32_t** state_trace, int32_t* n_levels, int32_t* launches, cudaStream_t st);

namespace {

constexpr int BC_THREADS = 256;
constexpr int BC_GROUP_DEFAULT = 32;   // lanes per vertex in bc_level (8 and 16 measured within 6 % on LJ-shaped: profiles/r1ag)
constexpr int BC_HEAVY = 1024;         // rows with at least this many edge slots go to bc_heavy (a CTA per row)
constexpr int BCH_THREADS = 1024;      // bc_heavy: threads per CTA (32 warps: the cross-warp scan is one warp wide)
constexpr int BC_PER_THREAD = 8;       // bc_heavy: consecutive slots per thread and chunk
constexpr int BC_CHUNK = BCH_THREADS * BC_PER_THREAD;   // slots per chunk (32 KB of terms in shared memory)
static_assert(BCH_THREADS == 1024, "bc_heavy scans 32 warp totals with one warp");
constexpr unsigned FULLM = 0xffffffffu;

// order[level_begin[l] + k] = the k-th vertex of level l (k in arrival order: vertices of one level do not depend on each other)
__global__ void bc_bucket(const int16_t* __restrict__ level, int32_t n, const int32_t* __restrict__ level_begin,
                          int32_t* __restrict__ cursor, int32_t* __restrict__ order) {
    const int32_t v = blockIdx.x * blockDim.x + threadIdx.x;
    const int lane = threadIdx.x & 31;
    const int l = v < n ? (int)level[v] : -1;
    const unsigned active = __ballot_sync(FULLM, l >= 0);
    if (l < 0) return;
    const unsigned peers = __match_any_sync(active, l);      // one atomic per level present in the warp
    const int leader = __ffs(peers) - 1;
    int32_t base = 0;
    if (lane == leader) base = atomicAdd(&cursor[l], __popc(peers));
    base = __shfl_sync(peers, base, leader);
    order[level_begin[l] + base + __popc(peers & ((1u << lane) - 1u))] = v;
}

__global__ void bc_seed(float* __restrict__ sigma, int32_t s) { sigma[s] = 1.0f; }

// One term of a callback's sum.  REV = false: visit_fw, sigma[w] of an in-neighbour on level `want` (= level - 1);
// REV = true: visit_rv, sigma[v] / sigma[w] * (1 + delta[w]) of an out-neighbour on level `want` (= level + 1).
template <bool REV>
__device__ __forceinline__ float bc_term(int32_t w, const int16_t* __restrict__ level, int want, const float* __restrict__ sigma,
                                         const float* __restrict__ delta, float sv, bool& match) {
    match = (int)level[w] == want;
    if (!match) return 0.0f;
    if (REV) return __fmul_rn(__fdiv_rn(sv, sigma[w]), __fadd_rn(1.0f, delta[w]));
    return sigma[w];
}

template <bool REV>
__device__ __forceinline__ void bc_store(int32_t v, float S, float* __restrict__ sigma, float* __restrict__ delta, float* __restrict__ bc) {
    if (REV) { delta[v] = S; bc[v] = __fadd_rn(bc[v], S); }      // v.delta = Sum ...; v.BC += v.delta
    else sigma[v] = S;                                           // v.sigma = Sum ...
}

// The callback of every vertex of one level (order[lo, hi)), BC_GROUP lanes per vertex (32 / BC_GROUP vertices per warp):
// BC_GROUP edge slots per step, the matching terms added in slot order by every lane of the group.  Rows of heavy_min
// slots or more are queued for bc_heavy instead.
template <bool REV, int BC_GROUP>
__global__ void __launch_bounds__(BC_THREADS)
bc_level(const int32_t* __restrict__ row_begin, const int32_t* __restrict__ row_idx, const int16_t* __restrict__ level,
         const int32_t* __restrict__ order, int32_t lo, int32_t hi, int want, float* __restrict__ sigma, float* __restrict__ delta,
         float* __restrict__ bc, int32_t* __restrict__ heavy_list, int32_t* __restrict__ heavy_count, int32_t heavy_min) {
    const int lane = threadIdx.x & 31, gl = lane & (BC_GROUP - 1), grp = lane / BC_GROUP;
    constexpr int GPW = 32 / BC_GROUP;                       // groups per warp
    const int32_t nw = (gridDim.x * BC_THREADS) >> 5;
    for (int32_t base = lo + ((blockIdx.x * BC_THREADS + threadIdx.x) >> 5) * GPW; base < hi; base += nw * GPW) {   // warp-uniform
        const int32_t i = base + grp;
        bool mine = i < hi;
        int32_t v = 0, b = 0, e = 0;
        if (mine) {
            v = order[i];
            b = row_begin[v];
            e = row_begin[v + 1];
            if (e - b >= heavy_min) {
                if (gl == 0 && heavy_count) heavy_list[atomicAdd(heavy_count, 1)] = v;      // NULL: the long rows are already listed
                mine = false;
                e = b;
            }
        }
        const float sv = (REV && mine) ? sigma[v] : 0.0f;
        int steps = (e - b + BC_GROUP - 1) / BC_GROUP;
#pragma unroll
        for (int o = BC_GROUP; o < 32; o <<= 1) steps = max(steps, __shfl_xor_sync(FULLM, steps, o));
        float S = 0.0f;
        for (int t = 0; t < steps; t++) {
            const int32_t q = b + t * BC_GROUP + gl;
            float x = 0.0f;
            bool m = false;
            if (q < e) x = bc_term<REV>(ld_stream(row_idx + q), level, want, sigma, delta, sv, m);
            unsigned gm = (__ballot_sync(FULLM, m) >> (grp * BC_GROUP)) & (BC_GROUP == 32 ? FULLM : (1u << (BC_GROUP & 31)) - 1u);   // the group's matches
            while (__any_sync(FULLM, gm != 0u)) {            // slot order = the reference's loop order
                const int j = gm ? __ffs(gm) - 1 : 0;
                const float xj = __shfl_sync(FULLM, x, grp * BC_GROUP + j);
                if (gm) S = __fadd_rn(S, xj);
                gm &= gm - 1u;
            }
        }
        if (mine && gl == 0) bc_store<REV>(v, S, sigma, delta, bc);
    }
}

// The long rows of the level, a CTA per row.  Per chunk of BC_CHUNK slots every thread turns 8 consecutive slots into
// terms (all loads of a chunk in flight at once), the matching ones are compacted in slot order into shared memory (one
// block-wide exclusive scan of the per-thread match counts), then the first warp adds them in that order: one fadd (4
// clocks) per matching term instead of a memory round trip per 32 slots.  The chain itself is what bit-identical float
// sums cost; everything around it is parallel.
template <bool REV>
__global__ void __launch_bounds__(BCH_THREADS)
bc_heavy(const int32_t* __restrict__ row_begin, const int32_t* __restrict__ row_idx, const int16_t* __restrict__ level, int want,
         float* __restrict__ sigma, float* __restrict__ delta, float* __restrict__ bc, const int32_t* __restrict__ heavy_list,
         const int32_t* __restrict__ heavy_count, int only_level) {
    __shared__ __align__(16) float xs[BC_CHUNK + 4];
    __shared__ int32_t wsum[BCH_THREADS / 32];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int32_t n = *heavy_count;
    for (int32_t h = blockIdx.x; h < n; h += gridDim.x) {
        const int32_t v = heavy_list[h];
        if (only_level >= 0 && (int)level[v] != only_level) continue;      // list of the long rows of all levels (overlap mode)
        const int32_t b = row_begin[v], e = row_begin[v + 1];
        const float sv = REV ? sigma[v] : 0.0f;
        float S = 0.0f;
        for (int32_t p = b; p < e; p += BC_CHUNK) {
            float x[BC_PER_THREAD];
            unsigned mbits = 0u;
#pragma unroll
            for (int k = 0; k < BC_PER_THREAD; k++) {
                const int32_t q = p + tid * BC_PER_THREAD + k;
                bool m = false;
                x[k] = q < e ? bc_term<REV>(ld_stream(row_idx + q), level, want, sigma, delta, sv, m) : 0.0f;
                if (m) mbits |= 1u << k;
            }
            const int c = __popc(mbits);
            int incl = c;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) { const int t = __shfl_up_sync(FULLM, incl, o); if (lane >= o) incl += t; }
            if (lane == 31) wsum[warp] = incl;
            __syncthreads();
            const int wv = wsum[lane];                        // BCH_THREADS / 32 == 32 warps
            int wincl = wv;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) { const int t = __shfl_up_sync(FULLM, wincl, o); if (lane >= o) wincl += t; }
            const int total = __shfl_sync(FULLM, wincl, 31);
            int pos = __shfl_sync(FULLM, wincl - wv, warp) + incl - c;
#pragma unroll
            for (int k = 0; k < BC_PER_THREAD; k++)
                if ((mbits >> k) & 1u) xs[pos++] = x[k];
            if (tid < 4) xs[total + tid] = 0.0f;              // the chain reads whole float4 groups: + (+0) changes nothing
            __syncthreads();
            if (warp == 0) {
                const float4* x4 = reinterpret_cast<const float4*>(xs);
#pragma unroll 4
                for (int k = 0; k < (total + 3) >> 2; k++) {
                    const float4 t = x4[k];
                    S = __fadd_rn(__fadd_rn(__fadd_rn(__fadd_rn(S, t.x), t.y), t.z), t.w);
                }
            }
            __syncthreads();
        }
        if (tid == 0) bc_store<REV>(v, S, sigma, delta, bc);
    }
}

inline unsigned level_grid(int32_t count, int group) {
    const int64_t want = div_up((int64_t)count * group, (int64_t)BC_THREADS);
    const int64_t cap = (int64_t)sm_count() * 8;
    return (unsigned)(want < cap ? (want > 0 ? want : 1) : cap);
}

// overlap mode: the long rows by direction (in-degree for visit_fw, out-degree for visit_rv), once per call
__global__ void bc_classify(const int32_t* __restrict__ begin, const int32_t* __restrict__ r_begin, int32_t n, int32_t heavy_min,
                            int32_t* __restrict__ fw_list, int32_t* __restrict__ rv_list, int32_t* __restrict__ counts) {
    const int32_t v = blockIdx.x * blockDim.x + threadIdx.x;
    if (v >= n) return;
    if (r_begin[v + 1] - r_begin[v] >= heavy_min) fw_list[atomicAdd(&counts[0], 1)] = v;
    if (begin[v + 1] - begin[v] >= heavy_min) rv_list[atomicAdd(&counts[1], 1)] = v;
}

template <bool REV>
void launch_level(int group, unsigned grid, cudaStream_t st, const int32_t* rb, const int32_t* ri, const int16_t* level, const int32_t* order,
                  int32_t lo, int32_t hi, int want, float* sigma, float* delta, float* bc, int32_t* hl, int32_t* hc, int32_t heavy_min) {
    if (group == 32) bc_level<REV, 32><<<grid, BC_THREADS, 0, st>>>(rb, ri, level, order, lo, hi, want, sigma, delta, bc, hl, hc, heavy_min);
    else if (group == 16) bc_level<REV, 16><<<grid, BC_THREADS, 0, st>>>(rb, ri, level, order, lo, hi, want, sigma, delta, bc, hl, hc, heavy_min);
    else bc_level<REV, 8><<<grid, BC_THREADS, 0, st>>>(rb, ri, level, order, lo, hi, want, sigma, delta, bc, hl, hc, heavy_min);
}

struct BcBuffers {
    float *bc = nullptr, *sigma = nullptr, *delta = nullptr;
    int32_t *order = nullptr, *level_begin = nullptr, *cursor = nullptr, *heavy_list = nullptr, *heavy_count = nullptr;
    int32_t *rv_list = nullptr, *cls_count = nullptr;       // overlap mode
    cudaStream_t s1 = nullptr, s2 = nullptr;
    cudaEvent_t join = nullptr, heavy_done = nullptr;
    ~BcBuffers() {
        cudaFree(bc); cudaFree(sigma); cudaFree(delta); cudaFree(order); cudaFree(level_begin); cudaFree(cursor);
        cudaFree(heavy_list); cudaFree(heavy_count); cudaFree(rv_list); cudaFree(cls_count);
        if (join) cudaEventDestroy(join);
        if (heavy_done) cudaEventDestroy(heavy_done);
        if (s1) cudaStreamDestroy(s1);
        if (s2) cudaStreamDestroy(s2);
    }
};

}  // namespace

// first_level = 0: bc.gm; 1: bc_adj.gm, whose (v != s) filters take the root -- the only vertex of level 0 -- out of both passes
static int comp_bc(gmc_graph* g, const int32_t* seeds, int32_t n_seeds, float* bc_host, gmc_bc_info* info, int32_t first_level) {
    GMC_REQUIRE(g && (seeds || n_seeds == 0) && n_seeds >= 0 && (bc_host || g->N == 0), GMC_ERR_ARG, "gmc_comp_bc: bad args");
    if (info) memset(info, 0, sizeof(*info));
    if (g->N == 0) return GMC_OK;
    GMC_REQUIRE(g->has_forward && g->has_reverse, GMC_ERR_STATE, "gmc_comp_bc: needs the forward and the reverse CSR (gmc_graph_make_reverse)");
    for (int32_t k = 0; k < n_seeds; k++)
        GMC_REQUIRE(seeds[k] >= 0 && seeds[k] < g->N, GMC_ERR_ARG, "gmc_comp_bc: seed %d out of range [0,%d)", seeds[k], g->N);
    const size_t n = (size_t)g->N;
    constexpr int32_t MAX_LEVELS = 32768;                 // level_t = short
    cudaStream_t st = nullptr;
    BcBuffers B;
    GMC_TRY(gmc_malloc((void**)&B.bc, sizeof(float) * n));
    GMC_TRY(gmc_malloc((void**)&B.sigma, sizeof(float) * n));
    GMC_TRY(gmc_malloc((void**)&B.delta, sizeof(float) * n));
    GMC_TRY(gmc_malloc((void**)&B.order, sizeof(int32_t) * n));
    GMC_TRY(gmc_malloc((void**)&B.level_begin, sizeof(int32_t) * (MAX_LEVELS + 1)));
    GMC_TRY(gmc_malloc((void**)&B.cursor, sizeof(int32_t) * MAX_LEVELS));
    GMC_TRY(gmc_malloc((void**)&B.heavy_list, sizeof(int32_t) * n));
    GMC_TRY(gmc_malloc((void**)&B.heavy_count, sizeof(int32_t) * 2 * MAX_LEVELS));      // one counter per level and direction
    const unsigned heavy_grid = (unsigned)sm_count() * 2;
    // experiment knobs (tools/bc_bench.py): lanes per vertex (8 | 16 | 32) and the row length from which bc_heavy takes over
    int group = BC_GROUP_DEFAULT, heavy_min = BC_HEAVY;
    if (const char* t = getenv("GMC_BC_GROUP")) { const int v = atoi(t); if (v == 8 || v == 16 || v == 32) group = v; }
    if (const char* t = getenv("GMC_BC_HEAVY")) { const int v = atoi(t); if (v >= 32) heavy_min = v; }
    // GMC_BC_OVERLAP=1 (experimental, not yet measured): the long rows are listed once per call by degree, and bc_level /
    // bc_heavy of a level -- independent of each other -- run on two streams that join after every level
    const bool overlap = getenv("GMC_BC_OVERLAP") != nullptr && atoi(getenv("GMC_BC_OVERLAP")) != 0;
    if (overlap) {
        GMC_CUDA(cudaStreamCreateWithFlags(&B.s1, cudaStreamNonBlocking));
        GMC_CUDA(cudaStreamCreateWithFlags(&B.s2, cudaStreamNonBlocking));
        GMC_CUDA(cudaEventCreateWithFlags(&B.join, cudaEventDisableTiming));
        GMC_CUDA(cudaEventCreateWithFlags(&B.heavy_done, cudaEventDisableTiming));
        GMC_TRY(gmc_malloc((void**)&B.rv_list, sizeof(int32_t) * n));
        GMC_TRY(gmc_malloc((void**)&B.cls_count, sizeof(int32_t) * 2));
        st = B.s1;
    }
    std::vector<int32_t> lb;                          // level offsets of the current seed
    const int16_t* level_cur = nullptr;               // ... its levels and level counts (run_level reads them)
    const int32_t* level_count_cur = nullptr;
    cudaEvent_t ev0, ev1;
    GMC_CUDA(cudaEventCreate(&ev0));
    GMC_CUDA(cudaEventCreate(&ev1));
    GMC_CUDA(cudaEventRecord(ev0, st));
    GMC_CUDA(cudaMemsetAsync(B.bc, 0, sizeof(float) * n, st));
    int32_t launches = 0, max_levels = 0;
    if (overlap) {
        GMC_CUDA(cudaMemsetAsync(B.cls_count, 0, sizeof(int32_t) * 2, st));
        bc_classify<<<(unsigned)div_up(n, (size_t)BC_THREADS), BC_THREADS, 0, st>>>(g->begin, g->r_begin, g->N, heavy_min, B.heavy_list,
                                                                                   B.rv_list, B.cls_count);
        launches++;
    }
    // one level of one direction: the short rows on `st`; the long rows behind them (default) or beside them (overlap)
    auto run_level = [&](bool rev, int32_t l, int want) -> bool {
        const int32_t* rb = rev ? g->begin : g->r_begin;
        const int32_t* ri = rev ? g->node_idx : g->r_node_idx;
        const unsigned grid = level_grid(level_count_cur[l], group);
        int32_t* queue_count = overlap ? nullptr : B.heavy_count + 2 * l + (rev ? 1 : 0);
        const int32_t* list = overlap ? (rev ? B.rv_list : B.heavy_list) : B.heavy_list;
        const int32_t* list_count = overlap ? B.cls_count + (rev ? 1 : 0) : queue_count;
        cudaStream_t hs = st;
        if (overlap) {
            hs = B.s2;
            if (cudaEventRecord(B.join, st) != cudaSuccess || cudaStreamWaitEvent(hs, B.join, 0) != cudaSuccess) return false;
        }
        if (rev) launch_level<true>(group, grid, st, rb, ri, level_cur, B.order, lb[l], lb[l + 1], want, B.sigma, B.delta, B.bc, B.heavy_list, queue_count, heavy_min);
        else launch_level<false>(group, grid, st, rb, ri, level_cur, B.order, lb[l], lb[l + 1], want, B.sigma, B.delta, B.bc, B.heavy_list, queue_count, heavy_min);
        if (rev) bc_heavy<true><<<heavy_grid, BCH_THREADS, 0, hs>>>(rb, ri, level_cur, want, B.sigma, B.delta, B.bc, list, list_count, overlap ? l : -1);
        else bc_heavy<false><<<heavy_grid, BCH_THREADS, 0, hs>>>(rb, ri, level_cur, want, B.sigma, B.delta, B.bc, list, list_count, overlap ? l : -1);
        if (overlap && (cudaEventRecord(B.heavy_done, hs) != cudaSuccess || cudaStreamWaitEvent(st, B.heavy_done, 0) != cudaSuccess)) return false;
        launches += 2;
        return true;
    };
    int rc = GMC_OK;
    for (int32_t k = 0; k < n_seeds && rc == GMC_OK; k++) {
        const int16_t* level = nullptr;
        const int32_t *level_count = nullptr, *state = nullptr;
        int32_t nl = 0, bfs_launches = 0;
        rc = gmc_bfs_levels_resident(g, seeds[k], true, &level, &level_count, &state, &nl, &bfs_launches, st);
        if (rc != GMC_OK) break;
        launches += bfs_launches;
        level_cur = level;
        level_count_cur = level_count;
        if (nl > max_levels) max_levels = nl;
        lb.assign((size_t)nl + 1, 0);
        for (int32_t l = 0; l < nl; l++) lb[l + 1] = lb[l] + level_count[l];
        // the stream is idle here (the traversal ends with a synchronisation), so the pageable copy is safe to reuse per seed
        if (cudaMemcpyAsync(B.level_begin, lb.data(), sizeof(int32_t) * (size_t)(nl + 1), cudaMemcpyHostToDevice, st) != cudaSuccess ||
            cudaMemsetAsync(B.cursor, 0, sizeof(int32_t) * (size_t)nl, st) != cudaSuccess ||
            cudaMemsetAsync(B.heavy_count, 0, sizeof(int32_t) * 2 * (size_t)nl, st) != cudaSuccess ||
            cudaMemsetAsync(B.sigma, 0, sizeof(float) * n, st) != cudaSuccess) {
            set_error("gmc_comp_bc: %s", cudaGetErrorString(cudaGetLastError()));
            rc = GMC_ERR_CUDA;
            break;
        }
        bc_seed<<<1, 1, 0, st>>>(B.sigma, seeds[k]);
        bc_bucket<<<(unsigned)div_up(n, (size_t)BC_THREADS), BC_THREADS, 0, st>>>(level, g->N, B.level_begin, B.cursor, B.order);
        launches += 2;
        bool ok = true;
        for (int32_t l = first_level; l < nl && ok; l++) ok = run_level(false, l, l - 1);
        for (int32_t l = nl - 1; l >= first_level && ok; l--) {
            // a level the traversal expanded bottom-up has no down edges in the reference: check_parent_rrd
            // (gm_bfs_template.h:689-708) does not mark them; its vertices get delta = 0 (level -1 matches nothing)
            const bool marked = state[l] != GMC_BFS_ST_RRD && state[l] != GMC_BFS_ST_RR2Q;
            ok = run_level(true, l, marked ? l + 1 : -1);
        }
        if (!ok) { set_error("gmc_comp_bc: %s", cudaGetErrorString(cudaGetLastError())); rc = GMC_ERR_CUDA; break; }
        cudaError_t ce = cudaGetLastError();
        if (ce == cudaSuccess) ce = cudaStreamSynchronize(st);           // lb is rewritten for the next seed
        if (ce != cudaSuccess) { set_error("gmc_comp_bc: %s", cudaGetErrorString(ce)); rc = GMC_ERR_CUDA; }
    }
    if (rc == GMC_OK && cudaEventRecord(ev1, st) != cudaSuccess) rc = GMC_ERR_CUDA;
    if (rc == GMC_OK) rc = gmc_d2h(bc_host, B.bc, sizeof(float) * n, st);
    if (rc == GMC_OK && info) {
        info->seeds = n_seeds;
        info->launches = launches;
        info->max_levels = max_levels;
        cudaEventElapsedTime(&info->device_ms, ev0, ev1);
    }
    cudaEventDestroy(ev0);
    cudaEventDestroy(ev1);
    return rc;
}

extern "C" int gmc_comp_bc(gmc_graph* g, const int32_t* seeds, int32_t n_seeds, float* bc_host, gmc_bc_info* info) {
    return comp_bc(g, seeds, n_seeds, bc_host, info, 0);
}
extern "C" int gmc_comp_bc_adj(gmc_graph* g, const int32_t* seeds, int32_t n_seeds, float* bc_host, gmc_bc_info* info) {
    return comp_bc(g, seeds, n_seeds, bc_host, info, 1);
}
