// Level-synchronous, direction-optimizing, frontier-compacted BFS (K6/K7/K8).
//
// One engine serves both reference entry points:
//   * generated hop_dist() (apps/src/hop_dist.gm:3-31): dist int32, INT_MAX when unreachable.  The
//     reference re-scans all N vertices twice per level; the device keeps a compacted
//     frontier instead -- the resulting dist[] is the same BFS distance.
//   * gm_bfs_template<short,...>::do_bfs_forward (inc/gm_bfs_template.h:69-271): level int16, -2 unvisited.
// The per-level state machine is the reference's (get_next_state, gm_bfs_template.h:352-421, thresholds
// :359,:720-721) so the direction switches happen on the same levels as on the CPU.  Device realisation
// of the seven states:
//   SMALL, QUE, Q2R, RD, R2Q  -> top-down: expand the current level's queue, test-and-set the visited
//                                bitmap with atomicOr (the byte fetch_and_or of gm_bitmap.cc:20-27 on
//                                32-bit words), write the level, append to the next queue
//                                (warp-aggregated).  The CPU-only "marks without queue" variants
//                                (Q2R/RD) still emit the queue here: one aggregated atomic per warp.
//   RRD, RR2Q                 -> bottom-up: one lane per vertex, one warp per visited word; an
//                                unvisited vertex scans its parents until one is in the frontier
//                                BITMAP of the current level (check_parent_rrd :689-708 reads
//                                visited_level[u]==curr; the bitmap carries the same predicate in N/8
//                                bytes).  The warp publishes its 32 results with one plain store:
//                                no atomics.  RR2Q additionally emits the queue.
#include "gmc_common.cuh"
#include <limits.h>
#include <vector>

using namespace gmc;

enum { ST_SMALL = 0, ST_QUE = 1, ST_Q2R = 2, ST_RD = 3, ST_R2Q = 4, ST_RRD = 5, ST_RR2Q = 6 };
constexpr int BFS_THREADS = 256;

struct gmc_bfs_ws {
    int32_t N = 0;
    int32_t words = 0;
    uint32_t* visited = nullptr;      // [words]
    uint32_t* frontier_bm = nullptr;  // [words] bitmap of the current level (bottom-up input)
    uint32_t* next_bm = nullptr;      // [words] bitmap of the next level (bottom-up output)
    int32_t* queue = nullptr;         // [N] all level queues back to back (global_vector)
    int32_t* counters = nullptr;      // [4]: 0 = next_count
    void* level = nullptr;            // [N] int32 (also used as int16)
    unsigned long long* edge_sum = nullptr;
    // host-side record of the last traversal
    std::vector<int32_t> level_count, level_queue_begin, state_trace;
    bool last_is_i16 = false;
    bool valid = false;
};

void gmc_bfs_ws_free(gmc_graph* g) {
    gmc_bfs_ws* w = g->bfs;
    if (!w) return;
    cudaFree(w->visited); cudaFree(w->frontier_bm); cudaFree(w->next_bm); cudaFree(w->queue);
    cudaFree(w->counters); cudaFree(w->level); cudaFree(w->edge_sum);
    delete w;
    g->bfs = nullptr;
}

namespace {

template <typename L>
__global__ void bfs_init(L* __restrict__ level, int32_t n, L invalid, int32_t root, uint32_t* __restrict__ visited,
                         int32_t words, int32_t* __restrict__ queue, int32_t* __restrict__ counters) {
    int32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) level[i] = (i == root) ? (L)0 : invalid;
    if (i < words) visited[i] = (i == (root >> 5)) ? (1u << (root & 31)) : 0u;
    if (i == 0) { queue[0] = root; counters[0] = 0; }
}

// append `v` (when `take`) to the next-level queue with one atomic per warp
__device__ __forceinline__ void warp_append(bool take, int32_t v, int32_t* __restrict__ out, int32_t* __restrict__ count) {
    unsigned mask = __ballot_sync(0xffffffffu, take);   // callers keep the warp converged here
    if (!take) return;
    int lane = threadIdx.x & 31;
    int leader = __ffs(mask) - 1;
    int32_t base = 0;
    if (lane == leader) base = atomicAdd(count, __popc(mask));
    base = __shfl_sync(mask, base, leader);
    out[base + __popc(mask & ((1u << lane) - 1))] = v;
}

// top-down step: CTA takes BFS_THREADS frontier vertices; their edges are expanded cooperatively
// (prefix sum of degrees + binary search) so a hub does not serialise on one thread.
template <typename L>
__global__ void __launch_bounds__(BFS_THREADS)
bfs_topdown(const int32_t* __restrict__ begin, const int32_t* __restrict__ node_idx, const int32_t* __restrict__ cur_queue,
            int32_t cur_count, L next_level, L* __restrict__ level, uint32_t* __restrict__ visited,
            int32_t* __restrict__ next_queue, int32_t* __restrict__ next_count, const uint8_t* __restrict__ edge_mask) {
    __shared__ int32_t s_start[BFS_THREADS];
    __shared__ int32_t s_off[BFS_THREADS + 1];
    __shared__ int32_t s_warp[BFS_THREADS / 32 + 1];
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    const int32_t qi = blockIdx.x * BFS_THREADS + tid;
    int32_t deg = 0, st = 0;
    if (qi < cur_count) {
        int32_t v = cur_queue[qi];
        st = begin[v];
        deg = begin[v + 1] - st;
    }
    // block exclusive scan of deg
    int32_t inc = deg;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        int32_t t = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= o) inc += t;
    }
    if (lane == 31) s_warp[wid] = inc;
    __syncthreads();
    if (tid == 0) {
        int32_t run = 0;
        for (int w = 0; w < BFS_THREADS / 32; w++) { int32_t t = s_warp[w]; s_warp[w] = run; run += t; }
        s_warp[BFS_THREADS / 32] = run;
    }
    __syncthreads();
    s_start[tid] = st;
    s_off[tid] = inc - deg + s_warp[wid];
    const int32_t total = s_warp[BFS_THREADS / 32];
    if (tid == 0) s_off[BFS_THREADS] = total;
    __syncthreads();

    for (int32_t j0 = 0; j0 < total; j0 += BFS_THREADS) {
        int32_t j = j0 + tid;
        bool won = false;
        int32_t u = 0;
        if (j < total) {
            int lo = 0, hi = BFS_THREADS;   // s_off[lo] <= j < s_off[hi]
            while (hi - lo > 1) {
                int mid = (lo + hi) >> 1;
                if (s_off[mid] <= j) lo = mid; else hi = mid;
            }
            const size_t eidx = (size_t)s_start[lo] + (j - s_off[lo]);
            u = ld_stream(node_idx + eidx);
            const uint32_t bit = 1u << (u & 31);
            // navigator (check_navigator, gm_bfs_template.h:488-492,596-598): filtered edges are not followed
            if (!(visited[u >> 5] & bit) && (!edge_mask || edge_mask[eidx])) {                     // test ...
                uint32_t old = atomicOr(&visited[u >> 5], bit);     // ... and test-and-set
                won = !(old & bit);
                if (won) level[u] = next_level;
            }
        }
        warp_append(won, u, next_queue, next_count);
    }
}

// bottom-up step: warp w owns vertices [32w, 32w+32) = one word of the bitmaps
template <typename L, bool EMIT>
__global__ void __launch_bounds__(BFS_THREADS)
bfs_bottomup(const int32_t* __restrict__ r_begin, const int32_t* __restrict__ r_node_idx, int32_t n,
             const uint32_t* __restrict__ frontier_bm, uint32_t* __restrict__ visited, uint32_t* __restrict__ next_bm,
             L next_level, L* __restrict__ level, int32_t* __restrict__ next_queue, int32_t* __restrict__ next_count) {
    const int lane = threadIdx.x & 31;
    const int32_t w = (blockIdx.x * BFS_THREADS + threadIdx.x) >> 5;
    const int32_t words = (n + 31) >> 5;
    if (w >= words) return;
    const uint32_t vis = visited[w];
    const int32_t t = (w << 5) + lane;
    bool found = false;
    if (t < n && !((vis >> lane) & 1u)) {
        const int32_t b = r_begin[t], e = r_begin[t + 1];
        for (int32_t nx = b; nx < e; nx++) {
            int32_t u = r_node_idx[nx];
            if ((frontier_bm[u >> 5] >> (u & 31)) & 1u) { found = true; break; }
        }
        if (found) level[t] = next_level;
    }
    const uint32_t got = __ballot_sync(0xffffffffu, found);
    if (lane == 0) {
        next_bm[w] = got;
        if (got) visited[w] = vis | got;
    }
    if (EMIT) {
        warp_append(found, t, next_queue, next_count);
    } else if (lane == 0 && got) {
        atomicAdd(next_count, __popc(got));
    }
}

__global__ void queue_to_bitmap(const int32_t* __restrict__ queue, int32_t count, uint32_t* __restrict__ bm) {
    int32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < count) { int32_t v = queue[i]; atomicOr(&bm[v >> 5], 1u << (v & 31)); }
}

template <typename L>
__global__ void sum_reached_degree(const int32_t* __restrict__ begin, const L* __restrict__ level, L invalid, int32_t n,
                                   unsigned long long* __restrict__ out) {
    int32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    int v = 0;
    if (i < n && level[i] != invalid) v = begin[i + 1] - begin[i];
    v = warp_sum(v);
    if ((threadIdx.x & 31) == 0 && v) atomicAdd(out, (unsigned long long)v);
}

int ensure_ws(gmc_graph* g) {
    if (g->bfs) return GMC_OK;
    gmc_bfs_ws* w = new gmc_bfs_ws();
    w->N = g->N;
    w->words = (g->N + 31) / 32;
    size_t wb = sizeof(uint32_t) * (size_t)(w->words ? w->words : 1);
    int rc = gmc_malloc((void**)&w->visited, wb);
    if (rc == GMC_OK) rc = gmc_malloc((void**)&w->frontier_bm, wb);
    if (rc == GMC_OK) rc = gmc_malloc((void**)&w->next_bm, wb);
    if (rc == GMC_OK) rc = gmc_malloc((void**)&w->queue, sizeof(int32_t) * (size_t)(g->N ? g->N : 1));
    if (rc == GMC_OK) rc = gmc_malloc((void**)&w->counters, sizeof(int32_t) * 4);
    if (rc == GMC_OK) rc = gmc_malloc((void**)&w->level, sizeof(int32_t) * (size_t)(g->N ? g->N : 1));
    if (rc == GMC_OK) rc = gmc_malloc((void**)&w->edge_sum, sizeof(unsigned long long));
    g->bfs = w;
    if (rc != GMC_OK) { gmc_bfs_ws_free(g); return rc; }
    return GMC_OK;
}

// The traversal.  fwd_* = arrays followed top-down, rev_* = arrays followed bottom-up (NULL: no bottom-up).
template <typename L>
int run_bfs(gmc_graph* g, const int32_t* fwd_begin, const int32_t* fwd_idx, const int32_t* rev_begin, const int32_t* rev_idx,
            int32_t root, L invalid, L* level_out_dev, gmc_bfs_info* info, cudaStream_t st,
            const uint8_t* edge_mask = nullptr) {
    GMC_REQUIRE(root >= 0 && root < g->N, GMC_ERR_ARG, "bfs: root %d out of range [0,%d)", root, g->N);
    GMC_TRY(ensure_ws(g));
    gmc_bfs_ws* w = g->bfs;
    const int32_t n = g->N;
    const bool has_reverse = rev_begin != nullptr;
    L* level = level_out_dev ? level_out_dev : (L*)w->level;
    w->level_count.clear(); w->level_queue_begin.clear(); w->state_trace.clear();
    w->valid = false;

    cudaEvent_t ev0, ev1;
    GMC_CUDA(cudaEventCreate(&ev0));
    GMC_CUDA(cudaEventCreate(&ev1));
    int launches = 0, bottom_up = 0;
    GMC_CUDA(cudaEventRecord(ev0, st));
    bfs_init<L><<<div_up(n, 256), 256, 0, st>>>(level, n, invalid, root, w->visited, w->words, w->queue, w->counters);
    launches++;

    int state = ST_SMALL;
    int32_t curr_level = 0, curr_begin = 0, next_begin = 1, curr_count = 1, next_count = 0;
    bool have_queue = true;       // queue[curr_begin, +curr_count) holds the current level
    bool have_bitmap = false;     // frontier_bm holds the current level
    const float thr = (float)n * 0.05f;   // RRD_THRESHOLD, gm_bfs_template.h:359 (float arithmetic as in the reference)
    for (;;) {
        w->state_trace.push_back(state);
        w->level_count.push_back(curr_count);
        w->level_queue_begin.push_back(have_queue ? curr_begin : -1);
        GMC_REQUIRE(sizeof(L) == 4 || curr_level + 1 <= 32767, GMC_ERR_STATE, "bfs: more levels than level_t can hold");
        const L nl = (L)(curr_level + 1);
        if (state == ST_RRD || state == ST_RR2Q) {
            if (!have_bitmap) {   // previous level was top-down: build its bitmap from the queue
                GMC_CUDA(cudaMemsetAsync(w->frontier_bm, 0, sizeof(uint32_t) * (size_t)w->words, st));
                queue_to_bitmap<<<div_up(curr_count, 256), 256, 0, st>>>(w->queue + curr_begin, curr_count, w->frontier_bm);
                launches++;
            }
            const unsigned grid = (unsigned)div_up((int64_t)w->words * 32, (int64_t)BFS_THREADS);
            if (state == ST_RR2Q)
                bfs_bottomup<L, true><<<grid, BFS_THREADS, 0, st>>>(rev_begin, rev_idx, n, w->frontier_bm, w->visited, w->next_bm, nl, level, w->queue + next_begin, w->counters);
            else
                bfs_bottomup<L, false><<<grid, BFS_THREADS, 0, st>>>(rev_begin, rev_idx, n, w->frontier_bm, w->visited, w->next_bm, nl, level, w->queue + next_begin, w->counters);
            launches++; bottom_up++;
        } else {
            bfs_topdown<L><<<div_up(curr_count, BFS_THREADS), BFS_THREADS, 0, st>>>(fwd_begin, fwd_idx, w->queue + curr_begin, curr_count, nl, level, w->visited, w->queue + next_begin, w->counters, edge_mask);
            launches++;
        }
        GMC_CUDA(cudaGetLastError());
        GMC_CUDA(cudaMemcpyAsync(&next_count, w->counters, sizeof(int32_t), cudaMemcpyDeviceToHost, st));
        GMC_CUDA(cudaMemsetAsync(w->counters, 0, sizeof(int32_t), st));
        GMC_CUDA(cudaStreamSynchronize(st));

        // get_next_state (gm_bfs_template.h:352-421)
        if (next_count == 0) break;
        int next_state = state;
        switch (state) {
            case ST_SMALL: if (next_count >= 128) next_state = ST_QUE; break;
            case ST_QUE:
                if ((float)next_count >= thr) next_state = has_reverse ? ST_RRD : ST_RD;
                else if (next_count >= 1024 && next_count >= curr_count * 5) next_state = ST_Q2R;
                break;
            case ST_Q2R: next_state = (has_reverse && (float)next_count >= thr) ? ST_RRD : ST_RD; break;
            case ST_RD:
                if (has_reverse && (float)next_count >= thr) next_state = ST_RRD;
                else if (next_count <= 2 * curr_count) next_state = ST_R2Q;
                break;
            case ST_R2Q: next_state = ST_QUE; break;
            case ST_RRD: if (next_count <= 2 * curr_count) next_state = ST_RR2Q; break;
            case ST_RR2Q: next_state = ST_QUE; break;
        }
        // finish_level (:423-447).  Device: every top-down state and RR2Q emitted the queue; RRD did not.
        const bool emitted_queue = (state != ST_RRD);
        const bool emitted_bitmap = (state == ST_RRD || state == ST_RR2Q);
        if (emitted_queue) { curr_begin = next_begin; next_begin += next_count; }
        have_queue = emitted_queue;
        have_bitmap = emitted_bitmap;
        if (emitted_bitmap) { uint32_t* t = w->frontier_bm; w->frontier_bm = w->next_bm; w->next_bm = t; }
        curr_count = next_count;
        next_count = 0;
        curr_level++;
        state = next_state;
    }
    GMC_CUDA(cudaEventRecord(ev1, st));
    GMC_CUDA(cudaStreamSynchronize(st));
    w->valid = true;
    w->last_is_i16 = sizeof(L) == 2;
    if (info) {
        info->levels = curr_level + 1;
        int32_t reached = 0;
        for (int32_t c : w->level_count) reached += c;
        info->reached = reached;
        info->launches = launches;
        info->bottom_up_levels = bottom_up;
        cudaEventElapsedTime(&info->kernel_ms, ev0, ev1);
        GMC_CUDA(cudaMemsetAsync(w->edge_sum, 0, sizeof(unsigned long long), st));
        sum_reached_degree<L><<<div_up(n, 256), 256, 0, st>>>(fwd_begin, level, invalid, n, w->edge_sum);
        unsigned long long es = 0;
        GMC_CUDA(cudaMemcpyAsync(&es, w->edge_sum, sizeof(es), cudaMemcpyDeviceToHost, st));
        GMC_CUDA(cudaStreamSynchronize(st));
        info->edges_traversed = (int64_t)es;
    }
    cudaEventDestroy(ev0); cudaEventDestroy(ev1);
    return GMC_OK;
}

}  // namespace

extern "C" {

int gmc_hop_dist_dev(gmc_graph* g, int32_t root, int32_t* dist_dev, gmc_bfs_info* info, void* stream) {
    GMC_REQUIRE(g && dist_dev, GMC_ERR_ARG, "gmc_hop_dist_dev: NULL");
    GMC_REQUIRE(g->N > 0, GMC_ERR_ARG, "gmc_hop_dist_dev: empty graph");
    return run_bfs<int32_t>(g, g->begin, g->node_idx, g->has_reverse ? g->r_begin : nullptr,
                            g->has_reverse ? g->r_node_idx : nullptr, root, INT_MAX, dist_dev, info, as_stream(stream));
}

int gmc_hop_dist(gmc_graph* g, int32_t root, int32_t* dist_host, gmc_bfs_info* info) {
    GMC_REQUIRE(g && dist_host, GMC_ERR_ARG, "gmc_hop_dist: NULL");
    GMC_REQUIRE(g->N > 0, GMC_ERR_ARG, "gmc_hop_dist: empty graph");
    GMC_TRY(ensure_ws(g));
    GMC_TRY(gmc_hop_dist_dev(g, root, (int32_t*)g->bfs->level, info, nullptr));
    return gmc_d2h(dist_host, g->bfs->level, sizeof(int32_t) * (size_t)g->N, nullptr);
}

int gmc_bfs_levels_masked(gmc_graph* g, int32_t root, int use_reverse_edge, const uint8_t* edge_mask_host, int16_t* level_host,
                          int32_t* state_trace, int32_t* level_count, int32_t trace_cap, gmc_bfs_info* info) {
    GMC_REQUIRE(g && level_host, GMC_ERR_ARG, "gmc_bfs_levels: NULL");
    GMC_REQUIRE(g->N > 0, GMC_ERR_ARG, "gmc_bfs_levels: empty graph");
    GMC_REQUIRE(!use_reverse_edge || g->has_reverse, GMC_ERR_STATE, "gmc_bfs_levels: use_reverse_edge needs the reverse CSR");
    GMC_TRY(ensure_ws(g));
    const int32_t *fb = g->begin, *fi = g->node_idx;
    const int32_t *rb = g->has_reverse ? g->r_begin : nullptr, *ri = g->has_reverse ? g->r_node_idx : nullptr;
    if (use_reverse_edge) { fb = g->r_begin; fi = g->r_node_idx; rb = g->begin; ri = g->node_idx; }
    uint8_t* mask_dev = nullptr;
    if (edge_mask_host) {
        // a filtered traversal stays top-down: the bottom-up parent scan (check_parent_rrd, :689-708)
        // has no navigator hook in the reference either
        rb = ri = nullptr;
        GMC_TRY(gmc_malloc((void**)&mask_dev, (size_t)(g->M ? g->M : 1)));
        int rc = gmc_h2d(mask_dev, edge_mask_host, (size_t)g->M, nullptr);
        if (rc != GMC_OK) { cudaFree(mask_dev); return rc; }
    }
    int rc = run_bfs<int16_t>(g, fb, fi, rb, ri, root, (int16_t)-2, (int16_t*)g->bfs->level, info, nullptr, mask_dev);
    if (mask_dev) cudaFree(mask_dev);
    GMC_TRY(rc);
    gmc_bfs_ws* w = g->bfs;
    for (int32_t i = 0; i < trace_cap && i < (int32_t)w->state_trace.size(); i++) {
        if (state_trace) state_trace[i] = w->state_trace[i];
        if (level_count) level_count[i] = w->level_count[i];
    }
    return gmc_d2h(level_host, w->level, sizeof(int16_t) * (size_t)g->N, nullptr);
}

int gmc_bfs_levels(gmc_graph* g, int32_t root, int use_reverse_edge, int16_t* level_host, int32_t* state_trace,
                   int32_t* level_count, int32_t trace_cap, gmc_bfs_info* info) {
    return gmc_bfs_levels_masked(g, root, use_reverse_edge, nullptr, level_host, state_trace, level_count, trace_cap, info);
}

// Vertices of every level in level order (the reference's global_vector / level_queue_begin,
// gm_bfs_template.h:423-447).  Bottom-up levels keep no device queue, so the order is rebuilt from the
// level array: ascending vertex id inside a level (the reference's order inside a level is unspecified).
int gmc_bfs_level_order(gmc_graph* g, int32_t* order_host, int32_t* level_begin_host, int32_t cap_levels) {
    GMC_REQUIRE(g && order_host && level_begin_host, GMC_ERR_ARG, "gmc_bfs_level_order: NULL");
    GMC_REQUIRE(g->bfs && g->bfs->valid && g->bfs->last_is_i16, GMC_ERR_STATE, "gmc_bfs_level_order: run gmc_bfs_levels first");
    gmc_bfs_ws* w = g->bfs;
    const int32_t nl = (int32_t)w->level_count.size();
    GMC_REQUIRE(cap_levels >= nl + 1, GMC_ERR_ARG, "gmc_bfs_level_order: need room for %d level offsets", nl + 1);
    std::vector<int16_t> lev((size_t)g->N);
    GMC_TRY(gmc_d2h(lev.data(), w->level, sizeof(int16_t) * (size_t)g->N, nullptr));
    level_begin_host[0] = 0;
    for (int32_t l = 0; l < nl; l++) level_begin_host[l + 1] = level_begin_host[l] + w->level_count[l];
    std::vector<int32_t> cur(level_begin_host, level_begin_host + nl);
    for (int32_t v = 0; v < g->N; v++)
        if (lev[v] >= 0) order_host[cur[lev[v]]++] = v;
    return GMC_OK;
}

}  // extern "C"
