// Level-synchronous, direction-optimizing, frontier-compacted BFS (K6/K7/K8), single GPU and sharded.
//
// One engine serves both reference entry points:
//   * generated hop_dist() (apps/src/hop_dist.gm:3-31): dist int32, INT_MAX when unreachable.  The
//     reference re-scans all N vertices twice per level; the device keeps a compacted
//     frontier instead -- the resulting dist[] is the same BFS distance.
//   * gm_bfs_template<short,...>::do_bfs_forward (inc/gm_bfs_template.h:69-271): level int16, -2 unvisited.
// The per-level state machine is the reference's (get_next_state, gm_bfs_template.h:352-421, thresholds
// :359,:720-721) so the direction switches happen on the same levels as on the CPU.
//
// Device realisation.  State: visited / frontier / candidate BITMAPS (N/8 bytes each, L1/L2 resident), the
// queue of the OWNED vertices of the current level, the level array.  Every level is
//   top-down  (SMALL, QUE, Q2R, RD, R2Q):
//       td_prepare  frontier degrees -> CTA-local offsets + fixed-size edge chunks (TD_CHUNK edge slots each),
//                   so a hub of degree 10^6 becomes 4096 chunks spread over the whole grid;
//       td_expand   one warp per chunk: coalesced neighbour loads, visited/candidate bitmap probe, atomicOr
//                   into the candidate bitmap (the byte fetch_and_or of gm_bitmap.cc:20-27 on 32-bit words);
//   bottom-up (RRD, RR2Q):
//       bu_scan     one lane per unvisited owned vertex scans its parents until one is in the frontier bitmap
//                   (check_parent_rrd :689-708 reads visited_level[u]==curr; the bitmap carries the same
//                   predicate in N/8 bytes); the warp publishes its 32 results with one plain store;
//   then  finish    new = candidates & ~visited; visited |= new; level[v] = l+1 for the new bits; the owned new
//                   vertices are compacted into the next queue; the level's vertex count is reduced.
// Sharded: rank r owns the vertex range [32*r*wpr, 32*(r+1)*wpr); it expands / scans only
// owned vertices and the candidate bitmaps are exchanged once per level with ncclAllGather (top-down: the
// full candidate bitmaps, OR-ed by `finish`; bottom-up: the owned slices, in place).  visited, frontier and
// level[] are replicated, so every rank drives the identical state machine without a count reduction.
#include "gmc_common.cuh"
#include "gmc_comm.h"
#include <limits.h>
#include <chrono>
#include <vector>

using namespace gmc;

enum { ST_SMALL = 0, ST_QUE = 1, ST_Q2R = 2, ST_RD = 3, ST_R2Q = 4, ST_RRD = 5, ST_RR2Q = 6 };
// words [BFS_MF_WORD, BFS_MF_WORD + 2) of the pinned host block: out-edges of the frontier a level leaves (64 bits), for the
// direction policy of hop_dist (see run_bfs)
constexpr int BFS_MF_WORD = 60;
constexpr int BFS_THREADS = 256;
constexpr int TD_CHUNK = 256;        // edge slots per warp chunk (8 per lane)
constexpr int TD_BATCHED_MAX_FRONTIER = 8192;   // owned frontier vertices up to which td_expand batches a lane's 8 loads
constexpr unsigned FULLMASK = 0xffffffffu;
constexpr int BFS_SPREAD = 64;          // counters a bottom-up level spreads its per-warp counts over (bu_scan_mark)

// Device-side traversal control (single-GPU runs): the reference's level-to-level state machine (get_next_state,
// gm_bfs_template.h:352-421) is evaluated by the last block of bfs_finish, so the host enqueues the kernels of several
// levels at once -- each kernel looks at the state and leaves at once when the level is not its kind -- instead of waiting
// for the counts of every level (~25 us of launch + PCIe round trip per level, 8 levels per traversal on the LJ-shaped graph).
constexpr int BFS_TRACE_CAP = 1024;
struct BfsCtl {
    int32_t state, curr_level, curr_count, own_count, qi, done, n_trace, overflow;
    float thr;
    int32_t has_reverse, rrd_direct, pad;
    int32_t trace_state[BFS_TRACE_CAP], trace_count[BFS_TRACE_CAP];
};
__host__ __device__ __forceinline__ bool ctl_bottom_up(int st) { return st == ST_RRD || st == ST_RR2Q; }

// get_next_state (gm_bfs_template.h:352-421): the state of the next level from the size of the frontier it found
// (next_count > 0).  One definition for the host loop and the device-side control blocks.
__host__ __device__ __forceinline__ int bfs_next_state(int state, int32_t next_count, int32_t curr_count, float thr, bool has_reverse,
                                                       bool rrd_direct) {
    switch (state) {
        case ST_SMALL: return next_count >= 128 ? ST_QUE : ST_SMALL;
        case ST_QUE:
            if ((float)next_count >= thr) return rrd_direct ? ST_RRD : ST_RD;
            if (next_count >= 1024 && next_count >= curr_count * 5) return ST_Q2R;
            return ST_QUE;
        case ST_Q2R: return (rrd_direct && (float)next_count >= thr) ? ST_RRD : ST_RD;
        case ST_RD:
            if (has_reverse && (float)next_count >= thr) return ST_RRD;
            return next_count <= 2 * curr_count ? ST_R2Q : ST_RD;
        case ST_R2Q: return ST_QUE;
        case ST_RRD: return next_count <= 2 * curr_count ? ST_RR2Q : ST_RRD;
        case ST_RR2Q: return ST_QUE;
    }
    return state;
}

struct gmc_bfs_ws {
    BfsCtl* ctl = nullptr;            // device
    int32_t N = 0;
    int32_t words = 0;                // ceil(N/32)
    int world = 1;
    int32_t wpr = 0;                  // bitmap words owned per rank
    int32_t wp = 0;                   // padded bitmap length = wpr * world
    uint32_t* visited = nullptr;      // [wp]
    uint32_t* frontier_bm = nullptr;  // 2 x [wp]: bitmap of the current level and the one being made (host-driven loop: ping-pong)
    uint32_t* cand = nullptr;         // [wp] candidates of the next level (top-down) / next level (bottom-up)
    uint32_t* allcand = nullptr;      // [world*wp] gathered candidate bitmaps (sharded top-down without peer mapping)
    // sharded runs: cand is the first of TWO buffers (level parity) in one exportable allocation, mapped by the peers, so that
    // a top-down level can OR its candidates straight into the OWNER's slice over NVLink (owner-directed exchange)
    bool peer_tried = false, peer_ok = false, peer_local = false;
    void* peer_base[8] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
    int rank = 0;
    int32_t* queue[2] = {nullptr, nullptr};   // owned vertices of the current / next level
    int32_t* blk_off = nullptr;       // [N] CTA-relative exclusive degree offsets of the queue entries
    int32_t* blk_total = nullptr;     // [ceil(N/256)]
    int2* chunks = nullptr;           // [M/TD_CHUNK + N/256 + 2] (queue block, chunk-in-block)
    size_t chunk_cap = 0;
    int32_t* best_parent = nullptr;   // [N] bottom-up shortcut: the in-neighbour of largest out-degree (-1: no in-edges), see bp_rows
    const int32_t* bp_key = nullptr;  // the reverse row array it was built from (NULL: not built)
    int32_t* bp_list = nullptr;       // [1 + BP_LIST_CAP] rows too long for a warp
    int32_t* spread = nullptr;        // [BFS_SPREAD] per-warp counts of a bottom-up level (bu_scan_mark), summed by bfs_publish
    int32_t* counters = nullptr;      // [8]: 0 = level count (global), 1 = owned queue count, 2 = nchunks, 3 = blocks of bfs_finish that are done, 4-6 = bfs_persistent, 7 = queue_from_bitmap
    int32_t* h_counters = nullptr;    // pinned + mapped [BFS_HOST_WORDS]: 0 = level count, 1 = owned count, 2 = sequence number of the level that wrote
                                      // them; bfs_small_levels: 0 = levels done, 1 = traversal ended, 2 = sequence number, 3 = the level after those is too big for it, 4.. = count found by each level
    int32_t seq = 0;
    void* level = nullptr;            // [N] int32 (also used as int16)
    unsigned long long* edge_sum = nullptr;
    // host-side record of the last traversal
    std::vector<int32_t> level_count, state_trace;
    bool last_is_i16 = false;
    bool valid = false;
    int last_launches = 0;
    int traversals = 0;               // traversals completed on this workspace (best_parent is built for graphs that are reused)
};

void gmc_bfs_ws_free(gmc_graph* g) {
    gmc_bfs_ws* w = g->bfs;
    if (!w) return;
    if (w->peer_ok) gmc_comm_ipc_close_all(w->world, w->rank, w->peer_base, w->peer_local);
    cudaFree(w->visited); cudaFree(w->frontier_bm); cudaFree(w->cand); cudaFree(w->allcand);
    cudaFree(w->queue[0]); cudaFree(w->queue[1]); cudaFree(w->blk_off); cudaFree(w->blk_total); cudaFree(w->chunks);
    cudaFree(w->best_parent); cudaFree(w->bp_list);
    cudaFree(w->counters); cudaFree(w->spread); cudaFree(w->level); cudaFree(w->edge_sum); cudaFree(w->ctl);
    if (w->h_counters) cudaFreeHost(w->h_counters);
    delete w;
    g->bfs = nullptr;
}

namespace {

// level[] = invalid (16-byte stores when the array allows it: 4.85 M one-element threads were 14 us of CTA launches), the bitmaps,
// the counters.  seen != 0: `cand` is the SEEN bitmap of the host-driven single-GPU loop (visited | candidates) and starts as visited.
template <typename L>
__global__ void bfs_init(L* __restrict__ level, int32_t n, L invalid, int32_t root, uint32_t* __restrict__ visited,
                         uint32_t* __restrict__ frontier, uint32_t* __restrict__ cand, int32_t wp,
                         int32_t* __restrict__ queue, int root_owned, int32_t* __restrict__ counters, int two_cand = 0, int seen = 0,
                         int32_t* __restrict__ spread = nullptr, unsigned long long* __restrict__ mf = nullptr) {
    constexpr int PER = 16 / (int)sizeof(L);
    const int32_t tid = blockIdx.x * blockDim.x + threadIdx.x, nthr = gridDim.x * blockDim.x;
    int32_t done = 0;
    if ((reinterpret_cast<uintptr_t>(level) & 15) == 0) {
        union { int4 v; L e[PER]; } fill;
#pragma unroll
        for (int k = 0; k < PER; k++) fill.e[k] = invalid;
        const int32_t nvec = n / PER;
        for (int32_t i = tid; i < nvec; i += nthr) {
            auto x = fill;
            if (root >= i * PER && root < (i + 1) * PER) x.e[root - i * PER] = (L)0;
            reinterpret_cast<int4*>(level)[i] = x.v;
        }
        done = nvec * PER;
    }
    for (int32_t i = done + tid; i < n; i += nthr) level[i] = (i == root) ? (L)0 : invalid;
    for (int32_t i = tid; i < wp; i += nthr) {
        const uint32_t b = (i == (root >> 5)) ? (1u << (root & 31)) : 0u;
        visited[i] = b; frontier[i] = b; cand[i] = seen ? b : 0u;
        if (two_cand) cand[wp + i] = 0u;
    }
    if (spread && tid < BFS_SPREAD) spread[tid] = 0;
    if (tid == 0) { if (root_owned) queue[0] = root; counters[0] = 0; counters[1] = 0; counters[2] = 0; counters[3] = 0; if (mf) *mf = 0ull; }
}

__global__ void bfs_ctl_init(BfsCtl* ctl, float thr, int has_reverse, int rrd_direct) {
    ctl->state = ST_SMALL; ctl->curr_level = 0; ctl->curr_count = 1; ctl->own_count = 1; ctl->qi = 0; ctl->done = 0;
    ctl->n_trace = 1; ctl->overflow = 0; ctl->thr = thr; ctl->has_reverse = has_reverse; ctl->rrd_direct = rrd_direct;
    ctl->trace_state[0] = ST_SMALL; ctl->trace_count[0] = 1;
}

// ---- top-down, step 1: degrees of the queued vertices -> CTA-relative offsets + chunk descriptors ----------
__device__ __forceinline__ void
td_prepare_body(const int32_t* __restrict__ begin, const int32_t* queue, int32_t qcount, int32_t* blk_off,
           int32_t* blk_total, int2* chunks, int32_t* nchunks, const int32_t* queue_alt, const BfsCtl* ctl,
           int32_t v_lo = 0, int32_t v_hi = INT_MAX) {
    __shared__ int32_t s_warp[BFS_THREADS / 32];
    __shared__ int32_t s_c0;
    if (ctl) {     // device-driven traversal: a top-down level? then the owned frontier is queue[qi], own_count long
        if (ctl->done || ctl_bottom_up(ctl->state)) return;
        qcount = ctl->own_count;
        if (ctl->qi) queue = queue_alt;
    }
    for (int32_t qblk = blockIdx.x; qblk * BFS_THREADS < qcount; qblk += gridDim.x) {
    __syncthreads();
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    const int32_t qi = qblk * BFS_THREADS + tid;
    int32_t deg = 0;
    // (sharded runs whose queue holds the WHOLE frontier -- the levels before were run redundantly by every rank -- expand only
    // the vertices in [v_lo, v_hi): the others count as rows without edges)
    if (qi < qcount) { const int32_t v = queue[qi]; if (v >= v_lo && v < v_hi) deg = begin[v + 1] - begin[v]; }
    int32_t inc = deg;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { const int32_t t = __shfl_up_sync(FULLMASK, inc, o); if (lane >= o) inc += t; }
    if (lane == 31) s_warp[wid] = inc;
    __syncthreads();
    int32_t woff = 0, total = 0;
#pragma unroll
    for (int k = 0; k < BFS_THREADS / 32; k++) { const int32_t t = s_warp[k]; if (k < wid) woff += t; total += t; }
    if (qi < qcount) blk_off[qi] = inc - deg + woff;
    const int32_t nch = (total + TD_CHUNK - 1) / TD_CHUNK;
    if (tid == 0) { blk_total[qblk] = total; s_c0 = nch ? atomicAdd(nchunks, nch) : 0; }
    __syncthreads();
    const int32_t c0 = s_c0;
    for (int32_t k = tid; k < nch; k += BFS_THREADS) chunks[c0 + k] = make_int2((int)qblk, k);
    }
}

// first k in [0,32] with e_k > pos (lane k holds e_k, ascending)
__device__ __forceinline__ int window_search(int32_t e, int32_t pos) {
    int k = 0;
#pragma unroll
    for (int s = 16; s > 0; s >>= 1) {
        const int32_t ek = __shfl_sync(FULLMASK, e, k + s - 1);
        if (ek <= pos) k += s;
    }
    if (__shfl_sync(FULLMASK, e, 31) <= pos) k = 32;      // beyond the 32-row window
    return k;
}

// Owner-directed exchange of the sharded top-down levels: base[r] = rank r's candidate buffer of this level (peer memory).
// A candidate is ORed into the local buffer -- which doubles as the "already sent" filter -- and, when another rank owns
// its word, into that rank's buffer over NVLink: at most one remote atomic per (rank, vertex) and level.
struct PeerCand {
    uint32_t* base[8];
    int32_t wpr, world, rank;
};
// pc == nullptr: no owner-directed exchange.  (pc points at a __grid_constant__ kernel parameter: indexing a by-value copy of the
// struct made every thread of the grid spill it to its stack -- 24 MB of local stores per launch, +8 us on every level)
__device__ __forceinline__ void mark_candidate(uint32_t* cand, const PeerCand* __restrict__ pc, int32_t uw, uint32_t bit) {
    if (cand[uw] & bit) return;
    atomicOr(&cand[uw], bit);
    if (pc && pc->world > 1) {
        const int o = uw / pc->wpr;
        if (o != pc->rank) atomicOr_system(pc->base[o] + uw, bit);       // performed at the owner's L2, system scope
    }
}

// ---- top-down, step 2: one warp per chunk of TD_CHUNK edge slots ---------------------------------------------
template <bool BATCHED>
__device__ __forceinline__ void
td_expand_body(const int32_t* __restrict__ begin, const int32_t* __restrict__ node_idx, const int32_t* queue, int32_t qcount,
          const int32_t* blk_off, const int32_t* blk_total, const int2* chunks, const int32_t* nchunks_p, const uint32_t* visited,
          uint32_t* cand, const uint8_t* __restrict__ edge_mask, const int32_t* queue_alt, const BfsCtl* ctl,
          const PeerCand* __restrict__ pc) {
    const int lane = threadIdx.x & 31;
    if (ctl) {     // device-driven traversal: both forms are enqueued, the one that fits the frontier runs
        if (ctl->done || ctl_bottom_up(ctl->state)) return;
        qcount = ctl->own_count;
        if ((qcount <= TD_BATCHED_MAX_FRONTIER) != BATCHED) return;
        if (ctl->qi) queue = queue_alt;
    }
    const int32_t nchunks = *nchunks_p;
    const int32_t nwarps = (gridDim.x * BFS_THREADS) >> 5;
    for (int32_t c = (blockIdx.x * BFS_THREADS + threadIdx.x) >> 5; c < nchunks; c += nwarps) {
        const int2 ch = chunks[c];
        const int32_t qbase = ch.x * BFS_THREADS;
        const int32_t nq = min(BFS_THREADS, qcount - qbase);
        const int32_t T = blk_total[ch.x];
        const int32_t s0 = ch.y * TD_CHUNK, s1 = min(s0 + TD_CHUNK, T);
        // row holding slot s0: last i with off[i] <= s0 (32-ary search over <= 256 offsets)
        int32_t r0;
        {
            const int i1 = lane * 8;
            const bool p1 = i1 < nq && blk_off[qbase + i1] <= s0;
            const int seg = 31 - __clz(__ballot_sync(FULLMASK, p1));      // lane 0 always true (off[0] = 0)
            const int i2 = seg * 8 + lane;
            const bool p2 = lane < 8 && i2 < nq && blk_off[qbase + i2] <= s0;
            r0 = seg * 8 + (31 - __clz(__ballot_sync(FULLMASK, p2)));
        }
        int32_t w_start = INT_MAX, w_end = INT_MAX, w_bv = 0;
        auto load_window = [&](int32_t r) {
            const int32_t row = r + lane;
            if (row < nq) {
                w_start = blk_off[qbase + row];
                w_end = (row + 1 < nq) ? blk_off[qbase + row + 1] : T;
                w_bv = begin[queue[qbase + row]];
            } else { w_start = INT_MAX; w_end = INT_MAX; w_bv = 0; }
        };
        load_window(r0);
        if (!BATCHED) {
            // large frontiers: one 32-slot step at a time; 64 resident warps per SM hide the two dependent loads of a step
            for (int32_t p0 = s0; p0 < s1; p0 += 32) {
                const int32_t pos = p0 + lane;
                bool todo = pos < s1;
                while (__any_sync(FULLMASK, todo)) {
                    const int kk = window_search(w_end, pos);
                    const int32_t st = __shfl_sync(FULLMASK, w_start, kk & 31);
                    const int32_t bv = __shfl_sync(FULLMASK, w_bv, kk & 31);
                    if (todo && kk < 32) {
                        const int32_t eidx = bv + (pos - st);
                        const int32_t u = ld_stream(node_idx + eidx);
                        const uint32_t bit = 1u << (u & 31);
                        const int32_t uw = u >> 5;
                        // navigator (check_navigator, gm_bfs_template.h:488-492,596-598): filtered edges are not followed
                        if (!(visited[uw] & bit) && (!edge_mask || edge_mask[eidx])) mark_candidate(cand, pc, uw, bit);
                        todo = false;
                    }
                    if (__any_sync(FULLMASK, todo)) { r0 += 32; load_window(r0); }     // slots beyond the 32-row window
                }
            }
        } else {
        // small frontiers (few chunks, the level is a latency chain): three phases per chunk so that the 8 loads of a lane
        // are in flight together: (1) edge index of each of the lane's 8 slots (row lookup inside the 32-row window,
        // shuffles only), (2) the 8 neighbour ids, (3) the 8 bitmap probes
        int32_t eidx[TD_CHUNK / 32];
#pragma unroll
        for (int j = 0; j < TD_CHUNK / 32; j++) {
            const int32_t pos = s0 + j * 32 + lane;
            bool todo = pos < s1;
            eidx[j] = -1;
            while (__any_sync(FULLMASK, todo)) {
                const int kk = window_search(w_end, pos);
                const int32_t st = __shfl_sync(FULLMASK, w_start, kk & 31);
                const int32_t bv = __shfl_sync(FULLMASK, w_bv, kk & 31);
                if (todo && kk < 32) { eidx[j] = bv + (pos - st); todo = false; }
                if (__any_sync(FULLMASK, todo)) { r0 += 32; load_window(r0); }     // slots beyond the 32-row window
            }
        }
        int32_t u[TD_CHUNK / 32];
#pragma unroll
        for (int j = 0; j < TD_CHUNK / 32; j++) u[j] = eidx[j] >= 0 ? ld_stream(node_idx + eidx[j]) : -1;
        uint32_t vis[TD_CHUNK / 32];
#pragma unroll
        for (int j = 0; j < TD_CHUNK / 32; j++) vis[j] = u[j] >= 0 ? visited[u[j] >> 5] : 0xffffffffu;
#pragma unroll
        for (int j = 0; j < TD_CHUNK / 32; j++) {
            if (u[j] < 0) continue;
            const uint32_t bit = 1u << (u[j] & 31);
            const int32_t uw = u[j] >> 5;
            if (!(vis[j] & bit) && (!edge_mask || edge_mask[eidx[j]])) mark_candidate(cand, pc, uw, bit);
        }
        }
    }
}

// ---- bottom-up: warp w owns vertices [32w, 32w+32) = one word of the bitmaps ------------------------------------
__device__ __forceinline__ void
bu_scan_body(const int32_t* __restrict__ r_begin, const int32_t* __restrict__ r_node_idx, int32_t n, int32_t w_lo, int32_t w_hi,
        const uint32_t* frontier_bm, const uint32_t* visited, uint32_t* cand, const BfsCtl* ctl, const int32_t* __restrict__ best_parent = nullptr) {
    const int lane = threadIdx.x & 31;
    if (ctl && (ctl->done || !ctl_bottom_up(ctl->state))) return;
    const int32_t wstep = (gridDim.x * BFS_THREADS) >> 5;
    for (int32_t w = w_lo + ((blockIdx.x * BFS_THREADS + threadIdx.x) >> 5); w < w_hi; w += wstep) {
    const uint32_t vis = visited[w];
    const int32_t t = (w << 5) + lane;
    bool found = false;
    if (t < n && !((vis >> lane) & 1u)) {
        // best parent first (bp_rows): one coalesced load instead of the row, and rows without in-edges end here
        const int32_t bp = best_parent ? best_parent[t] : -2;
        if (bp >= 0) found = (frontier_bm[bp >> 5] >> (bp & 31)) & 1u;
        const int32_t b = (found || bp == -1) ? 0 : r_begin[t], e = (found || bp == -1) ? 0 : r_begin[t + 1];
        // four parents per step: their ids, then their frontier words, are independent loads (the one-at-a-time loop was a
        // chain of two dependent loads per parent); any hit ends the scan as before
        for (int32_t nx = b; nx < e && !found; nx += 4) {
            int32_t u[4];
#pragma unroll
            for (int k = 0; k < 4; k++) u[k] = nx + k < e ? r_node_idx[nx + k] : -1;
            uint32_t fw[4];
#pragma unroll
            for (int k = 0; k < 4; k++) fw[k] = u[k] >= 0 ? frontier_bm[u[k] >> 5] : 0u;
#pragma unroll
            for (int k = 0; k < 4; k++) found = found || (u[k] >= 0 && ((fw[k] >> (u[k] & 31)) & 1u));
        }
    }
    const uint32_t got = __ballot_sync(FULLMASK, found);
    // vis | got: bfs_finish takes cand & ~visited, and where cand is the SEEN bitmap the word has to be visited | candidates
    if (lane == 0) cand[w] = vis | got;
    }
}

// ---- the kernels: one per body for the host-driven loop, and two MERGED ones for the device-driven loop, where a level is
// three launches whatever its direction (bfs_level_a: bottom-up scan OR top-down prepare; td_expand_any: either form of the
// expansion, nothing in a bottom-up level; bfs_finish)
__global__ void __launch_bounds__(BFS_THREADS)
td_prepare(const int32_t* __restrict__ begin, const int32_t* __restrict__ queue, int32_t qcount, int32_t* __restrict__ blk_off,
           int32_t* __restrict__ blk_total, int2* __restrict__ chunks, int32_t* __restrict__ nchunks, int32_t v_lo, int32_t v_hi) {
    td_prepare_body(begin, queue, qcount, blk_off, blk_total, chunks, nchunks, nullptr, nullptr, v_lo, v_hi);
}
// own[i] |= OR over the other ranks r of slices[r * wpr + i]: the owner's end of the all-to-all exchange of a top-down level
__global__ void __launch_bounds__(BFS_THREADS)
or_slices(const uint32_t* __restrict__ slices, int world, int rank, int32_t wpr, uint32_t* __restrict__ own) {
    for (int32_t i = blockIdx.x * BFS_THREADS + threadIdx.x; i < wpr; i += gridDim.x * BFS_THREADS) {
        uint32_t v = own[i];
        for (int r = 0; r < world; r++) if (r != rank) v |= slices[(size_t)r * wpr + i];
        own[i] = v;
    }
}
// the whole frontier as a queue, from the (replicated) frontier bitmap: sharded runs going back to levels that every rank runs
// redundantly (bfs_small_levels).  At most SMALL_Q bits are set; their order in the queue does not matter.  *count starts at 0.
__global__ void __launch_bounds__(BFS_THREADS)
queue_from_bitmap(const uint32_t* __restrict__ frontier, int32_t words, int32_t* __restrict__ queue, int32_t* __restrict__ count, int32_t cap) {
    for (int32_t w = blockIdx.x * BFS_THREADS + threadIdx.x; w < words; w += gridDim.x * BFS_THREADS) {
        uint32_t b = frontier[w];
        if (!b) continue;
        int32_t at = atomicAdd(count, __popc(b));
        while (b) {
            const int k = __ffs(b) - 1;
            b &= b - 1;
            if (at < cap) queue[at] = (w << 5) + k;
            at++;
        }
    }
}
template <bool BATCHED>
__global__ void __launch_bounds__(BFS_THREADS)
td_expand(const int32_t* __restrict__ begin, const int32_t* __restrict__ node_idx, const int32_t* __restrict__ queue, int32_t qcount,
          const int32_t* __restrict__ blk_off, const int32_t* __restrict__ blk_total, const int2* __restrict__ chunks,
          const int32_t* __restrict__ nchunks_p, const uint32_t* __restrict__ visited, uint32_t* __restrict__ cand,
          const uint8_t* __restrict__ edge_mask, const __grid_constant__ PeerCand pc) {
    td_expand_body<BATCHED>(begin, node_idx, queue, qcount, blk_off, blk_total, chunks, nchunks_p, visited, cand, edge_mask, nullptr, nullptr,
                            pc.world > 1 ? &pc : nullptr);
}
// seen mode: ONE bitmap (visited | candidates) is probed and marked; it changes under the kernel, so no __restrict__ / read-only loads
template <bool BATCHED>
__global__ void __launch_bounds__(BFS_THREADS)
td_expand_seen(const int32_t* __restrict__ begin, const int32_t* __restrict__ node_idx, const int32_t* __restrict__ queue, int32_t qcount,
               const int32_t* __restrict__ blk_off, const int32_t* __restrict__ blk_total, const int2* __restrict__ chunks,
               const int32_t* __restrict__ nchunks_p, uint32_t* seen, const uint8_t* __restrict__ edge_mask) {
    td_expand_body<BATCHED>(begin, node_idx, queue, qcount, blk_off, blk_total, chunks, nchunks_p, seen, seen, edge_mask, nullptr, nullptr, nullptr);
}
__global__ void __launch_bounds__(BFS_THREADS)
bu_scan(const int32_t* __restrict__ r_begin, const int32_t* __restrict__ r_node_idx, int32_t n, int32_t w_lo, int32_t w_hi,
        const uint32_t* __restrict__ frontier_bm, const uint32_t* __restrict__ visited, uint32_t* __restrict__ cand,
        const int32_t* __restrict__ best_parent) {
    bu_scan_body(r_begin, r_node_idx, n, w_lo, w_hi, frontier_bm, visited, cand, nullptr, best_parent);
}
__global__ void __launch_bounds__(BFS_THREADS)
bfs_level_a(const int32_t* __restrict__ r_begin, const int32_t* __restrict__ r_node_idx, int32_t n, int32_t w_lo, int32_t w_hi,
            const uint32_t* __restrict__ frontier_bm, const uint32_t* __restrict__ visited, uint32_t* __restrict__ cand,
            const int32_t* __restrict__ begin, const int32_t* __restrict__ queue0, const int32_t* __restrict__ queue1,
            int32_t* __restrict__ blk_off, int32_t* __restrict__ blk_total, int2* __restrict__ chunks, int32_t* __restrict__ nchunks,
            const BfsCtl* __restrict__ ctl) {
    if (ctl->done) return;
    if (ctl_bottom_up(ctl->state)) { if (r_begin) bu_scan_body(r_begin, r_node_idx, n, w_lo, w_hi, frontier_bm, visited, cand, ctl); }
    else td_prepare_body(begin, queue0, 0, blk_off, blk_total, chunks, nchunks, queue1, ctl);
}
__global__ void __launch_bounds__(BFS_THREADS)
td_expand_any(const int32_t* __restrict__ begin, const int32_t* __restrict__ node_idx, const int32_t* __restrict__ queue0,
              const int32_t* __restrict__ queue1, const int32_t* __restrict__ blk_off, const int32_t* __restrict__ blk_total,
              const int2* __restrict__ chunks, const int32_t* __restrict__ nchunks_p, const uint32_t* __restrict__ visited,
              uint32_t* __restrict__ cand, const uint8_t* __restrict__ edge_mask, const BfsCtl* __restrict__ ctl) {
    if (ctl->done || ctl_bottom_up(ctl->state)) return;
    if (ctl->own_count <= TD_BATCHED_MAX_FRONTIER)
        td_expand_body<true>(begin, node_idx, queue0, 0, blk_off, blk_total, chunks, nchunks_p, visited, cand, edge_mask, queue1, ctl, nullptr);
    else
        td_expand_body<false>(begin, node_idx, queue0, 0, blk_off, blk_total, chunks, nchunks_p, visited, cand, edge_mask, queue1, ctl, nullptr);
}

// ---- end of level: new = candidates & ~visited -> visited, frontier, level[], owned queue, count ------------------
// One warp, 32 consecutive bitmap words (lane = word w).
template <typename L>
__device__ __forceinline__ void
finish_warp(int32_t w, int32_t words, int32_t wp, int world, int use_all, uint32_t* cand, const uint32_t* allcand, uint32_t* visited,
            uint32_t* frontier, L next_level, L* __restrict__ level, int32_t own_lo, int32_t own_hi, int32_t* next_queue,
            int32_t* counters, int keep_cand = 0, const int32_t* __restrict__ deg_begin = nullptr, unsigned long long* mf = nullptr,
            int32_t n_vertices = 0) {
    const int lane = threadIdx.x & 31;
    uint32_t nw = 0;
    unsigned long long dsum = 0;       // deg_begin != null: out-edges of the new frontier vertices this lane stored
    if (w < words) {
        uint32_t c;
        if (use_all) {
            c = 0;
            for (int r = 0; r < world; r++) c |= allcand[(size_t)r * wp + w];
        } else c = cand[w];
        if (!keep_cand) cand[w] = 0u;          // (keep_cand: cand is the SEEN bitmap, visited | candidates = the new visited)
        const uint32_t vis = visited[w];
        nw = c & ~vis;
        if (nw) visited[w] = vis | nw;
        frontier[w] = nw;
    }
    const int pc = __popc(nw);
    const bool owned = w >= own_lo && w < own_hi;
    int tot = pc, own = owned ? pc : 0, pre = own;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { const int t = __shfl_up_sync(FULLMASK, pre, o); if (lane >= o) pre += t; }
    const int own_tot = __shfl_sync(FULLMASK, pre, 31);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) tot += __shfl_xor_sync(FULLMASK, tot, o);
    int32_t base = 0;
    if (lane == 0) {
        if (tot) atomicAdd(&counters[0], tot);
        if (own_tot) base = atomicAdd(&counters[1], own_tot);
    }
    base = __shfl_sync(FULLMASK, base, 0) + pre - own;
    int32_t db[33];
    if (deg_begin && nw) {
        const int32_t v0 = w << 5;
        if (v0 + 32 <= n_vertices && (reinterpret_cast<uintptr_t>(deg_begin) & 15) == 0) {
            const int4* __restrict__ p4 = reinterpret_cast<const int4*>(deg_begin + v0);
#pragma unroll
            for (int k = 0; k < 8; k++) { const int4 x = __ldg(p4 + k); db[4 * k] = x.x; db[4 * k + 1] = x.y; db[4 * k + 2] = x.z; db[4 * k + 3] = x.w; }
            db[32] = __ldg(deg_begin + v0 + 32);
        } else {
#pragma unroll
            for (int k = 0; k < 33; k++) db[k] = __ldg(deg_begin + min(v0 + k, n_vertices));       // (the last word: vertices >= n have no edges)
        }
    }
    // the warp goes through its non-empty words together: lane b takes bit b of the word, so the level[] and queue
    // stores of one word are one coalesced 128-byte store each (a lane per word wrote 32 scattered elements in turn)
    unsigned has = __ballot_sync(FULLMASK, nw != 0u);
    while (has) {
        const int src = __ffs(has) - 1;
        has &= has - 1;
        const uint32_t bits = __shfl_sync(FULLMASK, nw, src);
        const int32_t qb = __shfl_sync(FULLMASK, base, src);
        const int own_s = __shfl_sync(FULLMASK, (int)owned, src);
        const int32_t v = ((w - lane + src) << 5) + lane;
        if ((bits >> lane) & 1u) {
            level[v] = next_level;
            if (own_s) next_queue[qb + __popc(bits & ((1u << lane) - 1u))] = v;
        }
    }
    if (deg_begin) {
        // out-edges of the lane's own word: begin[32 w .. 32 w + 32] was requested before the loop above (33 independent loads, 128
        // contiguous bytes per lane), so the loop hid their latency.  (Two dependent loads per set bit inside the loop made it a
        // chain of L2 round trips: +15 us on the 405 K-vertex level of the LJ-shaped graph.)
        if (nw) {
#pragma unroll
            for (int k = 0; k < 32; k++) if ((nw >> k) & 1u) dsum += (unsigned long long)(db[k + 1] - db[k]);
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) dsum += __shfl_xor_sync(FULLMASK, dsum, o);
        if (lane == 0 && dsum) atomicAdd(mf, dsum);
    }
}

// get_next_state (gm_bfs_template.h:352-421) on the device, same arithmetic as the host loop of run_bfs: one thread, after
// every count of the level has landed.  Returns `done`.
__device__ __forceinline__ int ctl_advance(BfsCtl* ctl, int32_t next_count, int32_t own) {
    const int done_in = next_count == 0;
    const int next_state = done_in ? ctl->state
                                   : bfs_next_state(ctl->state, next_count, ctl->curr_count, ctl->thr, ctl->has_reverse != 0, ctl->rrd_direct != 0);
    int done = done_in;
    if (!done) {
        const int32_t k = ctl->n_trace;
        if (k >= BFS_TRACE_CAP) { ctl->overflow = 1; done = 1; }
        else {
            ctl->trace_state[k] = next_state; ctl->trace_count[k] = next_count; ctl->n_trace = k + 1;
            ctl->state = next_state; ctl->curr_count = next_count; ctl->own_count = own;
            ctl->curr_level = ctl->curr_level + 1; ctl->qi = ctl->qi ^ 1;
        }
    }
    ctl->done = done;
    return done;
}

template <typename L>
__global__ void __launch_bounds__(BFS_THREADS)
bfs_finish(int32_t words, int32_t wp, int world, int use_all, uint32_t* __restrict__ cand, const uint32_t* __restrict__ allcand,
           uint32_t* __restrict__ visited, uint32_t* __restrict__ frontier, L next_level, L* __restrict__ level,
           int32_t own_lo, int32_t own_hi, int32_t* __restrict__ next_queue, int32_t* __restrict__ counters,
           volatile int32_t* __restrict__ host_out, int32_t seq, int32_t* __restrict__ queue_alt = nullptr, BfsCtl* ctl = nullptr,
           int keep_cand = 0, const int32_t* __restrict__ deg_begin = nullptr, unsigned long long* mf = nullptr, int32_t n_vertices = 0) {
    if (ctl) {     // device-driven traversal: level number and queue come from the control block
        if (ctl->done) {      // the traversal ended earlier in this batch of launches: only tell the host
            if (blockIdx.x == 0 && threadIdx.x == 0) { host_out[1] = 1; __threadfence_system(); host_out[2] = seq; }
            return;
        }
        next_level = (L)(ctl->curr_level + 1);
        if (!ctl->qi) next_queue = queue_alt;       // next = queue[qi ^ 1]; next_queue = queue[0], queue_alt = queue[1]
    }
    finish_warp<L>(blockIdx.x * BFS_THREADS + threadIdx.x, words, wp, world, use_all, cand, allcand, visited, frontier, next_level, level,
                   own_lo, own_hi, next_queue, counters, keep_cand, deg_begin, mf, n_vertices);
    // the last block to finish publishes the two level counts straight into the host's (mapped, pinned) memory and
    // resets the device counters: the host polls the sequence number instead of a memcpy + memset + stream sync per level
    __shared__ int s_last;
    __syncthreads();
    if (threadIdx.x == 0) {
        __threadfence();
        s_last = (atomicAdd(&counters[3], 1) == (int)gridDim.x - 1);
    }
    __syncthreads();
    if (s_last && threadIdx.x == 0) {
        __threadfence();
        const int32_t next_count = atomicAdd(&counters[0], 0), own = atomicAdd(&counters[1], 0);
        counters[0] = 0; counters[1] = 0; counters[2] = 0; counters[3] = 0;
        if (mf) {
            const unsigned long long e = atomicAdd(mf, 0ull);
            *mf = 0ull;
            host_out[BFS_MF_WORD] = (int32_t)(uint32_t)e; host_out[BFS_MF_WORD + 1] = (int32_t)(uint32_t)(e >> 32);
        }
        if (!ctl) { host_out[0] = next_count; host_out[1] = own; }
        else {
            const int done = ctl_advance(ctl, next_count, own);
            __threadfence();
            host_out[0] = next_count; host_out[1] = done;
        }
        __threadfence_system();
        host_out[2] = seq;
    }
}

// ---- bottom-up level in ONE launch (single GPU): scan + end-of-level work of the same word ------------------------------------
// The warp that scans word w is the only one that reads visited[w], and what it finds IS the new part of the word, so it can
// finish the word itself: visited |= got, level[] of the found lanes, next-frontier word (into the OTHER frontier buffer: the
// current one is still being probed by every warp), queue entries.  One atomicAdd per CTA (8 words) hands out queue space -- per
// warp it would be 150 K atomics on one address.  Saves the bitmap pass and the launch of bfs_finish on the 2-3 widest levels.
template <typename L>
__global__ void __launch_bounds__(BFS_THREADS)
bu_scan_finish(const int32_t* __restrict__ r_begin, const int32_t* __restrict__ r_node_idx, int32_t n, int32_t words,
               const uint32_t* __restrict__ frontier_cur, uint32_t* __restrict__ frontier_next, uint32_t* __restrict__ visited,
               L next_level, L* __restrict__ level, int32_t* __restrict__ next_queue, int32_t* __restrict__ counters,
               volatile int32_t* __restrict__ host_out, int32_t seq) {
    __shared__ int32_t s_cnt[BFS_THREADS / 32];
    __shared__ int32_t s_base;
    __shared__ int s_last;
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    const int32_t w = (blockIdx.x * BFS_THREADS + threadIdx.x) >> 5;
    bool found = false;
    uint32_t vis = 0xffffffffu;
    const int32_t t = (w << 5) + lane;
    if (w < words) {
        vis = visited[w];
        if (t < n && !((vis >> lane) & 1u)) {
            const int32_t b = r_begin[t], e = r_begin[t + 1];
            for (int32_t nx = b; nx < e && !found; nx += 4) {       // four parents per step, as in bu_scan_body
                int32_t u[4];
#pragma unroll
                for (int k = 0; k < 4; k++) u[k] = nx + k < e ? r_node_idx[nx + k] : -1;
                uint32_t fw[4];
#pragma unroll
                for (int k = 0; k < 4; k++) fw[k] = u[k] >= 0 ? frontier_cur[u[k] >> 5] : 0u;
#pragma unroll
                for (int k = 0; k < 4; k++) found = found || (u[k] >= 0 && ((fw[k] >> (u[k] & 31)) & 1u));
            }
        }
    }
    const uint32_t got = __ballot_sync(FULLMASK, found);
    if (w < words && lane == 0) {
        if (got) visited[w] = vis | got;
        frontier_next[w] = got;
    }
    if (found) level[t] = next_level;
    if (lane == 0) s_cnt[wid] = __popc(got);
    __syncthreads();
    if (threadIdx.x == 0) {
        int total = 0;
#pragma unroll
        for (int k = 0; k < BFS_THREADS / 32; k++) total += s_cnt[k];
        s_base = total ? atomicAdd(&counters[1], total) : 0;
    }
    __syncthreads();
    if (found) {
        int32_t base = s_base;
        for (int k = 0; k < wid; k++) base += s_cnt[k];
        next_queue[base + __popc(got & ((1u << lane) - 1u))] = t;
    }
    // last CTA: publish the level's count (single GPU: all of it is owned) and reset the counters, as bfs_finish does
    __syncthreads();
    if (threadIdx.x == 0) {
        __threadfence();
        s_last = (atomicAdd(&counters[3], 1) == (int)gridDim.x - 1);
    }
    __syncthreads();
    if (s_last && threadIdx.x == 0) {
        __threadfence();
        const int32_t next_count = atomicAdd(&counters[1], 0);
        counters[0] = 0; counters[1] = 0; counters[2] = 0; counters[3] = 0;
        host_out[0] = next_count; host_out[1] = next_count;
        __threadfence_system();
        host_out[2] = seq;
    }
}

// ---- bottom-up level whose successor is bottom-up too (state RRD: the next state is RRD or RR2Q) ---------------------------------
// No top-down level reads its result, so no queue is needed, and then the scanning warp can finish its word alone: visited |= got,
// next-frontier word (the OTHER frontier buffer: this level's is being probed), level[] of the found lanes, seen word.  The level's
// count goes into BFS_SPREAD counters (a warp adds to counter w mod BFS_SPREAD: no block-wide step, a finished warp leaves) and a
// one-warp kernel adds them up for the host: scan + publish instead of scan + a bfs_finish pass over three bitmaps.
template <typename L>
__global__ void __launch_bounds__(BFS_THREADS)
bu_scan_mark(const int32_t* __restrict__ r_begin, const int32_t* __restrict__ r_node_idx, int32_t n, int32_t words,
             const uint32_t* __restrict__ frontier_cur, uint32_t* __restrict__ frontier_next, uint32_t* __restrict__ visited,
             uint32_t* __restrict__ seen, L next_level, L* __restrict__ level, int32_t* __restrict__ spread,
             const int32_t* __restrict__ best_parent) {
    // the CTA's count: shared-memory atomics, flushed by whichever warp finishes last (no barrier after the scan: a finished warp
    // leaves).  A global atomic per warp -- 150 K of them on 64 neighbouring counters -- ran at the L2's one-per-clock rate for a
    // cache line and cost 30 us per level (gpurun r3l).
    __shared__ int s_cnt, s_arrived;
    if (threadIdx.x == 0) { s_cnt = 0; s_arrived = 0; }
    __syncthreads();
    const int lane = threadIdx.x & 31;
    const int32_t w = (blockIdx.x * BFS_THREADS + threadIdx.x) >> 5;
    const uint32_t vis = w < words ? visited[w] : 0xffffffffu;
    const int32_t t = (w << 5) + lane;
    bool found = false;
    if (t < n && !((vis >> lane) & 1u)) {
        const int32_t bp = best_parent ? best_parent[t] : -2;           // best parent first, as in bu_scan_body
        if (bp >= 0) found = (frontier_cur[bp >> 5] >> (bp & 31)) & 1u;
        const int32_t b = (found || bp == -1) ? 0 : r_begin[t], e = (found || bp == -1) ? 0 : r_begin[t + 1];
        for (int32_t nx = b; nx < e && !found; nx += 4) {       // four parents per step, as in bu_scan_body
            int32_t u[4];
#pragma unroll
            for (int k = 0; k < 4; k++) u[k] = nx + k < e ? r_node_idx[nx + k] : -1;
            uint32_t fw[4];
#pragma unroll
            for (int k = 0; k < 4; k++) fw[k] = u[k] >= 0 ? frontier_cur[u[k] >> 5] : 0u;
#pragma unroll
            for (int k = 0; k < 4; k++) found = found || (u[k] >= 0 && ((fw[k] >> (u[k] & 31)) & 1u));
        }
    }
    const uint32_t got = __ballot_sync(FULLMASK, found);
    if (found) level[t] = next_level;
    if (lane == 0) {
        if (w < words) frontier_next[w] = got;
        if (got) {
            visited[w] = vis | got;
            if (seen) seen[w] = vis | got;
            atomicAdd_block(&s_cnt, __popc(got));
        }
        __threadfence_block();
        if (atomicAdd_block(&s_arrived, 1) == BFS_THREADS / 32 - 1) {
            const int c = atomicAdd_block(&s_cnt, 0);
            if (c) atomicAdd(&spread[blockIdx.x & (BFS_SPREAD - 1)], c);
        }
    }
}

// ---- best parent of every vertex: the in-neighbour with the most out-edges (ties: the first in the row), -1 without in-edges ------
// A bottom-up level asks "is ANY in-neighbour in the frontier"; on a power-law graph the in-neighbour of largest out-degree is the
// likeliest one at every level, so the scan probes it first -- one coalesced load per vertex instead of r_begin + a sector of the
// row -- and walks the row only when that fails.  Built once per graph (one pass over the reverse rows), kept in the workspace.
constexpr int BP_LIST_CAP = 4096, BP_WARP_MAX = 32768;
__device__ __forceinline__ unsigned long long bp_key_of(const int32_t* __restrict__ begin, int32_t src, int32_t pos) {
    return ((unsigned long long)(uint32_t)(begin[src + 1] - begin[src]) << 32) | (unsigned long long)(0xffffffffu - (uint32_t)pos);
}
__global__ void __launch_bounds__(BFS_THREADS)
bp_rows(const int32_t* __restrict__ begin, const int32_t* __restrict__ r_begin, const int32_t* __restrict__ r_node_idx, int32_t n,
        int32_t* __restrict__ best_parent, int32_t* __restrict__ list) {
    const int lane = threadIdx.x & 31;
    const int32_t nwarps = (gridDim.x * BFS_THREADS) >> 5;
    for (int32_t w = (blockIdx.x * BFS_THREADS + threadIdx.x) >> 5; w * 32 < n; w += nwarps) {
        const int32_t t = w * 32 + lane;
        int32_t b = 0, d = 0;
        if (t < n) { b = r_begin[t]; d = r_begin[t + 1] - b; }
        if (t < n && d <= 32) {          // short row: its lane alone
            unsigned long long best = 0ull;
            for (int32_t k = 0; k < d; k++) { const unsigned long long key = bp_key_of(begin, r_node_idx[b + k], k); best = key > best ? key : best; }
            best_parent[t] = d ? r_node_idx[b + (int32_t)(0xffffffffu - (uint32_t)best)] : -1;
        }
        unsigned longs = __ballot_sync(FULLMASK, t < n && d > 32);
        while (longs) {                  // long rows: the warp together, four entries per lane and step
            const int src = __ffs(longs) - 1;
            longs &= longs - 1;
            const int32_t rb = __shfl_sync(FULLMASK, b, src), rd = __shfl_sync(FULLMASK, d, src), row = w * 32 + src;
            if (rd > BP_WARP_MAX) {      // a hub: one CTA per row (bp_huge), unless the list is full
                int32_t at = 0;
                if (lane == 0) { at = atomicAdd(list, 1); if (at < BP_LIST_CAP) list[1 + at] = row; }
                if (__shfl_sync(FULLMASK, at, 0) < BP_LIST_CAP) continue;
            }
            unsigned long long best = 0ull;
            for (int32_t k0 = 0; k0 < rd; k0 += 128) {
                int32_t sv[4];
#pragma unroll
                for (int j = 0; j < 4; j++) { const int32_t k = k0 + j * 32 + lane; sv[j] = k < rd ? r_node_idx[rb + k] : -1; }
#pragma unroll
                for (int j = 0; j < 4; j++) if (sv[j] >= 0) { const unsigned long long key = bp_key_of(begin, sv[j], k0 + j * 32 + lane); best = key > best ? key : best; }
            }
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) { const unsigned long long x = __shfl_xor_sync(FULLMASK, best, o); best = x > best ? x : best; }
            if (lane == 0) best_parent[row] = r_node_idx[rb + (int32_t)(0xffffffffu - (uint32_t)best)];
        }
    }
}
__global__ void __launch_bounds__(1024)
bp_huge(const int32_t* __restrict__ begin, const int32_t* __restrict__ r_begin, const int32_t* __restrict__ r_node_idx,
        int32_t* __restrict__ best_parent, const int32_t* __restrict__ list) {
    __shared__ unsigned long long s_best[32];
    const int32_t cnt = min(list[0], BP_LIST_CAP);
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    for (int32_t i = blockIdx.x; i < cnt; i += gridDim.x) {
        const int32_t row = list[1 + i], rb = r_begin[row], rd = r_begin[row + 1] - rb;
        unsigned long long best = 0ull;
        for (int32_t k0 = 0; k0 < rd; k0 += 4096) {
            int32_t sv[4];
#pragma unroll
            for (int j = 0; j < 4; j++) { const int32_t k = k0 + j * 1024 + (int32_t)threadIdx.x; sv[j] = k < rd ? r_node_idx[rb + k] : -1; }
#pragma unroll
            for (int j = 0; j < 4; j++) if (sv[j] >= 0) { const unsigned long long key = bp_key_of(begin, sv[j], k0 + j * 1024 + (int32_t)threadIdx.x); best = key > best ? key : best; }
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) { const unsigned long long x = __shfl_xor_sync(FULLMASK, best, o); best = x > best ? x : best; }
        __syncthreads();
        if (lane == 0) s_best[wid] = best;
        __syncthreads();
        if (wid == 0) {
            best = s_best[lane];
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) { const unsigned long long x = __shfl_xor_sync(FULLMASK, best, o); best = x > best ? x : best; }
            if (lane == 0) best_parent[row] = r_node_idx[rb + (int32_t)(0xffffffffu - (uint32_t)best)];
        }
    }
}

// the count of a bu_scan_mark level -> the host's pinned block (same slots as bfs_finish writes); the spread counters start over
__global__ void bfs_publish(int32_t* __restrict__ spread, volatile int32_t* __restrict__ host_out, int32_t seq) {
    int v = 0;
    for (int i = threadIdx.x; i < BFS_SPREAD; i += 32) { v += spread[i]; spread[i] = 0; }
    v = warp_sum(v);
    if (threadIdx.x == 0) {
        host_out[0] = v; host_out[1] = v;
        __threadfence_system();
        host_out[2] = seq;
    }
}

// ---- runs of SMALL top-down levels (single GPU): one CTA, no grid-wide step at all ----------------------------------------------
// A level whose frontier has at most SMALL_Q vertices and SMALL_E out-edges is a latency chain, not a bandwidth problem: as
// prepare / expand / finish launches it costs ~35 us of launch, drain and bitmap passes for a few thousand edges.  Here one CTA of
// 1024 threads does the whole level between two __syncthreads: frontier degrees -> offsets in shared memory -> one edge slot per
// thread and step (row by binary search in shared memory) -> atomicOr straight into `visited`; the thread that sets the bit owns the
// discovery, writes level[u] and appends u to the next queue.  It keeps going, level after level (the reference's state machine
// evaluated in place), until the frontier needs the grid (more vertices / edges than the limits), the next level is bottom-up, or
// the traversal ends; the per-level counts go to the host's pinned block, which replays the state machine over them.
// The candidate and frontier bitmaps are not used by these levels; the frontier bitmap is rebuilt from the queue when the level
// that follows is bottom-up (only graphs of at most 20 x SMALL_E vertices can get there from a small level: RRD_THRESHOLD is N/20).
constexpr int SMALL_THREADS = 1024, SMALL_Q = 4096, SMALL_E = 32768, SMALL_LEVELS = 56, SMALL_HDR = 4;
constexpr int BFS_HOST_WORDS = SMALL_HDR + SMALL_LEVELS + 4;

template <typename L>
__global__ void __launch_bounds__(SMALL_THREADS, 1)
bfs_small_levels(const int32_t* __restrict__ begin, const int32_t* __restrict__ node_idx, const uint8_t* __restrict__ edge_mask,
                 int32_t* queue0, int32_t* queue1, uint32_t* visited, uint32_t* seen, uint32_t* frontier, int32_t words, L* level, int state,
                 int32_t curr_level, int32_t curr_count, int32_t qi, float thr, int has_reverse, int rrd_direct, int max_levels,
                 volatile int32_t* host_out, int32_t seq) {
    __shared__ int32_t s_off[SMALL_Q + 1];
    __shared__ int32_t s_beg[SMALL_Q];
    __shared__ int32_t s_warp[SMALL_THREADS / 32];
    __shared__ int32_t s_next;
    constexpr int PER = SMALL_Q / SMALL_THREADS;
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    int n_done = 0, ended = 0, refused = 0;
    int32_t own = curr_count;
    for (;;) {
        const int32_t* q = qi ? queue1 : queue0;
        int32_t* qn = qi ? queue0 : queue1;
        // degrees of the frontier: thread t takes the PER consecutive entries t*PER ...
        int32_t d[PER], b[PER], mine = 0;
#pragma unroll
        for (int k = 0; k < PER; k++) {
            const int32_t i = tid * PER + k;
            d[k] = 0; b[k] = 0;
            if (i < own) { const int32_t v = q[i]; b[k] = begin[v]; d[k] = begin[v + 1] - b[k]; }
            mine += d[k];
        }
        int32_t inc = mine;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) { const int32_t t = __shfl_up_sync(FULLMASK, inc, o); if (lane >= o) inc += t; }
        if (lane == 31) s_warp[wid] = inc;
        if (tid == 0) s_next = 0;
        __syncthreads();
        int32_t woff = 0, total = 0;
#pragma unroll
        for (int k = 0; k < SMALL_THREADS / 32; k++) { const int32_t t = s_warp[k]; if (k < wid) woff += t; total += t; }
        if (total > SMALL_E) { refused = 1; break; }       // (uniform) this level needs the grid
        int32_t run = inc - mine + woff;
#pragma unroll
        for (int k = 0; k < PER; k++) {
            const int32_t i = tid * PER + k;
            if (i < own) { s_off[i] = run; s_beg[i] = b[k]; }
            run += d[k];
        }
        if (tid == 0) s_off[own] = total;
        __syncthreads();
        const L nl = (L)(curr_level + 1);
        // four edge slots of a thread per step: rows first (shared memory only), then the four neighbour ids, then the four probes
        for (int32_t s0 = tid; s0 < total; s0 += 4 * SMALL_THREADS) {
            int32_t eidx[4], u[4];
            uint32_t vis[4];
#pragma unroll
            for (int j = 0; j < 4; j++) {
                const int32_t sl = s0 + j * SMALL_THREADS;
                eidx[j] = -1;
                if (sl < total) {
                    int32_t lo = 0, hi = own;            // last row with s_off[row] <= sl
                    while (hi - lo > 1) { const int32_t mid = (lo + hi) >> 1; if (s_off[mid] <= sl) lo = mid; else hi = mid; }
                    eidx[j] = s_beg[lo] + (sl - s_off[lo]);
                }
            }
#pragma unroll
            for (int j = 0; j < 4; j++) u[j] = eidx[j] >= 0 ? ld_stream(node_idx + eidx[j]) : -1;
#pragma unroll
            for (int j = 0; j < 4; j++) vis[j] = u[j] >= 0 ? visited[u[j] >> 5] : 0xffffffffu;
#pragma unroll
            for (int j = 0; j < 4; j++) {
                if (u[j] < 0) continue;
                const uint32_t bit = 1u << (u[j] & 31);
                // navigator (check_navigator, gm_bfs_template.h:488-492,596-598): filtered edges are not followed
                if ((vis[j] & bit) || (edge_mask && !edge_mask[eidx[j]])) continue;
                const uint32_t old = atomicOr(&visited[u[j] >> 5], bit);
                if (!(old & bit)) {
                    level[u[j]] = nl; qn[atomicAdd(&s_next, 1)] = u[j];
                    if (seen) atomicOr(&seen[u[j] >> 5], bit);          // seen = visited | candidates stays equal to visited
                }
            }
        }
        __syncthreads();
        const int32_t next_count = s_next;
        if (tid == 0) host_out[SMALL_HDR + n_done] = next_count;
        n_done++;
        if (next_count == 0) { ended = 1; break; }
        state = bfs_next_state(state, next_count, curr_count, thr, has_reverse != 0, rrd_direct != 0);
        curr_count = next_count; own = next_count; curr_level++; qi ^= 1;
        if (ctl_bottom_up(state)) {        // the scan of the next level probes the frontier bitmap: make it from the queue
            for (int32_t w = tid; w < words; w += SMALL_THREADS) frontier[w] = 0u;
            __syncthreads();
            for (int32_t i = tid; i < own; i += SMALL_THREADS) { const int32_t v = qn[i]; atomicOr(&frontier[v >> 5], 1u << (v & 31)); }
            break;
        }
        if (own > SMALL_Q || n_done >= max_levels) break;
        __syncthreads();
    }
    __syncthreads();
    // out-edges of the frontier the kernel leaves to the grid (queue[qi], `own` entries): the host's direction policy wants them
    __shared__ unsigned long long s_mf;
    if (tid == 0) s_mf = 0ull;
    __syncthreads();
    if (!ended) {
        const int32_t* q = qi ? queue1 : queue0;
        unsigned long long mine = 0;
        for (int32_t i = tid; i < own; i += SMALL_THREADS) { const int32_t v = q[i]; mine += (unsigned long long)(begin[v + 1] - begin[v]); }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) mine += __shfl_xor_sync(FULLMASK, mine, o);
        if (lane == 0 && mine) atomicAdd(&s_mf, mine);
    }
    __syncthreads();
    if (tid == 0) {
        host_out[0] = n_done; host_out[1] = ended; host_out[3] = refused;
        host_out[BFS_MF_WORD] = (int32_t)(uint32_t)s_mf; host_out[BFS_MF_WORD + 1] = (int32_t)(uint32_t)(s_mf >> 32);
        __threadfence_system();
        host_out[2] = seq;
    }
}
static_assert(BFS_MF_WORD == SMALL_HDR + SMALL_LEVELS && BFS_MF_WORD + 2 <= BFS_HOST_WORDS, "layout of the pinned host block");

// frontier bitmap from a (whole-frontier) queue; the bitmap was cleared before
__global__ void __launch_bounds__(BFS_THREADS)
bitmap_from_queue(const int32_t* __restrict__ queue, int32_t count, uint32_t* __restrict__ frontier) {
    for (int32_t i = blockIdx.x * BFS_THREADS + threadIdx.x; i < count; i += gridDim.x * BFS_THREADS) {
        const int32_t v = queue[i];
        atomicOr(&frontier[v >> 5], 1u << (v & 31));
    }
}

// ---- one launch per traversal (single GPU): the levels are separated by grid barriers instead of launches ------------------------
// counters[4] = arrivals of the current barrier, [5] = barrier generation, [6] = error (a barrier timed out).
struct BfsPersist {
    const int32_t *fwd_begin, *fwd_idx, *rev_begin, *rev_idx;
    const uint8_t* edge_mask;
    int32_t n, words;
    uint32_t *visited, *frontier, *cand;
    int32_t *queue0, *queue1, *blk_off, *blk_total;
    int2* chunks;
    int32_t* counters;
    void* level;
    BfsCtl* ctl;
};
constexpr unsigned BFS_BAR_SPINS = 1u << 26;       // x >= 40 ns: a barrier nobody completes fails the call after a few seconds instead of hanging

// Every CTA arrives; the LAST one runs `last` (the level's bookkeeping) before it opens the barrier.  False: timed out / failed.
template <typename F>
__device__ __forceinline__ bool bfs_grid_sync(int32_t* counters, F&& last) {
    __shared__ int s_ok;
    __syncthreads();
    if (threadIdx.x == 0) {
        volatile int32_t* gen = counters + 5;
        const int32_t g = *gen;
        int ok = 1;
        __threadfence();
        if (atomicAdd(counters + 4, 1) == (int)gridDim.x - 1) {
            __threadfence();
            last();
            counters[4] = 0;
            __threadfence();
            atomicAdd(counters + 5, 1);
        } else {
            unsigned spins = 0;
            while (*gen == g) {
                if (++spins > 256u) __nanosleep(40);
                if (spins > BFS_BAR_SPINS || *(volatile int32_t*)(counters + 6)) { counters[6] = 1; ok = 0; break; }
            }
        }
        __threadfence();
        s_ok = ok;
    }
    __syncthreads();
    return s_ok != 0;
}

template <typename L>
__global__ void __launch_bounds__(BFS_THREADS, 6)
bfs_persistent(const __grid_constant__ BfsPersist A) {
    volatile BfsCtl* vc = A.ctl;
    const int32_t nwarps = (gridDim.x * BFS_THREADS) >> 5, gwarp = (blockIdx.x * BFS_THREADS + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    auto nothing = [] {};
    for (;;) {
        // the control block was written by the last arrival of the previous barrier, before it opened
        if (vc->done) break;
        const int state = vc->state;
        const int32_t own = vc->own_count, qi = vc->qi;
        const L next_level = (L)(vc->curr_level + 1);
        int32_t* const q = qi ? A.queue1 : A.queue0;
        int32_t* const qn = qi ? A.queue0 : A.queue1;
        if (ctl_bottom_up(state)) {
            bu_scan_body(A.rev_begin, A.rev_idx, A.n, 0, A.words, A.frontier, A.visited, A.cand, nullptr);
        } else {
            td_prepare_body(A.fwd_begin, q, own, A.blk_off, A.blk_total, A.chunks, A.counters + 2, nullptr, nullptr);
            if (!bfs_grid_sync(A.counters, nothing)) return;
            if (own <= TD_BATCHED_MAX_FRONTIER)
                td_expand_body<true>(A.fwd_begin, A.fwd_idx, q, own, A.blk_off, A.blk_total, A.chunks, A.counters + 2, A.visited, A.cand,
                                     A.edge_mask, nullptr, nullptr, nullptr);
            else
                td_expand_body<false>(A.fwd_begin, A.fwd_idx, q, own, A.blk_off, A.blk_total, A.chunks, A.counters + 2, A.visited, A.cand,
                                      A.edge_mask, nullptr, nullptr, nullptr);
        }
        if (!bfs_grid_sync(A.counters, nothing)) return;
        for (int32_t w0 = gwarp * 32; w0 < A.words; w0 += nwarps * 32)
            finish_warp<L>(w0 + lane, A.words, A.words, 1, 0, A.cand, nullptr, A.visited, A.frontier, next_level, (L*)A.level, 0, A.words, qn,
                           A.counters);
        int32_t* const counters = A.counters;
        BfsCtl* const ctl = A.ctl;
        if (!bfs_grid_sync(A.counters, [counters, ctl] {
                const int32_t next_count = atomicAdd(&counters[0], 0), own_next = atomicAdd(&counters[1], 0);
                counters[0] = 0; counters[1] = 0; counters[2] = 0;
                ctl_advance(ctl, next_count, own_next);
            })) return;
    }
}

__global__ void bfs_persist_init(BfsCtl* ctl, float thr, int has_reverse, int rrd_direct, int32_t* counters) {
    ctl->state = ST_SMALL; ctl->curr_level = 0; ctl->curr_count = 1; ctl->own_count = 1; ctl->qi = 0; ctl->done = 0;
    ctl->n_trace = 1; ctl->overflow = 0; ctl->thr = thr; ctl->has_reverse = has_reverse; ctl->rrd_direct = rrd_direct;
    ctl->trace_state[0] = ST_SMALL; ctl->trace_count[0] = 1;
    counters[4] = 0; counters[5] = 0; counters[6] = 0;
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

int ensure_ws(gmc_graph* g, int world) {
    if (g->bfs && g->bfs->world == world) return GMC_OK;
    if (g->bfs) gmc_bfs_ws_free(g);
    gmc_bfs_ws* w = new gmc_bfs_ws();
    g->bfs = w;
    w->N = g->N;
    w->world = world;
    w->words = (g->N + 31) / 32;
    w->wpr = div_up(w->words > 0 ? w->words : 1, world);
    w->wp = w->wpr * world;
    const size_t wb = sizeof(uint32_t) * (size_t)w->wp;
    const size_t nn = (size_t)(g->N ? g->N : 1);
    w->chunk_cap = (size_t)g->M / TD_CHUNK + nn / BFS_THREADS + 2;
    int rc = gmc_malloc((void**)&w->visited, wb);
    if (rc == GMC_OK) rc = gmc_malloc((void**)&w->frontier_bm, 2 * wb);      // current / next level (ping-pong)
    if (rc == GMC_OK) rc = world > 1 ? gmc_malloc_ipc((void**)&w->cand, 2 * wb) : gmc_malloc((void**)&w->cand, wb);
    if (rc == GMC_OK && world > 1) rc = gmc_malloc((void**)&w->allcand, wb * (size_t)world);
    if (rc == GMC_OK) rc = gmc_malloc((void**)&w->queue[0], sizeof(int32_t) * nn);
    if (rc == GMC_OK) rc = gmc_malloc((void**)&w->queue[1], sizeof(int32_t) * nn);
    if (rc == GMC_OK) rc = gmc_malloc((void**)&w->blk_off, sizeof(int32_t) * nn);
    if (rc == GMC_OK) rc = gmc_malloc((void**)&w->blk_total, sizeof(int32_t) * (nn / BFS_THREADS + 1));
    if (rc == GMC_OK) rc = gmc_malloc((void**)&w->chunks, sizeof(int2) * w->chunk_cap);
    if (rc == GMC_OK) rc = gmc_malloc((void**)&w->counters, sizeof(int32_t) * 8);
    if (rc == GMC_OK) rc = gmc_malloc((void**)&w->spread, sizeof(int32_t) * BFS_SPREAD);
    if (rc == GMC_OK) rc = gmc_malloc((void**)&w->level, sizeof(int32_t) * nn);
    if (rc == GMC_OK) rc = gmc_malloc((void**)&w->edge_sum, 2 * sizeof(unsigned long long));     // [1] = out-edges of the new frontier (bfs_finish)
    if (rc == GMC_OK) rc = gmc_malloc((void**)&w->ctl, sizeof(BfsCtl));
    if (rc == GMC_OK && cudaMallocHost((void**)&w->h_counters, sizeof(int32_t) * BFS_HOST_WORDS) != cudaSuccess) {
        set_error("bfs: cudaMallocHost failed"); rc = GMC_ERR_NOMEM;
    }
    if (rc == GMC_OK) memset(w->h_counters, 0, sizeof(int32_t) * BFS_HOST_WORDS);
    if (rc != GMC_OK) { gmc_bfs_ws_free(g); return rc; }
    return GMC_OK;
}

// best_parent of the (forward, reverse) pair a traversal runs over; built on first use, rebuilt when another pair is asked for
int ensure_best_parent(gmc_bfs_ws* w, int32_t n, const int32_t* fwd_begin, const int32_t* rev_begin, const int32_t* rev_idx, cudaStream_t st) {
    if (w->bp_key == rev_begin && w->best_parent) return GMC_OK;
    if (!w->best_parent) GMC_TRY(gmc_malloc((void**)&w->best_parent, sizeof(int32_t) * (size_t)(n ? n : 1)));
    if (!w->bp_list) GMC_TRY(gmc_malloc((void**)&w->bp_list, sizeof(int32_t) * (1 + BP_LIST_CAP)));
    GMC_CUDA(cudaMemsetAsync(w->bp_list, 0, sizeof(int32_t), st));
    bp_rows<<<(unsigned)min(div_up(n, BFS_THREADS), sm_count() * 32), BFS_THREADS, 0, st>>>(fwd_begin, rev_begin, rev_idx, n, w->best_parent, w->bp_list);
    bp_huge<<<(unsigned)sm_count(), 1024, 0, st>>>(fwd_begin, rev_begin, rev_idx, w->best_parent, w->bp_list);
    GMC_CUDA(cudaGetLastError());
    w->bp_key = rev_begin;
    return GMC_OK;
}

// Wait until the kernel enqueued last has published its sequence number into the pinned block.
int wait_host_seq(gmc_bfs_ws* w, cudaStream_t st) {
    volatile int32_t* h = w->h_counters;
    unsigned spins = 0;
    while (h[2] != w->seq) {
        if ((++spins & 0x3fffu) == 0) {            // a failed launch or a fault would never publish: ask the stream now and then
            const cudaError_t q = cudaStreamQuery(st);
            if (q != cudaErrorNotReady && q != cudaSuccess) { set_error("bfs: %s", cudaGetErrorString(q)); return GMC_ERR_CUDA; }
            if (q == cudaSuccess && h[2] != w->seq) { set_error("bfs: level counts were not published"); return GMC_ERR_CUDA; }
        }
    }
    __sync_synchronize();
    return GMC_OK;
}

// The traversal.  fwd_* = arrays followed top-down, rev_* = arrays followed bottom-up (NULL: no bottom-up).
template <typename L>
int run_bfs(gmc_graph* g, gmc_comm* comm, const int32_t* fwd_begin, const int32_t* fwd_idx, const int32_t* rev_begin,
            const int32_t* rev_idx, int32_t root, L invalid, L* level_out_dev, gmc_bfs_info* info, cudaStream_t st,
            const uint8_t* edge_mask = nullptr, bool save_child = false, bool free_policy = false) {
    GMC_REQUIRE(root >= 0 && root < g->N, GMC_ERR_ARG, "bfs: root %d out of range [0,%d)", root, g->N);
    GMC_REQUIRE(g->has_forward, GMC_ERR_STATE, "bfs: the forward CSR (node_idx) was not uploaded for this graph");
    const int world = comm ? comm->world : 1, rank = comm ? comm->rank : 0;
    GMC_TRY(ensure_ws(g, world));
    gmc_bfs_ws* w = g->bfs;
    const int32_t n = g->N;
    const bool has_reverse = rev_begin != nullptr;
    // a save_child instantiation never switches QUE / Q2R -> RRD directly (gm_bfs_template.h:374,:389), only RD -> RRD (:398)
    const bool rrd_direct = has_reverse && !save_child;
    L* level = level_out_dev ? level_out_dev : (L*)w->level;
    w->level_count.clear(); w->state_trace.clear();
    w->valid = false;
    const int32_t own_lo = min(rank * w->wpr, w->words), own_hi = min((rank + 1) * w->wpr, w->words);   // owned words
    const bool root_owned = (root >> 5) >= own_lo && (root >> 5) < own_hi;

    cudaEvent_t ev0, ev1;
    GMC_CUDA(cudaEventCreate(&ev0));
    GMC_CUDA(cudaEventCreate(&ev1));
    int launches = 0, bottom_up = 0;
    // sharded: map the peers' candidate buffers once per workspace (collective: every rank is here on its first traversal);
    // not mappable (or GMC_BFS_NO_PEER=1): the levels fall back to allgathering `world` full bitmaps
    if (world > 1 && !w->peer_tried) {
        w->peer_tried = true;
        w->rank = rank;
        // Measured on the TW-shaped graph (profiles/r2t_bfs_exchange_8gpu.txt): owner-directed 1.81 ms per traversal on 8 GPUs
        // (2.99 ms on 2) against 1.52 ms (2.30 ms) for the allgather of full bitmaps -- the barrier per top-down level (an NCCL
        // allreduce) and the remote atomics cost more than the 36 MB per level they save.  So it is opt-in: GMC_BFS_PEER=1.
        if (getenv("GMC_BFS_PEER") && !getenv("GMC_BFS_NO_PEER")) {
            void* bases[8];
            const int prc = world <= 8 ? gmc_comm_ipc_open_all(comm, w->cand, bases, st) : GMC_ERR_STATE;
            if (prc == GMC_OK) { for (int r = 0; r < world; r++) w->peer_base[r] = bases[r]; w->peer_ok = true; w->peer_local = comm->local; }
            else if (prc != GMC_ERR_STATE) return prc;
        }
    }
    const bool peer = world > 1 && w->peer_ok;
    GMC_CUDA(cudaEventRecord(ev0, st));
    const unsigned init_grid = (unsigned)min(div_up(max(n / (16 / (int)sizeof(L)), w->wp), 256), sm_count() * 16);
    // Host-driven loop on a single GPU: `cand` is kept as the SEEN bitmap = visited | candidates of the level being expanded, so
    // td_expand probes ONE bitmap per edge instead of visited and then cand (the hub root's 105 K-vertex top-down level: 68.8 M L2
    // sectors in 221 us, profiles/r3k_bfs_launches_hub.md -- one random sector per probe); bfs_finish takes new = seen & ~visited
    // and leaves seen = the new visited.  GMC_BFS_NO_SEEN=1: separate bitmaps, cleared every level (what the sharded paths use).
    const bool seen_mode = world == 1 && getenv("GMC_BFS_NO_SEEN") == nullptr && getenv("GMC_BFS_DEVICE_LOOP") == nullptr &&
                           getenv("GMC_BFS_PERSISTENT") == nullptr && getenv("GMC_BFS_FUSED_BU") == nullptr;
    int qi = 0;     // queue[qi] = current level
    // Sharded runs: the CSR, the bitmaps and level[] are replicated, so a level whose frontier is small is run by EVERY rank in
    // full (bfs_small_levels, the single-GPU kernel) instead of expanding 1/world of it and allgathering `world` full bitmaps: no
    // exchange at all for the first and the last levels of a traversal.  queue_full: queue[qi] holds the whole frontier (identical
    // sets on every rank), not the owned part.  GMC_BFS_NO_SMALL_SHARDED=1: round-1 behaviour.
    const bool small_sharded = world > 1 && getenv("GMC_BFS_NO_SMALL") == nullptr && getenv("GMC_BFS_NO_SMALL_SHARDED") == nullptr;
    bool queue_full = small_sharded;
    // GMC_BFS_A2A=1 (experiment): exchange of the sharded top-down levels that stay as all-to-all of the candidate slices + OR at
    // the owner + allgather of the owned slices instead of allgathering `world` whole bitmaps.  Measured on 4 GPUs, TW-shaped
    // (profiles/r3o): 1.424 against 1.406 ms per traversal -- two collectives and a kernel cost what the smaller transfer saves.
    const bool a2a = world > 1 && (getenv("GMC_BFS_A2A") ? atoi(getenv("GMC_BFS_A2A")) != 0 : false);
    bfs_init<L><<<init_grid, 256, 0, st>>>(level, n, invalid, root, w->visited, w->frontier_bm, w->cand, w->wp,
                                           w->queue[0], (root_owned || small_sharded) ? 1 : 0, w->counters, world > 1 ? 1 : 0, seen_mode ? 1 : 0, w->spread,
                                           w->edge_sum + 1);
    launches++;
    // nobody sends candidates into a buffer its owner has not cleared yet
    if (peer) GMC_TRY(gmc_comm_stream_barrier(comm, st));

    int state = ST_SMALL;
    int32_t curr_level = 0, curr_count = 1, own_count = root_owned ? 1 : 0;
    // Direction policy.  gm_bfs_template traversals (levels + state trace are compared with the reference) switch on the
    // reference's rule: bottom-up once a level finds >= 5 % of the vertices (RRD_THRESHOLD).  hop_dist is not a template instance in
    // the reference (hop_dist.gm is a frontier-filtered sweep), only its distances are defined, so it is free to switch on EDGES
    // (Beamer et al., direction-optimizing BFS): bottom-up as soon as the out-edges of the new frontier, m_f, exceed 1/alpha of the
    // edges not yet explored, while the frontier grows.  On the TW-shaped graph the reference's rule runs the widest level top-down
    // (e.g. 318 K frontier vertices, > 10^8 out-edges, 16 M vertices found: 1.17 ms of a 1.8 ms traversal on 4 GPUs); leaving
    // bottom-up follows the reference's rule (RRD -> RR2Q -> QUE) either way.  GMC_BFS_REF_POLICY=1: the reference's rule everywhere.
    const bool beamer = free_policy && has_reverse && getenv("GMC_BFS_REF_POLICY") == nullptr && getenv("GMC_BFS_DEVICE_LOOP") == nullptr &&
                        getenv("GMC_BFS_PERSISTENT") == nullptr && getenv("GMC_BFS_FUSED_BU") == nullptr;
    const double alpha = getenv("GMC_BFS_ALPHA") ? atof(getenv("GMC_BFS_ALPHA")) : 14.0;
    double explored = 0.0;                // out-edges of the frontiers so far (as far as the levels reported them)
    // RRD_THRESHOLD, gm_bfs_template.h:359 (float arithmetic as in the reference); beamer: the vertex-count rule never fires
    const float thr = beamer ? 3.0e38f : (float)n * 0.05f;
    const int32_t* const deg_begin = (beamer || world > 1) ? fwd_begin : nullptr;     // the levels report the out-edges of the frontier they leave
    unsigned long long* const mf_dev = w->edge_sum + 1;
    // Sharded runs: a top-down level whose frontier has at most red_max out-edges is expanded IN FULL by every rank (the CSR and
    // the bitmaps are replicated): no exchange, where expanding 1/world of it costs an allgather of `world` whole bitmaps
    // (TW-shaped, 4 GPUs: 150-180 us for a 51-vertex level that one GPU does in 105 us).  GMC_BFS_RED_EDGES=0: never.
    // The limit grows with the world size (2 M out-edges per rank, at most 16 M): what a sharded level saves is (1 - 1/world) of the
    // one-GPU time, what it costs is the exchange of `world` bitmaps -- on 8 GPUs a 133 K-vertex level near the end of a traversal took
    // 162 us sharded against 55 us on one GPU (profiles/r3r_bfs_levels_8gpu.txt).
    const double red_max = getenv("GMC_BFS_RED_EDGES") ? atof(getenv("GMC_BFS_RED_EDGES")) : (double)(1 << 21) * (double)min(world, 8);
    double last_mf = 0.0;                 // out-edges of the current frontier ...
    bool mf_known = false;                // ... when the level that left it reported them
    // next state of a level that would run top-down: bottom-up instead when the frontier's out-edges say so
    auto policy = [&](int next_state, int32_t next_count, int32_t prev_count) {
        mf_known = false;
        if (ctl_bottom_up(next_state)) return next_state;
        last_mf = (double)(((unsigned long long)(uint32_t)w->h_counters[BFS_MF_WORD + 1] << 32) | (uint32_t)w->h_counters[BFS_MF_WORD]);
        mf_known = true;
        explored += last_mf;
        const double mu = (double)g->M - explored;
        return (beamer && next_count > prev_count && last_mf * alpha > mu) ? (int)ST_RRD : next_state;
    };
    const unsigned fin_grid = (unsigned)div_up(w->words, BFS_THREADS);
    const unsigned exp_grid = (unsigned)sm_count() * 8;
    // Measured on the LJ-shaped graph (gpurun r2e): device-driven 0.43 ms per traversal, host-driven 0.37 ms -- five launches per
    // level, two or three of them leaving at once (~3 us each), and the levels enqueued past the end of the traversal cost more
    // than the host round trip they replace.  So the host-driven loop stays the default; GMC_BFS_DEVICE_LOOP=1 selects this one.
    bool device_driven = world == 1 && getenv("GMC_BFS_DEVICE_LOOP") != nullptr && getenv("GMC_BFS_HOST_LOOP") == nullptr;
    // GMC_BFS_PERSISTENT=1: ONE cooperative launch per traversal (bfs_persistent), the levels separated by grid barriers instead of
    // launches.  Measured SLOWER than the host-driven loop (gpurun r3g, LJ-shaped: 0.361 against 0.336 ms from root 3, 0.568 against
    // 0.447 ms from the hub; 58 against 37 us for a one-level traversal): a software grid barrier (~3-4 us with the bookkeeping of the
    // last arrival) is no cheaper than a kernel boundary on this part, and the static word striding of the bottom-up scan balances
    // worse than 19 K CTAs handed out by the hardware.  Opt-in.
    bool persistent = world == 1 && !device_driven && getenv("GMC_BFS_PERSISTENT") != nullptr && getenv("GMC_BFS_HOST_LOOP") == nullptr;
    static thread_local int blocks_per_sm[2] = {0, 0};
    if (persistent) {
        int& bps = blocks_per_sm[sizeof(L) == 2];
        if (!bps) {
            int coop = 0, dev = 0;
            GMC_CUDA(cudaGetDevice(&dev));
            GMC_CUDA(cudaDeviceGetAttribute(&coop, cudaDevAttrCooperativeLaunch, dev));
            int nb = 0;
            if (coop) GMC_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, bfs_persistent<L>, BFS_THREADS, 0));
            bps = nb > 0 ? nb : -1;
        }
        if (bps < 0) persistent = false;
    }
    if (persistent) {
        BfsCtl* ctl = w->ctl;
        bfs_persist_init<<<1, 1, 0, st>>>(ctl, thr, has_reverse ? 1 : 0, rrd_direct ? 1 : 0, w->counters);
        BfsPersist A;
        A.fwd_begin = fwd_begin; A.fwd_idx = fwd_idx; A.rev_begin = rev_begin; A.rev_idx = rev_idx; A.edge_mask = edge_mask;
        A.n = n; A.words = w->words; A.visited = w->visited; A.frontier = w->frontier_bm; A.cand = w->cand;
        A.queue0 = w->queue[0]; A.queue1 = w->queue[1]; A.blk_off = w->blk_off; A.blk_total = w->blk_total; A.chunks = w->chunks;
        A.counters = w->counters; A.level = level; A.ctl = ctl;
        void* kargs[] = {&A};
        const unsigned grid = (unsigned)(sm_count() * blocks_per_sm[sizeof(L) == 2]);
        GMC_CUDA(cudaLaunchCooperativeKernel((const void*)bfs_persistent<L>, dim3(grid), dim3(BFS_THREADS), kargs, 0, st));
        launches += 2;
        GMC_CUDA(cudaEventRecord(ev1, st));
        static thread_local BfsCtl hc;
        static thread_local int32_t hcnt[8];
        GMC_CUDA(cudaMemcpyAsync(&hc, ctl, sizeof(BfsCtl), cudaMemcpyDeviceToHost, st));
        GMC_CUDA(cudaMemcpyAsync(hcnt, w->counters, sizeof(hcnt), cudaMemcpyDeviceToHost, st));
        GMC_CUDA(cudaStreamSynchronize(st));
        GMC_REQUIRE(hcnt[6] == 0 && hc.done, GMC_ERR_CUDA, "bfs: a grid barrier of the traversal kernel timed out");
        if (hc.overflow) {       // more levels than the trace holds: the host-driven loop below starts over
            persistent = false;
            GMC_CUDA(cudaEventRecord(ev0, st));
            bfs_init<L><<<init_grid, 256, 0, st>>>(level, n, invalid, root, w->visited, w->frontier_bm, w->cand, w->wp,
                                                   w->queue[0], root_owned ? 1 : 0, w->counters, 0, 0, w->spread);
            launches++;
        } else {
            for (int k = 0; k < hc.n_trace; k++) {
                w->state_trace.push_back(hc.trace_state[k]); w->level_count.push_back(hc.trace_count[k]);
                if (hc.trace_state[k] == ST_RRD || hc.trace_state[k] == ST_RR2Q) bottom_up++;
            }
            curr_level = hc.n_trace - 1;
            GMC_REQUIRE(sizeof(L) == 4 || curr_level + 1 <= 32767, GMC_ERR_STATE, "bfs: more levels than level_t can hold");
        }
    }
    if (device_driven) {
        // Single GPU: the kernels of several levels are enqueued back to back, the state machine runs on the device
        // (BfsCtl); the host only looks at the `done` word the last bfs_finish of a batch publishes.
        BfsCtl* ctl = w->ctl;
        bfs_ctl_init<<<1, 1, 0, st>>>(ctl, thr, has_reverse ? 1 : 0, rrd_direct ? 1 : 0);
        launches++;
        const unsigned prep_grid = (unsigned)sm_count() * 4, scan_grid = (unsigned)sm_count() * 16;
        int batch = 5;          // levels per batch: most traversals of a power-law graph end within 6..9 levels
        bool done = false;
        for (int guard = 0; !done && guard < BFS_TRACE_CAP + 8; guard += batch, batch = 2) {
            for (int b = 0; b < batch; b++) {
                bfs_level_a<<<scan_grid, BFS_THREADS, 0, st>>>(has_reverse ? rev_begin : nullptr, rev_idx, n, own_lo, own_hi, w->frontier_bm, w->visited, w->cand,
                                                               fwd_begin, w->queue[0], w->queue[1], w->blk_off, w->blk_total, w->chunks, w->counters + 2, ctl);
                td_expand_any<<<exp_grid, BFS_THREADS, 0, st>>>(fwd_begin, fwd_idx, w->queue[0], w->queue[1], w->blk_off, w->blk_total, w->chunks,
                                                                w->counters + 2, w->visited, w->cand, edge_mask, ctl);
                bfs_finish<L><<<fin_grid, BFS_THREADS, 0, st>>>(w->words, w->wp, 1, 0, w->cand, w->allcand, w->visited, w->frontier_bm, (L)0, level,
                                                                own_lo, own_hi, w->queue[0], w->counters, w->h_counters, ++w->seq, w->queue[1], ctl);
                launches += 3;
            }
            GMC_CUDA(cudaGetLastError());
            volatile int32_t* h = w->h_counters;
            unsigned spins = 0;
            while (h[2] != w->seq) {
                if ((++spins & 0x3fffu) == 0) {
                    const cudaError_t q = cudaStreamQuery(st);
                    if (q != cudaErrorNotReady && q != cudaSuccess) { set_error("bfs: %s", cudaGetErrorString(q)); return GMC_ERR_CUDA; }
                    if (q == cudaSuccess && h[2] != w->seq) { set_error("bfs: level counts were not published"); return GMC_ERR_CUDA; }
                }
            }
            __sync_synchronize();
            done = h[1] != 0;
        }
        GMC_CUDA(cudaEventRecord(ev1, st));
        static thread_local BfsCtl hc;
        GMC_CUDA(cudaMemcpyAsync(&hc, ctl, sizeof(BfsCtl), cudaMemcpyDeviceToHost, st));
        GMC_CUDA(cudaStreamSynchronize(st));
        if (!done || hc.overflow) device_driven = false;       // more levels than the trace holds: the host-driven loop below takes over
        else {
            for (int k = 0; k < hc.n_trace; k++) {
                w->state_trace.push_back(hc.trace_state[k]); w->level_count.push_back(hc.trace_count[k]);
                if (hc.trace_state[k] == ST_RRD || hc.trace_state[k] == ST_RR2Q) bottom_up++;
            }
            curr_level = hc.n_trace - 1;
            GMC_REQUIRE(sizeof(L) == 4 || curr_level + 1 <= 32767, GMC_ERR_STATE, "bfs: more levels than level_t can hold");
        }
        if (!device_driven) {       // start over on the host-driven path
            GMC_CUDA(cudaEventRecord(ev0, st));
            bfs_init<L><<<init_grid, 256, 0, st>>>(level, n, invalid, root, w->visited, w->frontier_bm, w->cand, w->wp,
                                                   w->queue[0], root_owned ? 1 : 0, w->counters, 0, 0, w->spread);
            launches++; bottom_up = 0;
        }
    }
    // runs of small top-down levels go to bfs_small_levels (one CTA, several levels per launch); GMC_BFS_NO_SMALL=1: never
    const bool small_ok = (world == 1 && getenv("GMC_BFS_NO_SMALL") == nullptr) || small_sharded;
    // GMC_BFS_FUSED_BU=1: bottom-up levels in one launch (bu_scan_finish) on a single GPU.  Measured SLOWER than scan + finish as
    // two launches (gpurun r3i, LJ-shaped, root 3: 0.324 against 0.281 ms; hub root 0.452 against 0.430): the block-wide queue
    // allocation keeps the 8 warps of a CTA resident until its slowest row scan ends, which costs more than the bitmap pass and
    // the launch it saves.  Opt-in.
    const bool fused_bu = world == 1 && getenv("GMC_BFS_FUSED_BU") != nullptr;
    // RRD levels (the next level is bottom-up again: no queue needed) as scan-and-mark + a one-warp publish; GMC_BFS_NO_BU_MARK=1: scan + finish
    const bool bu_mark = world == 1 && !fused_bu && getenv("GMC_BFS_NO_BU_MARK") == nullptr;
    uint32_t *fcur = w->frontier_bm, *fnext = w->frontier_bm + w->wp;      // every level writes fnext, then they swap
    int32_t small_refused = -1;       // level whose frontier had more than SMALL_E edges: it takes the grid path
    // GMC_BFS_LEVEL_TIMING=1 (profiling aid): host wall time of every level of the host-driven loop on stderr (rank 0)
    // bottom-up scans probe the in-neighbour of largest out-degree first (bp_rows).  The table costs one pass over all reverse
    // edges with a random out-degree lookup each -- more than a traversal -- so it is built for graphs that are REUSED: at the first
    // bottom-up level of the third traversal on (a one-shot driver such as bin/hop_dist never pays for it; bc / bc_random, many
    // roots, the bench do).  GMC_BFS_NO_BEST_PARENT=1: never; GMC_BFS_BEST_PARENT_NOW=1: from the first traversal.
    const bool bp_on = has_reverse && getenv("GMC_BFS_NO_BEST_PARENT") == nullptr &&
                       (w->traversals >= 2 || (w->best_parent && w->bp_key == rev_begin) || getenv("GMC_BFS_BEST_PARENT_NOW") != nullptr);
    const int32_t* best_parent = nullptr;
    const bool lvl_timing = getenv("GMC_BFS_LEVEL_TIMING") != nullptr && rank == 0;
    auto lvl_t0 = std::chrono::steady_clock::now();
    auto lvl_report = [&](const char* what, int st_, int32_t lvl, int32_t cnt) {
        if (!lvl_timing) return;
        const auto t1 = std::chrono::steady_clock::now();
        fprintf(stderr, "bfs level %d state %d count %d: %s %.1f us\n", lvl, st_, cnt, what, std::chrono::duration<double, std::micro>(t1 - lvl_t0).count());
        lvl_t0 = t1;
    };
    if (lvl_timing) { cudaStreamSynchronize(st); lvl_report("init", state, 0, 1); }
    for (; !device_driven && !persistent;) {
        if (small_ok && !ctl_bottom_up(state) && (world > 1 ? curr_count : own_count) > 0 && (world > 1 ? curr_count : own_count) <= SMALL_Q &&
            small_refused != curr_level) {
            if (world > 1 && !queue_full) {        // back from sharded levels: the whole frontier from the replicated bitmap
                GMC_CUDA(cudaMemsetAsync(w->counters + 7, 0, sizeof(int32_t), st));        // ([7]: a word no level kernel uses)
                queue_from_bitmap<<<(unsigned)min(div_up(w->words, BFS_THREADS), sm_count() * 8), BFS_THREADS, 0, st>>>(fcur, w->words, w->queue[qi],
                                                                                                                       w->counters + 7, SMALL_Q);
                launches++;
                queue_full = true;
            }
            int max_levels = SMALL_LEVELS;
            if (sizeof(L) == 2) max_levels = min(max_levels, 32767 - curr_level);
            GMC_REQUIRE(max_levels > 0, GMC_ERR_STATE, "bfs: more levels than level_t can hold");
            bfs_small_levels<L><<<1, SMALL_THREADS, 0, st>>>(fwd_begin, fwd_idx, edge_mask, w->queue[0], w->queue[1], w->visited,
                                                             seen_mode ? w->cand : nullptr, fcur, w->words, level, state, curr_level, curr_count, qi, thr, has_reverse ? 1 : 0,
                                                             rrd_direct ? 1 : 0, max_levels, w->h_counters, ++w->seq);
            launches++;
            GMC_CUDA(cudaGetLastError());
            GMC_TRY(wait_host_seq(w, st));
            const int k = w->h_counters[0];
            const bool refused = w->h_counters[3] != 0;
            lvl_report(refused ? "small levels (stopped: too many edges)" : "small levels", state, curr_level, curr_count);
            bool ended = false;
            for (int j = 0; j < k && !ended; j++) {      // replay the state machine over the counts of the levels it ran
                w->state_trace.push_back(state);
                w->level_count.push_back(curr_count);
                const int32_t next_count = w->h_counters[SMALL_HDR + j];
                if (next_count == 0) { ended = true; break; }
                state = bfs_next_state(state, next_count, curr_count, thr, has_reverse, rrd_direct);
                curr_count = next_count; own_count = next_count; curr_level++; qi ^= 1;
            }
            if (ended) break;
            if (refused || k == 0) small_refused = curr_level;     // the level it stopped at has too many edges for one CTA
            if (!ctl_bottom_up(state)) {
                // (prev_count 0: a level the small kernel hands over has more edges than it takes -- count it as growing)
                const int ns = policy(state, curr_count, 0);
                if (ctl_bottom_up(ns)) {        // the scan probes the frontier bitmap, which the small levels do not keep
                    GMC_CUDA(cudaMemsetAsync(fcur, 0, sizeof(uint32_t) * (size_t)w->wp, st));
                    bitmap_from_queue<<<(unsigned)min(div_up(curr_count, BFS_THREADS), sm_count() * 8), BFS_THREADS, 0, st>>>(w->queue[qi], curr_count, fcur);
                    launches++;
                    state = ns;
                }
            }
            continue;
        }
        w->state_trace.push_back(state);
        w->level_count.push_back(curr_count);
        GMC_REQUIRE(sizeof(L) == 4 || curr_level + 1 <= 32767, GMC_ERR_STATE, "bfs: more levels than level_t can hold");
        const L nl = (L)(curr_level + 1);
        int use_all = 0;
        // sharded: the candidate buffer alternates with the level (a rank one level ahead sends into the other buffer)
        uint32_t* const cand_l = w->cand + (world > 1 ? (size_t)(curr_level & 1) * w->wp : 0);
        PeerCand pc = PeerCand();
        if (peer) {
            for (int r = 0; r < world; r++) pc.base[r] = (uint32_t*)w->peer_base[r] + (size_t)(curr_level & 1) * w->wp;
            pc.wpr = w->wpr; pc.world = world; pc.rank = rank;
        }
        bool published = false;       // the level's counts are already on their way to the host (no bfs_finish)
        bool redundant = false;       // sharded: this (top-down) level is expanded in full by every rank
        if (ctl_bottom_up(state) && bp_on && !best_parent) {
            GMC_TRY(ensure_best_parent(w, n, fwd_begin, rev_begin, rev_idx, st));
            best_parent = w->best_parent;
        }
        if (state == ST_RRD && bu_mark) {
            bu_scan_mark<L><<<div_up(w->words * 32, BFS_THREADS), BFS_THREADS, 0, st>>>(rev_begin, rev_idx, n, w->words, fcur, fnext, w->visited,
                                                                                       seen_mode ? w->cand : nullptr, nl, level, w->spread, best_parent);
            bfs_publish<<<1, 32, 0, st>>>(w->spread, w->h_counters, ++w->seq);
            launches += 2;
            bottom_up++;
            published = true;
        } else if ((state == ST_RRD || state == ST_RR2Q) && fused_bu) {
            bu_scan_finish<L><<<div_up(w->words * 32, BFS_THREADS), BFS_THREADS, 0, st>>>(rev_begin, rev_idx, n, w->words, fcur, fnext, w->visited, nl,
                                                                                         level, w->queue[qi ^ 1], w->counters, w->h_counters, ++w->seq);
            launches++;
            bottom_up++;
            published = true;
        } else if (state == ST_RRD || state == ST_RR2Q) {
            if (own_hi > own_lo) {
                bu_scan<<<div_up((own_hi - own_lo) * 32, BFS_THREADS), BFS_THREADS, 0, st>>>(rev_begin, rev_idx, n, own_lo, own_hi,
                                                                                            fcur, w->visited, cand_l, best_parent);
                launches++;
            }
            bottom_up++;
            if (world > 1) GMC_TRY(gmc_comm_allgather(comm, cand_l + (size_t)rank * w->wpr, cand_l, sizeof(uint32_t) * (size_t)w->wpr, st));
        } else {
            redundant = world > 1 && !peer && mf_known && last_mf <= red_max && red_max > 0.0;
            if (redundant && !queue_full) {        // the whole frontier from the replicated bitmap
                GMC_CUDA(cudaMemsetAsync(w->counters + 7, 0, sizeof(int32_t), st));
                queue_from_bitmap<<<(unsigned)min(div_up(w->words, BFS_THREADS), sm_count() * 8), BFS_THREADS, 0, st>>>(fcur, w->words, w->queue[qi],
                                                                                                                       w->counters + 7, n);
                launches++;
                queue_full = true;
            }
            // queue_full (sharded, after levels every rank ran in full): the queue holds the whole frontier; unless this level is run
            // in full as well, td_prepare lets only the owned vertices count
            if (queue_full) own_count = curr_count;
            const bool filter = queue_full && !redundant;
            const int32_t v_lo = filter ? own_lo * 32 : 0, v_hi = filter ? (int32_t)min((long long)own_hi * 32, (long long)INT_MAX) : INT_MAX;
            if (own_count > 0) {
                td_prepare<<<div_up(own_count, BFS_THREADS), BFS_THREADS, 0, st>>>(fwd_begin, w->queue[qi], own_count, w->blk_off,
                                                                                  w->blk_total, w->chunks, w->counters + 2, v_lo, v_hi);
                // (full grid whatever the frontier size: the levels of few vertices AND few edges went to bfs_small_levels, and a
                // 500-vertex frontier of hubs has 10^6 edges -- on one block per SM that level took 43 us instead of 22)
                if (seen_mode && own_count <= TD_BATCHED_MAX_FRONTIER)
                    td_expand_seen<true><<<exp_grid, BFS_THREADS, 0, st>>>(fwd_begin, fwd_idx, w->queue[qi], own_count, w->blk_off, w->blk_total,
                                                                           w->chunks, w->counters + 2, w->cand, edge_mask);
                else if (seen_mode)
                    td_expand_seen<false><<<exp_grid, BFS_THREADS, 0, st>>>(fwd_begin, fwd_idx, w->queue[qi], own_count, w->blk_off, w->blk_total,
                                                                            w->chunks, w->counters + 2, w->cand, edge_mask);
                else if (own_count <= TD_BATCHED_MAX_FRONTIER)
                    td_expand<true><<<exp_grid, BFS_THREADS, 0, st>>>(fwd_begin, fwd_idx, w->queue[qi], own_count, w->blk_off, w->blk_total,
                                                                      w->chunks, w->counters + 2, w->visited, cand_l, edge_mask, pc);
                else
                    td_expand<false><<<exp_grid, BFS_THREADS, 0, st>>>(fwd_begin, fwd_idx, w->queue[qi], own_count, w->blk_off, w->blk_total,
                                                                       w->chunks, w->counters + 2, w->visited, cand_l, edge_mask, pc);
                launches += 2;
            }
            if (redundant) {
                // nothing to exchange: every rank found every candidate
            } else if (peer) {
                // every rank's candidates have been ORed into their owners' slices once all expansions are complete (barrier);
                // then the owned slices are allgathered in place: N/8 bytes per level in total instead of `world` full bitmaps
                GMC_TRY(gmc_comm_stream_barrier(comm, st));
                GMC_TRY(gmc_comm_allgather(comm, cand_l + (size_t)rank * w->wpr, cand_l, sizeof(uint32_t) * (size_t)w->wpr, st));
            } else if (world > 1 && a2a) {
                // slice o of every rank's candidates to its owner o (all-to-all), the owner ORs the `world` slices, the owned
                // slices are allgathered in place like a bottom-up level: a rank receives 2 (world - 1) / world bitmaps
                // instead of world - 1 whole ones
                GMC_TRY(gmc_comm_alltoall(comm, cand_l, w->allcand, sizeof(uint32_t) * (size_t)w->wpr, st));
                or_slices<<<(unsigned)min(div_up(w->wpr, BFS_THREADS), sm_count() * 8), BFS_THREADS, 0, st>>>(w->allcand, world, rank, w->wpr,
                                                                                                           cand_l + (size_t)rank * w->wpr);
                launches++;
                GMC_TRY(gmc_comm_allgather(comm, cand_l + (size_t)rank * w->wpr, cand_l, sizeof(uint32_t) * (size_t)w->wpr, st));
            } else if (world > 1) {
                GMC_TRY(gmc_comm_allgather(comm, cand_l, w->allcand, sizeof(uint32_t) * (size_t)w->wp, st));
                use_all = 1;
            }
        }
        if (!published) {
            // (a level every rank ran in full queues the whole new frontier: the levels after it are usually small as well; the level
            // after an RRD level is bottom-up whatever it finds -- RRD or RR2Q -- so it needs neither a queue nor the edge count)
            const bool no_queue = state == ST_RRD;
            const int32_t q_lo = no_queue ? 0 : (redundant ? 0 : own_lo), q_hi = no_queue ? 0 : (redundant ? w->words : own_hi);
            bfs_finish<L><<<fin_grid, BFS_THREADS, 0, st>>>(w->words, w->wp, world, use_all, cand_l, w->allcand, w->visited, fnext,
                                                            nl, level, q_lo, q_hi, w->queue[qi ^ 1], w->counters,
                                                            w->h_counters, ++w->seq, nullptr, nullptr, seen_mode ? 1 : 0, no_queue ? nullptr : deg_begin,
                                                            no_queue ? nullptr : mf_dev, n);
            launches++;
        }
        { uint32_t* t = fcur; fcur = fnext; fnext = t; }
        if (world > 1) queue_full = redundant;     // bfs_finish queued the owned part of the new frontier (all of it after a level run in full)
        GMC_CUDA(cudaGetLastError());
        GMC_TRY(wait_host_seq(w, st));    // the level's counts: the sequence number bfs_finish's last block writes into pinned memory
        const int32_t next_count = w->h_counters[0];
        own_count = w->h_counters[1];
        qi ^= 1;
        lvl_report(ctl_bottom_up(state) ? "bottom-up" : "top-down", state, curr_level, curr_count);

        // get_next_state (gm_bfs_template.h:352-421)
        if (next_count == 0) break;
        int next_state = bfs_next_state(state, next_count, curr_count, thr, has_reverse, rrd_direct);
        if (!published) next_state = policy(next_state, next_count, curr_count);       // (published: no bfs_finish, no edge count)
        else mf_known = false;
        curr_count = next_count;
        curr_level++;
        state = next_state;
    }
    if (!device_driven && !persistent) GMC_CUDA(cudaEventRecord(ev1, st));
    GMC_CUDA(cudaStreamSynchronize(st));
    w->valid = true;
    w->traversals++;
    w->last_is_i16 = sizeof(L) == 2;
    w->last_launches = launches;
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

// internal (gmc_bc.cu): forward traversal whose int16 levels stay in the workspace; level counts and states are the host record
int gmc_bfs_levels_resident(gmc_graph* g, int32_t root, bool save_child, const int16_t** level_dev, const int32_t** level_count,
                            const int32_t** state_trace, int32_t* n_levels, int32_t* launches, cudaStream_t st) {
    GMC_TRY(ensure_ws(g, 1));
    GMC_TRY(run_bfs<int16_t>(g, nullptr, g->begin, g->node_idx, g->has_reverse ? g->r_begin : nullptr,
                             g->has_reverse ? g->r_node_idx : nullptr, root, (int16_t)-2, (int16_t*)g->bfs->level, nullptr, st,
                             nullptr, save_child));
    gmc_bfs_ws* w = g->bfs;
    *level_dev = (const int16_t*)w->level;
    *level_count = w->level_count.data();
    *state_trace = w->state_trace.data();
    *n_levels = (int32_t)w->level_count.size();
    if (launches) *launches = w->last_launches;
    return GMC_OK;
}

extern "C" {

int gmc_hop_dist_sharded_dev(gmc_graph* g, gmc_comm* comm, int32_t root, int32_t* dist_dev, gmc_bfs_info* info, void* stream) {
    GMC_REQUIRE(g && dist_dev, GMC_ERR_ARG, "gmc_hop_dist_dev: NULL");
    GMC_REQUIRE(g->N > 0, GMC_ERR_ARG, "gmc_hop_dist_dev: empty graph");
    return run_bfs<int32_t>(g, comm, g->begin, g->node_idx, g->has_reverse ? g->r_begin : nullptr,
                            g->has_reverse ? g->r_node_idx : nullptr, root, INT_MAX, dist_dev, info, as_stream(stream), nullptr, false,
                            /*free_policy (hop_dist.gm defines distances, not a traversal order)*/ true);
}

int gmc_hop_dist_dev(gmc_graph* g, int32_t root, int32_t* dist_dev, gmc_bfs_info* info, void* stream) {
    return gmc_hop_dist_sharded_dev(g, nullptr, root, dist_dev, info, stream);
}

int gmc_hop_dist(gmc_graph* g, int32_t root, int32_t* dist_host, gmc_bfs_info* info) {
    GMC_REQUIRE(g && dist_host, GMC_ERR_ARG, "gmc_hop_dist: NULL");
    GMC_REQUIRE(g->N > 0, GMC_ERR_ARG, "gmc_hop_dist: empty graph");
    GMC_TRY(ensure_ws(g, 1));
    GMC_TRY(gmc_hop_dist_dev(g, root, (int32_t*)g->bfs->level, info, nullptr));
    return gmc_d2h(dist_host, g->bfs->level, sizeof(int32_t) * (size_t)g->N, nullptr);
}

int gmc_bfs_levels_sharded(gmc_graph* g, gmc_comm* comm, int32_t root, int use_reverse_edge, const uint8_t* edge_mask_host,
                           int16_t* level_host, int32_t* state_trace, int32_t* level_count, int32_t trace_cap, gmc_bfs_info* info) {
    GMC_REQUIRE(g && level_host, GMC_ERR_ARG, "gmc_bfs_levels: NULL");
    GMC_REQUIRE(g->N > 0, GMC_ERR_ARG, "gmc_bfs_levels: empty graph");
    const bool save_child = (use_reverse_edge & GMC_BFS_SAVE_CHILD) != 0;
    use_reverse_edge &= GMC_BFS_REVERSE_EDGE;
    GMC_REQUIRE(!use_reverse_edge || g->has_reverse, GMC_ERR_STATE, "gmc_bfs_levels: use_reverse_edge needs the reverse CSR");
    GMC_TRY(ensure_ws(g, comm ? comm->world : 1));
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
    int rc = run_bfs<int16_t>(g, comm, fb, fi, rb, ri, root, (int16_t)-2, (int16_t*)g->bfs->level, info, nullptr, mask_dev, save_child);
    if (mask_dev) cudaFree(mask_dev);
    GMC_TRY(rc);
    gmc_bfs_ws* w = g->bfs;
    for (int32_t i = 0; i < trace_cap && i < (int32_t)w->state_trace.size(); i++) {
        if (state_trace) state_trace[i] = w->state_trace[i];
        if (level_count) level_count[i] = w->level_count[i];
    }
    return gmc_d2h(level_host, w->level, sizeof(int16_t) * (size_t)g->N, nullptr);
}

int gmc_bfs_levels_masked(gmc_graph* g, int32_t root, int use_reverse_edge, const uint8_t* edge_mask_host, int16_t* level_host,
                          int32_t* state_trace, int32_t* level_count, int32_t trace_cap, gmc_bfs_info* info) {
    return gmc_bfs_levels_sharded(g, nullptr, root, use_reverse_edge, edge_mask_host, level_host, state_trace, level_count, trace_cap, info);
}

int gmc_bfs_levels(gmc_graph* g, int32_t root, int use_reverse_edge, int16_t* level_host, int32_t* state_trace,
                   int32_t* level_count, int32_t trace_cap, gmc_bfs_info* info) {
    return gmc_bfs_levels_sharded(g, nullptr, root, use_reverse_edge, nullptr, level_host, state_trace, level_count, trace_cap, info);
}

// Vertices of every level in level order (the reference's global_vector / level_queue_begin,
// gm_bfs_template.h:423-447), rebuilt from the level array: ascending vertex id inside a level (the
// reference's order inside a level is unspecified).
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
