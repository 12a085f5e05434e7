// PageRank pull sweep (K1) with the updater (K2) and init (K3) fused away -- one cooperative launch per iteration.
//
// Reference loop being replaced (generated from apps/src/pagerank.gm:8-19, shape in SURVEY.md 3.1):
//   foreach t: S = sum_{w in InNbrs(t), r_node_idx order} rank[w] / (double)outdeg(w)
//              val = (1-d)/N + d*S;  diff += |val - rank[t]|;  rank_nxt[t] = val
//   rank = rank_nxt; cnt++           while (diff > e && cnt < max)
//
// Device formulation.  contrib[w] = rank[w]/(double)outdeg(w) is produced once per vertex by the epilogue of the previous
// sweep (the same IEEE division the reference performs once per EDGE), so a sweep gathers one double per edge; rank/contrib
// are ping-ponged, which removes the copy-back loop.
//
// Layout: the plan of gmc_pr_plan.h (read its header first).  In short: the (T+1) * 16639 vertices with the largest
// out-degree own a slot in the hotpack; a CTA stages ONE 130 KB tier of it in shared memory (TMA bulk copies) and runs the
// entry stream that gathers from that tier -- stream 0 ("home": all rows; tier-0 sources from shared memory, everything
// that no tier stream serves from L2) or stream t (the (row, tier t) segments: shared-memory gathers only).  Both kinds of
// stream consist of the same two task classes:
//   long task   a fixed number of entries of row-major long rows whatever the row lengths (merge-path in the limit): a
//               warp streams it with 128-bit index loads, 8 gathers in flight per lane, lane-private accumulation and one
//               butterfly reduction per row end.  A row cut by a task boundary leaves per-task partials; the last task
//               to arrive (atomic counter) adds them in task order.
//   short task  sliced ELL, one lane per row, terms added strictly left to right in the original source order.
// A finished row either runs its epilogue (rows without segments: bit-identical to the CPU order when short) or stores its
// partial sum into the row's part slot; behind a grid barrier at the end of the launch the parts of every split row are added
// in slot order and their epilogues run, the diff partials are reduced in a fixed order and the device-side loop control
// advances (combine_rows; GMC_PR_TWO_KERNELS=1 runs it as a second kernel instead).  Nothing in the result depends on which
// CTA ran what.  Measurements, the variants tried and why the sweep stops at 0.35 of HBM peak: DESIGN.md 4.1.
#include "gmc_common.cuh"
#include "gmc_comm.h"
#include "gmc_pr_plan.h"
#include <cub/device/device_radix_sort.cuh>
#include <mutex>
#include <vector>
#include <stdlib.h>
#include <time.h>

using namespace gmc;
using prp::HT;
using prp::OD_HOT;
using prp::OD_MASK;
using prp::OD_SPLIT;
using prp::SCHED_W;
using prp::Stream;

#ifndef GMC_PR_THREADS
#define GMC_PR_THREADS 1024
#endif
constexpr int PR_THREADS = GMC_PR_THREADS;    // one CTA per SM, 32 independent warps
constexpr int CB_THREADS = 1024;              // pr_combine: one thread per split row
constexpr int ZT = prp::ZT;                   // ids per zero-in-degree task
// task sizes: shards with many entries run 4x / 2x larger tasks: the start and the end of a task (metadata chain, row
// hand-over, split-row arrival) are latency chains without gathers in flight; measured on the LJ-shaped graph
// 0.2149 -> 0.2015 ms per sweep (profiles/r1k_task_size_variants.txt), while graphs with few tasks per warp keep the small ones
constexpr int LT = 2048, ST_COST = 2048, SLICE_OVH = 68;
constexpr int LT_BIG = 8192, ST_COST_BIG = 4096, SLICE_OVH_BIG = 136;
constexpr int64_t BIG_TASK_ENTRIES = 24000000;   // per-shard entries from which the large tasks are used
// entries one row epilogue weighs when the shards of a multi-rank run are cut.  The epilogue stores the new contribution
// into every peer (world-1 NVLink stores per row), so the weight grows with the world size (profiles/r1p_shard_row_cost.txt)
static inline int shard_row_cost(int world) {
    if (const char* c = getenv("GMC_PR_ROWCOST")) return atoi(c);      // experiments
    // round 2 (tiered sweep, 8 ranks, TW-shaped; profiles/r2r_shard_row_cost_8gpu.txt): weight 12 -> 1605 GTEPS (sweep 0.68 .. 0.82 ms over
    // the ranks), 20 -> 1715 (0.63 .. 0.76), 30 -> 1639 (0.59 .. 0.80)
    return world > 2 ? (5 * world) / 2 : 0;
}
constexpr unsigned FULL = 0xffffffffu;
constexpr int MAX_STREAMS = prp::MAX_TIERS + 2;      // home + tiers 0..31

struct gmc_pr_ctrl {   // device-side loop control: lets the host enqueue sweeps without a sync per iteration
    double diff;
    int32_t cnt;
    int32_t done;
    unsigned int arrived;   // blocks that have finished their share of the split rows
    unsigned int bar;       // grid barrier of pr_sweep: arrivals of all launches since the control block was cleared
    int32_t error;          // 1: a peer's end-of-sweep flag did not arrive in time (sharded runs)
    unsigned int next[MAX_STREAMS];     // dynamic task counter of every stream of the running sweep (reset by pr_combine)
};
constexpr int MAX_PEERS = 7;            // peer-memory exchange up to 8 ranks (one NVSwitch domain); beyond: NCCL
constexpr int SYNC_RANKS = 64;          // slots of the sync area at the end of the state block
// doubles per contribution buffer inside the state block: N+2 used, rounded up to 256 bytes (TMA bulk source alignment)
static inline size_t state_stride(int32_t n) { return (((size_t)n + 4 + 31) / 32) * 32; }
constexpr long long PEER_TIMEOUT_CYCLES = 8000000000LL;    // ~4 s at 1.9 GHz: a dead peer fails the call, not the box

struct SweepArgs {
    // layout (gmc_pr_plan.h)
    const int32_t* __restrict__ E;
    const int32_t* __restrict__ eoff;
    const int32_t* __restrict__ sbeg;
    const int32_t* __restrict__ ltask_row;
    const int32_t* __restrict__ stask_slice;
    const int32_t* __restrict__ od_flags;    // [N+1] by id: out-degree | OD_SPLIT | OD_HOT
    const int32_t* __restrict__ hotslot;     // [N+1] by id (read for OD_HOT vertices only)
    const int32_t* __restrict__ home_slot;   // [N+1] by id (read for OD_SPLIT rows only)
    const int32_t* __restrict__ seg_slot;    // [n_seg] part slot of every segment, in stream order
    const int32_t* __restrict__ pbeg;        // [K+1] part slots of split row k
    const int32_t* __restrict__ prow;        // [K] its id
    const Stream* __restrict__ streams;      // [T+1]
    const int32_t* __restrict__ sched;       // [grid][SCHED_W]
    int32_t N, T, K, lt, lt_shift;
    int32_t long_id0, short_id0;              // id of the shard's first long row / of lane 0 of its first slice
    int32_t n_ztasks, z0, z_end, zt_on;       // [z0, z_end): the shard's ids without in-edges; zt_on: run their tasks this sweep
    int32_t dp_short, dp_zero, dp_cut, dp_total, n_cb;   // diff_parts layout; n_cb = pr_combine blocks (their partials follow dp_total)
    int32_t hot0_bytes;                       // bytes of tier 0 worth staging (small graphs do not fill it)
    int32_t split;                            // SPLIT mode (home tasks run with any tier staged)
    int32_t cold_s;                           // stream number of the cold stream (0: none)
    int32_t fused;                            // 1: the split rows are finished at the end of pr_sweep (grid barrier), no pr_combine launch
    unsigned int bar_target;                  // value of ctrl->bar when every CTA of this launch has arrived
    int32_t cold_warps;                       // warps of a CTA that prefer home tasks (L2 gathers) to their tier's tasks
    // cut rows of the long sections, parts of the split rows
    double* head_part;                       // [n_ltslots]
    double* tail_part;                       // [n_ltslots]
    unsigned int* split_cnt;                 // [n_ltslots]
    double* diff_parts;                      // [dp_total + n_cb]
    double* parts;                           // [n_parts]
    // state
    const double* __restrict__ contrib_in;   // [N+2]
    double* contrib_out;
    const double* __restrict__ hot_in;       // [(T+1) * HT] the hotpack
    double* hot_out;
    const double* __restrict__ rank_in;      // [N+1]
    double* rank_out;
    double d, base;
    // loop control.  loop == 0: a free-standing sweep (timing aid), cnt/done untouched.  diff_slot != null: sharded
    // run, the shard's diff goes there and the loop control is advanced by k_ctrl_combine once every shard's diff is known
    gmc_pr_ctrl* ctrl;
    double* diff_slot;
    int32_t loop;
    // peer-memory exchange (sharded, ranks mapped with CUDA IPC): every new contribution is also stored into the
    // peers' contrib_out / hot_out over NVLink from the row epilogue; pr_combine publishes diff + the sweep's sequence number
    int32_t npeers;
    unsigned int seq;
    unsigned int* my_flag;
    double* peer_out[MAX_PEERS];
    double* peer_hot[MAX_PEERS];
    double* peer_diff[MAX_PEERS];
    unsigned int* peer_flag[MAX_PEERS];
    double e;
    int32_t max_iter;
};

struct gmc_pr_plan {
    prp::Plan L;                              // the layout (arrays owned by this plan)
    int world = 1, shard = 0;
    bool owns_state = true;
    int32_t N = 0, M = 0;
    // state block (one allocation, so that one CUDA IPC handle maps it into the peers):
    //   contrib[0] [stride] | contrib[1] [stride] | hot[0] [NH] | hot[1] [NH] | diffs [2][SYNC_RANKS] doubles | flags [SYNC_RANKS] u32
    double* block = nullptr;
    double* diffs = nullptr;                  // per-shard diffs, double-buffered by sweep parity
    unsigned int* flags = nullptr;            // flags[r] = sequence number of rank r's last completed sweep
    bool p2p = false;                         // peers mapped: contributions are pushed from the sweep epilogue
    bool peer_local = false;                  // ... with plain pointers (all ranks in this process), nothing to unmap
    void* peer_base[MAX_PEERS + 1] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
    unsigned int seq = 0;                     // sweeps issued so far (identical on every rank)
    size_t block_bytes = 0;
    bool block_ipc = false;                   // block came from gmc_malloc_ipc (exportable)
    bool block_cached = false;                // block (and its peer mappings) taken over from the retired-block cache
    unsigned long long gen = 0;               // allocation number of the block inside this process
    unsigned long long peer_gen[MAX_PEERS + 1] = {0, 0, 0, 0, 0, 0, 0, 0};   // the peers' allocation numbers when they were mapped
    double *head_part = nullptr, *tail_part = nullptr, *diff_parts = nullptr, *parts = nullptr;
    unsigned int* split_cnt = nullptr;
    double* contrib[2] = {nullptr, nullptr};
    double* hot[2] = {nullptr, nullptr};
    double* rank[2] = {nullptr, nullptr};
    gmc_pr_ctrl* ctrl = nullptr;
    int grid = 0, n_cb = 1;
    bool lite = false;                        // built without tier streams for a short first call (see build_plan)
    unsigned int bar_arrivals = 0;            // state owner: CTAs that will have arrived at ctrl->bar after the launches enqueued so far
    size_t smem = 0;
};

namespace {

// ---------------------------------------------------------------------------------------------
// small PTX helpers: mbarrier + TMA bulk copy (global -> shared), streaming loads/stores
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(unsigned long long* bar, unsigned count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(unsigned long long* bar, unsigned bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(unsigned long long* bar, unsigned parity) {
    asm volatile(
        "{\n.reg .pred p;\nWAIT_%=:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra DONE_%=;\nbra WAIT_%=;\nDONE_%=:\n}" ::"r"(smem_u32(bar)), "r"(parity) : "memory");
}
__device__ __forceinline__ void tma_bulk_g2s(void* dst_smem, const void* src_gmem, unsigned bytes, unsigned long long* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(dst_smem)),
                 "l"(src_gmem), "r"(bytes), "r"(smem_u32(bar))
                 : "memory");
}
__device__ __forceinline__ int4 ld_idx4(const int4* p) { return __ldcs(p); }       // read-once index stream
__device__ __forceinline__ double butterfly(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(FULL, v, o);
    return v;
}

// ---------------------------------------------------------------------------------------------
// the sweep
// ---------------------------------------------------------------------------------------------
// home entries: value < HT = slot of tier 0 (shared memory), otherwise HT + id (L2; cold gathers bypass L1, whose capacity the
// hot block has taken); tier entries: slot inside the staged tier
template <bool SMEM>
__device__ __forceinline__ double gather(const double* hot, const double* cin_m, int32_t w) {
    if (SMEM) return hot[w];
    // (tried in round 2: two predicated loads into one register through inline PTX instead of this branch -- 0.159 -> 0.230 ms
    // per sweep, profiles/r2x_predicated_gather.txt: warps whose lanes are all hot or all cold skip the other load here)
    return (w < HT) ? hot[w] : __ldcg(cin_m + w);
}

// new contribution of vertex id: local copy + one copy per peer (fire-and-forget NVLink stores, coalesced across lanes)
__device__ __forceinline__ void put_contrib(const SweepArgs& A, int32_t id, double c) {
    A.contrib_out[id] = c;
    if (A.npeers == 0) return;                  // (uniform: one branch instead of seven predicates on a single GPU)
#pragma unroll
    for (int p = 0; p < MAX_PEERS; p++)
        if (p < A.npeers) A.peer_out[p][id] = c;
}
__device__ __forceinline__ void put_hot(const SweepArgs& A, int32_t id, double c) {
    const int32_t slot = __ldg(A.hotslot + id);
    A.hot_out[slot] = c;
    if (A.npeers == 0) return;
#pragma unroll
    for (int p = 0; p < MAX_PEERS; p++)
        if (p < A.npeers) A.peer_hot[p][slot] = c;
}
__device__ __forceinline__ void st_release_sys(unsigned int* p, unsigned int v) {
    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ unsigned int ld_acquire_sys(const unsigned int* p) {
    unsigned int v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}

// per-row epilogue; mul and add kept separate (no FMA contraction) to match the CPU build's rounding
__device__ __forceinline__ double finish_row(const SweepArgs& A, int32_t id, uint32_t of, double old, double S) {
    const double val = __dadd_rn(A.base, __dmul_rn(A.d, S));
    __stcs(A.rank_out + id, val);
    const int32_t od = (int32_t)(of & OD_MASK);
    const double c = od > 0 ? val / (double)od : 0.0;
    put_contrib(A, id, c);
    if (of & OD_HOT) put_hot(A, id, c);
    return fabs(__dsub_rn(val, old));
}
// a home row is finished: rows with segments hand their partial sum to pr_combine, the others run the epilogue
__device__ __forceinline__ double finish_home(const SweepArgs& A, int32_t id, double S) {
    const uint32_t of = (uint32_t)__ldcs(A.od_flags + id);
    if (of & OD_SPLIT) { __stcg(A.parts + __ldg(A.home_slot + id), S); return 0.0; }
    return finish_row(A, id, of, __ldcs(A.rank_in + id), S);
}

// first k in [0,32] with e_k > o, e_k ascending across lanes (lane k holds e_k)
__device__ __forceinline__ int row_search(int32_t e, int32_t o) {
    int k = 0;
#pragma unroll
    for (int s = 16; s > 0; s >>= 1) {
        int32_t ek = __shfl_sync(FULL, e, k + s - 1);
        if (ek <= o) k += s;
    }
    return k;
}

// the last task to deliver a piece of a cut row adds the pieces in task order and finishes the row.  t1..t2: the tasks
// (stream-local numbers) the row spans; slots of the cut-row arrays are st.lt_off + task
// MODE of a task: 0 = home stream (rows finished by their epilogue or the home part slot; gathers from the staged tier 0 and
// from L2), 1 = tier stream (segments: sums go to part slots; shared-memory gathers only), 2 = cold stream (segments of
// cold entries: part slots; L2 gathers -- runs with any tier staged, slot 0 of every tier being the padding's 0.0)
template <int MODE>
__device__ __forceinline__ void split_arrive(const SweepArgs& A, const Stream& st, int lane, int32_t t1, int32_t t2, int32_t row, double piece,
                                             double* slot) {
    int last = 0;
    if (lane == 0) {
        *slot = piece;
        __threadfence();
        unsigned old = atomicAdd(&A.split_cnt[st.lt_off + t1], 1u);
        last = (old == (unsigned)(t2 - t1));
    }
    last = __shfl_sync(FULL, last, 0);
    if (!last) return;
    __threadfence();
    double v = 0.0;
    for (int32_t p = t1 + lane; p <= t2; p += 32) v += (p == t1) ? __ldcg(A.tail_part + st.lt_off + p) : __ldcg(A.head_part + st.lt_off + p);
    const double S = butterfly(v);
    if (lane == 0) {
        constexpr bool TIER = MODE != 0;
        if (TIER) __stcg(A.parts + __ldg(A.seg_slot + st.meta_off + row), S);
        else A.diff_parts[A.dp_cut + t1] = finish_home(A, A.long_id0 + row, S);
        A.split_cnt[st.lt_off + t1] = 0u;
    }
}

template <int MODE>
__device__ __forceinline__ void long_task(const SweepArgs& A, const Stream& st, const double* hot, int lane, int32_t t) {
    constexpr bool TIER = MODE != 0, SMEM = MODE == 1;
    const double* cin_m = A.contrib_in - HT;
    const int32_t* __restrict__ lbeg = A.eoff + st.lbeg_off;
    const int32_t x0 = st.e_long + (t << A.lt_shift);
    const int32_t x_end = st.e_long + st.EL;
    const int32_t x1 = min(x0 + A.lt, x_end);
    const int32_t a = __ldg(A.ltask_row + st.lt_off + t);         // row holding entry x0
    const int32_t e = (a + lane < st.n_long) ? __ldg(lbeg + a + lane + 1) : INT32_MAX;   // end of row a+lane
    const int32_t rb0 = __ldg(lbeg + a);
    const int4* __restrict__ E4 = reinterpret_cast<const int4*>(A.E + x0);
    const int32_t span = x1 - x0;
    const int nst = (span + 127) >> 7;
    const int4 zero4 = make_int4(0, 0, 0, 0);
    int4 nx0 = (4 * lane < span) ? ld_idx4(E4 + lane) : zero4;
    int4 nx1 = (128 + 4 * lane < span) ? ld_idx4(E4 + 32 + lane) : zero4;
    const bool head_split = rb0 < x0;
    const int kl = row_search(e, x1 - 1);                          // last row of the task (relative)
    const int32_t e_kl = __shfl_sync(FULL, e, kl);
    const bool tail_split = e_kl > x1;

    double acc = 0.0, my_sum = 0.0, head_acc = 0.0, tail_acc = 0.0;
    int acc_row = -1;
    int32_t row_end = INT32_MIN;        // end of the lane's current row: forces a search at its first step

    // hand the finished rows of the lanes in `who` to the warp: one butterfly per distinct row
    auto hand_over = [&](bool who) {
        unsigned m = __ballot_sync(FULL, who);
        while (m) {
            const int R = __shfl_sync(FULL, acc_row, __ffs(m) - 1);
            const bool mem = who && acc_row == R;
            const unsigned mm = __ballot_sync(FULL, mem);
            const double tot = butterfly(mem ? acc : 0.0);
            if (R == 0 && head_split) head_acc += tot;
            else if (R == kl && tail_split) tail_acc += tot;
            else if (lane == R) my_sum += tot;
            m &= ~mm;
        }
    };
    // a lane's 4 entries lie in one row (rows are padded to 4); rows change rarely, so the row lookup
    // (5 shuffles) only runs in steps where some lane has run past the end of its row
    auto step = [&](int32_t o, bool in, double p) {
        const bool cross = in && o >= row_end;
        if (__any_sync(FULL, cross)) {
            const int row = row_search(e, o);
            hand_over(cross && acc_row >= 0);
            const int32_t re = __shfl_sync(FULL, e, row);
            if (cross) { acc_row = row; row_end = re; acc = 0.0; }
        }
        acc += p;
    };

    for (int i = 0; i < nst; i += 2) {
        const int4 c0 = nx0, c1 = nx1;
        const int32_t o0 = i * 128 + 4 * lane, o1 = o0 + 128;
        nx0 = (o0 + 256 < span) ? ld_idx4(E4 + (i + 2) * 32 + lane) : zero4;
        nx1 = (o1 + 256 < span) ? ld_idx4(E4 + (i + 3) * 32 + lane) : zero4;
        const double v0 = gather<SMEM>(hot, cin_m, c0.x), v1 = gather<SMEM>(hot, cin_m, c0.y), v2 = gather<SMEM>(hot, cin_m, c0.z), v3 = gather<SMEM>(hot, cin_m, c0.w);
        const double v4 = gather<SMEM>(hot, cin_m, c1.x), v5 = gather<SMEM>(hot, cin_m, c1.y), v6 = gather<SMEM>(hot, cin_m, c1.z), v7 = gather<SMEM>(hot, cin_m, c1.w);
        step(x0 + o0, o0 < span, (v0 + v1) + (v2 + v3));
        step(x0 + o1, o1 < span, (v4 + v5) + (v6 + v7));
    }
    hand_over(acc_row >= 0);
    // rows that lie completely inside the task: one lane per row, coalesced epilogue
    const int k_first = head_split ? 1 : 0;
    const int k_last = tail_split ? kl - 1 : kl;
    if (TIER) {
        if (lane >= k_first && lane <= k_last) __stcg(A.parts + __ldg(A.seg_slot + st.meta_off + a + lane), my_sum);
    } else {
        double diff = 0.0;
        if (lane >= k_first && lane <= k_last) diff = finish_home(A, A.long_id0 + a + lane, my_sum);
        diff = butterfly(diff);
        if (lane == 0) A.diff_parts[t] = diff;
    }
    if (head_split) {
        const int32_t e0 = __shfl_sync(FULL, e, 0);
        split_arrive<MODE>(A, st, lane, (rb0 - st.e_long) >> A.lt_shift, (e0 - 1 - st.e_long) >> A.lt_shift, a, head_acc, A.head_part + st.lt_off + t);
    }
    if (tail_split && !(kl == 0 && head_split))
        split_arrive<MODE>(A, st, lane, t, (e_kl - 1 - st.e_long) >> A.lt_shift, a + kl, tail_acc, A.tail_part + st.lt_off + t);
}

// rows / segments of up to S_MAX entries: one lane per row, terms added strictly left to right.  The loop runs over
// 8-column chunks; the indices of the next chunk (of this slice or the next one) are always in flight.
template <int MODE>
__device__ __forceinline__ void short_task(const SweepArgs& A, const Stream& st, const double* hot, int lane, int32_t u) {
    constexpr bool TIER = MODE != 0, SMEM = MODE == 1;
    const double* cin_m = A.contrib_in - HT;
    const int32_t* __restrict__ sbeg = A.sbeg + st.sbeg_off;
    const int32_t s_lo = __ldg(A.stask_slice + st.st_off + u), s_hi = __ldg(A.stask_slice + st.st_off + u + 1);
    const int ns = s_hi - s_lo;                                  // <= 31
    const int32_t sb = (lane <= ns) ? __ldg(sbeg + s_lo + lane) : 0;
    double diff = 0.0;
    if (ns <= 0) { if (!TIER && lane == 0) A.diff_parts[A.dp_short + u] = 0.0; return; }
    int q = 0, j = 0;
    int32_t base = __shfl_sync(FULL, sb, 0);
    int L = (__shfl_sync(FULL, sb, 1) - base) >> 5;
    int32_t w[8];
    {
        const int32_t* __restrict__ pe = A.E + base + lane;       // one pointer, constant offsets: 2 instructions per load
#pragma unroll
        for (int k = 0; k < 8; k++) w[k] = (k < L) ? __ldcs(pe + k * 32) : 0;
    }
    double S = 0.0, old = 0.0;
    uint32_t of = 0;          // home: od_flags of the lane's row; tier: part slot + 1 of the lane's segment (0: dead lane)
    while (q < ns) {
        if (j == 0) {
            if (TIER) {
                const int32_t loc = st.n_long + (s_lo + q) * 32 + lane;
                of = loc < st.n_rows ? (uint32_t)__ldcs(A.seg_slot + st.meta_off + loc) + 1u : 0u;
            } else {
                const int32_t id = A.short_id0 + (s_lo + q) * 32 + lane;
                if (id <= A.N) { old = __ldcs(A.rank_in + id); of = (uint32_t)__ldcs(A.od_flags + id); }
            }
        }
        // where the next chunk lives
        int nq = q, nj = j + 8;
        int32_t nbase = base;
        int nL = L;
        if (nj >= L) {
            nq = q + 1; nj = 0;
            nbase = __shfl_sync(FULL, sb, nq < ns ? nq : 0);
            nL = nq < ns ? (__shfl_sync(FULL, sb, nq + 1 <= ns ? nq + 1 : 0) - nbase) >> 5 : 0;
        }
        int32_t wn[8];
        {
            // the 8 index loads of the next chunk: one pointer and one remaining-column count, so that each load is a
            // predicate + LDG with an immediate offset (the per-load address arithmetic was 25 % of the kernel's instructions)
            const int32_t* __restrict__ pe = A.E + nbase + lane + nj * 32;
            const int rem = nL - nj;
#pragma unroll
            for (int k = 0; k < 8; k++) wn[k] = (k < rem) ? __ldcs(pe + k * 32) : 0;
        }
        double v[8];
#pragma unroll
        for (int k = 0; k < 8; k++) v[k] = gather<SMEM>(hot, cin_m, w[k]);
#pragma unroll
        for (int k = 0; k < 8; k++) S = S + v[k];                 // left to right: the reference's order (+0.0 for padding)
        if (nj == 0) {                                            // slice finished
            if (TIER) {
                if (of) __stcg(A.parts + (of - 1u), S);
            } else {
                const int32_t id = A.short_id0 + (s_lo + q) * 32 + lane;
                if (id <= A.N) {
                    if (of & OD_SPLIT) __stcg(A.parts + __ldg(A.home_slot + id), S);
                    else diff += finish_row(A, id, of, old, S);
                }
            }
            S = 0.0;
        }
#pragma unroll
        for (int k = 0; k < 8; k++) w[k] = wn[k];
        q = nq; j = nj; base = nbase; L = nL;
    }
    if (!TIER) {
        diff = butterfly(diff);
        if (lane == 0) A.diff_parts[A.dp_short + u] = diff;
    }
}

// rows without in-edges (the tail of the id space): val = (1-d)/N, a pure streaming update
__device__ __forceinline__ void zero_task(const SweepArgs& A, int lane, int32_t z) {
    const int32_t id0 = A.z0 + z * ZT;
    double diff = 0.0;
#pragma unroll 1
    for (int b = 0; b < ZT; b += 256) {
        double old[8];
        uint32_t of[8];
#pragma unroll
        for (int k = 0; k < 8; k++) {
            const int32_t id = id0 + b + k * 32 + lane;
            old[k] = id < A.z_end ? __ldcs(A.rank_in + id) : 0.0;
            of[k] = id < A.z_end ? (uint32_t)__ldcs(A.od_flags + id) : 0u;
        }
#pragma unroll
        for (int k = 0; k < 8; k++) {
            const int32_t id = id0 + b + k * 32 + lane;
            if (id < A.z_end) diff += finish_row(A, id, of[k], old[k], 0.0);       // (1-d)/N + d*0
        }
    }
    diff = butterfly(diff);
    if (lane == 0) A.diff_parts[A.dp_zero + z] = diff;
}

// The split rows (parts added in slot order + epilogue), then the last block to finish adds the diff partials of the whole
// sweep in a fixed order, advances the loop control and resets the task counters.  Runs behind a grid barrier at the end of
// pr_sweep (default) or as the kernel pr_combine (GMC_PR_TWO_KERNELS=1).
template <int NT>
__device__ __forceinline__ void combine_rows(const SweepArgs& A, double* s_red, int* s_last) {
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    double diff = 0.0;
    for (int32_t k = blockIdx.x * NT + tid; k < A.K; k += gridDim.x * NT) {
        const int32_t p0 = __ldg(A.pbeg + k), p1 = __ldg(A.pbeg + k + 1), id = __ldg(A.prow + k);
        // the row's flags and old rank are fetched beside its parts, the parts four at a time (independent loads; added in slot
        // order as before): a thread has 1-2 rows here, so this phase is a latency chain, not a bandwidth problem
        const uint32_t of = (uint32_t)__ldcs(A.od_flags + id);
        const double old = __ldcs(A.rank_in + id);
        double S = 0.0;                                                     // home part, then the segments in tier order
        for (int32_t p = p0; p < p1; p += 4) {
            double v[4];
#pragma unroll
            for (int j = 0; j < 4; j++) v[j] = p + j < p1 ? __ldcg(A.parts + p + j) : 0.0;
#pragma unroll
            for (int j = 0; j < 4; j++) if (p + j < p1) S += v[j];
        }
        diff += finish_row(A, id, of, old, S);
    }
    diff = butterfly(diff);
    if (lane == 0) s_red[warp] = diff;
    __syncthreads();
    if (warp == 0) {
        const double t = butterfly(lane < NT / 32 ? s_red[lane] : 0.0);
        if (lane == 0) A.diff_parts[A.dp_total + blockIdx.x] = t;
    }
    __syncthreads();
    if (tid == 0) {
        if (A.npeers > 0) __threadfence_system(); else __threadfence();
        *s_last = (atomicAdd(&A.ctrl->arrived, 1u) == gridDim.x - 1);
    }
    __syncthreads();
    if (!*s_last) return;
    __threadfence();
    const int32_t nparts = A.dp_total + A.n_cb;
    double v = 0.0;
    for (int32_t i = tid; i < nparts; i += NT) v += __ldcg(A.diff_parts + i);
    v = butterfly(v);
    if (lane == 0) s_red[warp] = v;
    __syncthreads();
    if (warp == 0) {
        const double t = butterfly(lane < NT / 32 ? s_red[lane] : 0.0);
        for (int i2 = lane; i2 < MAX_STREAMS; i2 += 32) A.ctrl->next[i2] = 0u;
        if (lane == 0) A.ctrl->arrived = 0u;
        if (lane == 0 && A.diff_slot) {
            *A.diff_slot = t;
            for (int p = 0; p < A.npeers; p++) *A.peer_diff[p] = t;
        } else if (lane == 0 && A.loop) {
            const int32_t cnt = A.ctrl->cnt + 1;
            A.ctrl->diff = t;
            A.ctrl->cnt = cnt;
            A.ctrl->done = !((t > A.e) && (cnt < A.max_iter));       // } While ((diff > e) && (cnt < max))
        }
        if (lane == 0 && A.my_flag) {      // every store of this sweep (both kernels, all blocks) is ordered before the flags
            __threadfence_system();
            *A.my_flag = A.seq;
            for (int p = 0; p < A.npeers; p++) st_release_sys(A.peer_flag[p], A.seq);
        }
    }
}

__global__ void __launch_bounds__(PR_THREADS, 1) pr_sweep(const SweepArgs A) {
    extern __shared__ __align__(16) double hot[];
    __shared__ __align__(8) unsigned long long bar;
    __shared__ Stream s_st, s_home, s_cold;
    __shared__ double s_red[PR_THREADS / 32];
    __shared__ int s_last;
    __shared__ int s_skip;
    const int tid = threadIdx.x, lane = tid & 31;

    if (tid == 0) mbar_init(&bar, 1);
    __syncthreads();
    // programmatic dependent launch: the CTAs of this sweep take their SMs as the blocks of the previous combine leave them
    // and wait here (everything above touches no global memory) until it has completed and its stores are visible
    asm volatile("griddepcontrol.wait;" ::: "memory");
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    if (A.loop && A.ctrl->done) return;

    // Every CTA stages the tiers of its list one after the other.  While a tier is staged its warps draw tasks from two
    // device-wide counters: the tier's stream (shared-memory gathers) and the home stream (L2 gathers; runs with any tier
    // staged in SPLIT mode, with tier 0 -- the only list entry -- in MIXED mode).  The first cold_warps warps prefer home
    // tasks, the others the tier's, and each falls back to the other kind: both gather paths of the SM stay busy.
    if (tid == 0) { s_home = A.streams[0]; s_cold = A.streams[A.cold_s]; }      // cold_s == 0: no cold stream (its task count is forced to 0 below)
    unsigned phase = 0;
    const bool cold_first = (tid >> 5) < A.cold_warps;
    bool home_left = true, cold_left = A.cold_s != 0;
    for (int i = 0; i < SCHED_W; i++) {
        const int s = __ldg(A.sched + blockIdx.x * SCHED_W + i);
        if (s == 0) break;
        if (i > 0) __syncthreads();                  // every warp has left the previous tier: hot[] and s_st may change
        if (tid == 0) {
            s_st = A.streams[s];
            // a later list entry (a place to help) is staged only while it has tasks left
            // (MIXED mode: tier 0 is staged for the home tasks, its own stream is empty)
            const bool tier_done = *(volatile unsigned int*)&A.ctrl->next[s] >= (unsigned)(s_st.n_ltasks + s_st.n_stasks);
            const bool home_done = *(volatile unsigned int*)&A.ctrl->next[0] >= (unsigned)(s_home.n_ltasks + s_home.n_stasks + (A.zt_on ? A.n_ztasks : 0));
            const bool cold_done = !A.cold_s || *(volatile unsigned int*)&A.ctrl->next[A.cold_s] >= (unsigned)(s_cold.n_ltasks + s_cold.n_stasks);
            s_skip = i > 0 && tier_done && (home_done || !(A.split || s == 1)) && cold_done;
        }
        __syncthreads();
        const Stream& st = s_st;
        const Stream& hm = s_home;
        const int32_t h_l = hm.n_ltasks, h_s = hm.n_stasks, h_total = h_l + h_s + (A.zt_on ? A.n_ztasks : 0);
        const int32_t t_l = st.n_ltasks, t_total = t_l + st.n_stasks;
        const int32_t c_l = A.cold_s ? s_cold.n_ltasks : 0, c_total = A.cold_s ? c_l + s_cold.n_stasks : 0;
        if (s_skip) continue;
        {   // stage the tier: hot_in[hot_off, hot_off + HT) -> shared memory, one TMA bulk copy stream
            if (tid == 0) {
                const unsigned bytes = s == 1 ? (unsigned)A.hot0_bytes : (unsigned)(HT * 8);
                mbar_expect_tx(&bar, bytes);
                const char* src = reinterpret_cast<const char*>(A.hot_in + st.hot_off);
                char* dst = reinterpret_cast<char*>(hot);
                // CTAs staging the same tier start at different 16 KB pieces so they do not queue on the same L2 lines
                const unsigned pieces = (bytes + 16383u) / 16384u;
                for (unsigned k = 0; k < pieces; k++) {
                    const unsigned off = ((k + blockIdx.x) % pieces) * 16384u;
                    tma_bulk_g2s(dst + off, src + off, min(16384u, bytes - off), &bar);
                }
            }
            mbar_wait(&bar, phase);
            phase ^= 1u;
        }
        bool tier_left = t_total > 0;
        home_left = home_left && h_total > 0;
        cold_left = cold_left && c_total > 0;
        const bool home_ok = A.split || s == 1;        // MIXED mode: the home rows gather from tier 0
        // task sources in this warp's order of preference.  Warps of a tier CTA: the first cold_warps start with the
        // cold stream (L2 gathers), the others with the tier's tasks (shared memory) -- both gather paths of the SM stay
        // busy; warps of a home CTA: home tasks first, the cold stream when home is exhausted.  0 = home, 1 = tier, 2 = cold
        int order[3];
        if (A.split) { order[0] = cold_first ? 0 : 1; order[1] = cold_first ? 1 : 0; order[2] = 2; }
        else if (s == 1) { order[0] = 0; order[1] = 2; order[2] = 1; }
        else { order[0] = cold_first ? 2 : 1; order[1] = cold_first ? 1 : 2; order[2] = 0; }
        for (;;) {
            bool did = false;
#pragma unroll 1
            for (int pass = 0; pass < 3 && !did; pass++) {
                const int src = order[pass];
                if (src == 0 ? !(home_left && home_ok) : (src == 1 ? !tier_left : !cold_left)) continue;
                int32_t u = 0;
                if (lane == 0) u = (int32_t)atomicAdd(&A.ctrl->next[src == 0 ? 0 : (src == 1 ? s : A.cold_s)], 1u);
                u = __shfl_sync(FULL, u, 0);
                if (src == 0) {
                    if (u >= h_total) { home_left = false; continue; }
                    if (u < h_l) long_task<0>(A, hm, hot, lane, u);
                    else if (u < h_l + h_s) short_task<0>(A, hm, hot, lane, u - h_l);
                    else zero_task(A, lane, u - h_l - h_s);
                } else if (src == 1) {
                    if (u >= t_total) { tier_left = false; continue; }
                    if (u < t_l) long_task<1>(A, st, hot, lane, u);
                    else short_task<1>(A, st, hot, lane, u - t_l);
                } else {
                    if (u >= c_total) { cold_left = false; continue; }
                    if (u < c_l) long_task<2>(A, s_cold, hot, lane, u);
                    else short_task<2>(A, s_cold, hot, lane, u - c_l);
                }
                did = true;
            }
            if (!did) break;
        }
    }
    if (!A.fused) return;
    // ---- every task of the sweep is done once all CTAs have passed this grid barrier (cooperative launch: they are all
    // resident); then the split rows.  bar counts arrivals of all launches of the call: this launch ends at bar_target.
    // (sharded: the epilogues' stores into the peers must be performed system-wide before this CTA counts as arrived -- the
    // last CTA publishes the sweep's sequence number to the peers on the strength of these arrivals)
    if (A.npeers > 0) __threadfence_system(); else __threadfence();
    __syncthreads();
    if (tid == 0) {
        atomicAdd(&A.ctrl->bar, 1u);
        const long long t0 = clock64();
        while ((int)(*(volatile unsigned int*)&A.ctrl->bar - A.bar_target) < 0) {
            if (clock64() - t0 > PEER_TIMEOUT_CYCLES) { A.ctrl->error = 2; break; }       // never hang the device
            __nanosleep(64);
        }
        __threadfence();
    }
    __syncthreads();
    combine_rows<PR_THREADS>(A, s_red, &s_last);
}

__global__ void __launch_bounds__(CB_THREADS, 2) pr_combine(const SweepArgs A) {
    __shared__ double s_red[CB_THREADS / 32];
    __shared__ int s_last;
    asm volatile("griddepcontrol.wait;" ::: "memory");
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    if (A.loop && A.ctrl->done) return;
    combine_rows<CB_THREADS>(A, s_red, &s_last);
}

// ---------------------------------------------------------------------------------------------
// state kernels
// ---------------------------------------------------------------------------------------------
__global__ void k_init(SweepArgs A, double* __restrict__ contrib0, double* __restrict__ contrib1, double* __restrict__ hot0, double* __restrict__ rank0) {
    int32_t id = blockIdx.x * blockDim.x + threadIdx.x;
    if (id > A.N + 1) return;
    if (id == 0 || id == A.N + 1) { contrib0[id] = 0.0; contrib1[id] = 0.0; if (id == 0) rank0[0] = 0.0; return; }
    const double v = 1 / (double)A.N;                       // G.pg_rank = 1 / N  (pagerank.gm:6-7)
    rank0[id] = v;
    const uint32_t of = (uint32_t)A.od_flags[id];
    const int32_t od = (int32_t)(of & OD_MASK);
    const double c = od > 0 ? v / (double)od : 0.0;
    contrib0[id] = c;
    if (of & OD_HOT) hot0[A.hotslot[id]] = c;
}
// EXPERIMENT (GMC_PR_FAKE_BANKS, results invalid): rewrites the bank residue of the shared-memory entries so that every LDS.64 of
// a warp is conflict-free -- the upper bound of what a bank-aware entry order inside segments could buy.  ell = 0: row-major
// section (lane = (p / 4) % 32), ell = 1: sliced section (lane = p % 32, column = p / 32)
__global__ void k_fake_banks(int32_t* __restrict__ E, int32_t e0, int32_t e1, int ell, int only_below) {
    const int32_t p = e0 + blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= e1) return;
    const int32_t x = E[p];
    if (x <= 0 || x >= only_below) return;
    const int32_t q = p - e0;
    const int res = ell ? ((q & 31) + (q >> 5)) & 15 : (q >> 2) & 15;
    int32_t y = (x & ~15) | res;
    if (y <= 0) y = 16 | res;
    if (y >= HT) y -= 16;
    E[p] = y;
}
__global__ void k_fake_allhot(int32_t* __restrict__ E, int32_t e0, int32_t e1) {
    const int32_t p = e0 + blockIdx.x * blockDim.x + threadIdx.x;
    if (p < e1 && E[p] >= HT) E[p] = 1 + (int32_t)(((uint32_t)E[p] * 2654435761u) % 16000u);
}
// hotpack from the contribution vector (NCCL exchange path: the peers' contributions arrive by allgather, not by remote stores)
__global__ void k_hot_refresh(const int32_t* __restrict__ idslot, int32_t nh, const double* __restrict__ contrib, double* __restrict__ hot) {
    int32_t s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s < nh) hot[s] = contrib[idslot[s]];               // dummy slots hold id 0, whose contribution is 0.0
}
__global__ void k_unpermute(const double* __restrict__ rank_new, const int32_t* __restrict__ new_of_old, int32_t n, double* __restrict__ out) {
    int32_t v = blockIdx.x * blockDim.x + threadIdx.x;
    if (v < n) out[v] = rank_new[new_of_old[v]];
}
// sharded loop control: diff = sum of the shards' diffs in rank order; cnt++; done = !(diff > e && cnt < max)
// wait until every rank has published sequence number `seq` (its sweep, with all its remote stores, is complete)
__device__ bool wait_peers(const unsigned int* flags, int world, unsigned int seq) {
    for (int r = 0; r < world; r++) {
        const long long t0 = clock64();
        while ((int)(ld_acquire_sys(flags + r) - seq) < 0) {
            if (clock64() - t0 > PEER_TIMEOUT_CYCLES) return false;
            __nanosleep(200);
        }
    }
    return true;
}
__global__ void k_wait_peers(gmc_pr_ctrl* ctrl, const unsigned int* flags, int world, unsigned int seq) {
    if (threadIdx.x != 0 || blockIdx.x != 0 || ctrl->error) return;
    if (!wait_peers(flags, world, seq)) ctrl->error = 1;
}
__global__ void k_ctrl_combine(gmc_pr_ctrl* ctrl, const double* diffs, int world, double e, int32_t max_iter,
                               const unsigned int* flags, unsigned int seq) {
    if (threadIdx.x != 0 || blockIdx.x != 0 || ctrl->done) return;
    if (flags && !wait_peers(flags, world, seq)) { ctrl->error = 1; ctrl->done = 1; return; }
    double t = 0.0;
    for (int r = 0; r < world; r++) t += __ldcg(diffs + r);
    const int32_t cnt = ctrl->cnt + 1;
    ctrl->diff = t;
    ctrl->cnt = cnt;
    ctrl->done = !((t > e) && (cnt < max_iter));
}

// ---------------------------------------------------------------------------------------------
// the plan builder's backend: its steps as CUDA kernels on stream `st`
// ---------------------------------------------------------------------------------------------
template <class F>
__global__ void k_pfor(int64_t n, const F f) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) f(i);
}
struct CudaBackend {
    cudaStream_t st;
    gmc_graph* g;
    std::string err;
    int fail(const char* what) { err = std::string("pagerank plan: ") + what + ": " + cudaGetErrorString(cudaGetLastError()); return GMC_ERR_CUDA; }
    int alloc(void** p, size_t bytes) { int rc = gmc_malloc(p, bytes ? bytes : 8); if (rc != GMC_OK) err = gmc_last_error(); return rc; }
    void free(void* p) { if (p) cudaFree(p); }
    int zero(void* p, size_t b) { return cudaMemsetAsync(p, 0, b, st) == cudaSuccess ? 0 : fail("memset"); }
    int d2h(void* d, const void* s, size_t b) {
        if (cudaMemcpyAsync(d, s, b, cudaMemcpyDeviceToHost, st) != cudaSuccess || cudaStreamSynchronize(st) != cudaSuccess) return fail("copy to host");
        return 0;
    }
    int h2d(void* d, const void* s, size_t b) {       // pageable source: the copy has been staged when the call returns
        return cudaMemcpyAsync(d, s, b, cudaMemcpyHostToDevice, st) == cudaSuccess ? 0 : fail("copy to device");
    }
    int sync() { return cudaStreamSynchronize(st) == cudaSuccess ? 0 : fail("synchronize"); }
    // GMC_PR_PLAN_TIMING=1: wall-clock time of every step of the builder (with a stream sync: the steps no longer overlap)
    double t_last = -1.0;
    void mark(const char* what) {
        static const bool on = getenv("GMC_PR_PLAN_TIMING") != nullptr;
        if (!on) return;
        cudaStreamSynchronize(st);
        struct timespec ts;
        clock_gettime(CLOCK_MONOTONIC, &ts);
        const double now = ts.tv_sec * 1e3 + ts.tv_nsec * 1e-6;
        if (t_last >= 0) fprintf(stderr, "pagerank plan: %-40s %8.3f ms\n", what, now - t_last);
        t_last = now;
    }
    // copy-in still in flight (gmc_pagerank_host_csr): the offsets arrive first, the edges in chunks
    int begin_offsets_ready() {
        if (g->pending && cudaStreamWaitEvent(st, g->pending->small, 0) != cudaSuccess) return fail("wait for offsets");
        return 0;
    }
    int edges_ready() {
        if (g->pending)
            for (int c = 0; c < g->pending->nchunks; c++)
                if (cudaStreamWaitEvent(st, g->pending->chunk[c], 0) != cudaSuccess) return fail("wait for edges");
        return 0;
    }
    int edge_chunks() { return g->pending ? g->pending->nchunks : 1; }
    int edge_chunk_ready(int c, int32_t* e_end) {
        if (!g->pending) { *e_end = g->M; return 0; }
        if (cudaStreamWaitEvent(st, g->pending->chunk[c], 0) != cudaSuccess) return fail("wait for edges");
        *e_end = g->pending->end[c];
        return 0;
    }
    std::string last_error() { return err.empty() ? std::string(gmc_last_error()) : err; }
    template <class F>
    void pfor(int64_t n, const F& f) {
        if (n <= 0) return;
        k_pfor<F><<<(unsigned)((n + 255) / 256), 256, 0, st>>>(n, f);
    }
    int scan_i32(int32_t* a, size_t n) { int rc = exclusive_scan_i32(a, a, n, nullptr, st); if (rc != GMC_OK) err = gmc_last_error(); return rc; }
    template <typename K>
    int sort_pairs(const K* kin, K* kout, const int32_t* vin, int32_t* vout, int32_t n, int end_bit) {
        void* tmp = nullptr;
        size_t bytes = 0;
        cub::DeviceRadixSort::SortPairs(nullptr, bytes, kin, kout, vin, vout, (int64_t)n, 0, end_bit, st);
        int rc = alloc(&tmp, bytes);
        if (rc != GMC_OK) return rc;
        cudaError_t e = cub::DeviceRadixSort::SortPairs(tmp, bytes, kin, kout, vin, vout, (int64_t)n, 0, end_bit, st);
        cudaStreamSynchronize(st);
        cudaFree(tmp);
        if (e != cudaSuccess) { err = std::string("pagerank plan: sort: ") + cudaGetErrorString(e); return GMC_ERR_CUDA; }
        return 0;
    }
    int sort_u32(const uint32_t* kin, uint32_t* kout, const int32_t* vin, int32_t* vout, int32_t n, int end_bit) { return sort_pairs(kin, kout, vin, vout, n, end_bit); }
    int sort_u64(const unsigned long long* kin, unsigned long long* kout, const int32_t* vin, int32_t* vout, int32_t n, int end_bit) { return sort_pairs(kin, kout, vin, vout, n, end_bit); }
};

// One retired state block per process, kept allocated AND mapped by the peers.  Exporting a block with CUDA IPC costs
// a plain cudaMalloc plus one cudaIpcOpenMemHandle per peer (~20 ms each for a 666 MB block: 147 ms of plan time and 44 ms
// of teardown per call at 8 ranks, profiles/r1x_bench_8gpu_tw.json), so a caller that runs pagerank() on graph after graph of
// the same size -- every step of the e2e loop -- pays it once.  Reuse is validated collectively (allocation numbers of
// all ranks must be the ones recorded when the mappings were made), see plan_map_peers.
struct SymCache {
    double* block = nullptr;
    size_t bytes = 0;
    int world = 0, rank = 0;
    bool p2p = false, local = false;
    void* peer_base[MAX_PEERS + 1] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
    unsigned long long gen = 0, peer_gen[MAX_PEERS + 1] = {0, 0, 0, 0, 0, 0, 0, 0};
};
// per thread: the ranks of a device group (gmc_group.cu) are threads of one process, each with its own device and block
thread_local SymCache g_sym;
thread_local unsigned long long g_block_gen = 0;

void sym_cache_release() {
    if (!g_sym.block) return;
    if (g_sym.p2p) gmc_comm_ipc_close_all(g_sym.world, g_sym.rank, g_sym.peer_base, g_sym.local);
    cudaFree(g_sym.block);
    g_sym = SymCache();
}

void free_plan(gmc_pr_plan* p) {
    if (!p) return;
    if (p->owns_state && p->block_ipc && p->block && !getenv("GMC_NO_BLOCK_CACHE")) {      // retire the block instead of unmapping it
        sym_cache_release();
        g_sym.block = p->block; g_sym.bytes = p->block_bytes; g_sym.world = p->world; g_sym.rank = p->shard; g_sym.p2p = p->p2p;
        g_sym.gen = p->gen; g_sym.local = p->peer_local;
        for (int r = 0; r <= MAX_PEERS; r++) { g_sym.peer_base[r] = p->peer_base[r]; g_sym.peer_gen[r] = p->peer_gen[r]; }
        p->block = nullptr; p->p2p = false;
    }
    prp::Plan& L = p->L;
    cudaFree(L.new_of_old); cudaFree(L.old_of_new); cudaFree(L.od_flags); cudaFree(L.hotslot); cudaFree(L.home_slot); cudaFree(L.idslot);
    cudaFree(L.E); cudaFree(L.eoff); cudaFree(L.sbeg); cudaFree(L.ltask_row); cudaFree(L.stask_slice); cudaFree(L.seg_slot);
    cudaFree(L.pbeg); cudaFree(L.prow); cudaFree(L.sched_dev); cudaFree(L.streams_dev);
    cudaFree(p->head_part); cudaFree(p->tail_part); cudaFree(p->diff_parts); cudaFree(p->split_cnt); cudaFree(p->parts);
    if (p->owns_state) {
        if (p->p2p) gmc_comm_ipc_close_all(p->world, p->shard, p->peer_base, p->peer_local);
        cudaFree(p->block); cudaFree(p->rank[0]); cudaFree(p->rank[1]); cudaFree(p->ctrl);
    }
    delete p;
}

static int env_int(const char* name, int dflt) { const char* v = getenv(name); return v && *v ? atoi(v) : dflt; }

#define PLAN_TRY(expr) do { rc = (expr); if (rc != GMC_OK) goto fail; } while (0)
#define PLAN_CUDA(expr) do { if ((expr) != cudaSuccess) { rc = GMC_ERR_CUDA; set_error("pagerank plan: %s", cudaGetErrorString(cudaGetLastError())); goto fail; } } while (0)

// Plan of shard `rank` of `world` (world == 1: the whole graph).  `state_of`: borrow the state buffers of
// another plan of the same graph (emulated shards on one device) instead of allocating them.
int build_plan(gmc_graph* g, int world, int rank, const gmc_pr_plan* state_of, gmc_pr_plan** out, cudaStream_t st) {
    const int32_t n = g->N, m = g->M;
    gmc_pr_plan* p = new gmc_pr_plan();
    p->N = n; p->M = m; p->world = world; p->shard = rank;
    p->owns_state = state_of == nullptr;
    int rc = GMC_OK;
    const size_t nn = (size_t)n;
    prp::Plan& L = p->L;
    {
        prp::Config cfg;
        // experiment / test switches; the defaults are the product
        cfg.tiers = env_int("GMC_PR_TIERS", cfg.tiers);
        cfg.kmin = env_int("GMC_PR_KMIN", cfg.kmin);
        cfg.split = env_int("GMC_PR_SPLIT", cfg.split);
        cfg.cold_stream = env_int("GMC_PR_COLD_STREAM", cfg.cold_stream);
        cfg.hot_first = env_int("GMC_PR_HOT_FIRST", cfg.hot_first);
        cfg.kcold = env_int("GMC_PR_KCOLD", cfg.kcold);
        if (cfg.kmin < 1) cfg.kmin = 1;
        // The FIRST pagerank() call on a graph may be its only one (the reference's drivers: load, one call, exit; every
        // gmc_pagerank_host_csr call).  The tier streams cost a second pass over the edges after the last byte of the graph has
        // landed (+2.1 ms on the LJ-shaped graph) and save ~0.025 ms per sweep, so they pay from ~100 sweeps on: a first call
        // shorter than that gets the plan without tiers ("lite": its entries are written behind the copy-in), and the second
        // call on the same graph -- now known to be reused -- rebuilds the tiered plan once (graph_plan).
        if (!getenv("GMC_PR_TIERS") && !getenv("GMC_PR_NO_LITE") && g->pr_iters_hint > 0 && g->pr_iters_hint < 100 && cfg.tiers > 0) {
            cfg.tiers = 0;
            p->lite = true;
        }
        if (const char* v = getenv("GMC_PR_TIER_MIN_EDGES")) cfg.tier_min_edges = atoll(v);
        if (const char* v = getenv("GMC_PR_TIER_SCALE")) cfg.tier_scale = atof(v);
        if (const char* v = getenv("GMC_PR_RATES")) sscanf(v, "%lf,%lf", &cfg.r_smem, &cfg.r_miss);
        bool big = (int64_t)m / world >= BIG_TASK_ENTRIES;
        if (const char* b = getenv("GMC_PR_BIGTASKS")) big = atoi(b) != 0;           // tests / experiments: force either task size
        cfg.lt = big ? LT_BIG : LT; cfg.st_cost = big ? ST_COST_BIG : ST_COST; cfg.slice_ovh = big ? SLICE_OVH_BIG : SLICE_OVH;
        if (const char* t = getenv("GMC_PR_TAIL")) sscanf(t, "%d,%d", &cfg.tail_big16, &cfg.tail_shift);      // "<sixteenths in big tasks>,<shift>"
        if (getenv("GMC_PR_NO_SMALL_TAIL")) cfg.tail_shift = 0;
        cfg.grid = sm_count();
        cfg.row_cost = shard_row_cost(world);
        CudaBackend be{st, g, std::string()};
        be.mark("start");
        prp::Input in;
        in.n = n; in.m = m; in.begin = g->begin; in.r_begin = g->r_begin; in.r_node_idx = g->r_node_idx; in.world = world; in.rank = rank;
        rc = prp::build(be, in, cfg, L);
        if (rc != 0) { set_error("%s", L.error.empty() ? "pagerank plan: failed" : L.error.c_str()); if (rc > 0) rc = GMC_ERR_CUDA; goto fail; }
        if (cudaGetLastError() != cudaSuccess) { rc = GMC_ERR_CUDA; set_error("pagerank plan: a plan kernel failed to launch"); goto fail; }
    }
    if (getenv("GMC_PR_FAKE_ALLHOT")) {       // experiment (results invalid): every L2 gather of the home stream becomes a tier-0 slot
        const Stream& stx = L.streams[0];
        int32_t b1 = 0;
        cudaMemcpy(&b1, L.sbeg + stx.sbeg_off + stx.n_slices, 4, cudaMemcpyDeviceToHost);
        const int32_t e1 = stx.n_slices > 0 ? b1 : stx.e_long + stx.EL;
        if (e1 > stx.e_long) k_fake_allhot<<<div_up(e1 - stx.e_long, 256), 256, 0, st>>>(L.E, stx.e_long, e1);
    }
    if (const char* fb = getenv("GMC_PR_FAKE_BANKS")) {     // experiment: 1 = tier streams, 2 = + the tier-0 entries of the home stream
        const int mode = atoi(fb);
        for (int s2 = mode >= 2 ? 0 : 2; s2 <= L.thi + 1; s2++) {
            if (s2 == 1 && !L.split) continue;
            const Stream& stx = L.streams[s2];
            std::vector<int32_t> hb(2);
            if (stx.EL > 0) k_fake_banks<<<div_up(stx.EL, 256), 256, 0, st>>>(L.E, stx.e_long, stx.e_long + stx.EL, 0, HT);
            if (stx.n_slices > 0) {
                int32_t b0 = 0, b1 = 0;
                cudaMemcpy(&b0, L.sbeg + stx.sbeg_off, 4, cudaMemcpyDeviceToHost);
                cudaMemcpy(&b1, L.sbeg + stx.sbeg_off + stx.n_slices, 4, cudaMemcpyDeviceToHost);
                if (b1 > b0) k_fake_banks<<<div_up(b1 - b0, 256), 256, 0, st>>>(L.E, b0, b1, 1, HT);
            }
        }
    }
    // cut-row arrays of the long sections, diff partials, part slots
    {
        p->n_cb = L.K > 0 ? (L.K + CB_THREADS - 1) / CB_THREADS : 1;
        if (p->n_cb < sm_count()) p->n_cb = sm_count();          // fused mode: one partial per CTA of pr_sweep
        const size_t nl = (size_t)(L.n_ltslots ? L.n_ltslots : 1), nparts = (size_t)L.dp_total + (size_t)p->n_cb + 1;
        PLAN_TRY(gmc_malloc((void**)&p->head_part, 8 * nl)); PLAN_TRY(gmc_malloc((void**)&p->tail_part, 8 * nl));
        PLAN_TRY(gmc_malloc((void**)&p->split_cnt, 4 * nl)); PLAN_TRY(gmc_malloc((void**)&p->diff_parts, 8 * nparts));
        PLAN_TRY(gmc_malloc((void**)&p->parts, 8 * ((size_t)L.n_parts + 1)));
        cudaMemsetAsync(p->head_part, 0, 8 * nl, st); cudaMemsetAsync(p->tail_part, 0, 8 * nl, st);
        cudaMemsetAsync(p->split_cnt, 0, 4 * nl, st); cudaMemsetAsync(p->diff_parts, 0, 8 * nparts, st);
        cudaMemsetAsync(p->parts, 0, 8 * ((size_t)L.n_parts + 1), st);
    }
    // state (replicated on every rank: the gather needs the whole contribution vector)
    if (state_of) {
        for (int k = 0; k < 2; k++) { p->contrib[k] = state_of->contrib[k]; p->hot[k] = state_of->hot[k]; p->rank[k] = state_of->rank[k]; }
        p->ctrl = state_of->ctrl; p->diffs = state_of->diffs; p->flags = state_of->flags;
    } else {
        if (world > SYNC_RANKS) { rc = GMC_ERR_ARG; set_error("pagerank: at most %d shards", SYNC_RANKS); goto fail; }
        const size_t stride = state_stride(n), nh = (size_t)L.NH, sync_doubles = 2 * SYNC_RANKS + SYNC_RANKS / 2;
        p->block_bytes = 8 * (2 * stride + 2 * nh + sync_doubles);
        if (world > 1 && g_sym.block && g_sym.bytes == p->block_bytes && g_sym.world == world && g_sym.rank == rank) {
            p->block = g_sym.block; p->block_ipc = true; p->block_cached = true; p->gen = g_sym.gen; p->peer_local = g_sym.local;
            for (int r = 0; r <= MAX_PEERS; r++) { p->peer_base[r] = g_sym.p2p ? g_sym.peer_base[r] : nullptr; p->peer_gen[r] = g_sym.peer_gen[r]; }
            p->peer_base[rank] = nullptr;                 // plan_map_peers validates the mappings before they are used
            g_sym = SymCache();
        } else if (world > 1) {
            PLAN_TRY(gmc_malloc_ipc((void**)&p->block, p->block_bytes));   // exported to the peers
            p->block_ipc = true; p->gen = ++g_block_gen;
        } else PLAN_TRY(gmc_malloc((void**)&p->block, p->block_bytes));
        p->contrib[0] = p->block; p->contrib[1] = p->block + stride;
        p->hot[0] = p->block + 2 * stride; p->hot[1] = p->hot[0] + nh;
        p->diffs = p->block + 2 * stride + 2 * nh;
        p->flags = reinterpret_cast<unsigned int*>(p->diffs + 2 * SYNC_RANKS);
        PLAN_CUDA(cudaMemsetAsync(p->diffs, 0, 8 * sync_doubles, st));
        for (int k = 0; k < 2; k++) PLAN_TRY(gmc_malloc((void**)&p->rank[k], 8 * (nn + 4)));
        PLAN_TRY(gmc_malloc((void**)&p->ctrl, sizeof(gmc_pr_ctrl)));
    }
    p->grid = sm_count();
    p->smem = (size_t)HT * 8;
    PLAN_CUDA(cudaFuncSetAttribute(pr_sweep, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)p->smem));
    PLAN_CUDA(cudaStreamSynchronize(st));
fail:
    if (rc != GMC_OK) { free_plan(p); return rc; }
    *out = p;
    return GMC_OK;
}

// sweep: index of the sweep inside the call (0-based).  From the third sweep on the rows without in-edges are
// skipped: their value is the constant (1-d)/N, both ping-pong buffers hold it (and its contribution) after two
// sweeps, and their |val - old| is exactly 0 -- the stored diff partials of sweep 1 stay valid.
SweepArgs make_args(const gmc_pr_plan* p, double d, int sweep, gmc_pr_ctrl* ctrl, int loop, double e, int32_t max_iter) {
    const int parity = sweep & 1;
    const prp::Plan& L = p->L;
    SweepArgs A;
    A.E = L.E; A.eoff = L.eoff; A.sbeg = L.sbeg; A.ltask_row = L.ltask_row; A.stask_slice = L.stask_slice;
    A.od_flags = L.od_flags; A.hotslot = L.hotslot; A.home_slot = L.home_slot; A.seg_slot = L.seg_slot; A.pbeg = L.pbeg; A.prow = L.prow;
    A.streams = L.streams_dev; A.sched = L.sched_dev;
    A.N = L.N; A.T = L.T; A.K = L.K; A.lt = L.lt; A.lt_shift = L.lt_shift;
    A.long_id0 = L.jl0 + 1; A.short_id0 = L.g_n_long + 1 + 32 * L.s0;
    A.n_ztasks = L.n_ztasks; A.z0 = L.z0; A.z_end = L.z_end; A.zt_on = sweep >= 2 ? 0 : 1;
    A.dp_short = L.dp_short; A.dp_zero = L.dp_zero; A.dp_cut = L.dp_cut; A.dp_total = L.dp_total; A.n_cb = p->n_cb;
    {
        const int32_t want = L.n_src + 1 < HT ? L.n_src + 1 : HT;         // slot 0 + the tier-0 sources that exist
        A.hot0_bytes = ((want + 1) & ~1) * 8;                              // multiple of 16 bytes
    }
    {
        static const int cw = env_int("GMC_PR_COLD_WARPS", (PR_THREADS / 32) * 3 / 8);      // 12 of 32 warps
        A.cold_warps = L.split ? cw : PR_THREADS / 32;
        A.split = L.split;
        A.cold_s = L.thi > L.T ? L.thi + 1 : 0;
        static const int two = env_int("GMC_PR_TWO_KERNELS", 0);
        A.fused = two ? 0 : 1;
        A.bar_target = 0;
        if (!L.split) A.cold_warps = L.thi > L.T ? env_int("GMC_PR_COLD_WARPS", (PR_THREADS / 32) * 3 / 8) : 0;
    }
    A.head_part = p->head_part; A.tail_part = p->tail_part; A.split_cnt = p->split_cnt; A.diff_parts = p->diff_parts; A.parts = p->parts;
    A.contrib_in = p->contrib[parity]; A.contrib_out = p->contrib[parity ^ 1];
    A.hot_in = p->hot[parity]; A.hot_out = p->hot[parity ^ 1];
    A.rank_in = p->rank[parity]; A.rank_out = p->rank[parity ^ 1];
    A.d = d;
    A.base = (1 - d) / (double)p->N;     // (1-d) / N, pagerank.gm:11
    A.ctrl = ctrl; A.loop = loop; A.diff_slot = nullptr; A.e = e; A.max_iter = max_iter;
    A.npeers = 0; A.seq = 0; A.my_flag = nullptr;
    for (int q = 0; q < MAX_PEERS; q++) { A.peer_out[q] = nullptr; A.peer_hot[q] = nullptr; A.peer_diff[q] = nullptr; A.peer_flag[q] = nullptr; }
    return A;
}

// sharded sweep number `sweep` of the call: where the shard's diff goes, and (peers mapped) where the copies go
void set_exchange_args(const gmc_pr_plan* p0, const gmc_pr_plan* p, int sweep, unsigned int seq, SweepArgs& A) {
    const size_t stride = state_stride(p0->N), nh = (size_t)p0->L.NH;
    const int par = sweep & 1;
    A.diff_slot = p0->diffs + par * SYNC_RANKS + p->shard;
    if (!p0->p2p) return;
    A.seq = seq;
    A.my_flag = p0->flags + p->shard;
    int q = 0;
    for (int r = 0; r < p0->world; r++) {
        if (r == p->shard) continue;
        double* base = static_cast<double*>(p0->peer_base[r]);
        A.peer_out[q] = base + (par ^ 1) * stride;
        A.peer_hot[q] = base + 2 * stride + (par ^ 1) * nh;
        A.peer_diff[q] = base + 2 * stride + 2 * nh + par * SYNC_RANKS + p->shard;
        A.peer_flag[q] = reinterpret_cast<unsigned int*>(base + 2 * stride + 2 * nh + 2 * SYNC_RANKS) + p->shard;
        q++;
    }
    A.npeers = q;
}

int launch_init(const gmc_pr_plan* p, cudaStream_t st) {
    SweepArgs A = make_args(p, 0.0, 0, p->ctrl, 0, 0.0, 0);
    GMC_CUDA(cudaMemsetAsync(p->hot[0], 0, 16 * (size_t)p->L.NH, st));       // both hotpacks: dummy and unused slots stay 0.0
    k_init<<<div_up(p->N + 2, 256), 256, 0, st>>>(A, p->contrib[0], p->contrib[1], p->hot[0], p->rank[0]);
    GMC_CUDA(cudaGetLastError());
    return GMC_OK;
}

// One sweep = one launch of pr_sweep: cooperative (every CTA resident: the grid barrier before the split rows) and with
// programmatic stream serialization (the next sweep's CTAs take their SMs while this one's tail drains; they order
// themselves behind it with griddepcontrol.wait).  GMC_PR_NO_PDL=1 = plain cooperative launches; GMC_PR_TWO_KERNELS=1 = no
// barrier, the split rows in a second kernel (pr_combine).  `state` = the plan that owns the control block.
int launch_sweep(gmc_pr_plan* state, const gmc_pr_plan* p, SweepArgs A, cudaStream_t st) {
    static const bool pdl = getenv("GMC_PR_NO_PDL") == nullptr;
    static int coop_pdl_ok = 1;              // cleared when the driver refuses the two attributes together
    cudaLaunchAttribute at[3];
    at[0].id = cudaLaunchAttributeCooperative;
    at[0].val.cooperative = 1;
    at[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[1].val.programmaticStreamSerializationAllowed = 1;
    // GMC_PR_L2_WINDOW_MB=W (experiment): the first W MB of the contribution vector this sweep gathers from (the ids of the
    // largest in-degree, which on a power-law graph are the most gathered sources) are read with the persisting L2 property, so
    // that the cold gathers of a vector larger than L2 (TW-shaped: 333 MB) do not evict them.  The two buffers alternate, so the
    // set-aside is sized for both windows.
    static const int l2_window_mb = env_int("GMC_PR_L2_WINDOW_MB", 0);
    int n_extra = 0;
    if (l2_window_mb > 0) {
        static thread_local int set_aside_dev = -1;
        int dev = 0;
        GMC_CUDA(cudaGetDevice(&dev));
        if (set_aside_dev != dev) {
            int max_persist = 0;
            GMC_CUDA(cudaDeviceGetAttribute(&max_persist, cudaDevAttrMaxPersistingL2CacheSize, dev));
            const size_t want = std::min<size_t>((size_t)max_persist, 2 * ((size_t)l2_window_mb << 20));
            GMC_CUDA(cudaDeviceSetLimit(cudaLimitPersistingL2CacheSize, want));
            if (getenv("GMC_PR_L2_WINDOW_VERBOSE")) fprintf(stderr, "pr_sweep: persisting L2 set-aside %zu of max %d bytes\n", want, max_persist);
            set_aside_dev = dev;
        }
        int max_win = 0;
        GMC_CUDA(cudaDeviceGetAttribute(&max_win, cudaDevAttrMaxAccessPolicyWindowSize, dev));
        size_t bytes = std::min<size_t>((size_t)l2_window_mb << 20, 8 * ((size_t)p->N + 2));
        bytes = std::min<size_t>(bytes, (size_t)max_win);
        static const double hit = getenv("GMC_PR_L2_WINDOW_HIT") ? atof(getenv("GMC_PR_L2_WINDOW_HIT")) : 1.0;
        at[2].id = cudaLaunchAttributeAccessPolicyWindow;
        at[2].val.accessPolicyWindow.base_ptr = const_cast<double*>(A.contrib_in);
        at[2].val.accessPolicyWindow.num_bytes = bytes;
        at[2].val.accessPolicyWindow.hitRatio = (float)hit;
        at[2].val.accessPolicyWindow.hitProp = cudaAccessPropertyPersisting;
        at[2].val.accessPolicyWindow.missProp = cudaAccessPropertyNormal;
        n_extra = 1;
    }
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)p->grid);
    cfg.blockDim = dim3(PR_THREADS);
    cfg.dynamicSmemBytes = p->smem;
    cfg.stream = st;
    if (A.fused) {
        state->bar_arrivals += (unsigned)p->grid;
        A.bar_target = state->bar_arrivals;
        A.n_cb = p->grid;
        cfg.attrs = at;
        const bool two = pdl && coop_pdl_ok;
        if (!two && n_extra) at[1] = at[2];
        cfg.numAttrs = (two ? 2 : 1) + n_extra;
        cudaError_t e = cudaLaunchKernelEx(&cfg, pr_sweep, A);
        if (e != cudaSuccess && two) {       // cooperative + programmatic serialization not accepted: cooperative only
            cudaGetLastError();
            coop_pdl_ok = 0;
            if (n_extra) at[1] = at[2];
            cfg.numAttrs = 1 + n_extra;
            e = cudaLaunchKernelEx(&cfg, pr_sweep, A);
        }
        GMC_CUDA(e);
        return GMC_OK;
    }
    cfg.attrs = at + 1;
    cfg.numAttrs = pdl ? 1 : 0;
    GMC_CUDA(cudaLaunchKernelEx(&cfg, pr_sweep, A));
    cfg.gridDim = dim3((unsigned)(p->L.K > 0 ? (p->L.K + CB_THREADS - 1) / CB_THREADS : 1));
    A.n_cb = (int32_t)cfg.gridDim.x;
    cfg.blockDim = dim3(CB_THREADS);
    cfg.dynamicSmemBytes = 0;
    GMC_CUDA(cudaLaunchKernelEx(&cfg, pr_combine, A));
    return GMC_OK;
}
constexpr int LAUNCHES_PER_SWEEP = 1;

}  // namespace

// called by gmc_comm_destroy: the retired state block (and the peers' mappings of it) does not outlive the communicator
extern "C" void gmc_pr_release_cached_state(void) { sym_cache_release(); }

void gmc_pr_plan_free(gmc_graph* g) {
    if (g->pr) { free_plan(g->pr); g->pr = nullptr; }
}

// iters_hint: max_iter of the call the plan is wanted for (0: a full plan whatever the call)
static int graph_plan(gmc_graph* g, int world, int rank, cudaStream_t st, int32_t iters_hint = 0) {
    if (g->pr && g->pr->world == world && g->pr->shard == rank) {
        if (!g->pr->lite) return GMC_OK;
        iters_hint = 0;                       // second use of a lite plan: the graph is reused, build the tiered plan now
    }
    g->pr_iters_hint = iters_hint;
    GMC_REQUIRE(g->has_reverse, GMC_ERR_STATE, "pagerank needs the reverse CSR: call gmc_graph_make_reverse first");
    gmc_pr_plan_free(g);
    gmc_pr_plan* p = nullptr;
    GMC_TRY(build_plan(g, world, rank, nullptr, &p, st));
    g->pr = p;
    return GMC_OK;
}

// Map the peers' state blocks (once per plan).  Every rank takes the same path: all mapped, or all on NCCL.
static int plan_map_peers(gmc_pr_plan* p, gmc_comm* comm, cudaStream_t st) {
    if (!comm || comm->world == 1 || p->p2p || p->peer_base[p->shard]) return GMC_OK;
    p->peer_base[p->shard] = p->block;        // marks "tried"
    if (comm->world > MAX_PEERS + 1) return GMC_OK;
    const int W = comm->world;
    // every rank publishes (allocation number, "my block and mappings come from the cache"); the cached mappings are used
    // only if EVERY rank reuses the very allocation its peers mapped last time -- otherwise all ranks map afresh
    unsigned long long mine[2] = {p->gen, p->block_cached ? 1ull : 0ull}, all[2 * (MAX_PEERS + 1)];
    unsigned long long* stage = nullptr;
    GMC_TRY(gmc_malloc((void**)&stage, 16 * (size_t)W));
    int rc = GMC_OK;
    if (cudaMemcpyAsync(stage + 2 * p->shard, mine, 16, cudaMemcpyHostToDevice, st) != cudaSuccess) rc = GMC_ERR_CUDA;
    if (rc == GMC_OK) rc = gmc_comm_allgather(comm, stage + 2 * p->shard, stage, 16, st);
    if (rc == GMC_OK && (cudaMemcpyAsync(all, stage, 16 * (size_t)W, cudaMemcpyDeviceToHost, st) != cudaSuccess || cudaStreamSynchronize(st) != cudaSuccess)) rc = GMC_ERR_CUDA;
    cudaFree(stage);
    if (rc != GMC_OK) { set_error("pagerank: peer block numbers could not be exchanged"); return rc; }
    bool reuse = true;                        // the gathered vector is the same on every rank, and so is the recorded one
    for (int r = 0; r < W; r++) reuse = reuse && all[2 * r + 1] == 1ull && all[2 * r] == (r == p->shard ? p->gen : p->peer_gen[r]) && (r == p->shard || p->peer_base[r]);
    if (reuse) { p->p2p = true; return GMC_OK; }
    for (int r = 0; r < W; r++)               // stale mappings of a cached block (a peer has a new allocation)
        if (r != p->shard && p->peer_base[r]) { if (!p->peer_local) cudaIpcCloseMemHandle(p->peer_base[r]); p->peer_base[r] = nullptr; }
    void* bases[MAX_PEERS + 1];
    rc = gmc_comm_ipc_open_all(comm, p->block, bases, st);
    p->peer_local = comm->local;
    if (rc == GMC_OK) {
        for (int r = 0; r < W; r++) { p->peer_base[r] = bases[r]; p->peer_gen[r] = all[2 * r]; }
        p->p2p = true;
    } else if (rc != GMC_ERR_STATE) return rc;
    return GMC_OK;
}

// The iteration loop over a set of shard plans living on this device: one plan and a communicator (a real
// rank), or `world` plans and no communicator (emulated shards: the exchange is the shared state itself).
static int run_pagerank(gmc_graph* g, gmc_pr_plan* const* plans, int nplans, gmc_comm* comm, double e, double d, int32_t max_iter,
                        double* rank_dev, gmc_pagerank_info* info, cudaStream_t st) {
    gmc_pr_plan* p0 = plans[0];
    const int world = p0->world;
    const bool sharded = world > 1;
    GMC_CUDA(cudaMemsetAsync(p0->ctrl, 0, sizeof(gmc_pr_ctrl), st));
    p0->bar_arrivals = 0;
    // peer-memory exchange: the ranks enter the loop together (device-side barrier), so that the flag timeout of the sweeps
    // bounds waits inside the loop only -- not the skew between the ranks' host threads before a repeat call on a kept plan
    if (sharded && comm && p0->p2p) GMC_TRY(gmc_comm_stream_barrier(comm, st));
    cudaEvent_t ev0, ev1;
    GMC_CUDA(cudaEventCreate(&ev0));
    GMC_CUDA(cudaEventCreate(&ev1));
    int launches = 0;
    GMC_CUDA(cudaEventRecord(ev0, st));
    GMC_TRY(launch_init(p0, st));
    launches++;
    // do { sweep; cnt++ } while (diff > e && cnt < max): sweeps are enqueued in batches; kernels of a
    // finished loop exit at once on ctrl->done, so the host only synchronises once per batch.
    gmc_pr_ctrl h;
    memset(&h, 0, sizeof(h));
    int32_t issued = 0;
    const int32_t BATCH = 32;
    do {
        int32_t todo = max_iter - issued;
        if (todo > BATCH) todo = BATCH;
        if (todo < 1) todo = 1;   // a do-while body runs at least once (max <= 0 still sweeps once)
        for (int32_t k = 0; k < todo; k++, issued++) {
            for (int i = 0; i < nplans; i++) {
                SweepArgs A = make_args(plans[i], d, issued, p0->ctrl, 1, e, max_iter);
                if (sharded) set_exchange_args(p0, plans[i], issued, p0->seq + 1, A);
                GMC_TRY(launch_sweep(p0, plans[i], A, st));
                launches += LAUNCHES_PER_SWEEP;
            }
            if (sharded) {
                double* dslot = p0->diffs + (issued & 1) * SYNC_RANKS;
                p0->seq++;
                if (comm && !p0->p2p) {
                    // exchange step over NCCL: every rank's new contributions (3 id ranges per rank) and its diff
                    GMC_TRY(gmc_comm_allgather_segments(comm, p0->contrib[(issued & 1) ^ 1], 8, p0->L.seg_off.data(), p0->L.seg_len.data(), 3, st));
                    k_hot_refresh<<<div_up(p0->L.NH, 256), 256, 0, st>>>(p0->L.idslot, p0->L.NH, p0->contrib[(issued & 1) ^ 1], p0->hot[(issued & 1) ^ 1]);
                    launches++;
                    GMC_TRY(gmc_comm_allgather(comm, dslot + p0->shard, dslot, 8, st));
                }
                // peers mapped: the contributions and diffs were pushed by the sweeps themselves; wait for the flags
                k_ctrl_combine<<<1, 32, 0, st>>>(p0->ctrl, dslot, world, e, max_iter, p0->p2p ? p0->flags : nullptr, p0->seq);
                launches++;
            }
        }
        GMC_CUDA(cudaMemcpyAsync(&h, p0->ctrl, sizeof(h), cudaMemcpyDeviceToHost, st));
        GMC_CUDA(cudaStreamSynchronize(st));
    } while (!h.done);
    if (h.error) { set_error(h.error == 2 ? "pagerank: the grid barrier of pr_sweep timed out" : "pagerank: a peer rank did not finish its sweep in time (peer-memory exchange)"); return GMC_ERR_STATE; }
    if (sharded && comm)      // the result vector: gather the owned slices of rank[] once
        GMC_TRY(gmc_comm_allgather_segments(comm, p0->rank[h.cnt & 1], 8, p0->L.seg_off.data(), p0->L.seg_len.data(), 3, st));
    k_unpermute<<<div_up(g->N, 256), 256, 0, st>>>(p0->rank[h.cnt & 1], p0->L.new_of_old, g->N, rank_dev);
    launches++;
    GMC_CUDA(cudaEventRecord(ev1, st));
    GMC_CUDA(cudaStreamSynchronize(st));
    if (info) {
        info->iterations = h.cnt;
        info->diff = h.diff;
        info->launches = launches;
        cudaEventElapsedTime(&info->kernel_ms, ev0, ev1);
    }
    cudaEventDestroy(ev0); cudaEventDestroy(ev1);
    return GMC_OK;
}

extern "C" {

int gmc_pagerank_sharded_dev(gmc_graph* g, gmc_comm* comm, double e, double d, int32_t max_iter, double* rank_dev,
                             gmc_pagerank_info* info, void* stream) {
    GMC_REQUIRE(g && rank_dev, GMC_ERR_ARG, "gmc_pagerank_dev: NULL");
    GMC_REQUIRE(g->N > 0, GMC_ERR_ARG, "gmc_pagerank_dev: empty graph");
    cudaStream_t st = as_stream(stream);
    const int world = comm ? comm->world : 1, rank = comm ? comm->rank : 0;
    static const bool timing = getenv("GMC_PR_CALL_TIMING") != nullptr;       // phases of the call, wall clock, to stderr
    auto now_ms = []() { struct timespec ts; clock_gettime(CLOCK_MONOTONIC, &ts); return ts.tv_sec * 1e3 + ts.tv_nsec * 1e-6; };
    const double t0 = timing ? now_ms() : 0.0;
    GMC_TRY(graph_plan(g, world, rank, st, max_iter > 0 ? max_iter : 1));
    const double t1 = timing ? now_ms() : 0.0;
    gmc_pr_plan* p = g->pr;
    GMC_TRY(plan_map_peers(p, comm, st));
    const double t2 = timing ? now_ms() : 0.0;
    const int rc = run_pagerank(g, &p, 1, world > 1 ? comm : nullptr, e, d, max_iter, rank_dev, info, st);
    if (timing) fprintf(stderr, "pagerank call (rank %d of %d): plan %.2f ms (block %s), peer mapping %.2f ms, loop + result %.2f ms\n", rank, world,
                        t1 - t0, p->block_cached ? "reused" : "new", t2 - t1, now_ms() - t2);
    return rc;
}

int gmc_pagerank_dev(gmc_graph* g, double e, double d, int32_t max_iter, double* rank_dev, gmc_pagerank_info* info, void* stream) {
    return gmc_pagerank_sharded_dev(g, nullptr, e, d, max_iter, rank_dev, info, stream);
}

int gmc_pagerank_emulated_shards_dev(gmc_graph* g, int world, double e, double d, int32_t max_iter, double* rank_dev,
                                     gmc_pagerank_info* info, void* stream) {
    GMC_REQUIRE(g && rank_dev && world >= 1 && world <= 64, GMC_ERR_ARG, "gmc_pagerank_emulated_shards_dev: bad args");
    GMC_REQUIRE(g->N > 0, GMC_ERR_ARG, "gmc_pagerank_emulated_shards_dev: empty graph");
    GMC_REQUIRE(g->has_reverse, GMC_ERR_STATE, "pagerank needs the reverse CSR: call gmc_graph_make_reverse first");
    cudaStream_t st = as_stream(stream);
    std::vector<gmc_pr_plan*> plans;
    int rc = GMC_OK;
    g->pr_iters_hint = 0;                 // full (tiered) shard plans
    for (int r = 0; r < world && rc == GMC_OK; r++) {
        gmc_pr_plan* p = nullptr;
        rc = build_plan(g, world, r, r ? plans[0] : nullptr, &p, st);
        if (rc == GMC_OK) plans.push_back(p);
    }
    if (rc == GMC_OK) rc = run_pagerank(g, plans.data(), world, nullptr, e, d, max_iter, rank_dev, info, st);
    for (size_t i = plans.size(); i-- > 0;) free_plan(plans[i]);      // plan 0 (the state owner) last
    return rc;
}

// measurement aid for the roofline figure: every sweep bracketed by its own pair of CUDA events
int gmc_pagerank_sharded_time_sweeps(gmc_graph* g, gmc_comm* comm, double d, int32_t warmup, int32_t n, float* sweep_ms,
                                     float* exchange_ms, void* stream) {
    GMC_REQUIRE(g && sweep_ms && n > 0 && warmup >= 0, GMC_ERR_ARG, "gmc_pagerank_time_sweeps: bad args");
    GMC_REQUIRE(g->N > 0, GMC_ERR_ARG, "gmc_pagerank_time_sweeps: empty graph");
    cudaStream_t st = as_stream(stream);
    const int world = comm ? comm->world : 1, rank = comm ? comm->rank : 0;
    GMC_TRY(graph_plan(g, world, rank, st));
    gmc_pr_plan* p = g->pr;
    GMC_TRY(plan_map_peers(p, comm, st));
    GMC_CUDA(cudaMemsetAsync(p->ctrl, 0, sizeof(gmc_pr_ctrl), st));
    p->bar_arrivals = 0;
    GMC_TRY(launch_init(p, st));
    std::vector<cudaEvent_t> ev(3 * (size_t)n);
    for (auto& e : ev) GMC_CUDA(cudaEventCreate(&e));
    for (int32_t k = 0; k < warmup + n; k++) {
        const bool timed = k >= warmup;
        if (timed) GMC_CUDA(cudaEventRecord(ev[3 * (k - warmup)], st));
        SweepArgs A = make_args(p, d, k, p->ctrl, 0, 0.0, 0);
        if (world > 1) set_exchange_args(p, p, k, p->seq + 1, A);
        GMC_TRY(launch_sweep(p, p, A, st));
        if (timed) GMC_CUDA(cudaEventRecord(ev[3 * (k - warmup) + 1], st));
        if (world > 1) {
            p->seq++;
            if (p->p2p) k_wait_peers<<<1, 32, 0, st>>>(p->ctrl, p->flags, world, p->seq);
            else {
                GMC_TRY(gmc_comm_allgather_segments(comm, p->contrib[(k & 1) ^ 1], 8, p->L.seg_off.data(), p->L.seg_len.data(), 3, st));
                k_hot_refresh<<<div_up(p->L.NH, 256), 256, 0, st>>>(p->L.idslot, p->L.NH, p->contrib[(k & 1) ^ 1], p->hot[(k & 1) ^ 1]);
            }
        }
        if (timed) GMC_CUDA(cudaEventRecord(ev[3 * (k - warmup) + 2], st));
    }
    GMC_CUDA(cudaStreamSynchronize(st));
    for (int32_t k = 0; k < n; k++) {
        GMC_CUDA(cudaEventElapsedTime(&sweep_ms[k], ev[3 * k], ev[3 * k + 1]));
        if (exchange_ms) GMC_CUDA(cudaEventElapsedTime(&exchange_ms[k], ev[3 * k + 1], ev[3 * k + 2]));
    }
    for (auto& e : ev) cudaEventDestroy(e);
    return GMC_OK;
}

int gmc_pagerank_time_sweeps(gmc_graph* g, double d, int32_t warmup, int32_t n, float* ms_out, int32_t* launches_per_sweep, void* stream) {
    if (launches_per_sweep) *launches_per_sweep = LAUNCHES_PER_SWEEP;
    return gmc_pagerank_sharded_time_sweeps(g, nullptr, d, warmup, n, ms_out, nullptr, stream);
}

int gmc_pagerank(gmc_graph* g, double e, double d, int32_t max_iter, double* rank_host, gmc_pagerank_info* info) {
    GMC_REQUIRE(g && (rank_host || g->N == 0), GMC_ERR_ARG, "gmc_pagerank: NULL");
    if (g->N == 0) { if (info) memset(info, 0, sizeof(*info)); return GMC_OK; }
    double* tmp = nullptr;
    GMC_TRY(gmc_malloc((void**)&tmp, sizeof(double) * (size_t)g->N));
    int rc = gmc_pagerank_dev(g, e, d, max_iter, tmp, info, nullptr);
    if (rc == GMC_OK) rc = gmc_d2h(rank_host, tmp, sizeof(double) * (size_t)g->N, nullptr);
    cudaFree(tmp);
    return rc;
}

// how the current plan exchanges contributions between ranks: 0 = single rank, 1 = NCCL allgather after the sweep,
// 2 = peer-memory stores from the sweep epilogue (CUDA IPC mapping over NVLink)
// The generated pagerank() procedure as ONE call on host arrays: Shoal copy-in of begin / r_begin / r_node_idx
// (gm_cpp_gen.cc:685-775), the do-while loop nest, copy_back of the rank array (:1480).  The copy-in runs on its own
// stream in chunks; the plan builder starts as soon as the two offset arrays have landed (degrees, sorts, relabelling
// need nothing else) and fills the entry array chunk by chunk behind the PCIe copy, so ~2 ms of plan building hide
// behind the 5.7 ms the LJ-shaped copy-in takes.  Host arrays should be pinned (gmc_host_register) for full PCIe rate.
int gmc_pagerank_host_csr(int32_t num_nodes, int32_t num_edges, const int32_t* begin, const int32_t* r_begin, const int32_t* r_node_idx,
                          double e, double d, int32_t max_iter, double* rank_host, gmc_pagerank_info* info) {
    GMC_REQUIRE(num_nodes >= 0 && num_edges >= 0 && begin && r_begin && (r_node_idx || num_edges == 0) && (rank_host || num_nodes == 0),
                GMC_ERR_ARG, "gmc_pagerank_host_csr: bad args");
    if (num_nodes == 0) { if (info) memset(info, 0, sizeof(*info)); return GMC_OK; }
    // the copy stream and the events are made once per process (creating them per call costs driver round trips whose
    // latency is not bounded); calls are serialised on them
    static std::mutex mtx;
    static cudaStream_t cs = nullptr;
    static gmc_pending_upload pu;
    std::lock_guard<std::mutex> lock(mtx);
    if (!cs) {
        // non-blocking: the allocator and the plan kernels use the legacy default stream, which must not wait for the copies
        cudaStream_t s2 = nullptr;
        bool ok = cudaStreamCreateWithFlags(&s2, cudaStreamNonBlocking) == cudaSuccess &&
                  cudaEventCreateWithFlags(&pu.small, cudaEventDisableTiming) == cudaSuccess;
        for (int c = 0; c < gmc_pending_upload::MAX_CHUNKS && ok; c++) ok = cudaEventCreateWithFlags(&pu.chunk[c], cudaEventDisableTiming) == cudaSuccess;
        if (!ok) { set_error("gmc_pagerank_host_csr: stream/event: %s", cudaGetErrorString(cudaGetLastError())); return GMC_ERR_CUDA; }
        cs = s2;
    }
    gmc_graph* g = nullptr;
    GMC_TRY(gmc_graph_create(&g, num_nodes, num_edges));
    double* tmp = nullptr;
    int rc = GMC_OK;
    const size_t nb = sizeof(int32_t) * ((size_t)num_nodes + 1);
    auto fail = [&](const char* what) { set_error("gmc_pagerank_host_csr: %s: %s", what, cudaGetErrorString(cudaGetLastError())); rc = GMC_ERR_CUDA; };
    do {
        g->has_forward = num_edges == 0;           // pull-only graph: node_idx is not copied in
        if ((rc = gmc_malloc((void**)&g->r_begin, nb)) != GMC_OK) break;
        if ((rc = gmc_malloc((void**)&g->r_node_idx, sizeof(int32_t) * (size_t)(num_edges ? num_edges : 1))) != GMC_OK) break;
        if ((rc = gmc_malloc((void**)&tmp, sizeof(double) * (size_t)num_nodes)) != GMC_OK) break;
        if (cudaMemcpyAsync(g->begin, begin, nb, cudaMemcpyHostToDevice, cs) != cudaSuccess ||
            cudaMemcpyAsync(g->r_begin, r_begin, nb, cudaMemcpyHostToDevice, cs) != cudaSuccess ||
            cudaEventRecord(pu.small, cs) != cudaSuccess) { fail("copy-in"); break; }
        pu.nchunks = num_edges >= (1 << 22) ? gmc_pending_upload::MAX_CHUNKS : 1;
        bool ok = true;
        for (int c = 0; c < pu.nchunks && ok; c++) {
            const int32_t lo = (int32_t)((int64_t)num_edges * c / pu.nchunks), hi = (int32_t)((int64_t)num_edges * (c + 1) / pu.nchunks);
            pu.end[c] = hi;
            if (hi > lo) ok = cudaMemcpyAsync(g->r_node_idx + lo, r_node_idx + lo, sizeof(int32_t) * (size_t)(hi - lo), cudaMemcpyHostToDevice, cs) == cudaSuccess;
            if (ok) ok = cudaEventRecord(pu.chunk[c], cs) == cudaSuccess;
        }
        if (!ok) { fail("copy-in"); break; }
        g->has_reverse = true;
        g->pending = &pu;
        rc = gmc_pagerank_dev(g, e, d, max_iter, tmp, info, nullptr);       // plan (waits on the events) + sweeps
        g->pending = nullptr;
        if (rc == GMC_OK) rc = gmc_d2h(rank_host, tmp, sizeof(double) * (size_t)num_nodes, nullptr);
    } while (0);
    cudaStreamSynchronize(cs);          // the host arrays are not read after the call returns, whatever happened
    cudaFree(tmp);
    g->pending = nullptr;
    gmc_graph_destroy(g);
    return rc;
}

int gmc_pagerank_exchange_path(gmc_graph* g, int* path) {
    GMC_REQUIRE(g && path, GMC_ERR_ARG, "gmc_pagerank_exchange_path: NULL");
    *path = !g->pr || g->pr->world == 1 ? 0 : (g->pr->p2p ? 2 : 1);
    return GMC_OK;
}

// layout facts of the plan, for tests and DESIGN.md tables
int gmc_pagerank_plan_info(gmc_graph* g, gmc_pagerank_plan_info_t* out) {
    GMC_REQUIRE(g && out, GMC_ERR_ARG, "gmc_pagerank_plan_info: NULL");
    GMC_REQUIRE(g->N > 0, GMC_ERR_ARG, "gmc_pagerank_plan_info: empty graph");
    if (!g->pr || g->pr->lite) GMC_TRY(graph_plan(g, g->pr ? g->pr->world : 1, g->pr ? g->pr->shard : 0, nullptr));
    gmc_pr_plan* p = g->pr;
    const prp::Plan& L = p->L;
    memset(out, 0, sizeof(*out));
    out->n_long = L.n_long; out->hot_n = L.n_src + 1 < HT ? L.n_src + 1 : HT; out->n_slices = L.n_slices; out->long_entries = L.streams[0].EL;
    out->total_entries = L.EN; out->n_long_tasks = L.streams[0].n_ltasks; out->n_short_tasks = L.streams[0].n_stasks;
    out->grid = p->grid; out->threads = PR_THREADS; out->smem_bytes = (int32_t)p->smem;
    out->n_pos = L.n_pos; out->n_src = L.n_src; out->tiers = L.T; out->segments = L.n_seg; out->split_rows = L.K; out->part_slots = L.n_parts;
    out->home_entries = L.streams[0].entries; out->hot0_entries = L.hot0_entries; out->cold_entries = L.cold_entries; out->segment_entries = L.seg_entries; out->cold_stream_entries = L.cold_stream_entries;
    {
        int c = 0;
        for (int b = 0; b < p->grid; b++) c += L.sched[(size_t)b * SCHED_W] > 1;          // CTAs that start in a warm tier
        out->tier_ctas = c;
    }
    return GMC_OK;
}

}  // extern "C"
