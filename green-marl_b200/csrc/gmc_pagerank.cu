// PageRank pull sweep (K1) with the updater (K2) and init (K3) fused away -- one launch per iteration.
//
// Reference loop being replaced (generated from apps/src/pagerank.gm:8-19, shape in SURVEY.md 3.1):
//   foreach t: S = sum_{w in InNbrs(t), r_node_idx order} rank[w] / (double)outdeg(w)
//              val = (1-d)/N + d*S;  diff += |val - rank[t]|;  rank_nxt[t] = val
//   rank = rank_nxt; cnt++           while (diff > e && cnt < max)
//
// Device formulation.  contrib[w] = rank[w]/(double)outdeg(w) is produced once per vertex by the
// epilogue of the previous sweep (the same IEEE division the reference performs once per EDGE), so a
// sweep gathers one double per edge; rank/contrib are ping-ponged, which removes the copy-back loop.
//
// Layout (built once per graph by build_plan, "the plan"): vertices are RELABELLED so that the kernel's
// two problems -- the random 8-byte gathers and the power-law row lengths -- become layout properties:
//   id 0            a dummy vertex whose contribution is always 0.0 (padding entries point at it)
//   ids [1,hot_n)   the hot block: among rows with in-degree > S_MAX, the hot_n-1 vertices with the largest
//                   OUT-degree (= the most gathered contributions).  Every CTA stages contrib[0,hot_n) in
//                   shared memory with one TMA bulk copy per sweep; ~half of all gathers of an R-MAT graph
//                   are then served from shared memory instead of L2 sectors.
//   ids [.., n_long] the other rows with in-degree > S_MAX, in-degree descending ("long" rows)
//   ids (n_long, N] rows with in-degree <= S_MAX, in-degree descending ("short" rows)
// The reverse CSR is rewritten over the new ids into one entry array E:
//   long rows   row-major, each row padded to a multiple of 4 entries (16-byte aligned, so a lane's int4
//               never straddles two rows).  Cut into fixed tasks of LT entries whatever the row lengths
//               (merge-path in the limit): a warp streams its task with 128-bit index loads, 8 gathers in
//               flight per lane, lane-private accumulation and one butterfly reduction per row end.  A row
//               cut by a task boundary leaves per-task partials; the last task to arrive (atomic counter)
//               adds them in task order -- deterministic, no spin-waiting.
//   short rows  sliced ELL: 32 consecutive rows per slice, column-major, padded to the slice's longest row
//               (rows are sorted by length, so padding is nil).  One lane per row adds its terms strictly
//               left to right in the ORIGINAL source order: the reference's summation order, so every row
//               of in-degree <= S_MAX is bit-identical to the CPU result given bit-identical inputs.
// HBM traffic per sweep: 4 B/edge (index stream) + 40 B/vertex (row pointer amortised, old rank, new rank,
// new contrib, out-degree, first touch of contrib); gathers are served by shared memory / L2.
// -DGMC_EXPERIMENT (tools/build_variant.sh) compiles the GMC_PR_MODE switches of the microbenchmark drivers into the sweep;
// the product build has none of them
#include "gmc_common.cuh"
#include "gmc_comm.h"
#include <cub/device/device_radix_sort.cuh>
#include <mutex>
#include <vector>
#include <stdlib.h>

using namespace gmc;

#ifndef GMC_PR_THREADS
#define GMC_PR_THREADS 1024
#endif
#ifndef GMC_PR_CTAS
#define GMC_PR_CTAS 1
#endif
constexpr int PR_THREADS = GMC_PR_THREADS;    // one CTA per SM, 32 independent warps (experiment builds: 512, 1 or 2 CTAs per SM)
constexpr int PR_CTAS = GMC_PR_CTAS;
constexpr int PR_WARPS = PR_THREADS / 32;
constexpr int S_MAX = 512;                // rows up to this in-degree are summed sequentially, one lane per row (bit-exact)
constexpr int LT = 2048;                  // entries per long-row task, default (16 warp steps of 128); power of two
constexpr int ST_COST = 2048;             // cost units per short-row task, default (slice cost = 32*len + slice_ovh)
constexpr int SLICE_OVH = 68;             // keeps a default task at <= 31 slices (one lane per slice offset)
// shards with many entries run 4x / 2x larger tasks: the start and the end of a task (metadata chain, row
// hand-over, split-row arrival) are latency chains without gathers in flight; measured on the LJ-shaped graph
// 0.2149 -> 0.2015 ms per sweep (gpurun_out/r1k_variants2.log), while graphs with few tasks per warp keep the small ones
constexpr int LT_BIG = 8192, ST_COST_BIG = 4096, SLICE_OVH_BIG = 136;
constexpr int64_t BIG_TASK_ENTRIES = 24000000;   // per-shard entries from which the large tasks are used
constexpr int ZT = 2048;                  // ids per zero-in-degree task
// entries one row epilogue weighs when the shards of a multi-rank run are cut.  The epilogue stores the new contribution
// into every peer (world-1 NVLink stores per row), so the weight grows with the world size.  Measured on the TW-shaped
// graph (max / min sweep time over ranks): 2 ranks: weight 0 -> 3.05 / 2.81 ms, 2 -> 3.12 / 2.82, 6 -> 3.23 / 2.65;
// 8 ranks: 6 -> 1.12 / 0.87, 14 -> 1.00 / 0.85, 26 -> 1.05 / 0.80 (profiles/r1p_shard_row_cost.txt)
static inline int shard_row_cost(int world) {
    if (const char* c = getenv("GMC_PR_ROWCOST")) return atoi(c);      // experiments
    return world > 2 ? 2 * (world - 2) : 0;
}
#ifndef GMC_HOT_CAP
#define GMC_HOT_CAP 16640
#endif
// doubles of shared memory for the hot block: 130 KB, the 132 KB carve-out step, so that 124 KB of the SM's 256 KB stay L1.
// The number of missing gathers an SM keeps in flight grows with the L1 left (tools/ubench), and the L1 itself serves
// the next-hottest sources; measured per sweep on the LJ-shaped graph (profiles/r1s_hot_cap.txt): 3900 doubles
// 0.2002 ms, 8000 0.1966, 12600 0.1914, 16640 0.1907, 20480 0.2000, 24576 0.2412
constexpr int HOT_CAP = GMC_HOT_CAP;
constexpr unsigned FULL = 0xffffffffu;

struct gmc_pr_ctrl {   // device-side loop control: lets the host enqueue sweeps without a sync per iteration
    double diff;
    int32_t cnt;
    int32_t done;
    unsigned int arrived;
    unsigned int next;      // dynamic task counter of the running sweep (reset by its last CTA)
    int32_t error;          // 1: a peer's end-of-sweep flag did not arrive in time (sharded runs)
    int32_t pad;
};
constexpr int MAX_PEERS = 7;            // peer-memory exchange up to 8 ranks (one NVSwitch domain); beyond: NCCL
constexpr int SYNC_RANKS = 64;          // slots of the sync area at the end of the state block
// doubles per contribution buffer inside the state block: N+2 used, rounded up to 256 bytes (TMA bulk source alignment)
static inline size_t state_stride(int32_t n) { return (((size_t)n + 4 + 31) / 32) * 32; }
constexpr long long PEER_TIMEOUT_CYCLES = 8000000000LL;    // ~4 s at 1.9 GHz: a dead peer fails the call, not the box

struct SweepArgs {
    // layout
    const int32_t* __restrict__ E;
    const int32_t* __restrict__ lbeg;        // [n_long+1]
    const int32_t* __restrict__ sbeg;        // [n_slices+1]
    const int32_t* __restrict__ ltask_row;   // [n_ltasks+1]
    const int32_t* __restrict__ stask_slice; // [n_stasks+1]
    const int32_t* __restrict__ outdeg;      // [N+1] by new id
    int32_t N, n_long, hot_n, EL;             // n_long, EL: rows / entries of THIS shard's long section
    int32_t LT, lt_shift;                     // entries per long-row task = 1 << lt_shift
    int32_t long_id0, short_id0;              // id of the shard's first long row / of lane 0 of its first slice
    int32_t n_ltasks, n_stasks, n_ztasks, z0, z_end;  // [z0, z_end): the shard's ids without in-edges
    int32_t lt_lo, lt_hi, st_lo, st_hi, zt_lo, zt_hi;   // task ranges executed by this launch (experiment switches)
    // split rows
    double* head_part;                       // [n_ltasks]
    double* tail_part;                       // [n_ltasks]
    unsigned int* split_cnt;                 // [n_ltasks]
    double* diff_parts;                      // [n_ltasks + n_stasks + n_ztasks + n_ltasks]
    // state
    const double* __restrict__ contrib_in;   // [N+2]
    double* contrib_out;
    const double* __restrict__ rank_in;      // [N+1]
    double* rank_out;
    double d, base;
    // loop control.  loop == 0: a free-standing sweep (timing aid), cnt/done untouched.  diff_slot != null: sharded
    // run, the shard's diff goes there and the loop control is advanced by k_ctrl_combine once every shard's
    // diff is known
    gmc_pr_ctrl* ctrl;
    double* diff_slot;
    int32_t loop;
    // peer-memory exchange (sharded, ranks mapped with CUDA IPC): every new contribution is also stored into the
    // peers' contrib_out over NVLink from the row epilogue; the last CTA publishes diff + the sweep's sequence number
    int32_t npeers;
    unsigned int seq;
    unsigned int* my_flag;
    double* peer_out[MAX_PEERS];
    double* peer_diff[MAX_PEERS];
    unsigned int* peer_flag[MAX_PEERS];
    double e;
    int32_t max_iter;
    int32_t mode;                            // experiment switches (GMC_EXPERIMENT builds only)
};

struct gmc_pr_plan {
    int world = 1, shard = 0;
    bool owns_state = true;
    // global layout facts
    int32_t N = 0, M = 0, g_n_long = 0, n_pos = 0, hot_n = 1, g_nz_slices = 0, g_z0 = 0;
    // this shard: long rows [jl0, jl0+n_long), slices [s0, s0+n_slices), zero ids [z0, z_end)
    int32_t lt = LT, lt_shift = 11, st_cost = ST_COST, slice_ovh = SLICE_OVH;     // task sizes of this plan
    int32_t jl0 = 0, n_long = 0, s0 = 0, n_slices = 0, EL = 0, EN = 0, n_ltasks = 0, n_stasks = 0, n_ztasks = 0, z0 = 0, z_end = 0;
    std::vector<int64_t> seg_off, seg_len;    // [world][3] owned id ranges of every rank (long, short, zero)
    // state block (one allocation, so that one CUDA IPC handle maps it into the peers):
    //   contrib[0] [stride] | contrib[1] [stride] | diffs [2][SYNC_RANKS] doubles | flags [SYNC_RANKS] u32
    double* block = nullptr;
    double* diffs = nullptr;                  // per-shard diffs, double-buffered by sweep parity
    unsigned int* flags = nullptr;            // flags[r] = sequence number of rank r's last completed sweep
    bool p2p = false;                         // peers mapped: contributions are pushed from the sweep epilogue
    void* peer_base[MAX_PEERS + 1] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
    unsigned int seq = 0;                     // sweeps issued so far (identical on every rank)
    size_t block_bytes = 0;
    bool block_ipc = false;                   // block came from gmc_malloc_ipc (exportable)
    bool block_cached = false;                // block (and its peer mappings) taken over from the retired-block cache
    unsigned long long gen = 0;               // allocation number of the block inside this process
    unsigned long long peer_gen[MAX_PEERS + 1] = {0, 0, 0, 0, 0, 0, 0, 0};   // the peers' allocation numbers when they were mapped
    int32_t *new_of_old = nullptr, *old_of_new = nullptr, *E = nullptr, *lbeg = nullptr, *sbeg = nullptr;
    int32_t *ltask_row = nullptr, *stask_slice = nullptr, *outdeg = nullptr;
    double *head_part = nullptr, *tail_part = nullptr, *diff_parts = nullptr;
    unsigned int* split_cnt = nullptr;
    double* contrib[2] = {nullptr, nullptr};
    double* rank[2] = {nullptr, nullptr};
    gmc_pr_ctrl* ctrl = nullptr;
    int grid = 0;
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
__device__ __forceinline__ double gather(const double* hot, const SweepArgs& A, int32_t w) {
#ifdef GMC_EXPERIMENT
    if (A.mode & 1) w &= 16383;
#endif
#ifdef GMC_COLD_LDG
    return (w < A.hot_n) ? hot[w] : __ldg(A.contrib_in + w);
#else
    return (w < A.hot_n) ? hot[w] : __ldcg(A.contrib_in + w);    // cold gathers bypass L1 (see HOT_CAP)
#endif
}

// new contribution of vertex id: local copy + one copy per peer (fire-and-forget NVLink stores, coalesced across lanes)
__device__ __forceinline__ void put_contrib(const SweepArgs& A, int32_t id, double c) {
    A.contrib_out[id] = c;
#pragma unroll
    for (int p = 0; p < MAX_PEERS; p++)
        if (p < A.npeers) A.peer_out[p][id] = c;
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
__device__ __forceinline__ double finish_row(const SweepArgs& A, int32_t id, double S) {
    const double val = __dadd_rn(A.base, __dmul_rn(A.d, S));
    const double old = __ldcs(A.rank_in + id);
    __stcs(A.rank_out + id, val);
    const int32_t od = __ldcs(A.outdeg + id);
    put_contrib(A, id, od > 0 ? val / (double)od : 0.0);
    return fabs(__dsub_rn(val, old));
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

// the last task to deliver a piece of a cut row adds the pieces in task order and finishes the row
__device__ __forceinline__ void split_arrive(const SweepArgs& A, int lane, int32_t t1, int32_t t2, int32_t row, double piece,
                                             double* slot, int32_t n_tasks_total) {
    int last = 0;
    if (lane == 0) {
        *slot = piece;
        __threadfence();
        unsigned old = atomicAdd(&A.split_cnt[t1], 1u);
        last = (old == (unsigned)(t2 - t1));
    }
    last = __shfl_sync(FULL, last, 0);
    if (!last) return;
    __threadfence();
    double v = 0.0;
    for (int32_t p = t1 + lane; p <= t2; p += 32) v += (p == t1) ? __ldcg(A.tail_part + p) : __ldcg(A.head_part + p);
    const double S = butterfly(v);
    if (lane == 0) {
        A.diff_parts[n_tasks_total + t1] = finish_row(A, A.long_id0 + row, S);
        A.split_cnt[t1] = 0u;
    }
}

__device__ __forceinline__ void long_task(const SweepArgs& A, const double* hot, int lane, int32_t t) {
    const int32_t x0 = t << A.lt_shift;
    const int32_t x1 = min(x0 + A.LT, A.EL);
    const int32_t a = __ldg(A.ltask_row + t);                     // row (0-based; id = row + 1) holding entry x0
    const int32_t e = (a + lane < A.n_long) ? __ldg(A.lbeg + a + lane + 1) : INT32_MAX;   // end of row a+lane
    const int32_t rb0 = __ldg(A.lbeg + a);
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
        bool cross = in && o >= row_end;
#ifdef GMC_EXPERIMENT
        if ((A.mode & 4) && acc_row >= 0) cross = false;
#endif
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
        const double v0 = gather(hot, A, c0.x), v1 = gather(hot, A, c0.y), v2 = gather(hot, A, c0.z), v3 = gather(hot, A, c0.w);
        const double v4 = gather(hot, A, c1.x), v5 = gather(hot, A, c1.y), v6 = gather(hot, A, c1.z), v7 = gather(hot, A, c1.w);
        step(x0 + o0, o0 < span, (v0 + v1) + (v2 + v3));
        step(x0 + o1, o1 < span, (v4 + v5) + (v6 + v7));
    }
    hand_over(acc_row >= 0);
    // rows that lie completely inside the task: one lane per row, coalesced epilogue
    const int k_first = head_split ? 1 : 0;
    const int k_last = tail_split ? kl - 1 : kl;
    double diff = 0.0;
    if (lane >= k_first && lane <= k_last) diff = finish_row(A, A.long_id0 + a + lane, my_sum);
    diff = butterfly(diff);
    if (lane == 0) A.diff_parts[t] = diff;
    const int32_t split_base = A.n_ltasks + A.n_stasks + A.n_ztasks;
    if (head_split) {
        const int32_t e0 = __shfl_sync(FULL, e, 0);
        split_arrive(A, lane, rb0 >> A.lt_shift, (e0 - 1) >> A.lt_shift, a, head_acc, A.head_part + t, split_base);
    }
    if (tail_split && !(kl == 0 && head_split))
        split_arrive(A, lane, t, (e_kl - 1) >> A.lt_shift, a + kl, tail_acc, A.tail_part + t, split_base);
}

// rows with 1..S_MAX in-edges: one lane per row, terms added strictly left to right.  The loop runs over
// 8-column chunks; the indices of the next chunk (of this slice or the next one) are always in flight.
__device__ __forceinline__ void short_task(const SweepArgs& A, const double* hot, int lane, int32_t u) {
    const int32_t s_lo = __ldg(A.stask_slice + u), s_hi = __ldg(A.stask_slice + u + 1);
    const int ns = s_hi - s_lo;                                  // <= 31
    const int32_t sb = (lane <= ns) ? __ldg(A.sbeg + s_lo + lane) : 0;
    double diff = 0.0;
    if (ns <= 0) { if (lane == 0) A.diff_parts[A.n_ltasks + u] = 0.0; return; }
    int q = 0, j = 0;
    int32_t base = __shfl_sync(FULL, sb, 0);
    int L = (__shfl_sync(FULL, sb, 1) - base) >> 5;
    int32_t w[8];
#pragma unroll
    for (int k = 0; k < 8; k++) w[k] = (k < L) ? __ldcs(A.E + base + lane + k * 32) : 0;
    double S = 0.0, old = 0.0;
    int32_t od = 0;
    while (q < ns) {
        const int32_t id = A.short_id0 + (s_lo + q) * 32 + lane;
        if (j == 0 && id <= A.N) { old = __ldcs(A.rank_in + id); od = __ldcs(A.outdeg + id); }
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
#pragma unroll
        for (int k = 0; k < 8; k++) wn[k] = (nj + k < nL) ? __ldcs(A.E + nbase + lane + (nj + k) * 32) : 0;
        double v[8];
#pragma unroll
        for (int k = 0; k < 8; k++) v[k] = gather(hot, A, w[k]);
#pragma unroll
        for (int k = 0; k < 8; k++) S = S + v[k];                 // left to right: the reference's order (+0.0 for padding)
        if (nj == 0) {                                            // slice finished
            if (id <= A.N) {
                const double val = __dadd_rn(A.base, __dmul_rn(A.d, S));
                __stcs(A.rank_out + id, val);
                put_contrib(A, id, od > 0 ? val / (double)od : 0.0);
                diff += fabs(__dsub_rn(val, old));
            }
            S = 0.0;
        }
#pragma unroll
        for (int k = 0; k < 8; k++) w[k] = wn[k];
        q = nq; j = nj; base = nbase; L = nL;
    }
    diff = butterfly(diff);
    if (lane == 0) A.diff_parts[A.n_ltasks + u] = diff;
}

// rows without in-edges (the tail of the id space): val = (1-d)/N, a pure streaming update
__device__ __forceinline__ void zero_task(const SweepArgs& A, int lane, int32_t z) {
    const int32_t id0 = A.z0 + z * ZT;
    double diff = 0.0;
#pragma unroll 1
    for (int b = 0; b < ZT; b += 256) {
        double old[8];
        int32_t od[8];
#pragma unroll
        for (int k = 0; k < 8; k++) {
            const int32_t id = id0 + b + k * 32 + lane;
            old[k] = id < A.z_end ? __ldcs(A.rank_in + id) : 0.0;
            od[k] = id < A.z_end ? __ldcs(A.outdeg + id) : 0;
        }
#pragma unroll
        for (int k = 0; k < 8; k++) {
            const int32_t id = id0 + b + k * 32 + lane;
            if (id < A.z_end) {
                const double val = A.base;                       // (1-d)/N + d*0
                __stcs(A.rank_out + id, val);
                put_contrib(A, id, od[k] > 0 ? val / (double)od[k] : 0.0);
                diff += fabs(__dsub_rn(val, old[k]));
            }
        }
    }
    diff = butterfly(diff);
    if (lane == 0) A.diff_parts[A.n_ltasks + A.n_stasks + z] = diff;
}

__global__ void __launch_bounds__(PR_THREADS, PR_CTAS) pr_sweep(const SweepArgs A) {
    extern __shared__ __align__(16) double hot[];
    __shared__ __align__(8) unsigned long long bar;
    __shared__ double s_red[PR_WARPS];
    __shared__ int s_last;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;

    if (tid == 0) mbar_init(&bar, 1);
    __syncthreads();
    // programmatic dependent launch: the CTAs of sweep k+1 take their SMs as the CTAs of sweep k leave them and wait
    // here (everything above touches no global memory) until sweep k has completed and its stores are visible
    asm volatile("griddepcontrol.wait;" ::: "memory");
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    if (A.loop && A.ctrl->done) return;

    // stage the hot block: contrib_in[0, hot_n) -> shared memory, one TMA bulk copy stream
    if (tid == 0) {
        const unsigned bytes = (unsigned)(((A.hot_n + 1) & ~1) * 8);     // multiple of 16
        mbar_expect_tx(&bar, bytes);
        const char* src = reinterpret_cast<const char*>(A.contrib_in);
        char* dst = reinterpret_cast<char*>(hot);
        // every CTA copies the same bytes: start at a different 16 KB piece per CTA so they do not queue on the same L2 lines
        const unsigned pieces = (bytes + 16383u) / 16384u;
        for (unsigned k = 0; k < pieces; k++) {
            const unsigned off = ((k + blockIdx.x) % pieces) * 16384u;
            tma_bulk_g2s(dst + off, src + off, min(16384u, bytes - off), &bar);
        }
    }
    mbar_wait(&bar, 0);

    // tasks, long-row tasks first: a warp's first task is static, the following ones come from a device-wide
    // counter (task costs vary with the row mix; a static round-robin left a tail of ~10 % at the final barrier)
    const int32_t n_l = A.lt_hi - A.lt_lo, n_s = A.st_hi - A.st_lo, n_z = A.zt_hi - A.zt_lo;
    const int32_t total = n_l + n_s + n_z, nwarps = gridDim.x * PR_WARPS;
    int32_t u = blockIdx.x * PR_WARPS + warp;
    while (u < total) {
        if (u < n_l) long_task(A, hot, lane, A.lt_lo + u);
        else if (u < n_l + n_s) short_task(A, hot, lane, A.st_lo + (u - n_l));
        else zero_task(A, lane, A.zt_lo + (u - n_l - n_s));
        if (lane == 0) u = nwarps + (int32_t)atomicAdd(&A.ctrl->next, 1u);
        u = __shfl_sync(FULL, u, 0);
    }

    // the last CTA adds the per-task diff partials in a fixed order and advances the loop control
    __syncthreads();
    if (tid == 0) {
        if (A.npeers > 0) __threadfence_system(); else __threadfence();
        s_last = (atomicAdd(&A.ctrl->arrived, 1u) == gridDim.x - 1);
    }
    __syncthreads();
    if (!s_last) return;
    __threadfence();
    const int32_t nparts = 2 * A.n_ltasks + A.n_stasks + A.n_ztasks;
    double v = 0.0;
    for (int32_t i = tid; i < nparts; i += PR_THREADS) v += __ldcg(A.diff_parts + i);
    v = butterfly(v);
    if (lane == 0) s_red[warp] = v;
    __syncthreads();
    if (warp == 0) {
        double t = butterfly(lane < PR_WARPS ? s_red[lane] : 0.0);
        if (lane == 0) { A.ctrl->arrived = 0u; A.ctrl->next = 0u; }
        if (lane == 0 && A.diff_slot) {
            *A.diff_slot = t;
            for (int p = 0; p < A.npeers; p++) *A.peer_diff[p] = t;
        } else if (lane == 0 && A.loop) {
            const int32_t cnt = A.ctrl->cnt + 1;
            A.ctrl->diff = t;
            A.ctrl->cnt = cnt;
            A.ctrl->done = !((t > A.e) && (cnt < A.max_iter));       // } While ((diff > e) && (cnt < max))
        }
        if (lane == 0 && A.my_flag) {      // every store of this sweep (all CTAs) is ordered before the flags
            __threadfence_system();
            *A.my_flag = A.seq;
            for (int p = 0; p < A.npeers; p++) st_release_sys(A.peer_flag[p], A.seq);
        }
    }
}

// ---------------------------------------------------------------------------------------------
// plan kernels
// ---------------------------------------------------------------------------------------------
__global__ void k_degrees(const int32_t* __restrict__ begin, const int32_t* __restrict__ r_begin, int32_t n,
                          int32_t* __restrict__ outdeg, int32_t* __restrict__ indeg, uint32_t* __restrict__ key_a,
                          int32_t* __restrict__ ids, int32_t* __restrict__ is_long, int32_t* __restrict__ is_pos, int32_t s_max) {
    int32_t v = blockIdx.x * blockDim.x + threadIdx.x;
    if (v >= n) return;
    const int32_t od = begin[v + 1] - begin[v], id = r_begin[v + 1] - r_begin[v];
    outdeg[v] = od;
    indeg[v] = id;
    const bool lg = id > s_max;
    is_long[v] = lg;
    is_pos[v] = id > 0;
    key_a[v] = lg ? (uint32_t)(0x7fffffff - od) : 0xffffffffu;    // long rows by out-degree descending, short rows last
    ids[v] = v;
}

// hot rank of the first `hot_ids` vertices of the out-degree order, then the final ordering key
__global__ void k_hot_rank(const int32_t* __restrict__ sorted_ids, int32_t n, int32_t hot_ids, int32_t* __restrict__ hot_rank) {
    int32_t p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p < n) hot_rank[sorted_ids[p]] = p < hot_ids ? p : -1;
}
__global__ void k_order_key(const int32_t* __restrict__ indeg, const int32_t* __restrict__ hot_rank, int32_t n,
                            unsigned long long* __restrict__ key, int32_t* __restrict__ ids, int32_t s_max) {
    int32_t v = blockIdx.x * blockDim.x + threadIdx.x;
    if (v >= n) return;
    const int32_t hr = hot_rank[v], id = indeg[v];
    unsigned long long k;
    if (hr >= 0) k = (unsigned long long)hr;
    else k = ((unsigned long long)(id > s_max ? 1 : 2) << 32) | (unsigned)(0x7fffffff - id);
    key[v] = k;
    ids[v] = v;
}
__global__ void k_relabel(const int32_t* __restrict__ sorted_ids, int32_t n, int32_t* __restrict__ old_of_new,
                          int32_t* __restrict__ new_of_old) {
    int32_t p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p == 0) old_of_new[0] = -1;
    if (p >= n) return;
    const int32_t v = sorted_ids[p];
    old_of_new[p + 1] = v;
    new_of_old[v] = p + 1;
}
// padded long-row lengths (n_long values + a trailing 0), slice lengths (n_slices values + trailing 0), outdeg by new id
__global__ void k_lengths(const int32_t* __restrict__ old_of_new, const int32_t* __restrict__ indeg, const int32_t* __restrict__ outdeg_old,
                          int32_t n, int32_t n_long, int32_t n_slices, int32_t* __restrict__ lbeg, int32_t* __restrict__ sbeg,
                          int32_t* __restrict__ scost, int32_t* __restrict__ outdeg_new, int32_t slice_ovh) {
    int32_t j = blockIdx.x * blockDim.x + threadIdx.x;     // 0-based row, id = j+1
    if (j == 0) { lbeg[n_long] = 0; sbeg[n_slices] = 0; scost[n_slices] = 0; outdeg_new[0] = 0; }
    if (j >= n) return;
    const int32_t v = old_of_new[j + 1];
    outdeg_new[j + 1] = outdeg_old[v];
    const int32_t len = indeg[v];
    if (j < n_long) lbeg[j] = (len + 3) & ~3;
    else if (((j - n_long) & 31) == 0) {
        const int32_t s = (j - n_long) >> 5;
        sbeg[s] = 32 * len;                                 // first row of the slice is its longest
        scost[s] = len > 0 ? 32 * len + slice_ovh : 0;          // empty slices belong to the zero tasks
    }
}
// dst[i] = src[lo + i] - src[lo] + add, i in [0, cnt): the shard's slice of a global offset array, rebased
__global__ void k_rebase(const int32_t* __restrict__ src, int32_t lo, int32_t cnt, int32_t add, int32_t* __restrict__ dst) {
    int32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < cnt) dst[i] = src[lo + i] - src[lo] + add;
}
// hint[b] = row holding reverse edge b * FILL_TILE (b <= ceil(m / FILL_TILE): the last one is n): narrows the per-edge
// row search of k_fill_entries from log2(N) global steps to the few rows one tile of edges spans
constexpr int FILL_TILE = 2048;
__global__ void k_row_hints(const int32_t* __restrict__ r_begin, int32_t n, int32_t m, int32_t n_hints, int32_t* __restrict__ hint) {
    int32_t b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= n_hints) return;
    const int64_t e64 = (int64_t)b * FILL_TILE;
    if (e64 >= m) { hint[b] = n; return; }
    const int32_t eidx = (int32_t)e64;
    int32_t lo = 0, hi = n;   // r_begin[lo] <= eidx < r_begin[hi]
    while (hi - lo > 1) {
        const int32_t mid = lo + ((hi - lo) >> 1);
        if (r_begin[mid] <= eidx) lo = mid; else hi = mid;
    }
    hint[b] = lo;
}
// scatter the reverse edges [e_lo, e_hi) whose destination row belongs to the shard into E, keeping the position inside
// the row (= the reference's summation order).  jl0/nl: the shard's long rows, s0/ns: its slices.
__global__ void k_fill_entries(const int32_t* __restrict__ r_begin, const int32_t* __restrict__ r_node_idx, int32_t n, int32_t e_lo, int32_t e_hi,
                               const int32_t* __restrict__ hint, const int32_t* __restrict__ new_of_old, int32_t g_n_long, int32_t jl0,
                               int32_t nl, int32_t s0, int32_t ns, const int32_t* __restrict__ lbeg, const int32_t* __restrict__ sbeg,
                               int32_t* __restrict__ E) {
    const int64_t e64 = (int64_t)e_lo + (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (e64 >= e_hi) return;
    const int32_t eidx = (int32_t)e64;
    int32_t lo = hint[eidx / FILL_TILE], hi = min(hint[eidx / FILL_TILE + 1] + 1, n);   // r_begin[lo] <= eidx < r_begin[hi]
    while (hi - lo > 1) {
        const int32_t mid = lo + ((hi - lo) >> 1);
        if (r_begin[mid] <= eidx) lo = mid; else hi = mid;
    }
    const int32_t k = eidx - r_begin[lo];
    const int32_t j = new_of_old[lo] - 1;
    if (j < g_n_long) {
        const int32_t jj = j - jl0;
        if (jj < 0 || jj >= nl) return;
        E[lbeg[jj] + k] = new_of_old[ld_stream(r_node_idx + eidx)];
    } else {
        const int32_t q = j - g_n_long;
        const int32_t sl = (q >> 5) - s0;
        if (sl < 0 || sl >= ns) return;
        E[sbeg[sl] + k * 32 + (q & 31)] = new_of_old[ld_stream(r_node_idx + eidx)];
    }
}
// Row-driven fill (every CSR array already resident): work proportional to the SHARD's entries instead of a scan over all
// M edges per rank.  Long rows: one warp per row, coalesced copy through the relabelling.  Slices: one warp per slice,
// lane = row of the slice (rows of a slice have similar lengths), column-major stores are coalesced.
__global__ void k_fill_long_rows(const int32_t* __restrict__ r_begin, const int32_t* __restrict__ r_node_idx, const int32_t* __restrict__ old_of_new,
                                 const int32_t* __restrict__ new_of_old, int32_t jl0, int32_t nl, int32_t el, const int32_t* __restrict__ lbeg,
                                 int32_t* __restrict__ E) {
    // one thread per entry slot of the long section (hub rows have 1e5 entries: a warp per row would leave a long tail);
    // the row is found by bisection over the shard's nl + 1 row offsets (a few thousand to a million: L1/L2 resident)
    const int64_t x64 = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (x64 >= el) return;
    const int32_t x = (int32_t)x64;
    int32_t lo = 0, hi = nl;             // lbeg[lo] <= x < lbeg[hi]
    while (hi - lo > 1) {
        const int32_t mid = lo + ((hi - lo) >> 1);
        if (__ldg(lbeg + mid) <= x) lo = mid; else hi = mid;
    }
    const int32_t v = old_of_new[jl0 + lo + 1];
    const int32_t rb = r_begin[v], len = r_begin[v + 1] - rb, k = x - __ldg(lbeg + lo);
    if (k < len) E[x] = new_of_old[ld_stream(r_node_idx + rb + k)];        // k >= len: padding of the row to 4 entries (stays 0)
}
__global__ void k_fill_slices(const int32_t* __restrict__ r_begin, const int32_t* __restrict__ r_node_idx, const int32_t* __restrict__ old_of_new,
                              const int32_t* __restrict__ new_of_old, int32_t n, int32_t g_n_long, int32_t s0, int32_t ns,
                              const int32_t* __restrict__ sbeg, int32_t* __restrict__ E) {
    const int32_t sl = (int32_t)(((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5);
    const int lane = threadIdx.x & 31;
    if (sl >= ns) return;
    const int64_t id = (int64_t)g_n_long + 1 + ((int64_t)(s0 + sl) << 5) + lane;
    if (id > n) return;
    const int32_t v = old_of_new[id];
    const int32_t rb = r_begin[v], len = r_begin[v + 1] - rb, base = sbeg[sl] + lane;
    for (int32_t k = 0; k < len; k++) E[base + (k << 5)] = new_of_old[__ldg(r_node_idx + rb + k)];
}
// first[t] = row holding entry t*LT: last j with lbeg[j] <= target   (task cut points)
__global__ void k_ltask_rows(const int32_t* __restrict__ lbeg, int32_t n_long, int32_t n_tasks, int32_t lt, int32_t* __restrict__ first) {
    int32_t t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t > n_tasks) return;
    if (t == n_tasks) { first[t] = n_long; return; }
    const int32_t target = t * lt;
    int32_t lo = 0, hi = n_long;             // lbeg[lo] <= target < lbeg[hi]
    while (hi - lo > 1) {
        const int32_t mid = lo + ((hi - lo) >> 1);
        if (lbeg[mid] <= target) lo = mid; else hi = mid;
    }
    first[t] = lo;
}
// short tasks: the first n_big ones hold st_cost cost units, the rest (the tail of the sweep: the last tasks handed out)
// a quarter of that, so that the warps finish closer together
__global__ void k_stask_slices(const int32_t* __restrict__ cpos, int32_t n_slices, int32_t n_tasks, int32_t st_cost, int32_t n_big, int32_t tail_shift,
                               int32_t* __restrict__ first) {
    int32_t u = blockIdx.x * blockDim.x + threadIdx.x;
    if (u > n_tasks) return;
    if (u == n_tasks) { first[u] = n_slices; return; }
    const int32_t target = u <= n_big ? u * st_cost : n_big * st_cost + (u - n_big) * (st_cost >> tail_shift);      // first slice with cpos >= target
    int32_t lo = 0, hi = n_slices;
    while (lo < hi) {
        const int32_t mid = lo + ((hi - lo) >> 1);
        if (cpos[mid] >= target) hi = mid; else lo = mid + 1;
    }
    first[u] = lo;
}
__global__ void k_init(SweepArgs A, double* __restrict__ contrib0, double* __restrict__ contrib1, double* __restrict__ rank0) {
    int32_t id = blockIdx.x * blockDim.x + threadIdx.x;
    if (id > A.N + 1) return;
    if (id == 0 || id == A.N + 1) { contrib0[id] = 0.0; contrib1[id] = 0.0; if (id == 0) rank0[0] = 0.0; return; }
    const double v = 1 / (double)A.N;                       // G.pg_rank = 1 / N  (pagerank.gm:6-7)
    rank0[id] = v;
    const int32_t od = A.outdeg[id];
    contrib0[id] = od > 0 ? v / (double)od : 0.0;
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

template <typename K>
int sort_pairs(K* keys_in, K* keys_out, int32_t* vals_in, int32_t* vals_out, int32_t n, int end_bit, cudaStream_t st) {
    void* tmp = nullptr;
    size_t bytes = 0;
    cub::DeviceRadixSort::SortPairs(nullptr, bytes, keys_in, keys_out, vals_in, vals_out, (int64_t)n, 0, end_bit, st);
    GMC_TRY(gmc_malloc(&tmp, bytes));
    cudaError_t e = cub::DeviceRadixSort::SortPairs(tmp, bytes, keys_in, keys_out, vals_in, vals_out, (int64_t)n, 0, end_bit, st);
    cudaStreamSynchronize(st);
    cudaFree(tmp);
    if (e != cudaSuccess) { set_error("plan sort: %s", cudaGetErrorString(e)); return GMC_ERR_CUDA; }
    return GMC_OK;
}

// One retired state block per process, kept allocated AND mapped by the peers.  Exporting a block with CUDA IPC costs
// a plain cudaMalloc plus one cudaIpcOpenMemHandle per peer (~20 ms each for a 666 MB block: 147 ms of plan time and 44 ms
// of teardown per call at 8 ranks, profiles/r1x_bench_8gpu_tw.json), so a caller that runs pagerank() on graph after graph of
// the same size -- every step of the e2e loop -- pays it once.  Reuse is validated collectively (allocation numbers of
// all ranks must be the ones recorded when the mappings were made), see plan_map_peers.
struct SymCache {
    double* block = nullptr;
    size_t bytes = 0;
    int world = 0, rank = 0;
    bool p2p = false;
    void* peer_base[MAX_PEERS + 1] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
    unsigned long long gen = 0, peer_gen[MAX_PEERS + 1] = {0, 0, 0, 0, 0, 0, 0, 0};
};
SymCache g_sym;
unsigned long long g_block_gen = 0;

void sym_cache_release() {
    if (!g_sym.block) return;
    if (g_sym.p2p) gmc_comm_ipc_close_all(g_sym.world, g_sym.rank, g_sym.peer_base);
    cudaFree(g_sym.block);
    g_sym = SymCache();
}

void free_plan(gmc_pr_plan* p) {
    if (!p) return;
    if (p->owns_state && p->block_ipc && p->block && !getenv("GMC_NO_BLOCK_CACHE")) {      // retire the block instead of unmapping it
        sym_cache_release();
        g_sym.block = p->block; g_sym.bytes = p->block_bytes; g_sym.world = p->world; g_sym.rank = p->shard; g_sym.p2p = p->p2p;
        g_sym.gen = p->gen;
        for (int r = 0; r <= MAX_PEERS; r++) { g_sym.peer_base[r] = p->peer_base[r]; g_sym.peer_gen[r] = p->peer_gen[r]; }
        p->block = nullptr; p->p2p = false;
    }
    cudaFree(p->new_of_old); cudaFree(p->old_of_new); cudaFree(p->E); cudaFree(p->lbeg); cudaFree(p->sbeg);
    cudaFree(p->ltask_row); cudaFree(p->stask_slice); cudaFree(p->outdeg);
    cudaFree(p->head_part); cudaFree(p->tail_part); cudaFree(p->diff_parts); cudaFree(p->split_cnt);
    if (p->owns_state) {
        if (p->p2p) gmc_comm_ipc_close_all(p->world, p->shard, p->peer_base);
        cudaFree(p->block); cudaFree(p->rank[0]); cudaFree(p->rank[1]); cudaFree(p->ctrl);
    }
    delete p;
}

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
    int32_t *outdeg_old = nullptr, *indeg = nullptr, *ids = nullptr, *ids2 = nullptr, *is_long = nullptr, *is_pos = nullptr, *hot_rank = nullptr;
    int32_t *g_lbeg = nullptr, *g_sbeg = nullptr, *g_scost = nullptr, *scost = nullptr;
    uint32_t *key_a = nullptr, *key_a2 = nullptr;
    unsigned long long *key_b = nullptr, *key_b2 = nullptr;
    int32_t* tot = nullptr;
    int32_t* hints = nullptr;
    const size_t nn = (size_t)n;
    const unsigned gb = (unsigned)div_up(n, 256);
    int32_t h[2] = {0, 0};
    int32_t g_n_slices = 0, jl1 = 0, s1 = 0, cost_total = 0, s_max = S_MAX, n_big = 0, tail_shift = 2;
    std::vector<int32_t> h_lbeg, h_sbeg, h_scost, cutsL((size_t)world + 1), cutsS((size_t)world + 1), cutsZ((size_t)world + 1);

    PLAN_TRY(gmc_malloc((void**)&outdeg_old, 4 * nn)); PLAN_TRY(gmc_malloc((void**)&indeg, 4 * nn));
    PLAN_TRY(gmc_malloc((void**)&ids, 4 * nn)); PLAN_TRY(gmc_malloc((void**)&ids2, 4 * nn));
    PLAN_TRY(gmc_malloc((void**)&is_long, 4 * nn)); PLAN_TRY(gmc_malloc((void**)&hot_rank, 4 * nn));
    PLAN_TRY(gmc_malloc((void**)&is_pos, 4 * nn));
    PLAN_TRY(gmc_malloc((void**)&key_a, 4 * nn)); PLAN_TRY(gmc_malloc((void**)&key_a2, 4 * nn));
    PLAN_TRY(gmc_malloc((void**)&key_b, 8 * nn)); PLAN_TRY(gmc_malloc((void**)&key_b2, 8 * nn));
    PLAN_TRY(gmc_malloc((void**)&tot, 16));
    PLAN_TRY(gmc_malloc((void**)&p->new_of_old, 4 * nn)); PLAN_TRY(gmc_malloc((void**)&p->old_of_new, 4 * (nn + 1)));
    PLAN_TRY(gmc_malloc((void**)&p->outdeg, 4 * (nn + 1)));

    if (g->pending) PLAN_CUDA(cudaStreamWaitEvent(st, g->pending->small, 0));       // begin / r_begin have landed
    // 1. degrees, number of long rows, the hot block, the relabelling (identical on every rank)
    if (const char* sm = getenv("GMC_PR_SMAX")) { s_max = atoi(sm); if (s_max < 64) s_max = 64; if (s_max > 4096) s_max = 4096; }   // experiments (>= 64: a long task holds at most 32 rows)
    k_degrees<<<gb, 256, 0, st>>>(g->begin, g->r_begin, n, outdeg_old, indeg, key_a, ids, is_long, is_pos, s_max);
    PLAN_TRY(exclusive_scan_i32(is_long, is_long, nn, tot, st));
    PLAN_TRY(exclusive_scan_i32(is_pos, is_pos, nn, tot + 1, st));
    PLAN_CUDA(cudaMemcpyAsync(h, tot, 8, cudaMemcpyDeviceToHost, st));
    PLAN_CUDA(cudaStreamSynchronize(st));
    p->g_n_long = h[0];
    p->n_pos = h[1];
    {
        const int32_t hot_ids = p->g_n_long < HOT_CAP - 2 ? p->g_n_long : HOT_CAP - 2;
        p->hot_n = hot_ids + 1;                             // ids [0, hot_n): the dummy + the hot block
        PLAN_TRY(sort_pairs<uint32_t>(key_a, key_a2, ids, ids2, n, 32, st));
        k_hot_rank<<<gb, 256, 0, st>>>(ids2, n, hot_ids, hot_rank);
        k_order_key<<<gb, 256, 0, st>>>(indeg, hot_rank, n, key_b, ids, s_max);
        PLAN_TRY(sort_pairs<unsigned long long>(key_b, key_b2, ids, ids2, n, 34, st));
        k_relabel<<<gb, 256, 0, st>>>(ids2, n, p->old_of_new, p->new_of_old);
    }
    // 2. global row offsets of both sections, then the shard cuts (host: gmc_shard_cuts)
    g_n_slices = div_up(n - p->g_n_long, 32);
    p->g_nz_slices = p->n_pos > p->g_n_long ? div_up(p->n_pos - p->g_n_long, 32) : 0;     // slices holding a row with in-edges
    p->g_z0 = p->g_n_long + 1 + 32 * p->g_nz_slices;
    PLAN_TRY(gmc_malloc((void**)&g_lbeg, 4 * ((size_t)p->g_n_long + 1)));
    PLAN_TRY(gmc_malloc((void**)&g_sbeg, 4 * ((size_t)g_n_slices + 1)));
    PLAN_TRY(gmc_malloc((void**)&g_scost, 4 * ((size_t)g_n_slices + 1)));
    // task sizes: large tasks for large shards (a long task may hold at most 32 rows: LT <= 31 * (s_max + 1))
    if ((int64_t)m / world >= BIG_TASK_ENTRIES && (int64_t)LT_BIG <= 31LL * (s_max + 1)) {
        p->lt = LT_BIG; p->st_cost = ST_COST_BIG; p->slice_ovh = SLICE_OVH_BIG;
    }
    if (const char* big = getenv("GMC_PR_BIGTASKS")) {            // tests / experiments: force either task size
        const bool on = atoi(big) != 0 && (int64_t)LT_BIG <= 31LL * (s_max + 1);
        p->lt = on ? LT_BIG : LT; p->st_cost = on ? ST_COST_BIG : ST_COST; p->slice_ovh = on ? SLICE_OVH_BIG : SLICE_OVH;
    }
    for (p->lt_shift = 0; (1 << p->lt_shift) < p->lt; p->lt_shift++) {}
    k_lengths<<<gb, 256, 0, st>>>(p->old_of_new, indeg, outdeg_old, n, p->g_n_long, g_n_slices, g_lbeg, g_sbeg, g_scost, p->outdeg, p->slice_ovh);
    PLAN_TRY(exclusive_scan_i32(g_lbeg, g_lbeg, (size_t)p->g_n_long + 1, nullptr, st));
    PLAN_TRY(exclusive_scan_i32(g_sbeg, g_sbeg, (size_t)g_n_slices + 1, nullptr, st));
    PLAN_TRY(exclusive_scan_i32(g_scost, g_scost, (size_t)g_n_slices + 1, nullptr, st));
    h_lbeg.resize((size_t)p->g_n_long + 1); h_sbeg.resize((size_t)g_n_slices + 1); h_scost.resize((size_t)g_n_slices + 1);
    PLAN_CUDA(cudaMemcpyAsync(h_lbeg.data(), g_lbeg, 4 * h_lbeg.size(), cudaMemcpyDeviceToHost, st));
    PLAN_CUDA(cudaMemcpyAsync(h_sbeg.data(), g_sbeg, 4 * h_sbeg.size(), cudaMemcpyDeviceToHost, st));
    PLAN_CUDA(cudaMemcpyAsync(h_scost.data(), g_scost, 4 * h_scost.size(), cudaMemcpyDeviceToHost, st));
    PLAN_CUDA(cudaStreamSynchronize(st));
    if (h_lbeg[p->g_n_long] < 0 || h_sbeg[g_n_slices] < 0 || (int64_t)h_lbeg[p->g_n_long] + h_sbeg[g_n_slices] > 0x7fffff00LL) {
        rc = GMC_ERR_ARG; set_error("pagerank plan: entry array exceeds int32 offsets"); goto fail;
    }
    if (world > 1) {
        // shard cuts balance entries + shard_row_cost(world) per row: a row's epilogue (28 B of HBM traffic, a division and, with
        // mapped peers, world-1 remote stores) weighs a few entries; with entry counts alone the ranks owning the short
        // tail rows were 2x slower than the others on the TW-shaped graph (profiles/r1e_bench_8gpu_tw_unbalanced.json)
        const int row_cost = shard_row_cost(world);
        std::vector<int64_t> wl((size_t)p->g_n_long + 1), ws((size_t)p->g_nz_slices + 1);
        for (int32_t j = 0; j <= p->g_n_long; j++) wl[j] = (int64_t)h_lbeg[j] + (int64_t)j * row_cost;
        ws[0] = 0;
        for (int32_t q = 0; q < p->g_nz_slices; q++) {          // a slice is processed in 8-column chunks
            const int32_t len = (h_sbeg[q + 1] - h_sbeg[q]) >> 5;
            ws[q + 1] = ws[q] + 32LL * (((len + 7) & ~7) + row_cost);
        }
        // gmc_shard_cuts takes int32 prefixes: scale the weights down when their total does not fit (only ratios matter)
        int sh = 0;
        while (((wl[p->g_n_long] > ws[p->g_nz_slices] ? wl[p->g_n_long] : ws[p->g_nz_slices]) >> sh) > 0x7fffff00LL) sh++;
        std::vector<int32_t> cl(wl.size()), cs(ws.size());
        for (size_t j = 0; j < wl.size(); j++) cl[j] = (int32_t)(wl[j] >> sh);
        for (size_t q = 0; q < ws.size(); q++) cs[q] = (int32_t)(ws[q] >> sh);
        PLAN_TRY(gmc_shard_cuts(cl.data(), p->g_n_long, world, 1, cutsL.data()));
        PLAN_TRY(gmc_shard_cuts(cs.data(), p->g_nz_slices, world, 1, cutsS.data()));
    } else {
        PLAN_TRY(gmc_shard_cuts(h_lbeg.data(), p->g_n_long, world, 1, cutsL.data()));
        PLAN_TRY(gmc_shard_cuts(h_scost.data(), p->g_nz_slices, world, 1, cutsS.data()));
    }
    {
        const int32_t nz = n + 1 - p->g_z0 > 0 ? n + 1 - p->g_z0 : 0;           // ids [g_z0, N] have no in-edges
        const int32_t per = div_up(div_up(nz, world), ZT) * ZT;
        for (int r = 0; r <= world; r++) { int64_t c = (int64_t)p->g_z0 + (int64_t)r * per; cutsZ[r] = (int32_t)(c < n + 1 ? c : n + 1); }
        cutsZ[world] = n + 1;
    }
    p->seg_off.assign((size_t)world * 3, 0); p->seg_len.assign((size_t)world * 3, 0);
    for (int r = 0; r < world; r++) {
        p->seg_off[r * 3 + 0] = 1 + cutsL[r];                         p->seg_len[r * 3 + 0] = cutsL[r + 1] - cutsL[r];
        const int64_t sa = (int64_t)p->g_n_long + 1 + 32LL * cutsS[r], sb = (int64_t)p->g_n_long + 1 + 32LL * cutsS[r + 1];
        p->seg_off[r * 3 + 1] = sa;                                   p->seg_len[r * 3 + 1] = (sb < n + 1 ? sb : n + 1) - sa;
        p->seg_off[r * 3 + 2] = cutsZ[r];                             p->seg_len[r * 3 + 2] = cutsZ[r + 1] - cutsZ[r];
    }
    p->jl0 = cutsL[rank]; jl1 = cutsL[rank + 1]; p->n_long = jl1 - p->jl0;
    p->s0 = cutsS[rank]; s1 = cutsS[rank + 1]; p->n_slices = s1 - p->s0;
    p->z0 = cutsZ[rank]; p->z_end = cutsZ[rank + 1];
    p->n_ztasks = div_up(p->z_end - p->z0, ZT);
    p->EL = h_lbeg[jl1] - h_lbeg[p->jl0];
    p->EN = p->EL + (h_sbeg[s1] - h_sbeg[p->s0]);
    cost_total = h_scost[s1] - h_scost[p->s0];
    PLAN_TRY(gmc_malloc((void**)&p->lbeg, 4 * ((size_t)p->n_long + 1)));
    PLAN_TRY(gmc_malloc((void**)&p->sbeg, 4 * ((size_t)p->n_slices + 1)));
    PLAN_TRY(gmc_malloc((void**)&scost, 4 * ((size_t)p->n_slices + 1)));
    k_rebase<<<div_up(p->n_long + 1, 256), 256, 0, st>>>(g_lbeg, p->jl0, p->n_long + 1, 0, p->lbeg);
    k_rebase<<<div_up(p->n_slices + 1, 256), 256, 0, st>>>(g_sbeg, p->s0, p->n_slices + 1, p->EL, p->sbeg);
    k_rebase<<<div_up(p->n_slices + 1, 256), 256, 0, st>>>(g_scost, p->s0, p->n_slices + 1, 0, scost);
    // 3. the shard's entry array
    PLAN_TRY(gmc_malloc((void**)&p->E, 4 * ((size_t)p->EN + 8)));
    PLAN_CUDA(cudaMemsetAsync(p->E, 0, 4 * ((size_t)p->EN + 8), st));
    if (m > 0 && !g->pending) {
        if (p->n_long > 0 && p->EL > 0)
            k_fill_long_rows<<<div_up(p->EL, 256), 256, 0, st>>>(g->r_begin, g->r_node_idx, p->old_of_new, p->new_of_old, p->jl0, p->n_long, p->EL, p->lbeg, p->E);
        if (p->n_slices > 0)
            k_fill_slices<<<div_up(p->n_slices, 8), 256, 0, st>>>(g->r_begin, g->r_node_idx, p->old_of_new, p->new_of_old, n, p->g_n_long, p->s0,
                                                                  p->n_slices, p->sbeg, p->E);
    } else if (m > 0) {
        const int32_t n_hints = div_up(m, FILL_TILE) + 1;
        PLAN_TRY(gmc_malloc((void**)&hints, 4 * (size_t)n_hints));
        k_row_hints<<<div_up(n_hints, 256), 256, 0, st>>>(g->r_begin, n, m, n_hints, hints);
        // r_node_idx may still be arriving over PCIe (gmc_pagerank_host_csr): fill chunk by chunk as the copies land
        const gmc_pending_upload* pu = g->pending;
        const int nch = pu ? pu->nchunks : 1;
        int32_t e_lo = 0;
        for (int c = 0; c < nch; c++) {
            const int32_t e_hi = pu ? pu->end[c] : m;
            if (pu) PLAN_CUDA(cudaStreamWaitEvent(st, pu->chunk[c], 0));
            if (e_hi > e_lo)
                k_fill_entries<<<div_up(e_hi - e_lo, 256), 256, 0, st>>>(g->r_begin, g->r_node_idx, n, e_lo, e_hi, hints, p->new_of_old, p->g_n_long,
                                                                         p->jl0, p->n_long, p->s0, p->n_slices, p->lbeg, p->sbeg, p->E);
            e_lo = e_hi;
        }
    }
    // 4. tasks
    p->n_ltasks = div_up(p->EL, p->lt);
    // ~7/8 of the short-row cost in full-size tasks, the last eighth in quarter-size ones (tail of the sweep)
    {
        int big16 = 14, sh = 2;                                                        // experiments: GMC_PR_TAIL="<sixteenths in big tasks>,<shift>"
        if (const char* t = getenv("GMC_PR_TAIL")) sscanf(t, "%d,%d", &big16, &sh);
        tail_shift = sh;
        n_big = (int32_t)((int64_t)cost_total * big16 / 16 / p->st_cost);
    }
    if (getenv("GMC_PR_NO_SMALL_TAIL")) n_big = div_up(cost_total, p->st_cost);       // experiments
    p->n_stasks = n_big + div_up(cost_total > n_big * p->st_cost ? cost_total - n_big * p->st_cost : 0, p->st_cost >> tail_shift);
    PLAN_TRY(gmc_malloc((void**)&p->ltask_row, 4 * ((size_t)p->n_ltasks + 1)));
    PLAN_TRY(gmc_malloc((void**)&p->stask_slice, 4 * ((size_t)p->n_stasks + 1)));
    k_ltask_rows<<<div_up(p->n_ltasks + 1, 256), 256, 0, st>>>(p->lbeg, p->n_long, p->n_ltasks, p->lt, p->ltask_row);
    k_stask_slices<<<div_up(p->n_stasks + 1, 256), 256, 0, st>>>(scost, p->n_slices, p->n_stasks, p->st_cost, n_big, tail_shift, p->stask_slice);
    {
        const size_t nl = (size_t)(p->n_ltasks ? p->n_ltasks : 1), nparts = 2 * nl + (size_t)p->n_stasks + (size_t)p->n_ztasks + 1;
        PLAN_TRY(gmc_malloc((void**)&p->head_part, 8 * nl)); PLAN_TRY(gmc_malloc((void**)&p->tail_part, 8 * nl));
        PLAN_TRY(gmc_malloc((void**)&p->split_cnt, 4 * nl)); PLAN_TRY(gmc_malloc((void**)&p->diff_parts, 8 * nparts));
        cudaMemsetAsync(p->head_part, 0, 8 * nl, st); cudaMemsetAsync(p->tail_part, 0, 8 * nl, st);
        cudaMemsetAsync(p->split_cnt, 0, 4 * nl, st); cudaMemsetAsync(p->diff_parts, 0, 8 * nparts, st);
    }
    // 5. state (replicated on every rank: the gather needs the whole contribution vector)
    if (state_of) {
        for (int k = 0; k < 2; k++) { p->contrib[k] = state_of->contrib[k]; p->rank[k] = state_of->rank[k]; }
        p->ctrl = state_of->ctrl; p->diffs = state_of->diffs; p->flags = state_of->flags;
    } else {
        if (world > SYNC_RANKS) { rc = GMC_ERR_ARG; set_error("pagerank: at most %d shards", SYNC_RANKS); goto fail; }
        const size_t stride = state_stride(n), sync_doubles = 2 * SYNC_RANKS + SYNC_RANKS / 2;
        p->block_bytes = 8 * (2 * stride + sync_doubles);
        if (world > 1 && g_sym.block && g_sym.bytes == p->block_bytes && g_sym.world == world && g_sym.rank == rank) {
            p->block = g_sym.block; p->block_ipc = true; p->block_cached = true; p->gen = g_sym.gen;
            for (int r = 0; r <= MAX_PEERS; r++) { p->peer_base[r] = g_sym.p2p ? g_sym.peer_base[r] : nullptr; p->peer_gen[r] = g_sym.peer_gen[r]; }
            p->peer_base[rank] = nullptr;                 // plan_map_peers validates the mappings before they are used
            g_sym = SymCache();
        } else if (world > 1) {
            PLAN_TRY(gmc_malloc_ipc((void**)&p->block, p->block_bytes));   // exported to the peers
            p->block_ipc = true; p->gen = ++g_block_gen;
        } else PLAN_TRY(gmc_malloc((void**)&p->block, p->block_bytes));
        p->contrib[0] = p->block; p->contrib[1] = p->block + stride;
        p->diffs = p->block + 2 * stride;
        p->flags = reinterpret_cast<unsigned int*>(p->block + 2 * stride + 2 * SYNC_RANKS);
        PLAN_CUDA(cudaMemsetAsync(p->diffs, 0, 8 * sync_doubles, st));
        for (int k = 0; k < 2; k++) PLAN_TRY(gmc_malloc((void**)&p->rank[k], 8 * (nn + 4)));
        PLAN_TRY(gmc_malloc((void**)&p->ctrl, sizeof(gmc_pr_ctrl)));
    }
    p->grid = sm_count() * PR_CTAS;
    p->smem = (size_t)HOT_CAP * 8;
    PLAN_CUDA(cudaFuncSetAttribute(pr_sweep, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)p->smem));
    PLAN_CUDA(cudaStreamSynchronize(st));
fail:
    cudaFree(outdeg_old); cudaFree(indeg); cudaFree(ids); cudaFree(ids2); cudaFree(is_long); cudaFree(is_pos); cudaFree(hot_rank); cudaFree(scost);
    cudaFree(g_lbeg); cudaFree(g_sbeg); cudaFree(g_scost);
    cudaFree(key_a); cudaFree(key_a2); cudaFree(key_b); cudaFree(key_b2); cudaFree(tot); cudaFree(hints);
    if (rc != GMC_OK) { free_plan(p); return rc; }
    *out = p;
    return GMC_OK;
}

// sweep: index of the sweep inside the call (0-based).  From the third sweep on the rows without in-edges are
// skipped: their value is the constant (1-d)/N, both ping-pong buffers hold it (and its contribution) after two
// sweeps, and their |val - old| is exactly 0 -- the stored diff partials of sweep 1 stay valid.
SweepArgs make_args(const gmc_pr_plan* p, double d, int sweep, gmc_pr_ctrl* ctrl, int loop, double e, int32_t max_iter) {
    const int parity = sweep & 1;
    SweepArgs A;
    A.E = p->E; A.lbeg = p->lbeg; A.sbeg = p->sbeg; A.ltask_row = p->ltask_row; A.stask_slice = p->stask_slice; A.outdeg = p->outdeg;
    A.N = p->N; A.n_long = p->n_long; A.hot_n = p->hot_n; A.EL = p->EL; A.LT = p->lt; A.lt_shift = p->lt_shift; A.n_ltasks = p->n_ltasks; A.n_stasks = p->n_stasks;
    A.long_id0 = p->jl0 + 1; A.short_id0 = p->g_n_long + 1 + 32 * p->s0;
    A.n_ztasks = p->n_ztasks; A.z0 = p->z0; A.z_end = p->z_end;
    A.lt_lo = 0; A.lt_hi = p->n_ltasks; A.st_lo = 0; A.st_hi = p->n_stasks; A.zt_lo = 0; A.zt_hi = sweep >= 2 ? 0 : p->n_ztasks;
    A.head_part = p->head_part; A.tail_part = p->tail_part; A.split_cnt = p->split_cnt; A.diff_parts = p->diff_parts;
    A.contrib_in = p->contrib[parity]; A.contrib_out = p->contrib[parity ^ 1];
    A.rank_in = p->rank[parity]; A.rank_out = p->rank[parity ^ 1];
    A.d = d;
    A.base = (1 - d) / (double)p->N;     // (1-d) / N, pagerank.gm:11
    A.ctrl = ctrl; A.loop = loop; A.diff_slot = nullptr; A.e = e; A.max_iter = max_iter; A.mode = 0;
    A.npeers = 0; A.seq = 0; A.my_flag = nullptr;
    for (int q = 0; q < MAX_PEERS; q++) { A.peer_out[q] = nullptr; A.peer_diff[q] = nullptr; A.peer_flag[q] = nullptr; }
    return A;
}

// sharded sweep number `sweep` of the call: where the shard's diff goes, and (peers mapped) where the copies go
void set_exchange_args(const gmc_pr_plan* p0, const gmc_pr_plan* p, int sweep, unsigned int seq, SweepArgs& A) {
    const size_t stride = state_stride(p0->N);
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
        A.peer_diff[q] = base + 2 * stride + par * SYNC_RANKS + p->shard;
        A.peer_flag[q] = reinterpret_cast<unsigned int*>(base + 2 * stride + 2 * SYNC_RANKS) + p->shard;
        q++;
    }
    A.npeers = q;
}

int launch_init(const gmc_pr_plan* p, cudaStream_t st) {
    SweepArgs A = make_args(p, 0.0, 0, p->ctrl, 0, 0.0, 0);
    k_init<<<div_up(p->N + 2, 256), 256, 0, st>>>(A, p->contrib[0], p->contrib[1], p->rank[0]);
    GMC_CUDA(cudaGetLastError());
    return GMC_OK;
}

int launch_sweep(const gmc_pr_plan* p, const SweepArgs& A, cudaStream_t st) {
    // launched with programmatic stream serialization: the next sweep's CTAs may become resident while this one's tail
    // drains (pr_sweep orders itself behind its predecessor with griddepcontrol.wait); GMC_PR_NO_PDL=1 = plain launch
    static const bool pdl = getenv("GMC_PR_NO_PDL") == nullptr;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)p->grid);
    cfg.blockDim = dim3(PR_THREADS);
    cfg.dynamicSmemBytes = p->smem;
    cfg.stream = st;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at;
    cfg.numAttrs = pdl ? 1 : 0;
    GMC_CUDA(cudaLaunchKernelEx(&cfg, pr_sweep, A));
    return GMC_OK;
}

}  // namespace

// called by gmc_comm_destroy: the retired state block (and the peers' mappings of it) does not outlive the communicator
extern "C" void gmc_pr_release_cached_state(void) { sym_cache_release(); }

void gmc_pr_plan_free(gmc_graph* g) {
    if (g->pr) { free_plan(g->pr); g->pr = nullptr; }
}

static int graph_plan(gmc_graph* g, int world, int rank, cudaStream_t st) {
    if (g->pr && g->pr->world == world && g->pr->shard == rank) return GMC_OK;
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
        if (r != p->shard && p->peer_base[r]) { cudaIpcCloseMemHandle(p->peer_base[r]); p->peer_base[r] = nullptr; }
    void* bases[MAX_PEERS + 1];
    rc = gmc_comm_ipc_open_all(comm, p->block, bases, st);
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
                GMC_TRY(launch_sweep(plans[i], A, st));
                launches++;
            }
            if (sharded) {
                double* dslot = p0->diffs + (issued & 1) * SYNC_RANKS;
                p0->seq++;
                if (comm && !p0->p2p) {
                    // exchange step over NCCL: every rank's new contributions (3 id ranges per rank) and its diff
                    GMC_TRY(gmc_comm_allgather_segments(comm, p0->contrib[(issued & 1) ^ 1], 8, p0->seg_off.data(), p0->seg_len.data(), 3, st));
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
    if (h.error) { set_error("pagerank: a peer rank did not finish its sweep in time (peer-memory exchange)"); return GMC_ERR_STATE; }
    if (sharded && comm)      // the result vector: gather the owned slices of rank[] once
        GMC_TRY(gmc_comm_allgather_segments(comm, p0->rank[h.cnt & 1], 8, p0->seg_off.data(), p0->seg_len.data(), 3, st));
    k_unpermute<<<div_up(g->N, 256), 256, 0, st>>>(p0->rank[h.cnt & 1], p0->new_of_old, g->N, rank_dev);
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
    GMC_TRY(graph_plan(g, world, rank, st));
    gmc_pr_plan* p = g->pr;
    GMC_TRY(plan_map_peers(p, comm, st));
    return run_pagerank(g, &p, 1, world > 1 ? comm : nullptr, e, d, max_iter, rank_dev, info, st);
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
    GMC_TRY(launch_init(p, st));
    std::vector<cudaEvent_t> ev(3 * (size_t)n);
    for (auto& e : ev) GMC_CUDA(cudaEventCreate(&e));
    for (int32_t k = 0; k < warmup + n; k++) {
        const bool timed = k >= warmup;
        if (timed) GMC_CUDA(cudaEventRecord(ev[3 * (k - warmup)], st));
        SweepArgs A = make_args(p, d, k, p->ctrl, 0, 0.0, 0);
        if (world > 1) set_exchange_args(p, p, k, p->seq + 1, A);
        if (const char* md = getenv("GMC_PR_MODE")) { A.mode = atoi(md); if (A.mode & 2) A.hot_n = 1; }
        if (const char* only = getenv("GMC_PR_ONLY")) {      // experiment switch: time one task class alone (results invalid)
            if (strcmp(only, "long") != 0) A.lt_hi = 0;
            if (strcmp(only, "short") != 0) A.st_hi = 0;
            if (strcmp(only, "zero") != 0) A.zt_hi = 0;
        }
        GMC_TRY(launch_sweep(p, A, st));
        if (timed) GMC_CUDA(cudaEventRecord(ev[3 * (k - warmup) + 1], st));
        if (world > 1) {
            p->seq++;
            if (p->p2p) k_wait_peers<<<1, 32, 0, st>>>(p->ctrl, p->flags, world, p->seq);
            else GMC_TRY(gmc_comm_allgather_segments(comm, p->contrib[(k & 1) ^ 1], 8, p->seg_off.data(), p->seg_len.data(), 3, st));
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
    if (launches_per_sweep) *launches_per_sweep = 1;
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
    if (!g->pr) GMC_TRY(graph_plan(g, 1, 0, nullptr));
    gmc_pr_plan* p = g->pr;
    out->n_long = p->n_long; out->hot_n = p->hot_n; out->n_slices = p->n_slices; out->long_entries = p->EL;
    out->total_entries = p->EN; out->n_long_tasks = p->n_ltasks; out->n_short_tasks = p->n_stasks;
    out->grid = p->grid; out->threads = PR_THREADS; out->smem_bytes = (int32_t)p->smem;
    return GMC_OK;
}

}  // extern "C"
