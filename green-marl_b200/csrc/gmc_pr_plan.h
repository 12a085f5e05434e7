// Builder of the PageRank sweep plan ("tiered" layout), written against a small backend interface so that the very same
// steps run as CUDA kernels inside libgm_cuda (gmc_pagerank.cu: CudaBackend) and as plain loops on the host
// (tests/cpp/pr_plan_check.cc: the CPU check of the host logic, no GPU needed).  No CUDA header is needed to read this file.
//
// Reference loop the plan serves: apps/src/pagerank.gm:10-17 (pull over InNbrs, emitted by
// src/backend_cpp/gm_cpp_gen_foreach.cc:297-311).  The sweep gathers contrib[w] = rank[w] / outdeg(w) once per reverse edge.
// A random 8-byte gather that misses the SM costs one 32-byte L2 sector, and the L2 sector rate -- not HBM -- bounds the sweep
// (DESIGN.md 4.1), so the plan's job is to have as many gathers as possible served by shared memory:
//
//   hot slots   the (T+1) * 16639 vertices with the largest OUT-degree get a slot in the "hotpack" array (tier t = slots
//               [t*HT, (t+1)*HT), slot 0 of every tier is a dummy whose value is always 0.0).  A CTA stages ONE tier (130 KB)
//               in shared memory with TMA bulk copies.  Every row epilogue also stores the new contribution of a hot vertex
//               into its slot, so the hotpack of the next sweep is complete when the sweep ends.
//   ids         rows are relabelled 1..N by (in-degree > S_MAX, in-degree > 0, none) and in-degree descending; rank[],
//               contrib[] and the multi-GPU exchange live in this id space (coalesced epilogues).
//   stream 0    "home": every row of the shard with the entries that no tier stream serves (entry value = HT + id, an L2
//               gather; 0 = padding).  Long rows (in-degree > S_MAX) row-major in fixed-size tasks, the others as sliced
//               ELL: the two task classes of round 1, original source order kept inside a row.
//   stream t+1  (t = 0..T) the SEGMENTS of tier t: the entries of one row whose sources lie in tier t, if there are at least
//               KMIN of them (fewer stay cold entries of the home row: a segment costs a partial sum written and read once,
//               ~1.25 sector operations, so it must replace at least ~2 cold gathers; tier 0: KMIN = 1).  Entry value = slot
//               inside the tier.  Segments sorted by length: the long ones row-major in tasks, the others sliced ELL -- the
//               same two task classes again.  No gather of a tier stream leaves the SM.
//   two modes   SPLIT (Config::split, off by default): tier 0 is a segment stream too, home holds cold entries only, so a home task runs on ANY CTA
//               whatever tier it has staged, and every SM keeps both of its gather paths busy at once: some warps on its
//               tier's stream (shared memory), the others on home tasks (L2 sectors).  MIXED (default): tier-0 entries stay
//               inside the home rows (entry value = slot < HT); home tasks run on CTAs that have tier 0 staged, the tier
//               CTAs run their warm tiers first and join them.  With T = 0 (small graphs, short first calls) no row is split
//               and every short row is added in exactly the reference's order (bit-identical to the CPU result).
//   parts       a row with segments ("split row") owns 1 + nseg consecutive part slots: slot 0 for the home part, then one
//               per segment in tier order.  Tasks store partial sums there; the combine kernel adds them in slot order and
//               runs the row epilogue: deterministic whatever the CTA schedule, no atomics or fences in the sweep.
// On the LJ-shaped R-MAT (tools/model_segments.py): T = 16, KMIN = 3 turns 24.5 M of the 39.8 M cold gathers into shared-memory
// gathers at the price of 1.9 M part slots.  Measured: MIXED mode with T = 12, KMIN = 6 is the fastest and the
// default (0.159 ms per sweep against 0.189 ms for T = 0); SPLIT mode, the cold stream (class T + 1) and hot_first are kept as
// switches of Config -- each was built, checked here on the CPU and on the GPU, and measured slower.
#pragma once
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <algorithm>
#include <string>
#include <vector>

#if defined(__CUDACC__)
#define PRP_HD __host__ __device__ __forceinline__
#else
#define PRP_HD inline
#endif

namespace prp {

#ifndef PRP_HT
#define PRP_HT 16640                   // the CPU check (tests/cpp) shrinks the tiers so that small graphs exercise them
#endif
#ifndef PRP_S_MAX
#define PRP_S_MAX 512
#endif
constexpr int HT = PRP_HT;             // slots per tier block (130 KB of doubles: the 132 KB shared-memory carve-out step)
constexpr int HT_REAL = HT - 1;        // slot 0 of a tier is the dummy
constexpr int S_MAX = PRP_S_MAX;            // rows / segments up to this length are summed by one lane, left to right
constexpr int PIECE = 512;             // entries of a long row one plan thread walks
constexpr int MAX_TIERS = 31;          // tiers 0..31 (segment mask = one 32-bit word, tier number = 5 key bits)
constexpr int ZT = 2048;               // ids per zero-in-degree task
constexpr int SCHED_W = 4;             // tier streams a CTA may stage one after the other
constexpr uint32_t OD_SPLIT = 0x80000000u, OD_HOT = 0x40000000u, OD_MASK = 0x3fffffffu;   // od_flags[id]

struct Config {
    int tiers = 12;                    // warm tiers wanted (0: the single hot block of round 1)
    int kmin = 6;                      // entries from which a (row, tier) group becomes a segment
    int64_t tier_min_edges = 4 << 20;  // smaller graphs keep T = 0 (every short row stays bit-identical to the CPU order)
    int lt = 2048, st_cost = 2048, slice_ovh = 68;     // task sizes (entries per long task, cost units per short task)
    int tail_big16 = 14, tail_shift = 2;               // home short tasks: 14/16 of the cost in full tasks, the rest in quarter ones
    int grid = 148;                    // CTAs of the sweep
    int row_cost = 0;                  // shard cuts: entries one row epilogue weighs (grows with the world size)
    double tier_scale = 1.25;          // multiplier on the CTAs given to the tier streams
    int hot_first = 0;                 // experiment (measured 1.7 % slower): tier-0 entries first inside a home row of a tiered plan
    int cold_stream = 0;               // MIXED mode with warm tiers: the cold entries of split rows get their own stream (class T + 1)
    int kcold = 4;                     // ... when the row has at least this many of them
    int split = 0;                     // 1: SPLIT mode (tier 0 as a segment stream, home = cold entries only); see the header
    double r_smem = 2.4, r_miss = 0.75;   // gathers per clock and SM: shared memory / L2 (sched model)
};

struct Stream {
    int32_t lbeg_off;      // eoff[lbeg_off + j] = entry offset of long row j (n_long + 1 values)
    int32_t n_long;
    int32_t e_long;        // entry offset of the long section (multiple of 4)
    int32_t EL;            // entries of the long section
    int32_t sbeg_off;      // sbeg[sbeg_off + q] = entry offset of slice q (n_slices + 1 values)
    int32_t n_slices;
    int32_t lt_off;        // ltask_row[lt_off + t] (n_ltasks + 1 values); also the task's slot in head/tail/split arrays
    int32_t n_ltasks;
    int32_t st_off;        // stask_slice[st_off + u] (n_stasks + 1 values)
    int32_t n_stasks;
    int32_t meta_off;      // tier streams: seg_slot[meta_off + local row]; long rows first, then the sliced ones
    int32_t n_rows;        // tier streams: segments (long + sliced)
    int32_t hot_off;       // first slot of the staged block in the hotpack: t * HT
    int32_t entries;       // entry slots (padding included), for the schedule and the statistics
    int32_t n_big;         // short tasks of full size (the rest are the small ones at the tail)
    int32_t i_mid;         // eoff index of the first slice's leader
};

struct Plan {
    int32_t N = 0, M = 0, world = 1, shard = 0;
    int32_t T = 0, NH = HT;
    int32_t thi = 0;                              // highest class with a stream: T, or T + 1 = the cold stream (stream thi + 1)
    int32_t g_n_long = 0, n_pos = 0, n_src = 0, g_nz_slices = 0, g_z0 = 0;
    int32_t lt = 2048, lt_shift = 11, st_cost = 2048, slice_ovh = 68, minpad = 72;
    int32_t jl0 = 0, n_long = 0, s0 = 0, n_slices = 0, z0 = 0, z_end = 0, n_ztasks = 0;
    int32_t EN = 0, n_eoff = 0, n_seg = 0, K = 0, n_parts = 0;
    int32_t n_sbeg = 0, n_ltslots = 0, n_stslots = 0;
    int32_t dp_short = 0, dp_zero = 0, dp_cut = 0, dp_total = 0;     // diff_parts: [home long | home short | zero | cut rows]
    int64_t hot0_entries = 0, cold_entries = 0, seg_entries = 0, cold_stream_entries = 0;
    int32_t split = 0;                            // 1: SPLIT mode (see the header)
    Stream streams[MAX_TIERS + 2];                // [0] home, [t + 1] tier t
    std::vector<int64_t> seg_off, seg_len;        // [world][3] id ranges owned by every rank (long, short, zero)
    std::vector<int32_t> sched;                   // [grid][SCHED_W] streams a CTA visits (0 = home ends the list)
    // device (backend) arrays
    int32_t *new_of_old = nullptr, *old_of_new = nullptr, *od_flags = nullptr, *hotslot = nullptr, *home_slot = nullptr, *idslot = nullptr;
    int32_t *E = nullptr, *eoff = nullptr, *sbeg = nullptr, *ltask_row = nullptr, *stask_slice = nullptr, *seg_slot = nullptr;
    int32_t *pbeg = nullptr, *prow = nullptr, *sched_dev = nullptr;
    Stream* streams_dev = nullptr;
    std::string error;
};

PRP_HD int32_t atomic_add_i32(int32_t* p, int32_t v) {
#if defined(__CUDA_ARCH__)
    return atomicAdd(p, v);
#else
    return __atomic_fetch_add(p, v, __ATOMIC_RELAXED);
#endif
}
PRP_HD void atomic_add_i64(long long* p, long long v) {
#if defined(__CUDA_ARCH__)
    atomicAdd(reinterpret_cast<unsigned long long*>(p), (unsigned long long)v);
#else
    __atomic_fetch_add(p, v, __ATOMIC_RELAXED);
#endif
}
PRP_HD int popc32(uint32_t x) {
#if defined(__CUDA_ARCH__)
    return __popc(x);
#else
    return __builtin_popcount(x);
#endif
}
PRP_HD int32_t roundup4(int32_t x) { return (x + 3) & ~3; }
PRP_HD unsigned long long seg_key(int t, int32_t len, int32_t row) {
    return ((unsigned long long)t << 59) | ((unsigned long long)(0xffffffffu - (uint32_t)len) << 27) | (unsigned long long)(uint32_t)row;
}
PRP_HD int key_tier(unsigned long long k) { return (int)(k >> 59); }
PRP_HD int32_t key_len(unsigned long long k) { return (int32_t)(0xffffffffu - (uint32_t)((k >> 27) & 0xffffffffu)); }
PRP_HD int32_t key_row(unsigned long long k) { return (int32_t)(k & 0x7ffffffu); }
// first i in [lo, hi) with a[i] >= x (a ascending)
template <typename T>
PRP_HD int32_t lower_bound_i(const T* a, int32_t lo, int32_t hi, T x) {
    while (lo < hi) {
        const int32_t mid = lo + ((hi - lo) >> 1);
        if (a[mid] >= x) hi = mid; else lo = mid + 1;
    }
    return lo;
}
// last j in [0, n] with a[j] <= x (a ascending, a[0] <= x)
PRP_HD int32_t last_le(const int32_t* a, int32_t n, int32_t x) {
    int32_t lo = 0, hi = n + 1;
    while (hi - lo > 1) {
        const int32_t mid = lo + ((hi - lo) >> 1);
        if (a[mid] <= x) lo = mid; else hi = mid;
    }
    return lo;
}

// ---------------------------------------------------------------------------------------------------------------------
// step functors: one call per item, no cooperation between items (so a host loop runs them as well as a CUDA grid)
// ---------------------------------------------------------------------------------------------------------------------
struct DegF {
    const int32_t *begin, *r_begin;
    int32_t *outdeg, *indeg, *ids;
    uint32_t* keyo;
    unsigned long long* keyr;
    int32_t* err;
    PRP_HD void operator()(int64_t v) const {
        const int32_t od = begin[v + 1] - begin[v], id = r_begin[v + 1] - r_begin[v];
        outdeg[v] = od; indeg[v] = id; ids[v] = (int32_t)v;
        if (od < 0 || id < 0 || (uint32_t)od > OD_MASK) *err = 1;
        keyo[v] = od > 0 ? (uint32_t)(0x7fffffff - od) : 0xffffffffu;            // out-degree descending, sinks last
        const int cls = id > S_MAX ? 0 : (id > 0 ? 1 : 2);
        keyr[v] = ((unsigned long long)cls << 32) | (uint32_t)(0x7fffffff - id);  // long rows, short rows, rows without in-edges
    }
};
struct BoundsF {       // one item: class boundaries of the two sorted orders
    const uint32_t* so_keys;
    const unsigned long long* sr_keys;
    int32_t n;
    int32_t* out;      // [0] n_long, [1] n_pos, [2] n_src
    PRP_HD void operator()(int64_t) const {
        out[0] = lower_bound_i<unsigned long long>(sr_keys, 0, n, 1ull << 32);
        out[1] = lower_bound_i<unsigned long long>(sr_keys, 0, n, 2ull << 32);
        out[2] = lower_bound_i<uint32_t>(so_keys, 0, n, 0xffffffffu);
    }
};
struct HotF {          // p-th vertex of the out-degree order -> hot slot (tier * HT + 1 + position inside the tier) or -1
    const int32_t *so_ids, *outdeg;
    int32_t* hotslot_old;
    int32_t nhot;
    PRP_HD void operator()(int64_t p) const {
        const int32_t v = so_ids[p];
        hotslot_old[v] = (p < nhot && outdeg[v] > 0) ? (int32_t)(p / HT_REAL) * HT + 1 + (int32_t)(p % HT_REAL) : -1;
    }
};
struct TierSumF {      // edges leaving every tier (64 items per tier)
    const int32_t *so_ids, *outdeg;
    int32_t nhot, n;
    long long* tier_out;
    PRP_HD void operator()(int64_t i) const {
        const int t = (int)(i >> 6), part = (int)(i & 63);
        const int32_t per = (HT_REAL + 63) / 64;
        int32_t p0 = t * HT_REAL + part * per, p1 = p0 + per;
        const int32_t lim = std::min(std::min((t + 1) * HT_REAL, nhot), n);
        if (p1 > lim) p1 = lim;
        long long s = 0;
        for (int32_t p = p0; p < p1; p++) s += outdeg[so_ids[p]];
        if (s) atomic_add_i64(tier_out + t, s);
    }
};
struct RelabelF {
    const int32_t* sr_ids;
    int32_t *new_of_old, *old_of_new;
    PRP_HD void operator()(int64_t p) const {
        const int32_t v = sr_ids[p];
        if (p == 0) old_of_new[0] = -1;
        old_of_new[p + 1] = v;
        new_of_old[v] = (int32_t)p + 1;
    }
};
struct ByIdF {         // arrays indexed by the new id, and the gather code of every vertex
    const int32_t *old_of_new, *outdeg, *hotslot_old;
    int32_t *od_flags, *hotslot, *idslot, *code;
    int32_t split;
    PRP_HD void operator()(int64_t p) const {
        const int32_t id = (int32_t)p + 1, v = old_of_new[id], slot = hotslot_old[v];
        if (p == 0) { od_flags[0] = 0; hotslot[0] = -1; }
        od_flags[id] = (int32_t)((uint32_t)outdeg[v] | (slot >= 0 ? OD_HOT : 0u));
        hotslot[id] = slot;
        if (slot >= 0) idslot[slot] = id;
        // entry code of a source: served by a tier stream -> -slot; tier 0 in MIXED mode -> its slot (< HT); no slot -> HT + id
        code[v] = slot < 0 ? HT + id : ((slot < HT && !split) ? slot : -slot);
    }
};
// global row offsets by in-degree (shard cuts only): padded long rows, slice lengths and costs
struct GlobalLenF {
    const int32_t *old_of_new, *indeg;
    int32_t n, g_n_long, g_n_slices, slice_ovh;
    int32_t *g_lbeg, *g_sbeg;
    PRP_HD void operator()(int64_t j) const {
        if (j == 0) { g_lbeg[g_n_long] = 0; g_sbeg[g_n_slices] = 0; }
        const int32_t len = indeg[old_of_new[j + 1]];
        if (j < g_n_long) g_lbeg[j] = roundup4(len);
        else if (((j - g_n_long) & 31) == 0) g_sbeg[(j - g_n_long) >> 5] = 32 * len;
    }
};

struct RowCtx {        // the shard's rows: r in [0, nl) long rows (id = jl0 + 1 + r), r >= nl short rows (id = short_id0 + r - nl)
    int32_t hot_first;     // MIXED mode with warm tiers: a home row holds its tier-0 entries first, then its L2 entries
    int32_t n, nl, jl0, short_id0, tlo, T, thi, kmin, kcold;   // thi = T, or T + 1 = the cold class when it has a stream       // tiers tlo..T have streams (tlo = 0 in SPLIT mode, 1 otherwise)
    const int32_t *old_of_new, *r_begin, *r_node_idx, *code;
    PRP_HD int32_t id_of(int32_t r) const { return r < nl ? jl0 + 1 + r : short_id0 + (r - nl); }
};

// emits the segment records of one row from its per-tier counts; returns the number of entries that move to segments
template <typename CNT>
PRP_HD int32_t emit_segments(const CNT* cnt, int tlo, int T, int thi, int kmin, int kcold, int32_t r, uint32_t* mask_out, int32_t* nrec, unsigned long long* rec_key,
                             int32_t* rec_j, int32_t rec_cap, int32_t* err) {
    uint32_t mask = 0;
    int32_t tot = 0;
    int ns = 0;
    for (int t = tlo; t <= T; t++)
        if ((int32_t)cnt[t] >= (t == 0 ? 1 : kmin)) { mask |= 1u << t; tot += (int32_t)cnt[t]; ns++; }
    // a row that is split anyway hands its cold entries to the cold stream (any CTA takes its tasks)
    if (thi > T && ns > 0 && (int32_t)cnt[thi] >= kcold) { mask |= 1u << thi; tot += (int32_t)cnt[thi]; ns++; }
    *mask_out = mask;
    if (ns) {
        int32_t base = atomic_add_i32(nrec, ns);
        if (base + ns > rec_cap) { *err = 2; return tot; }
        int j = 1;
        for (int t = tlo; t <= thi; t++)
            if (mask >> t & 1u) { rec_key[base] = seg_key(t, (int32_t)cnt[t], r); rec_j[base] = j++; base++; }
    }
    return tot;
}

// Walks the entries [a, b) of a reverse-CSR row in order, f(src, code[src]).  The sources and their codes are fetched eight at a
// time so that a thread has 8 + 8 independent loads in flight instead of a chain of two dependent loads per entry: the plan
// passes are latency chains of up to 512 entries per thread.
template <class F>
PRP_HD void walk_entries(const int32_t* __restrict__ idx, const int32_t* __restrict__ code, int32_t a, int32_t b, F&& f) {
    for (int32_t k = a; k < b; k += 8) {
        int32_t sr[8], cd[8];
#if defined(__CUDA_ARCH__)
#pragma unroll
#endif
        for (int i = 0; i < 8; i++) sr[i] = k + i < b ? idx[k + i] : -1;
#if defined(__CUDA_ARCH__)
#pragma unroll
#endif
        for (int i = 0; i < 8; i++) cd[i] = sr[i] >= 0 ? code[sr[i]] : 0;
#if defined(__CUDA_ARCH__)
#pragma unroll
#endif
        for (int i = 0; i < 8; i++) if (sr[i] >= 0) f(sr[i], cd[i]);
    }
}

struct CountShortF {   // item i: short row r = nl + i
    RowCtx c;
    uint32_t* segmask;
    int32_t *home_len, *home_hot, *nrec, *rec_j, *err;
    unsigned long long* rec_key;
    int32_t rec_cap;
    int32_t e_lo, e_hi;    // this pass handles the rows whose LAST edge lies in [e_lo, e_hi) (copy-in still in flight: the edges
                           // arrive in chunks); rows without edges go with e_lo == 0
    PRP_HD void operator()(int64_t i) const {
        const int32_t r = c.nl + (int32_t)i, id = c.short_id0 + (int32_t)i;
        if (id > c.n) { if (e_lo == 0) { segmask[r] = 0; home_len[r] = 0; home_hot[r] = -1; } return; }
        const int32_t v = c.old_of_new[id], rb = c.r_begin[v], len = c.r_begin[v + 1] - rb;
        if (len == 0 ? e_lo != 0 : (rb + len - 1 < e_lo || rb + len - 1 >= e_hi)) return;
        // (rows too short for a segment are not walked: they keep their source order, home_hot = -1)
        if (c.tlo > c.T || len < (c.tlo == 0 ? 1 : c.kmin)) { segmask[r] = 0; home_len[r] = len; home_hot[r] = -1; return; }
        uint16_t cnt[MAX_TIERS + 1];
        for (int t = 0; t <= c.thi; t++) cnt[t] = 0;
        const RowCtx& cc = c;
        int32_t hot0 = 0;
        walk_entries(c.r_node_idx, c.code, rb, rb + len, [&](int32_t, int32_t cd) {
            if (cd < 0) cnt[(-cd) / HT]++;
            else if (cd < HT) hot0++;
            else if (cc.thi > cc.T && cd >= HT) cnt[cc.thi]++;
        });
        uint32_t mask;
        const int32_t tot = emit_segments(cnt, c.tlo, c.T, c.thi, c.kmin, c.kcold, r, &mask, nrec, rec_key, rec_j, rec_cap, err);
        segmask[r] = mask;
        home_len[r] = len - tot;
        home_hot[r] = c.hot_first ? hot0 : -1;
    }
};
struct HomeLenDegF {   // no tier streams: every entry stays home, nothing needs the edges
    RowCtx c;
    int32_t nrows;
    uint32_t* segmask;
    int32_t *home_len, *home_hot;
    PRP_HD void operator()(int64_t r) const {
        const int32_t id = c.id_of((int32_t)r);
        segmask[r] = 0;
        home_hot[r] = -1;
        home_len[r] = id <= c.n ? c.r_begin[c.old_of_new[id] + 1] - c.r_begin[c.old_of_new[id]] : 0;
    }
};
struct PieceNumF {     // pieces of long row j (+ a trailing 0 for the scan)
    RowCtx c;
    int32_t* np;
    PRP_HD void operator()(int64_t j) const {
        if (j == c.nl) { np[j] = 0; return; }
        const int32_t v = c.old_of_new[c.jl0 + 1 + j];
        np[j] = (c.r_begin[v + 1] - c.r_begin[v] + PIECE - 1) / PIECE;
    }
};
struct PieceRowF {
    const int32_t* piece_first;   // [nl + 1]
    int32_t nl;
    int32_t* piece_row;
    PRP_HD void operator()(int64_t p) const {
        if (p >= piece_first[nl]) return;
        piece_row[p] = last_le(piece_first, nl, (int32_t)p);
    }
};
struct PieceCountF {   // per-tier counts of one piece -> G[t * nps + p]
    RowCtx c;
    const int32_t *piece_first, *piece_row;
    int32_t nps;
    int32_t* G;
    int32_t e_lo, e_hi;    // pieces whose last edge lies in [e_lo, e_hi)
    PRP_HD void operator()(int64_t p) const {
        if (p >= piece_first[c.nl]) return;
        const int32_t j = piece_row[p], v = c.old_of_new[c.jl0 + 1 + j], rb = c.r_begin[v], len = c.r_begin[v + 1] - rb;
        const int32_t k0 = ((int32_t)p - piece_first[j]) * PIECE, k1 = std::min(len, k0 + PIECE);
        if (rb + k1 - 1 < e_lo || rb + k1 - 1 >= e_hi) return;
        uint16_t cnt[MAX_TIERS + 1];
        for (int t = 0; t <= c.thi; t++) cnt[t] = 0;
        const RowCtx& cc = c;
        int32_t hot0 = 0;
        walk_entries(c.r_node_idx, c.code, rb + k0, rb + k1, [&](int32_t, int32_t cd) {
            if (cd < 0) cnt[(-cd) / HT]++;
            else if (cd < HT) hot0++;
            else if (cc.thi > cc.T && cd >= HT) cnt[cc.thi]++;
        });
        for (int t = c.tlo; t <= c.thi; t++) G[(size_t)t * nps + p] = cnt[t];
        if (c.hot_first) G[p] = hot0;                 // column 0 is free in MIXED mode: the tier-0 entries that stay home
    }
};
struct LongRowF {      // G scanned (exclusive): totals of long row j -> segments
    RowCtx c;
    const int32_t *piece_first, *G;
    int32_t nps;
    uint32_t* segmask;
    int32_t *home_len, *home_hot, *nrec, *rec_j, *err;
    unsigned long long* rec_key;
    int32_t rec_cap;
    PRP_HD void operator()(int64_t j) const {
        const int32_t v = c.old_of_new[c.jl0 + 1 + j], len = c.r_begin[v + 1] - c.r_begin[v];
        const int32_t p0 = piece_first[j], p1 = piece_first[j + 1];
        int32_t cnt[MAX_TIERS + 1];
        cnt[0] = 0;
        for (int t = 0; t <= c.thi; t++) cnt[t] = t >= c.tlo ? G[(size_t)t * nps + p1] - G[(size_t)t * nps + p0] : 0;
        uint32_t mask = 0;
        const int32_t tot = c.tlo <= c.T ? emit_segments(cnt, c.tlo, c.T, c.thi, c.kmin, c.kcold, (int32_t)j, &mask, nrec, rec_key, rec_j, rec_cap, err) : 0;
        segmask[j] = mask;
        home_len[j] = len - tot;
        home_hot[j] = c.hot_first ? G[p1] - G[p0] : -1;
    }
};
struct TierBoundF {    // item t = 0..T+1: first sorted segment of tier t and first one of it that is not long
    const unsigned long long* keys;
    int32_t nseg, T;
    int32_t *tier_q, *tier_mid;      // [T + 2]
    PRP_HD void operator()(int64_t i) const {
        const int t = (int)i;
        if (t > T) { tier_q[t] = nseg; tier_mid[t] = nseg; return; }
        tier_q[t] = lower_bound_i<unsigned long long>(keys, 0, nseg, (unsigned long long)t << 59);
        tier_mid[t] = lower_bound_i<unsigned long long>(keys, 0, nseg, ((unsigned long long)t << 59) | ((unsigned long long)(0xffffffffu - (uint32_t)S_MAX) << 27));
    }
};
struct PartsF {        // part slots and split-row numbers (before their scans); item nrows = the trailing 0
    const uint32_t* segmask;
    int32_t nrows;
    int32_t *nparts, *issplit;
    PRP_HD void operator()(int64_t r) const {
        const uint32_t m = r < nrows ? segmask[r] : 0u;
        nparts[r] = m ? 1 + popc32(m) : 0;
        issplit[r] = m ? 1 : 0;
    }
};
struct SplitF {        // pslot / kidx = the scanned arrays
    RowCtx c;
    const uint32_t* segmask;
    const int32_t *pslot, *kidx;
    int32_t nrows;
    int32_t *pbeg, *prow, *home_slot, *od_flags;
    PRP_HD void operator()(int64_t r) const {
        if (r == nrows) { pbeg[kidx[r]] = pslot[r]; return; }
        if (!segmask[r]) return;
        const int32_t id = c.id_of((int32_t)r), k = kidx[r];
        pbeg[k] = pslot[r]; prow[k] = id; home_slot[id] = pslot[r];
        od_flags[id] = (int32_t)((uint32_t)od_flags[id] | OD_SPLIT);
    }
};
struct HomeLongLenF {
    const int32_t* home_len;
    int32_t minpad;
    int32_t* elen;
    PRP_HD void operator()(int64_t j) const { const int32_t l = roundup4(home_len[j]); elen[j] = l > minpad ? l : minpad; }
};
struct HomeSliceLenF {
    const int32_t* home_len;
    int32_t nl;
    int32_t* elen;
    PRP_HD void operator()(int64_t q) const {
        int32_t mx = 0;
        for (int l = 0; l < 32; l++) mx = std::max(mx, home_len[nl + 32 * q + l]);
        elen[nl + q] = 32 * mx;
    }
};
struct SegLenF {       // entry slots of the q-th sorted segment (a slice's slots are booked by its first = longest segment)
    const unsigned long long* keys;
    const int32_t *tier_q, *tier_mid;
    int32_t* elen;     // already offset to the segment part of the array
    PRP_HD void operator()(int64_t q) const {
        const unsigned long long k = keys[q];
        const int t = key_tier(k);
        const int32_t len = key_len(k);
        if (q < tier_mid[t]) elen[q] = roundup4(len);
        else elen[q] = (((int32_t)q - tier_mid[t]) & 31) == 0 ? 32 * len : 0;
    }
};
struct SegDestF {      // where the entries of segment q go, and which part slot its sum goes to
    const unsigned long long* keys;
    const int32_t *rec_j, *tier_mid, *pslot, *eoff_seg;     // eoff_seg = eoff + nl + ns
    int32_t *seg_slot, *seg_ebase, *seg_estride;
    PRP_HD void operator()(int64_t q) const {
        const unsigned long long k = keys[q];
        const int t = key_tier(k);
        const int32_t slot = pslot[key_row(k)] + rec_j[q];
        seg_slot[q] = slot;
        if (q < tier_mid[t]) { seg_ebase[slot] = eoff_seg[q]; seg_estride[slot] = 1; }
        else {
            const int32_t lane = ((int32_t)q - tier_mid[t]) & 31;
            seg_ebase[slot] = eoff_seg[q - lane] + lane; seg_estride[slot] = 32;
        }
    }
};
struct SbegF {         // slice offsets of one stream: sbeg[q] = eoff[pos0 + q * step], q <= n (the last one = end of the stream)
    const int32_t* eoff;
    int32_t pos0, step, n, end_idx;
    int32_t* sbeg;
    PRP_HD void operator()(int64_t q) const { sbeg[q] = q < n ? eoff[pos0 + q * step] : eoff[end_idx]; }
};
struct SliceCostF {    // cost of a slice = its entry slots + the per-slice overhead; the scan gives the task cuts
    const int32_t* sbeg;
    int32_t n, ovh;
    int32_t* cost;
    PRP_HD void operator()(int64_t q) const { cost[q] = q < n ? sbeg[q + 1] - sbeg[q] + ovh : 0; }
};
struct LtaskF {        // first[t] = long row holding entry e_long + t * lt
    const int32_t* lbeg;  // eoff + lbeg_off
    int32_t n_long, n_tasks, lt, e_long;
    int32_t* first;
    PRP_HD void operator()(int64_t t) const {
        if (t >= n_tasks) { first[t] = n_long; return; }
        first[t] = last_le(lbeg, n_long, e_long + (int32_t)t * lt);
    }
};
struct StaskF {        // short tasks: the first n_big hold st_cost cost units, the others st_cost >> tail_shift
    const int32_t* cpos;  // scanned slice costs of the stream (n_slices + 1)
    int32_t n_slices, n_tasks, st_cost, n_big, tail_shift;
    int32_t* first;
    PRP_HD void operator()(int64_t u) const {
        if (u >= n_tasks) { first[u] = n_slices; return; }
        const int32_t target = cpos[0] + (u <= n_big ? (int32_t)u * st_cost : n_big * st_cost + ((int32_t)u - n_big) * (st_cost >> tail_shift));
        first[u] = lower_bound_i<int32_t>(cpos, 0, n_slices, target);
    }
};

// destination bookkeeping of one row part while its entries are walked in source order
struct SegCursor {
    int32_t pos[MAX_TIERS + 1], stride[MAX_TIERS + 1];
};
struct FillCtx {
    RowCtx c;
    const uint32_t* segmask;
    const int32_t *pslot, *seg_ebase, *seg_estride, *new_of_old, *eoff, *home_hot;
    int32_t* E;
};
// class of an entry with code cd: its tier (code < 0), the cold class (no slot, and the cold stream exists), or -1 = stays home
PRP_HD int entry_class(const RowCtx& c, int32_t cd) { return cd < 0 ? (-cd) / HT : ((c.thi > c.T && cd >= HT) ? c.thi : -1); }
PRP_HD int32_t home_value(const FillCtx& f, int32_t cd, int32_t src) { return cd < 0 ? HT + f.new_of_old[src] : cd; }

struct FillShortF {    // item i: short row r = nl + i; its home entries go to column-major slice slots, original order kept
    FillCtx f;
    int32_t ns;
    int32_t e_lo, e_hi;    // rows whose last edge lies in [e_lo, e_hi) (plans without tiers are filled behind the copy-in)
    PRP_HD void operator()(int64_t i) const {
        const RowCtx& c = f.c;
        const int32_t r = c.nl + (int32_t)i, id = c.short_id0 + (int32_t)i;
        if (id > c.n) return;
        const int32_t v = c.old_of_new[id], rb = c.r_begin[v], len = c.r_begin[v + 1] - rb;
        if (len == 0 || rb + len - 1 < e_lo || rb + len - 1 >= e_hi) return;
        const uint32_t mask = f.segmask[r];
        int32_t h = f.eoff[c.nl + (i >> 5)] + (int32_t)(i & 31);
        if (!mask) {
            const FillCtx& f0 = f;
            const int32_t nh0 = f.home_hot[r];
            if (nh0 < 0) walk_entries(c.r_node_idx, c.code, rb, rb + len, [&](int32_t src, int32_t cd) { f0.E[h] = home_value(f0, cd, src); h += 32; });
            else {              // tier-0 entries first, then the L2 ones (each group in source order)
                int32_t hc0 = h + 32 * nh0;
                walk_entries(c.r_node_idx, c.code, rb, rb + len, [&](int32_t src, int32_t cd) {
                    if (cd > 0 && cd < HT) { f0.E[h] = cd; h += 32; } else { f0.E[hc0] = home_value(f0, cd, src); hc0 += 32; }
                });
            }
            return;
        }
        SegCursor cur;
        int j = 1;
        for (int t = 0; t <= c.thi; t++)
            if (mask >> t & 1u) { const int32_t s = f.pslot[r] + j++; cur.pos[t] = f.seg_ebase[s]; cur.stride[t] = f.seg_estride[s]; }
        const FillCtx& ff = f;
        const int32_t nh = f.home_hot[r];
        int32_t hc = nh >= 0 ? h + 32 * nh : h;          // first slot of the L2 entries when the row is hot-first
        walk_entries(c.r_node_idx, c.code, rb, rb + len, [&](int32_t src, int32_t cd) {
            const int t = entry_class(ff.c, cd);
            if (t >= 0 && (mask >> t & 1u)) { ff.E[cur.pos[t]] = cd < 0 ? (-cd) - t * HT : cd; cur.pos[t] += cur.stride[t]; }
            else if (nh < 0 || (cd > 0 && cd < HT)) { ff.E[h] = home_value(ff, cd, src); h += 32; }
            else { ff.E[hc] = home_value(ff, cd, src); hc += 32; }
        });
    }
};
struct FillPieceF {    // one piece of a long row: the scanned piece table says where the piece starts in every destination
    FillCtx f;
    const int32_t *piece_first, *piece_row, *G;
    int32_t nps;
    int32_t e_lo, e_hi;    // pieces whose last edge lies in [e_lo, e_hi)
    PRP_HD void operator()(int64_t p) const {
        const RowCtx& c = f.c;
        if (p >= piece_first[c.nl]) return;
        const int32_t j = piece_row[p], v = c.old_of_new[c.jl0 + 1 + j], rb = c.r_begin[v], len = c.r_begin[v + 1] - rb;
        const int32_t p0 = piece_first[j];
        const int32_t k0 = ((int32_t)p - p0) * PIECE, k1 = std::min(len, k0 + PIECE);
        if (rb + k1 - 1 < e_lo || rb + k1 - 1 >= e_hi) return;
        const uint32_t mask = f.segmask[j];
        SegCursor cur;
        int32_t moved = 0;
        int jj = 1;
        for (int t = 0; t <= c.thi; t++)
            if (mask >> t & 1u) {
                const int32_t s = f.pslot[j] + jj++;
                const int32_t before = G[(size_t)t * nps + p] - G[(size_t)t * nps + p0];
                cur.stride[t] = f.seg_estride[s];
                cur.pos[t] = f.seg_ebase[s] + before * cur.stride[t];
                moved += before;
            }
        const int32_t nh = f.home_hot[j];
        const int32_t hot_before = nh >= 0 ? G[p] - G[p0] : 0;                 // column 0: tier-0 entries of the earlier pieces
        int32_t h = nh >= 0 ? f.eoff[j] + hot_before : f.eoff[j] + (k0 - moved);
        int32_t hc = nh >= 0 ? f.eoff[j] + nh + (k0 - moved - hot_before) : h;
        const FillCtx& ff = f;
        walk_entries(c.r_node_idx, c.code, rb + k0, rb + k1, [&](int32_t src, int32_t cd) {
            const int t = entry_class(ff.c, cd);
            if (t >= 0 && (mask >> t & 1u)) { ff.E[cur.pos[t]] = cd < 0 ? (-cd) - t * HT : cd; cur.pos[t] += cur.stride[t]; }
            else if (nh < 0 || (cd > 0 && cd < HT)) ff.E[h++] = home_value(ff, cd, src);
            else ff.E[hc++] = home_value(ff, cd, src);
        });
    }
};
struct GatherF {       // out[i] = a[idx[i]]: a handful of offsets the host needs
    const int32_t *a, *idx;
    int32_t* out;
    PRP_HD void operator()(int64_t i) const { out[i] = a[idx[i]]; }
};

// balanced contiguous partition of a prefix array (host): cuts[r] = first i with prefix[i] >= r * total / world
static inline void shard_cuts(const int64_t* prefix, int32_t n, int world, int32_t* cuts) {
    const int64_t total = prefix[n];
    cuts[0] = 0;
    for (int r = 1; r < world; r++) {
        const int64_t target = (total * r + world - 1) / world;
        int32_t lo = 0, hi = n;
        while (lo < hi) { const int32_t mid = lo + ((hi - lo) >> 1); if (prefix[mid] >= target) hi = mid; else lo = mid + 1; }
        cuts[r] = lo < cuts[r - 1] ? cuts[r - 1] : lo;
    }
    cuts[world] = n;
}


// Which tier streams every CTA stages, in order (sched[b][i] = stream number, 0 ends the list).  Home tasks need no tier in
// SPLIT mode and are taken by the warps of every CTA beside its tier's tasks, so only the tier streams are apportioned: CTAs in
// proportion to the entries of a tier (all shared-memory gathers, same cost each); tiers too small for one CTA are grouped, up
// to SCHED_W - 1 per CTA; every list ends with stream 1 (tier 0, the largest) as a place to help once the own work is done.
// MIXED mode: every CTA stages tier 0 (its stream is empty; the home rows gather from it).
static inline void make_schedule(const Config& cfg, Plan& P) {
    const int G = cfg.grid > 0 ? cfg.grid : 1;
    P.sched.assign((size_t)G * SCHED_W, 0);
    for (int b = 0; b < G; b++) P.sched[(size_t)b * SCHED_W] = 1;
    if (!P.split && P.T > 0) {
        // MIXED mode with warm tiers: home tasks need tier 0, so a CTA is either a home CTA (list = tier 0 only) or a tier CTA
        // (its warm tiers, then tier 0 to join the home tasks).  Tier CTAs in proportion to the modelled cost: shared-memory
        // gathers for the tiers; for home the slower of its shared-memory and its L2 gathers.
        std::vector<double> c((size_t)P.T + 2, 0.0);
        const double hot = (double)P.hot0_entries / cfg.r_smem, cold = (double)P.cold_entries / cfg.r_miss;
        double total = std::max(hot, cold) + 0.25 * std::min(hot, cold);
        for (int s = 2; s <= P.T + 1; s++) { c[s] = (double)P.streams[s].entries / cfg.r_smem * cfg.tier_scale; total += c[s]; }
        double per = total / G;
        if (per <= 0) per = 1;
        const int budget = G * 3 / 5 > 0 ? G * 3 / 5 : 1;
        for (int attempt = 0; attempt < 200; attempt++, per *= 1.05) {
            std::vector<int32_t> sc((size_t)G * SCHED_W, 0);
            int b = 0, s = 2;
            bool ok = true;
            while (s <= P.T + 1) {
                if (b >= budget) { ok = false; break; }
                if (c[s] >= 0.75 * per) {
                    int k = std::max(1, (int)(c[s] / per + 0.999));
                    if (b + k > budget) { ok = false; break; }
                    for (int i = 0; i < k; i++) { sc[(size_t)(b + i) * SCHED_W] = s; sc[(size_t)(b + i) * SCHED_W + 1] = 1; }
                    b += k; s++;
                } else {
                    double sum = 0;
                    int cnt = 0;
                    while (s <= P.T + 1 && cnt < SCHED_W - 1 && (cnt == 0 || sum + c[s] <= 1.1 * per)) { sc[(size_t)b * SCHED_W + cnt] = s; sum += c[s]; cnt++; s++; }
                    sc[(size_t)b * SCHED_W + cnt] = 1;
                    b++;
                }
            }
            if (!ok) continue;
            for (; b < G; b++) sc[(size_t)b * SCHED_W] = 1;
            P.sched = sc;
            return;
        }
        return;
    }
    if (!P.split) return;
    const int NTS = P.T + 1;                       // tier streams 1..NTS
    std::vector<double> c((size_t)NTS + 1, 0.0);
    double total = 0;
    for (int s = 1; s <= NTS; s++) { c[s] = (double)P.streams[s].entries * (s == 1 ? 1.0 : cfg.tier_scale) + 1.0; total += c[s]; }
    double per = total / G;
    for (int attempt = 0; attempt < 200; attempt++, per *= 1.03) {
        std::vector<int> k((size_t)NTS + 1, 0);    // CTAs of a big tier; 0 = grouped
        std::vector<std::vector<int>> groups;
        int used = 0;
        for (int s = 1; s <= NTS;) {
            if (c[s] >= 0.75 * per) { k[s] = std::max(1, (int)(c[s] / per + 0.5)); used += k[s]; s++; continue; }
            std::vector<int> g;
            double sum = 0;
            while (s <= NTS && c[s] < 0.75 * per && (int)g.size() < SCHED_W - 1 && (g.empty() || sum + c[s] <= 1.1 * per)) { g.push_back(s); sum += c[s]; s++; }
            groups.push_back(g); used++;
        }
        if (used > G) continue;
        for (; used < G; used++) {                 // leftover CTAs go where a CTA has the most entries to run
            int best = 0;
            double load = -1;
            for (int s = 1; s <= NTS; s++) if (k[s] > 0 && c[s] / k[s] > load) { load = c[s] / k[s]; best = s; }
            if (!best) break;
            k[best]++;
        }
        std::vector<int32_t> sc((size_t)G * SCHED_W, 0);
        int b = 0;
        for (int s = 1; s <= NTS; s++) for (int i = 0; i < k[s] && b < G; i++) sc[(size_t)(b++) * SCHED_W] = s;
        for (auto& g : groups) { if (b >= G) break; for (size_t i = 0; i < g.size(); i++) sc[(size_t)b * SCHED_W + i] = g[i]; b++; }
        for (; b < G; b++) sc[(size_t)b * SCHED_W] = 1;
        for (int bb = 0; bb < G; bb++) {           // the helper entry
            int32_t* row = &sc[(size_t)bb * SCHED_W];
            int len = 0;
            bool has1 = false;
            while (len < SCHED_W && row[len]) { has1 = has1 || row[len] == 1; len++; }
            if (!has1 && len < SCHED_W) row[len] = 1;
        }
        P.sched = sc;
        return;
    }
}

// ---------------------------------------------------------------------------------------------------------------------
// the builder
// ---------------------------------------------------------------------------------------------------------------------
struct Input {
    int32_t n = 0, m = 0;
    const int32_t *begin = nullptr, *r_begin = nullptr, *r_node_idx = nullptr;
    int world = 1, rank = 0;
};

#define PRP_TRY(expr) do { if ((rc = (expr)) != 0) goto done; } while (0)

template <class B>
struct Temps {
    B& be;
    std::vector<void*> v;
    explicit Temps(B& b) : be(b) {}
    ~Temps() { for (void* p : v) be.free(p); }
    template <typename T> int get(T** p, size_t n) { int rc = be.alloc((void**)p, sizeof(T) * (n ? n : 1)); if (rc == 0) v.push_back(*p); return rc; }
};

// Builds plan P of shard `rank` of `world` for the reverse CSR in `in`.  Returns 0 or the backend's error code (P.error set).
// The arrays listed in Plan are allocated with be.alloc and stay owned by the caller (free_plan_arrays).
template <class B>
int build(B& be, const Input& in, const Config& cfg, Plan& P) {
    int rc = 0;
    const int32_t n = in.n, m = in.m;
    const int world = in.world, rank = in.rank;
    Temps<B> tmp(be);
    int32_t *outdeg = nullptr, *indeg = nullptr, *ids = nullptr, *so_ids = nullptr, *sr_ids = nullptr, *hotslot_old = nullptr, *code = nullptr, *small = nullptr, *err = nullptr;
    uint32_t *keyo = nullptr, *so_keys = nullptr;
    unsigned long long *keyr = nullptr, *sr_keys = nullptr;
    long long* tier_out = nullptr;
    int32_t h3[4] = {0, 0, 0, 0};
    std::vector<int32_t> cutsL((size_t)world + 1), cutsS((size_t)world + 1), cutsZ((size_t)world + 1);
    std::vector<long long> h_tier_out(MAX_TIERS + 2, 0);
    int32_t tlo = 1, thi = 0, g_n_slices = 0, nl = 0, ns = 0, nrows = 0, short_id0 = 0, nps = 0, rec_cap = 0, nseg = 0, X = 0, herr = 0;
    uint32_t* segmask = nullptr;
    int32_t *home_len = nullptr, *home_hot = nullptr, *nrec = nullptr, *rec_j = nullptr, *rec_j2 = nullptr, *np = nullptr, *piece_row = nullptr, *G = nullptr;
    unsigned long long *rec_key = nullptr, *rec_key2 = nullptr;
    int32_t *tier_q = nullptr, *tier_mid = nullptr, *nparts = nullptr, *issplit = nullptr, *elen = nullptr, *seg_ebase = nullptr, *seg_estride = nullptr, *scost = nullptr;
    int32_t *gidx = nullptr, *gout = nullptr;
    std::vector<int32_t> h_tq(MAX_TIERS + 3, 0), h_tm(MAX_TIERS + 3, 0), h_idx, h_val;
    RowCtx rc_;
    P = Plan();
    P.N = n; P.M = m; P.world = world; P.shard = rank;
    const size_t nn = (size_t)n;

    // ---- 1. degrees, the out-degree order (hot slots) and the row order (ids): needs begin / r_begin only
    PRP_TRY(tmp.get(&outdeg, nn)); PRP_TRY(tmp.get(&indeg, nn)); PRP_TRY(tmp.get(&ids, nn)); PRP_TRY(tmp.get(&so_ids, nn)); PRP_TRY(tmp.get(&sr_ids, nn));
    PRP_TRY(tmp.get(&keyo, nn)); PRP_TRY(tmp.get(&so_keys, nn)); PRP_TRY(tmp.get(&keyr, nn)); PRP_TRY(tmp.get(&sr_keys, nn));
    PRP_TRY(tmp.get(&hotslot_old, nn)); PRP_TRY(tmp.get(&code, nn)); PRP_TRY(tmp.get(&small, 64)); PRP_TRY(tmp.get(&err, 1)); PRP_TRY(tmp.get(&tier_out, MAX_TIERS + 2));
    PRP_TRY(be.alloc((void**)&P.new_of_old, 4 * nn)); PRP_TRY(be.alloc((void**)&P.old_of_new, 4 * (nn + 1)));
    PRP_TRY(be.alloc((void**)&P.od_flags, 4 * (nn + 1))); PRP_TRY(be.alloc((void**)&P.hotslot, 4 * (nn + 1))); PRP_TRY(be.alloc((void**)&P.home_slot, 4 * (nn + 1)));
    PRP_TRY(be.zero(err, 4)); PRP_TRY(be.zero(tier_out, 8 * (MAX_TIERS + 2)));
    PRP_TRY(be.begin_offsets_ready());
    be.pfor(n, DegF{in.begin, in.r_begin, outdeg, indeg, ids, keyo, keyr, err});
    PRP_TRY(be.sort_u32(keyo, so_keys, ids, so_ids, n, 32));
    PRP_TRY(be.sort_u64(keyr, sr_keys, ids, sr_ids, n, 34));
    be.pfor(1, BoundsF{so_keys, sr_keys, n, small});
    PRP_TRY(be.d2h(h3, small, 12));
    PRP_TRY(be.d2h(&herr, err, 4));
    if (herr) { P.error = "pagerank plan: a degree is negative or >= 2^30"; rc = -1; goto done; }
    P.g_n_long = h3[0]; P.n_pos = h3[1]; P.n_src = h3[2];
    // warm tiers: as many as asked for and as there are sources to fill them
    {
        int T = cfg.tiers;
        if (T > MAX_TIERS - 1) T = MAX_TIERS - 1;            // class T + 1 may be the cold stream
        if ((int64_t)m < cfg.tier_min_edges) T = 0;
        const int full = (P.n_src + HT_REAL - 1) / HT_REAL;      // tiers needed to give every source a slot
        if (T > full - 1) T = full - 1 > 0 ? full - 1 : 0;
        P.T = T; P.NH = (T + 1) * HT;
        P.split = (T >= 1 && cfg.split) ? 1 : 0;
    }
    tlo = P.split ? 0 : 1;
    thi = (!P.split && P.T >= 1 && cfg.cold_stream) ? P.T + 1 : P.T;
    P.thi = thi;
    be.pfor(n, HotF{so_ids, outdeg, hotslot_old, (P.T + 1) * HT_REAL});
    be.pfor((int64_t)(P.T + 1) * 64, TierSumF{so_ids, outdeg, (P.T + 1) * HT_REAL, n, tier_out});
    PRP_TRY(be.alloc((void**)&P.idslot, 4 * (size_t)P.NH));
    PRP_TRY(be.zero(P.idslot, 4 * (size_t)P.NH));
    be.pfor(n, RelabelF{sr_ids, P.new_of_old, P.old_of_new});
    be.pfor(n, ByIdF{P.old_of_new, outdeg, hotslot_old, P.od_flags, P.hotslot, P.idslot, code, P.split});

    be.mark("1 degrees, orders, hot slots, ids");
    // ---- 2. task sizes, the shard's rows
    P.lt = cfg.lt; P.st_cost = cfg.st_cost; P.slice_ovh = cfg.slice_ovh;
    for (P.lt_shift = 0; (1 << P.lt_shift) < P.lt; P.lt_shift++) {}
    P.minpad = roundup4((P.lt + 29) / 30);                    // a long task then touches at most 32 rows
    g_n_slices = (n - P.g_n_long + 31) / 32;
    P.g_nz_slices = P.n_pos > P.g_n_long ? (P.n_pos - P.g_n_long + 31) / 32 : 0;
    P.g_z0 = P.g_n_long + 1 + 32 * P.g_nz_slices;
    if (world > 1) {
        // cuts balance in-degree entries + row_cost per row (the epilogue's remote stores), as in round 1
        int32_t *g_lbeg = nullptr, *g_sbeg = nullptr;
        PRP_TRY(tmp.get(&g_lbeg, (size_t)P.g_n_long + 1)); PRP_TRY(tmp.get(&g_sbeg, (size_t)g_n_slices + 1));
        be.pfor(n, GlobalLenF{P.old_of_new, indeg, n, P.g_n_long, g_n_slices, P.slice_ovh, g_lbeg, g_sbeg});
        std::vector<int32_t> hl((size_t)P.g_n_long + 1), hs((size_t)g_n_slices + 1);
        PRP_TRY(be.d2h(hl.data(), g_lbeg, 4 * hl.size())); PRP_TRY(be.d2h(hs.data(), g_sbeg, 4 * hs.size()));
        std::vector<int64_t> wl((size_t)P.g_n_long + 1), ws((size_t)P.g_nz_slices + 1);
        wl[0] = 0; ws[0] = 0;
        for (int32_t j = 0; j < P.g_n_long; j++) wl[j + 1] = wl[j] + hl[j] + cfg.row_cost;
        for (int32_t q = 0; q < P.g_nz_slices; q++) { const int32_t len = hs[q] >> 5; ws[q + 1] = ws[q] + 32LL * (((len + 7) & ~7) + cfg.row_cost); }
        shard_cuts(wl.data(), P.g_n_long, world, cutsL.data());
        shard_cuts(ws.data(), P.g_nz_slices, world, cutsS.data());
    } else { cutsL[0] = 0; cutsL[1] = P.g_n_long; cutsS[0] = 0; cutsS[1] = P.g_nz_slices; }
    {
        const int32_t nz = n + 1 - P.g_z0 > 0 ? n + 1 - P.g_z0 : 0;                    // ids [g_z0, N] have no in-edges
        const int32_t per = ((nz + world - 1) / world + ZT - 1) / ZT * ZT;
        for (int r = 0; r <= world; r++) { const int64_t c = (int64_t)P.g_z0 + (int64_t)r * per; cutsZ[r] = (int32_t)(c < n + 1 ? c : n + 1); }
        cutsZ[world] = n + 1;
    }
    P.seg_off.assign((size_t)world * 3, 0); P.seg_len.assign((size_t)world * 3, 0);
    for (int r = 0; r < world; r++) {
        P.seg_off[r * 3 + 0] = 1 + cutsL[r];                          P.seg_len[r * 3 + 0] = cutsL[r + 1] - cutsL[r];
        const int64_t sa = (int64_t)P.g_n_long + 1 + 32LL * cutsS[r], sb = (int64_t)P.g_n_long + 1 + 32LL * cutsS[r + 1];
        P.seg_off[r * 3 + 1] = sa;                                    P.seg_len[r * 3 + 1] = (sb < n + 1 ? sb : n + 1) - sa;
        P.seg_off[r * 3 + 2] = cutsZ[r];                              P.seg_len[r * 3 + 2] = cutsZ[r + 1] - cutsZ[r];
    }
    P.jl0 = cutsL[rank]; P.n_long = nl = cutsL[rank + 1] - cutsL[rank];
    P.s0 = cutsS[rank]; P.n_slices = ns = cutsS[rank + 1] - cutsS[rank];
    P.z0 = cutsZ[rank]; P.z_end = cutsZ[rank + 1];
    P.n_ztasks = (P.z_end - P.z0 + ZT - 1) / ZT;
    nrows = nl + 32 * ns;
    short_id0 = P.g_n_long + 1 + 32 * P.s0;
    if (nrows >= (1 << 27)) { P.error = "pagerank plan: more than 2^27 rows in one shard"; rc = -1; goto done; }
    rc_ = RowCtx{(!P.split && P.T >= 1 && cfg.hot_first) ? 1 : 0, n, nl, P.jl0, short_id0, tlo, P.T, thi, cfg.kmin, cfg.kcold, P.old_of_new, in.r_begin, in.r_node_idx, code};

    be.mark("2 shard cuts");
    // ---- 3. pass 1 over the shard's reverse edges: (row, tier) counts -> segments, home lengths
    PRP_TRY(tmp.get(&segmask, (size_t)nrows + 1)); PRP_TRY(tmp.get(&home_len, (size_t)nrows + 32)); PRP_TRY(tmp.get(&home_hot, (size_t)nrows + 32)); PRP_TRY(tmp.get(&nrec, 1));
    PRP_TRY(be.zero(nrec, 4)); PRP_TRY(be.zero(home_len, 4 * ((size_t)nrows + 32)));
    {
        // a row has at most one record per tier, and a record stands for >= 1 (tier 0) / >= kmin entries
        const int64_t cap = tlo <= P.T ? std::min<int64_t>((int64_t)(thi + 1 - tlo) * nrows, (int64_t)m / (tlo == 0 ? 1 : cfg.kmin) + 1) + 1 : 1;
        if (cap > 0x7fffff00LL) { P.error = "pagerank plan: too many segment records"; rc = -1; goto done; }
        rec_cap = (int32_t)cap;
    }
    PRP_TRY(tmp.get(&rec_key, (size_t)rec_cap)); PRP_TRY(tmp.get(&rec_key2, (size_t)rec_cap));
    PRP_TRY(tmp.get(&rec_j, (size_t)rec_cap)); PRP_TRY(tmp.get(&rec_j2, (size_t)rec_cap));
    nps = (int32_t)std::min<int64_t>((int64_t)m / PIECE + nl + 2, 0x7fffff00LL);
    PRP_TRY(tmp.get(&np, (size_t)nl + 1)); PRP_TRY(tmp.get(&piece_row, (size_t)nps));
    PRP_TRY(tmp.get(&G, (size_t)(thi + 1) * nps));
    be.pfor((int64_t)nl + 1, PieceNumF{rc_, np});
    PRP_TRY(be.scan_i32(np, (size_t)nl + 1));                               // np = piece_first from here on
    be.pfor(nps, PieceRowF{np, nl, piece_row});
    if (tlo <= P.T) PRP_TRY(be.zero(G, 4 * (size_t)(thi + 1) * nps));
    if (tlo > P.T) be.pfor(nrows, HomeLenDegF{rc_, nrows, segmask, home_len, home_hot});        // no tiers: the layout follows from the degrees
    else {
        // the reverse edges may still be arriving over PCIe (gmc_pagerank_host_csr): count chunk by chunk behind the copies
        const int nch = be.edge_chunks();
        int32_t e_lo = 0;
        for (int ch = 0; ch < nch; ch++) {
            int32_t e_hi = m;
            PRP_TRY(be.edge_chunk_ready(ch, &e_hi));
            if (ch == nch - 1) e_hi = m > 0 ? m : 1;
            if (e_hi > e_lo || ch == 0) {
                be.pfor((int64_t)32 * ns, CountShortF{rc_, segmask, home_len, home_hot, nrec, rec_j, err, rec_key, rec_cap, e_lo, e_hi > e_lo ? e_hi : e_lo + 1});
                if (tlo <= P.T) be.pfor(nps, PieceCountF{rc_, np, piece_row, nps, G, e_lo, e_hi});
            }
            if (e_hi > e_lo) e_lo = e_hi;
        }
    }
    if (tlo <= P.T) PRP_TRY(be.scan_i32(G, (size_t)(thi + 1) * nps));
    if (tlo <= P.T) be.pfor(nl, LongRowF{rc_, np, G, nps, segmask, home_len, home_hot, nrec, rec_j, err, rec_key, rec_cap});
    PRP_TRY(be.d2h(&nseg, nrec, 4));
    PRP_TRY(be.d2h(&herr, err, 4));
    if (herr) { P.error = "pagerank plan: segment record buffer overflow"; rc = -1; goto done; }
    P.n_seg = nseg;

    be.mark("3 pass 1 (counts, segments)");
    // ---- 4. segments sorted by (tier, length descending, row); split rows and their part slots
    PRP_TRY(tmp.get(&tier_q, MAX_TIERS + 3)); PRP_TRY(tmp.get(&tier_mid, MAX_TIERS + 3));
    PRP_TRY(be.zero(tier_q, 4 * (MAX_TIERS + 3))); PRP_TRY(be.zero(tier_mid, 4 * (MAX_TIERS + 3)));
    if (nseg > 0) PRP_TRY(be.sort_u64(rec_key, rec_key2, rec_j, rec_j2, nseg, 64));
    be.pfor(thi + 2, TierBoundF{rec_key2, nseg, thi, tier_q, tier_mid});
    PRP_TRY(tmp.get(&nparts, (size_t)nrows + 1)); PRP_TRY(tmp.get(&issplit, (size_t)nrows + 1));
    be.pfor((int64_t)nrows + 1, PartsF{segmask, nrows, nparts, issplit});
    PRP_TRY(be.scan_i32(nparts, (size_t)nrows + 1));                        // nparts = pslot from here on
    PRP_TRY(be.scan_i32(issplit, (size_t)nrows + 1));                       // issplit = kidx
    {
        int32_t two[2];
        PRP_TRY(be.d2h(&two[0], nparts + nrows, 4)); PRP_TRY(be.d2h(&two[1], issplit + nrows, 4));
        P.n_parts = two[0]; P.K = two[1];
    }
    PRP_TRY(be.d2h(h_tq.data(), tier_q, 4 * (MAX_TIERS + 3))); PRP_TRY(be.d2h(h_tm.data(), tier_mid, 4 * (MAX_TIERS + 3)));
    PRP_TRY(be.alloc((void**)&P.pbeg, 4 * ((size_t)P.K + 1))); PRP_TRY(be.alloc((void**)&P.prow, 4 * ((size_t)P.K + 1)));
    be.pfor((int64_t)nrows + 1, SplitF{rc_, segmask, nparts, issplit, nrows, P.pbeg, P.prow, P.home_slot, P.od_flags});

    be.mark("4 segment sort, split rows");
    // ---- 5. entry offsets of everything: [home long rows | home slices | sorted segments] + a trailing 0, one scan
    X = nl + ns + nseg;
    P.n_eoff = X + 1;
    PRP_TRY(be.alloc((void**)&P.eoff, 4 * ((size_t)X + 1)));
    elen = P.eoff;
    PRP_TRY(be.zero(elen + X, 4));
    be.pfor(nl, HomeLongLenF{home_len, P.minpad, elen});
    be.pfor(ns, HomeSliceLenF{home_len, nl, elen});
    be.pfor(nseg, SegLenF{rec_key2, tier_q, tier_mid, elen + nl + ns});
    PRP_TRY(be.scan_i32(elen, (size_t)X + 1));
    PRP_TRY(be.alloc((void**)&P.seg_slot, 4 * ((size_t)nseg + 1)));
    PRP_TRY(tmp.get(&seg_ebase, (size_t)P.n_parts + 1)); PRP_TRY(tmp.get(&seg_estride, (size_t)P.n_parts + 1));
    be.pfor(nseg, SegDestF{rec_key2, rec_j2, tier_mid, nparts, P.eoff + nl + ns, P.seg_slot, seg_ebase, seg_estride});

    be.mark("5 entry offsets, segment destinations");
    // ---- 6. stream table (host) -- needs the entry offsets at the stream boundaries
    {
        Stream* S = P.streams;
        memset(S, 0, sizeof(P.streams));
        // eoff indices: home long [0, nl], home slices [nl, nl + ns], tier t long [nl+ns+tq[t], nl+ns+tm[t]], sliced from there to tq[t+1]
        h_idx.clear();
        h_idx.push_back(0); h_idx.push_back(nl); h_idx.push_back(nl + ns);
        for (int t = 0; t <= thi; t++) { h_idx.push_back(nl + ns + h_tq[t]); h_idx.push_back(nl + ns + h_tm[t]); h_idx.push_back(nl + ns + h_tq[t + 1]); }      // stream t + 1
        h_idx.push_back(X);
        h_val.assign(h_idx.size(), 0);
        PRP_TRY(tmp.get(&gidx, h_idx.size())); PRP_TRY(tmp.get(&gout, h_idx.size()));
        PRP_TRY(be.h2d(gidx, h_idx.data(), 4 * h_idx.size()));
        be.pfor((int64_t)h_idx.size(), GatherF{P.eoff, gidx, gout});
        PRP_TRY(be.d2h(h_val.data(), gout, 4 * h_val.size()));
        P.EN = h_val.back();
        if (P.EN < 0) { P.error = "pagerank plan: entry array exceeds int32 offsets"; rc = -1; goto done; }
        int32_t sb = 0, lto = 0, sto = 0;
        for (int s = 0; s <= thi + 1; s++) {
            Stream& st = S[s];
            int32_t i_long, i_mid, i_end, v_long, v_mid, v_end;
            if (s == 0) { i_long = 0; i_mid = nl; i_end = nl + ns; v_long = h_val[0]; v_mid = h_val[1]; v_end = h_val[2]; }
            else { i_long = h_idx[3 * s]; i_mid = h_idx[3 * s + 1]; i_end = h_idx[3 * s + 2]; v_long = h_val[3 * s]; v_mid = h_val[3 * s + 1]; v_end = h_val[3 * s + 2]; }
            st.lbeg_off = i_long; st.n_long = i_mid - i_long; st.e_long = v_long; st.EL = v_mid - v_long;
            st.n_slices = s == 0 ? ns : (i_end - i_mid + 31) / 32;
            st.sbeg_off = sb; sb += st.n_slices + 1;
            st.n_ltasks = (st.EL + P.lt - 1) / P.lt;
            st.lt_off = lto; lto += st.n_ltasks + 1;
            const int64_t cost_total = (int64_t)(v_end - v_mid) + (int64_t)P.slice_ovh * st.n_slices;
            if (cost_total > 0x7fffff00LL) { P.error = "pagerank plan: slice costs exceed int32"; rc = -1; goto done; }
            int32_t n_big = (int32_t)((cost_total + P.st_cost - 1) / P.st_cost);
            int32_t n_st = n_big;
            if (s == 0 && cfg.tail_shift > 0) {      // home is the stream every CTA ends with: small tasks at its tail
                n_big = (int32_t)(cost_total * cfg.tail_big16 / 16 / P.st_cost);
                const int64_t rest = cost_total - (int64_t)n_big * P.st_cost;
                const int32_t small_cost = P.st_cost >> cfg.tail_shift;
                n_st = n_big + (int32_t)((rest + small_cost - 1) / small_cost);
            }
            st.n_stasks = n_st;
            st.n_big = n_big;
            st.st_off = sto; sto += st.n_stasks + 1;
            st.meta_off = s == 0 ? 0 : h_tq[s - 1];
            st.n_rows = s == 0 ? nrows : h_tq[s] - h_tq[s - 1];
            st.hot_off = s == 0 ? 0 : (s - 1) * HT;
            st.i_mid = i_mid;
        }
        P.n_sbeg = sb; P.n_ltslots = lto; P.n_stslots = sto;
        PRP_TRY(be.alloc((void**)&P.sbeg, 4 * (size_t)sb)); PRP_TRY(tmp.get(&scost, (size_t)sb));
        PRP_TRY(be.alloc((void**)&P.ltask_row, 4 * (size_t)lto)); PRP_TRY(be.alloc((void**)&P.stask_slice, 4 * (size_t)sto));
        for (int s = 0; s <= thi + 1; s++) {
            Stream& st = S[s];
            const int32_t i_end = s == 0 ? nl + ns : h_idx[3 * s + 2];
            be.pfor((int64_t)st.n_slices + 1, SbegF{P.eoff, st.i_mid, s == 0 ? 1 : 32, st.n_slices, i_end, P.sbeg + st.sbeg_off});
            be.pfor((int64_t)st.n_slices + 1, SliceCostF{P.sbeg + st.sbeg_off, st.n_slices, P.slice_ovh, scost + st.sbeg_off});
        }
        PRP_TRY(be.scan_i32(scost, (size_t)sb));
        for (int s = 0; s <= thi + 1; s++) {
            Stream& st = S[s];
            be.pfor((int64_t)st.n_ltasks + 1, LtaskF{P.eoff + st.lbeg_off, st.n_long, st.n_ltasks, P.lt, st.e_long, P.ltask_row + st.lt_off});
            be.pfor((int64_t)st.n_stasks + 1, StaskF{scost + st.sbeg_off, st.n_slices, st.n_stasks, P.st_cost, st.n_big, s == 0 ? cfg.tail_shift : 0, P.stask_slice + st.st_off});
        }
        // statistics for the schedule: real entries per stream
        PRP_TRY(be.d2h(h_tier_out.data(), tier_out, 8 * (MAX_TIERS + 2)));
    }

    be.mark("6 stream table, tasks");
    // ---- 7. pass 2: the entries
    PRP_TRY(be.alloc((void**)&P.E, 4 * ((size_t)P.EN + 8)));
    PRP_TRY(be.zero(P.E, 4 * ((size_t)P.EN + 8)));
    {
        FillCtx f{rc_, segmask, nparts, seg_ebase, seg_estride, P.new_of_old, P.eoff, home_hot, P.E};
        if (tlo > P.T) {
            // no tier streams: nothing above needed the edges, so the entries are written chunk by chunk behind their copy-in
            const int nch = be.edge_chunks();
            int32_t e_lo = 0;
            for (int ch = 0; ch < nch; ch++) {
                int32_t e_hi = m;
                PRP_TRY(be.edge_chunk_ready(ch, &e_hi));
                if (ch == nch - 1) e_hi = m;
                if (e_hi > e_lo) {
                    be.pfor((int64_t)32 * ns, FillShortF{f, ns, e_lo, e_hi});
                    be.pfor(nps, FillPieceF{f, np, piece_row, G, nps, e_lo, e_hi});
                    e_lo = e_hi;
                }
            }
        } else {
            be.pfor((int64_t)32 * ns, FillShortF{f, ns, 0, 0x7fffffff});
            be.pfor(nps, FillPieceF{f, np, piece_row, G, nps, 0, 0x7fffffff});
        }
    }
    // real entries per tier stream = sum of its segment lengths: long part from the offsets is padded, so walk the keys on the host
    // only for small plans; otherwise estimate from the padded slots (<= 3 % off)
    {
        Stream* S = P.streams;
        int64_t seg_total = 0;
        for (int s = 1; s <= thi + 1; s++) {
            const int64_t slots = (int64_t)h_val[3 * s + 2] - h_val[3 * s];
            S[s].entries = (int32_t)std::min<int64_t>(slots, 0x7fffffff);
            if (s == thi + 1 && thi > P.T) P.cold_stream_entries = slots; else seg_total += slots;
        }
        P.seg_entries = seg_total;
        const double frac = world > 1 ? (double)(nl + ns * 32) / (double)(P.n_pos > 0 ? P.n_pos : 1) : 1.0;
        P.hot0_entries = (int64_t)((double)h_tier_out[0] * (frac < 1.0 ? frac : 1.0));
        const int64_t home_slots = (int64_t)h_val[2] - h_val[0];
        S[0].entries = (int32_t)std::min<int64_t>(home_slots, 0x7fffffff);
        // SPLIT: home holds cold entries only; MIXED: the tier-0 entries are in the home rows too
        P.cold_entries = P.split ? home_slots : (home_slots > P.hot0_entries ? home_slots - P.hot0_entries : 0);
    }

    be.mark("7 pass 2 (entries)");
    // ---- 8. diff_parts layout and the CTA schedule
    P.dp_short = P.streams[0].n_ltasks;
    P.dp_zero = P.dp_short + P.streams[0].n_stasks;
    P.dp_cut = P.dp_zero + P.n_ztasks;
    P.dp_total = P.dp_cut + P.streams[0].n_ltasks + 1;
    make_schedule(cfg, P);
    PRP_TRY(be.alloc((void**)&P.sched_dev, 4 * P.sched.size()));
    PRP_TRY(be.h2d(P.sched_dev, P.sched.data(), 4 * P.sched.size()));
    PRP_TRY(be.alloc((void**)&P.streams_dev, sizeof(Stream) * (MAX_TIERS + 2)));
    PRP_TRY(be.h2d(P.streams_dev, P.streams, sizeof(Stream) * (MAX_TIERS + 2)));
    PRP_TRY(be.sync());
    be.mark("8 schedule, end");
done:
    if (rc != 0 && P.error.empty()) P.error = be.last_error();
    return rc;
}

}  // namespace prp
