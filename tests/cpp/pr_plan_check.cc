// CPU check of the PageRank plan builder (green-marl_b200/csrc/gmc_pr_plan.h) -- test infrastructure, not product code.
// The builder's steps are backend-agnostic functors; here they run as host loops on a synthetic R-MAT graph, then
//   1. every reverse edge must appear exactly once across the streams (home + tier segments), in its row,
//   2. 20 sweeps executed THROUGH the plan (hotpack staging, part slots, combine in slot order, hot-slot updates) must agree
//      with the plain pull loop of apps/src/pagerank.gm:8-19 to 1e-12.
// usage: pr_plan_check <log2 nodes> <edges> <tiers> <kmin> <world> [seed]
//   g++ -O2 -fopenmp -std=c++17 -I green-marl_b200/csrc tests/cpp/pr_plan_check.cc -o pr_plan_check
#include "gmc_pr_plan.h"
#include <math.h>
#include <stdio.h>
#include <numeric>
#include <random>

struct CpuBackend {
    std::string err;
    int alloc(void** p, size_t bytes) { *p = malloc(bytes ? bytes : 8); if (!*p) { err = "out of memory"; return -1; } return 0; }
    void free(void* p) { ::free(p); }
    int zero(void* p, size_t b) { memset(p, 0, b); return 0; }
    int d2h(void* d, const void* s, size_t b) { memcpy(d, s, b); return 0; }
    int h2d(void* d, const void* s, size_t b) { memcpy(d, s, b); return 0; }
    int sync() { return 0; }
    void mark(const char*) {}
    int begin_offsets_ready() { return 0; }
    int edges_ready() { return 0; }
    int n_chunks = 1, m_edges = 0;      // the check also runs the chunked pass 1 (edges arriving in pieces)
    int edge_chunks() { return n_chunks; }
    int edge_chunk_ready(int c, int32_t* e_end) { *e_end = (int32_t)((int64_t)m_edges * (c + 1) / n_chunks); return 0; }
    std::string last_error() { return err; }
    template <class F>
    void pfor(int64_t n, const F& f) {
#pragma omp parallel for schedule(dynamic, 256)
        for (int64_t i = 0; i < n; i++) f(i);
    }
    int scan_i32(int32_t* a, size_t n) {
        int32_t run = 0;
        for (size_t i = 0; i < n; i++) { const int32_t x = a[i]; a[i] = run; run += x; }
        return 0;
    }
    template <typename K>
    int sort_any(const K* kin, K* kout, const int32_t* vin, int32_t* vout, int32_t n) {
        std::vector<int32_t> perm((size_t)n);
        std::iota(perm.begin(), perm.end(), 0);
        std::stable_sort(perm.begin(), perm.end(), [&](int32_t a, int32_t b) { return kin[a] < kin[b]; });
        for (int32_t i = 0; i < n; i++) { kout[i] = kin[perm[i]]; vout[i] = vin[perm[i]]; }
        return 0;
    }
    int sort_u32(const uint32_t* kin, uint32_t* kout, const int32_t* vin, int32_t* vout, int32_t n, int) { return sort_any(kin, kout, vin, vout, n); }
    int sort_u64(const unsigned long long* kin, unsigned long long* kout, const int32_t* vin, int32_t* vout, int32_t n, int) { return sort_any(kin, kout, vin, vout, n); }
};

static void free_plan(prp::Plan& P) {
    void* a[] = {P.new_of_old, P.old_of_new, P.od_flags, P.hotslot, P.home_slot, P.idslot, P.E, P.eoff, P.sbeg, P.ltask_row, P.stask_slice,
                 P.seg_slot, P.pbeg, P.prow, P.sched_dev, P.streams_dev};
    for (void* p : a) free(p);
}

int main(int argc, char** argv) {
    const int ln = argc > 1 ? atoi(argv[1]) : 14;
    const int64_t m_req = argc > 2 ? atoll(argv[2]) : (1 << 18);
    const int tiers = argc > 3 ? atoi(argv[3]) : 3, kmin = argc > 4 ? atoi(argv[4]) : 3, world = argc > 5 ? atoi(argv[5]) : 1;
    const unsigned seed = argc > 6 ? (unsigned)atoi(argv[6]) : 7u;
    const int32_t n = 1 << ln;
    // R-MAT (a, b, c) = (0.57, 0.19, 0.19), vertex permutation, duplicates kept (the reference generator keeps them too)
    std::mt19937_64 rng(seed);
    std::uniform_real_distribution<double> U(0.0, 1.0);
    std::vector<int32_t> perm((size_t)n);
    std::iota(perm.begin(), perm.end(), 0);
    std::shuffle(perm.begin(), perm.end(), rng);
    std::vector<int32_t> src((size_t)m_req), dst((size_t)m_req);
    for (int64_t e = 0; e < m_req; e++) {
        int32_t u = 0, v = 0;
        for (int l = 0; l < ln; l++) {
            const double x = U(rng);
            const int q = x < 0.57 ? 0 : (x < 0.76 ? 1 : (x < 0.95 ? 2 : 3));
            u = (u << 1) | (q >> 1); v = (v << 1) | (q & 1);
        }
        src[e] = perm[u]; dst[e] = perm[v];
    }
    const int32_t m = (int32_t)m_req;
    std::vector<int32_t> begin((size_t)n + 1, 0), r_begin((size_t)n + 1, 0), r_idx((size_t)m);
    for (int32_t e = 0; e < m; e++) { begin[src[e] + 1]++; r_begin[dst[e] + 1]++; }
    for (int32_t v = 0; v < n; v++) { begin[v + 1] += begin[v]; r_begin[v + 1] += r_begin[v]; }
    {
        std::vector<int32_t> pos(r_begin.begin(), r_begin.end() - 1);
        for (int32_t e = 0; e < m; e++) r_idx[pos[dst[e]]++] = src[e];
        for (int32_t v = 0; v < n; v++) std::sort(r_idx.begin() + r_begin[v], r_idx.begin() + r_begin[v + 1]);     // semi-sorted
    }
    // plain PageRank (apps/src/pagerank.gm)
    const double d = 0.85;
    const int iters = 20;
    std::vector<double> want((size_t)n, 1.0 / n), nxt((size_t)n);
    for (int it = 0; it < iters; it++) {
#pragma omp parallel for schedule(dynamic, 512)
        for (int32_t t = 0; t < n; t++) {
            double S = 0.0;
            for (int32_t k = r_begin[t]; k < r_begin[t + 1]; k++) { const int32_t w = r_idx[k]; S += want[w] / (double)(begin[w + 1] - begin[w]); }
            nxt[t] = (1 - d) / n + d * S;
        }
        want.swap(nxt);
    }

    CpuBackend be;
    be.m_edges = m; be.n_chunks = getenv("PRP_CHUNKS") ? atoi(getenv("PRP_CHUNKS")) : 1;
    prp::Config cfg;
    cfg.tiers = tiers; cfg.kmin = kmin; cfg.tier_min_edges = 0;
    cfg.split = getenv("PRP_SPLIT") ? atoi(getenv("PRP_SPLIT")) : 1;
    cfg.cold_stream = getenv("PRP_COLD") ? atoi(getenv("PRP_COLD")) : 1;
    cfg.kcold = 2;
    cfg.hot_first = getenv("PRP_HOT_FIRST") ? atoi(getenv("PRP_HOT_FIRST")) : 1;       // the check exercises the optional layouts too
    const bool hot_first = cfg.hot_first != 0;
    cfg.row_cost = world > 2 ? 2 * (world - 2) : 0;
    if (const char* big = getenv("PRP_BIGTASKS")) if (atoi(big)) { cfg.lt = 8192; cfg.st_cost = 4096; cfg.slice_ovh = 136; }
    std::vector<prp::Plan> plans((size_t)world);
    for (int r = 0; r < world; r++) {
        prp::Input in; in.n = n; in.m = m; in.begin = begin.data(); in.r_begin = r_begin.data(); in.r_node_idx = r_idx.data(); in.world = world; in.rank = r;
        if (prp::build(be, in, cfg, plans[r]) != 0) { printf("FAIL: build rank %d: %s\n", r, plans[r].error.c_str()); return 1; }
    }
    const prp::Plan& P0 = plans[0];
    const int HT = prp::HT;
    int sched_bad = 0;
    printf("N=%d M=%d T=%d world=%d: n_long %d n_pos %d n_src %d | rank0: EN %d (%.3f x M/world) segs %d split rows %d parts %d; sched:", n, m, P0.T, world,
           P0.g_n_long, P0.n_pos, P0.n_src, P0.EN, (double)P0.EN * world / m, P0.n_seg, P0.K, P0.n_parts);
    {   // every tier stream must be on some CTA's list; print CTAs per first stream
        std::vector<int> first(prp::MAX_TIERS + 3, 0), seen(prp::MAX_TIERS + 3, 0);
        for (int b = 0; b < 148; b++) { first[P0.sched[(size_t)b * prp::SCHED_W]]++; for (int i = 0; i < prp::SCHED_W; i++) seen[P0.sched[(size_t)b * prp::SCHED_W + i]] = 1; }
        for (int s2 = 1; s2 <= P0.T + 1; s2++) { printf(" s%d:%d", s2, first[s2]); if (!seen[s2] && (P0.split || s2 == 1)) { printf("(UNSCHEDULED)"); sched_bad++; } }
        printf(" split=%d\n", P0.split);
    }

    // ---- check 1: edge coverage.  got[id] = multiset of source ids found in the streams
    int bad = sched_bad;
    std::vector<std::vector<int32_t>> got((size_t)n + 1);
    std::vector<int32_t> slot_row;      // part slot -> row id (through pbeg / prow)
    for (int r = 0; r < world; r++) {
        const prp::Plan& P = plans[r];
        slot_row.assign((size_t)P.n_parts + 1, -1);
        for (int32_t k = 0; k < P.K; k++) for (int32_t s = P.pbeg[k]; s < P.pbeg[k + 1]; s++) slot_row[s] = P.prow[k];
        auto decode = [&](int s, int32_t x) -> int32_t { return (s == 0 || (s == P.thi + 1 && P.thi > P.T)) ? (x < HT ? P.idslot[x] : x - HT) : P.idslot[(s - 1) * HT + x]; };
        for (int s = 0; s <= P.thi + 1; s++) {
            const prp::Stream& st = P.streams[s];
            if (st.e_long & 3) { printf("FAIL: stream %d long section not 16-byte aligned\n", s); bad++; }
            for (int32_t j = 0; j < st.n_long; j++) {
                const int32_t id = s == 0 ? P.jl0 + 1 + j : slot_row[P.seg_slot[st.meta_off + j]];
                if (id < 1 || id > n) { printf("FAIL: stream %d long row %d -> id %d\n", s, j, id); bad++; continue; }
                bool seen_cold = false;       // MIXED mode with warm tiers: a home row holds its tier-0 entries first (hot_first)
                for (int32_t x = P.eoff[st.lbeg_off + j]; x < P.eoff[st.lbeg_off + j + 1]; x++) {
                    const int32_t w = decode(s, P.E[x]); if (w) got[id].push_back(w);
                    if (s == 0 && hot_first && !P.split && P.T >= 1 && P.E[x]) {
                        if (P.E[x] >= HT) seen_cold = true;
                        else if (seen_cold) { printf("FAIL: home long row %d: a tier-0 entry after an L2 entry\n", j); bad++; seen_cold = false; }
                    }
                }
            }
            for (int32_t q = 0; q < st.n_slices; q++) {
                const int32_t b0 = P.sbeg[st.sbeg_off + q], L = (P.sbeg[st.sbeg_off + q + 1] - b0) >> 5;
                for (int l = 0; l < 32; l++) {
                    int32_t id;
                    if (s == 0) { id = P.g_n_long + 1 + 32 * (P.s0 + q) + l; if (id > n) continue; }
                    else { const int32_t loc = st.n_long + 32 * q + l; if (loc >= st.n_rows) { for (int32_t k = 0; k < L; k++) if (P.E[b0 + 32 * k + l]) { printf("FAIL: entry in a dead lane\n"); bad++; } continue; } id = slot_row[P.seg_slot[st.meta_off + loc]]; }
                    if (id < 1 || id > n) { printf("FAIL: stream %d slice %d lane %d -> id %d\n", s, q, l, id); bad++; continue; }
                    for (int32_t k = 0; k < L; k++) { const int32_t w = decode(s, P.E[b0 + 32 * k + l]); if (w) got[id].push_back(w); }
                }
            }
            // tasks: cuts must be monotone and cover the stream
            for (int32_t t = 0; t < st.n_ltasks; t++) if (P.ltask_row[st.lt_off + t] > P.ltask_row[st.lt_off + t + 1]) { printf("FAIL: long task rows not monotone\n"); bad++; }
            if (st.n_ltasks && P.ltask_row[st.lt_off + st.n_ltasks] != st.n_long) { printf("FAIL: long task sentinel\n"); bad++; }
            for (int32_t t = 0; t < st.n_ltasks; t++) {        // a task spans at most 32 rows
                const int32_t x1 = std::min(st.e_long + (t + 1) * P.lt, st.e_long + st.EL) - 1;
                const int32_t a = P.ltask_row[st.lt_off + t];
                int32_t b = a; while (b < st.n_long && P.eoff[st.lbeg_off + b + 1] <= x1) b++;
                if (b - a + 1 > 32) { printf("FAIL: stream %d long task %d spans %d rows\n", s, t, b - a + 1); bad++; }
            }
            if (P.stask_slice[st.st_off] != 0 || P.stask_slice[st.st_off + st.n_stasks] != st.n_slices) { printf("FAIL: stream %d short task ends %d %d (%d slices, %d tasks)\n", s, P.stask_slice[st.st_off], P.stask_slice[st.st_off + st.n_stasks], st.n_slices, st.n_stasks); bad++; }
            for (int32_t u = 0; u < st.n_stasks; u++) {
                const int32_t a = P.stask_slice[st.st_off + u], b = P.stask_slice[st.st_off + u + 1];
                if (b < a || b - a > 31) { printf("FAIL: stream %d short task %d holds %d slices\n", s, u, b - a); bad++; }
            }
        }
    }
    int64_t miss = 0;
    for (int32_t v = 0; v < n; v++) {
        const int32_t id = P0.new_of_old[v];
        std::vector<int32_t> exp;
        for (int32_t k = r_begin[v]; k < r_begin[v + 1]; k++) exp.push_back(P0.new_of_old[r_idx[k]]);
        std::sort(exp.begin(), exp.end());
        std::sort(got[id].begin(), got[id].end());
        if (exp != got[id]) { if (miss < 5) printf("FAIL: row id %d (old %d): %zu entries expected, %zu found\n", id, v, exp.size(), got[id].size()); miss++; }
    }
    if (miss) { printf("FAIL: %lld rows with wrong entries\n", (long long)miss); bad++; }

    // ---- check 2: 20 sweeps through the plan(s): shared state, every shard's streams, combine in slot order
    const int NH = P0.NH;
    std::vector<double> contrib[2] = {std::vector<double>((size_t)n + 2, 0.0), std::vector<double>((size_t)n + 2, 0.0)};
    std::vector<double> hotp[2] = {std::vector<double>((size_t)NH, 0.0), std::vector<double>((size_t)NH, 0.0)};
    std::vector<double> rk[2] = {std::vector<double>((size_t)n + 2, 0.0), std::vector<double>((size_t)n + 2, 0.0)};
    for (int32_t id = 1; id <= n; id++) {
        rk[0][id] = 1.0 / n;
        const uint32_t of = (uint32_t)P0.od_flags[id];
        const int32_t od = (int32_t)(of & prp::OD_MASK);
        contrib[0][id] = od > 0 ? rk[0][id] / (double)od : 0.0;
        if (of & prp::OD_HOT) hotp[0][P0.hotslot[id]] = contrib[0][id];
    }
    const double base = (1 - d) / n;
    for (int it = 0; it < iters; it++) {
        const int par = it & 1;
        const std::vector<double>&cin = contrib[par], &hin = hotp[par], &rin = rk[par];
        std::vector<double>&cout = contrib[par ^ 1], &hout = hotp[par ^ 1], &rout = rk[par ^ 1];
        (void)rin;
        for (int r = 0; r < world; r++) {
            const prp::Plan& P = plans[r];
            std::vector<double> part((size_t)P.n_parts + 1, 0.0);
            auto finish = [&](int32_t id, double S) {
                const double val = base + d * S;
                rout[id] = val;
                const uint32_t of = (uint32_t)P.od_flags[id];
                const int32_t od = (int32_t)(of & prp::OD_MASK);
                const double c = od > 0 ? val / (double)od : 0.0;
                cout[id] = c;
                if (of & prp::OD_HOT) hout[P.hotslot[id]] = c;
            };
            auto home_finish = [&](int32_t id, double S) { if ((uint32_t)P.od_flags[id] & prp::OD_SPLIT) part[P.home_slot[id]] = S; else finish(id, S); };
            for (int s = 0; s <= P.thi + 1; s++) {
                const prp::Stream& st = P.streams[s];
                auto val = [&](int32_t x) -> double { return (s == 0 || (s == P.thi + 1 && P.thi > P.T)) ? (x < HT ? hin[x] : cin[x - HT]) : hin[(s - 1) * HT + x]; };
                for (int32_t j = 0; j < st.n_long; j++) {
                    double S = 0.0;
                    for (int32_t x = P.eoff[st.lbeg_off + j]; x < P.eoff[st.lbeg_off + j + 1]; x++) S += val(P.E[x]);
                    if (s == 0) home_finish(P.jl0 + 1 + j, S); else part[P.seg_slot[st.meta_off + j]] = S;
                }
                for (int32_t q = 0; q < st.n_slices; q++) {
                    const int32_t b0 = P.sbeg[st.sbeg_off + q], L = (P.sbeg[st.sbeg_off + q + 1] - b0) >> 5;
                    for (int l = 0; l < 32; l++) {
                        double S = 0.0;
                        for (int32_t k = 0; k < L; k++) S += val(P.E[b0 + 32 * k + l]);
                        if (s == 0) { const int32_t id = P.g_n_long + 1 + 32 * (P.s0 + q) + l; if (id <= n) home_finish(id, S); }
                        else { const int32_t loc = st.n_long + 32 * q + l; if (loc < st.n_rows) part[P.seg_slot[st.meta_off + loc]] = S; }
                    }
                }
            }
            for (int32_t id = P.z0; id < P.z_end; id++) finish(id, 0.0);
            for (int32_t k = 0; k < P.K; k++) {
                double S = 0.0;
                for (int32_t s2 = P.pbeg[k]; s2 < P.pbeg[k + 1]; s2++) S += part[s2];
                finish(P.prow[k], S);
            }
        }
    }
    double err = 0.0;
    for (int32_t v = 0; v < n; v++) err = std::max(err, fabs(rk[iters & 1][P0.new_of_old[v]] - want[v]));
    printf("max |plan sweep - plain sweep| after %d iterations: %.3e\n", iters, err);
    if (!(err <= 1e-12)) { printf("FAIL: sweep through the plan differs\n"); bad++; }
    for (auto& P : plans) free_plan(P);
    printf(bad ? "FAILED (%d)\n" : "OK\n", bad);
    return bad ? 1 : 0;
}
