/* TEST INFRASTRUCTURE ONLY -- see gm_oracle.h for the rules and the parity pin.
 * Plain-C restatement of the reference algorithms; every function cites what it follows
 * (paths relative to /root/reference/). */
#define _GNU_SOURCE
#include "gm_oracle.h"
#include <limits.h>
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <arpa/inet.h>
#include <omp.h>

void orc_set_threads(int n) { if (n > 0) omp_set_num_threads(n); }

/* Shared tail of every generator: degree count -> prefix sum -> fill each row FROM THE BACK
 * (pos = degree[u]--), graph_gen.cc:85-97 / :263-279. */
static void edges_to_csr(int32_t N, int32_t M, const int32_t* src, const int32_t* dst,
                         int32_t* begin, int32_t* node_idx) {
    int32_t* degree = (int32_t*)calloc((size_t)N, sizeof(int32_t));
    for (int32_t i = 0; i < M; i++) degree[src[i]]++;
    begin[0] = 0;
    for (int32_t i = 1; i <= N; i++) begin[i] = begin[i - 1] + degree[i - 1];
    for (int32_t i = 0; i < M; i++) {
        int32_t u = src[i];
        int32_t pos = degree[u]--;
        node_idx[begin[u] + pos - 1] = dst[i];
    }
    free(degree);
}

/* graph_gen.cc:58-104 with use_xorshift_rng=false: srand(seed); per edge src=rand()%N then dst=rand()%N. */
int orc_uniform_random(int32_t N, int32_t M, long seed, int32_t* begin, int32_t* node_idx) {
    if (N <= 0 || M < 0) return -1;
    int32_t* src = (int32_t*)malloc(sizeof(int32_t) * (size_t)(M ? M : 1));
    int32_t* dst = (int32_t*)malloc(sizeof(int32_t) * (size_t)(M ? M : 1));
    srand((unsigned)seed);
    for (int32_t i = 0; i < M; i++) {
        int32_t r = rand();
        src[i] = r % N;
        r = rand();
        dst[i] = r % N;
    }
    edges_to_csr(N, M, src, dst, begin, node_idx);
    free(src); free(dst);
    return 0;
}

/* graph_gen.cc:159-287.  SCALE=(node_t)log2(N) :185; first quadrant pick :196-206; per further level the
 * four probabilities are jittered by 0.95+0.1*drand48() and renormalised :207-220; self edges are
 * re-drawn :232-235; permutation :240-258 swaps P[i] with P[(node_t)(N*drand48())] for i ascending. */
int orc_rmat(int32_t N, int32_t M, long seed, double a, double b, double c, int permute,
             int32_t* begin, int32_t* node_idx) {
    if (N <= 0 || M < 0 || !(a + b + c < 1)) return -1;
    double d = 1 - (a + b + c);
    srand48(seed);
    int32_t* src = (int32_t*)malloc(sizeof(int32_t) * (size_t)(M ? M : 1));
    int32_t* dst = (int32_t*)malloc(sizeof(int32_t) * (size_t)(M ? M : 1));
    int32_t SCALE = (int32_t)log2((double)N);
    for (int32_t i = 0; i < M; i++) {
        int32_t u = 1, v = 1, step = N / 2;
        double av = a, bv = b, cv = c, dv = d;
        double p = drand48();
        if (p < av) { }
        else if (p < (av + bv)) { v += step; }
        else if (p < (av + bv + cv)) { u += step; }
        else { v += step; u += step; }
        for (int32_t j = 1; j < SCALE; j++) {
            step = step / 2;
            double var = 0.1;
            av *= 0.95 + var * drand48();
            bv *= 0.95 + var * drand48();
            cv *= 0.95 + var * drand48();
            dv *= 0.95 + var * drand48();
            double S = av + bv + cv + dv;
            av = av / S; bv = bv / S; cv = cv / S; dv = dv / S;
            p = drand48();
            if (p < av) { }
            else if (p < (av + bv)) { v += step; }
            else if (p < (av + bv + cv)) { u += step; }
            else { v += step; u += step; }
        }
        src[i] = u - 1;
        dst[i] = v - 1;
        if (src[i] == dst[i]) { i = i - 1; continue; }
    }
    if (permute) {
        int32_t* P = (int32_t*)malloc(sizeof(int32_t) * (size_t)N);
        for (int32_t i = 0; i < N; i++) P[i] = i;
        for (int32_t i = 0; i < N; i++) {
            int32_t j = (int32_t)(N * drand48());
            int32_t t = P[j]; P[j] = P[i]; P[i] = t;
        }
        for (int32_t i = 0; i < M; i++) { src[i] = P[src[i]]; dst[i] = P[dst[i]]; }
        free(P);
    }
    edges_to_csr(N, M, src, dst, begin, node_idx);
    free(src); free(dst);
    return 0;
}

/* semi_sort_main, gm_graph.cc:380-423: every row sorted ascending by neighbour id, aux permuted
 * alongside.  The reference sorts an index vector with std::sort (unstable); ties here keep their
 * original order (a legal outcome of the reference's). */
typedef struct { int32_t key, aux, pos; } sort_rec;
static int rec_cmp(const void* x, const void* y) {
    const sort_rec* a = (const sort_rec*)x; const sort_rec* b = (const sort_rec*)y;
    if (a->key != b->key) return a->key < b->key ? -1 : 1;
    return a->pos < b->pos ? -1 : (a->pos > b->pos);
}
void orc_semi_sort(int32_t N, const int32_t* begin, int32_t* node_idx, int32_t* aux) {
    #pragma omp parallel
    {
        sort_rec* buf = NULL; int32_t cap = 0;
        #pragma omp for schedule(static, 4096) nowait
        for (int32_t i = 0; i < N; i++) {
            int32_t sz = begin[i + 1] - begin[i];
            if (sz < 2) continue;
            if (sz > cap) { cap = sz * 2; buf = (sort_rec*)realloc(buf, sizeof(sort_rec) * (size_t)cap); }
            int32_t* row = node_idx + begin[i];
            int32_t* arow = aux ? aux + begin[i] : NULL;
            for (int32_t j = 0; j < sz; j++) { buf[j].key = row[j]; buf[j].aux = arow ? arow[j] : 0; buf[j].pos = j; }
            qsort(buf, (size_t)sz, sizeof(sort_rec), rec_cmp);
            for (int32_t j = 0; j < sz; j++) { row[j] = buf[j].key; if (arow) arow[j] = buf[j].aux; }
        }
        free(buf);
    }
}

/* make_reverse_edges, gm_graph.cc:205-304: in-degree count, exclusive scan (:249-255), scatter
 * (:268-280), then do_semi_sort_reverse (:293).  Visiting sources in ascending order makes every
 * reverse row come out already sorted by source id, which is the post-sort state of the reference. */
void orc_make_reverse(int32_t N, int32_t M, const int32_t* begin, const int32_t* node_idx,
                      int32_t* r_begin, int32_t* r_node_idx, int32_t* e_rev2idx) {
    for (int32_t i = 0; i <= N; i++) r_begin[i] = 0;
    for (int32_t e = 0; e < M; e++) r_begin[node_idx[e]]++;
    int32_t sum = 0;
    for (int32_t i = 0; i < N; i++) { int32_t old = sum; sum += r_begin[i]; r_begin[i] = old; }
    r_begin[N] = sum;
    int32_t* cursor = (int32_t*)malloc(sizeof(int32_t) * (size_t)(N ? N : 1));
    if (N > 0) memcpy(cursor, r_begin, sizeof(int32_t) * (size_t)(uint32_t)N);
    for (int32_t i = 0; i < N; i++)
        for (int32_t e = begin[i]; e < begin[i + 1]; e++) {
            int32_t slot = cursor[node_idx[e]]++;
            r_node_idx[slot] = i;
            if (e_rev2idx) e_rev2idx[slot] = e;
        }
    free(cursor);
}

/* store_binary_internal, gm_graph_binary_loader.cc:222-252: big-endian words
 * [magic 0x03939999][sizeof node_t][sizeof edge_t][N][M][begin[N+1]][node_idx[M]]. */
int orc_store_binary(const char* path, int32_t N, int32_t M, const int32_t* begin, const int32_t* node_idx) {
    FILE* f = fopen(path, "wb");
    if (!f) return 0;
    uint32_t hdr[5] = { htonl(0x03939999u), htonl(4), htonl(4), htonl((uint32_t)N), htonl((uint32_t)M) };
    fwrite(hdr, 4, 5, f);
    for (int32_t i = 0; i <= N; i++) { uint32_t w = htonl((uint32_t)begin[i]); fwrite(&w, 4, 1, f); }
    for (int32_t i = 0; i < M; i++) { uint32_t w = htonl((uint32_t)node_idx[i]); fwrite(&w, 4, 1, f); }
    fclose(f);
    return 1;
}
/* load_binary_internal, gm_graph_binary_loader.cc:57-205 (32/32 files only; the "old flipped"
 * host-endian variant :78-80 is accepted the same way the reference does). */
static int read_header(FILE* f, int32_t* N, int32_t* M, int* flipped) {
    uint32_t hdr[5];
    if (fread(hdr, 4, 5, f) != 5) return 0;
    if (ntohl(hdr[0]) == 0x03939999u) *flipped = 0;
    else if (hdr[0] == 0x03939999u) *flipped = 1;
    else return 0;
    uint32_t ns = *flipped ? hdr[1] : ntohl(hdr[1]);
    uint32_t es = *flipped ? hdr[2] : ntohl(hdr[2]);
    if (ns != 4 || es != 4) return 0;
    *N = (int32_t)(*flipped ? hdr[3] : ntohl(hdr[3]));
    *M = (int32_t)(*flipped ? hdr[4] : ntohl(hdr[4]));
    return 1;
}
int orc_load_binary_header(const char* path, int32_t* N, int32_t* M) {
    FILE* f = fopen(path, "rb");
    if (!f) return 0;
    int flipped, ok = read_header(f, N, M, &flipped);
    fclose(f);
    return ok;
}
int orc_load_binary(const char* path, int32_t* begin, int32_t* node_idx) {
    FILE* f = fopen(path, "rb");
    if (!f) return 0;
    int32_t N, M; int flipped;
    if (!read_header(f, &N, &M, &flipped)) { fclose(f); return 0; }
    int ok = fread(begin, 4, (size_t)N + 1, f) == (size_t)N + 1 && fread(node_idx, 4, (size_t)M, f) == (size_t)M;
    fclose(f);
    if (!ok) return 0;
    if (!flipped) {
        for (int32_t i = 0; i <= N; i++) begin[i] = (int32_t)ntohl((uint32_t)begin[i]);
        for (int32_t i = 0; i < M; i++) node_idx[i] = (int32_t)ntohl((uint32_t)node_idx[i]);
    }
    return 1;
}

/* Generated pagerank(): apps/src/pagerank.gm:6-19.  Loop shape per the emitters cited in
 * ref_glue/pagerank.cc: rank=1/N; do { diff=0; foreach t (dynamic,128): S=sum over in-nbrs w in
 * r_node_idx order of rank[w]/(double)(begin[w+1]-begin[w]); val=(1-d)/N+d*S; diff_prv+=|val-rank[t]|;
 * rank_nxt[t]=val; ATOMIC_ADD(diff,diff_prv); rank=rank_nxt; cnt++ } while (diff>e && cnt<max).
 * The per-row sum is strictly left-to-right, so rank is bit-identical for any thread count. */
int orc_pagerank(int32_t N, const int32_t* begin, const int32_t* r_begin, const int32_t* r_node_idx,
                 double e, double d, int32_t max_iter, double* rank, double* diff_out) {
    double* nxt = (double*)malloc(sizeof(double) * (size_t)(N ? N : 1));
    double Nd = (double)N;
    double diff = 0.0;
    int32_t cnt = 0;
    #pragma omp parallel for
    for (int32_t t = 0; t < N; t++) rank[t] = 1 / Nd;
    do {
        diff = 0.0;
        #pragma omp parallel
        {
            double diff_prv = 0.0;
            #pragma omp for nowait schedule(dynamic, 128)
            for (int32_t t = 0; t < N; t++) {
                double S = 0.0;
                for (int32_t w_idx = r_begin[t]; w_idx < r_begin[t + 1]; w_idx++) {
                    int32_t w = r_node_idx[w_idx];
                    S = S + rank[w] / ((double)(begin[w + 1] - begin[w]));
                }
                double val = (1 - d) / Nd + d * S;
                diff_prv = diff_prv + fabs(val - rank[t]);
                nxt[t] = val;
            }
            #pragma omp atomic
            diff += diff_prv;
        }
        #pragma omp parallel for
        for (int32_t i = 0; i < N; i++) rank[i] = nxt[i];
        cnt = cnt + 1;
    } while ((diff > e) && (cnt < max_iter));
    free(nxt);
    if (diff_out) *diff_out = diff;
    return cnt;
}

/* Generated hop_dist(): apps/src/hop_dist.gm:9-30.  The reference's `min=` under a hashed spinlock
 * (gm_cpp_gen.cc:1563-1741, gm_lock.cc:48-58) is an atomic min with a flag; a CAS-min gives the same
 * dist_nxt/updated_nxt because every writer of one sweep writes the same value dist[n]+1 = sweep no. */
int orc_hop_dist(int32_t N, const int32_t* begin, const int32_t* node_idx, int32_t root, int32_t* dist) {
    unsigned char* updated = (unsigned char*)malloc((size_t)(N ? N : 1));
    unsigned char* updated_nxt = (unsigned char*)malloc((size_t)(N ? N : 1));
    int32_t* dist_nxt = (int32_t*)malloc(sizeof(int32_t) * (size_t)(N ? N : 1));
    int sweeps = 0, fin = 0;
    #pragma omp parallel for
    for (int32_t t = 0; t < N; t++) {
        dist[t] = (t == root) ? 0 : INT_MAX;
        updated[t] = (t == root);
        dist_nxt[t] = dist[t];
        updated_nxt[t] = updated[t];
    }
    while (!fin) {
        int any = 0;
        fin = 1;
        #pragma omp parallel for schedule(dynamic, 128)
        for (int32_t n = 0; n < N; n++) {
            if (!updated[n]) continue;
            for (int32_t s_idx = begin[n]; s_idx < begin[n + 1]; s_idx++) {
                int32_t s = node_idx[s_idx];
                int32_t cand = dist[n] + 1;
                int32_t cur = __atomic_load_n(&dist_nxt[s], __ATOMIC_RELAXED);
                while (cur > cand) {
                    if (__atomic_compare_exchange_n(&dist_nxt[s], &cur, cand, 0, __ATOMIC_RELAXED, __ATOMIC_RELAXED)) {
                        updated_nxt[s] = 1;
                        break;
                    }
                }
            }
        }
        #pragma omp parallel for reduction(|:any)
        for (int32_t t = 0; t < N; t++) {
            dist[t] = dist_nxt[t];
            updated[t] = updated_nxt[t];
            updated_nxt[t] = 0;
            any |= updated[t];
        }
        fin = !any;
        sweeps++;
    }
    free(updated); free(updated_nxt); free(dist_nxt);
    return sweeps;
}

/* gm_bfs_template::do_bfs_forward, inc/gm_bfs_template.h:69-271, serial restatement of all seven
 * states with the reference's transition rules (get_next_state :352-421, finish_level :423-447).
 * The level of a vertex does not depend on which state discovered it; the state trace is exported
 * so the device state machine can be checked against the reference's direction switches. */
enum { ST_SMALL = 0, ST_QUE, ST_Q2R, ST_RD, ST_R2Q, ST_RRD, ST_RR2Q };
/* save_child (template parameter, :14): the direct QUE / Q2R -> RRD switches are disabled (:374,:389), the RD -> RRD
 * one is not (:398-400) */
static int bfs_levels_impl(int32_t N, const int32_t* begin, const int32_t* node_idx,
                           const int32_t* r_begin, const int32_t* r_node_idx, int has_reverse, int save_child,
                           int32_t root, int16_t* levels, int32_t* state_trace, int32_t* level_count,
                           int32_t trace_cap) {
    const int rrd_direct = has_reverse && !save_child;
    const int16_t INVALID = -2;                       /* __INVALID_LEVEL, :725 */
    const int THRESHOLD1 = 128, THRESHOLD2 = 1024;    /* :720-721 */
    const float RRD_THRESHOLD = 0.05f;                /* :359 (float arithmetic in the comparisons) */
    int32_t* queue = (int32_t*)malloc(sizeof(int32_t) * (size_t)(N ? N : 1));
    for (int32_t i = 0; i < N; i++) levels[i] = INVALID;
    int state = ST_SMALL;
    int16_t curr_level = 0;
    int32_t curr_begin = 0, next_begin = 1, curr_count = 1, next_count = 0;
    levels[root] = 0;
    queue[0] = root;
    for (;;) {
        int16_t nl = (int16_t)(curr_level + 1);
        if (curr_level < trace_cap) {
            if (state_trace) state_trace[curr_level] = state;
            if (level_count) level_count[curr_level] = curr_count;
        }
        switch (state) {
        case ST_SMALL: case ST_QUE:        /* queue in, queue out (:98-133; iterate_neighbor_small/que) */
            for (int32_t i = 0; i < curr_count; i++) {
                int32_t t = queue[curr_begin + i];
                for (int32_t nx = begin[t]; nx < begin[t + 1]; nx++) {
                    int32_t u = node_idx[nx];
                    if (levels[u] == INVALID) { levels[u] = nl; queue[next_begin + next_count++] = u; }
                }
            }
            break;
        case ST_Q2R:                       /* queue in, marks only (:134-159; iterate_neighbor_rd) */
            for (int32_t i = 0; i < curr_count; i++) {
                int32_t t = queue[curr_begin + i];
                for (int32_t nx = begin[t]; nx < begin[t + 1]; nx++) {
                    int32_t u = node_idx[nx];
                    if (levels[u] == INVALID) { levels[u] = nl; next_count++; }
                }
            }
            break;
        case ST_RD: case ST_R2Q:           /* scan all vertices of the current level (:161-212) */
            for (int32_t t = 0; t < N; t++) {
                if (levels[t] != curr_level) continue;
                for (int32_t nx = begin[t]; nx < begin[t + 1]; nx++) {
                    int32_t u = node_idx[nx];
                    if (levels[u] == INVALID) {
                        levels[u] = nl;
                        if (state == ST_R2Q) queue[next_begin + next_count] = u;
                        next_count++;
                    }
                }
            }
            break;
        case ST_RRD: case ST_RR2Q:         /* bottom-up (:215-259; check_parent_rrd :689-708) */
            for (int32_t t = 0; t < N; t++) {
                if (levels[t] != INVALID) continue;
                for (int32_t nx = r_begin[t]; nx < r_begin[t + 1]; nx++) {
                    if (levels[r_node_idx[nx]] == curr_level) {
                        levels[t] = nl;
                        if (state == ST_RR2Q) queue[next_begin + next_count] = t;
                        next_count++;
                        break;
                    }
                }
            }
            break;
        }
        /* get_next_state, :352-421 */
        if (next_count == 0) break;
        int next_state = state;
        float thr = (float)N * RRD_THRESHOLD;
        switch (state) {
        case ST_SMALL: if (next_count >= THRESHOLD1) next_state = ST_QUE; break;
        case ST_QUE:
            if ((float)next_count >= thr) next_state = rrd_direct ? ST_RRD : ST_RD;
            else if (next_count >= THRESHOLD2 && next_count >= curr_count * 5) next_state = ST_Q2R;
            break;
        case ST_Q2R: next_state = (rrd_direct && (float)next_count >= thr) ? ST_RRD : ST_RD; break;
        case ST_RD:
            if ((float)next_count >= thr) next_state = ST_RRD;
            else if (next_count <= 2 * curr_count) next_state = ST_R2Q;
            break;
        case ST_R2Q: next_state = ST_QUE; break;
        case ST_RRD: if (next_count <= 2 * curr_count) next_state = ST_RR2Q; break;
        case ST_RR2Q: next_state = ST_QUE; break;
        }
        /* finish_level, :423-447 */
        if (!(state == ST_RD || state == ST_Q2R || state == ST_RRD)) {
            curr_begin = next_begin;
            next_begin = next_begin + next_count;
        }
        curr_count = next_count;
        next_count = 0;
        curr_level++;
        state = next_state;
    }
    free(queue);
    return curr_level + 1;
}
int orc_bfs_levels(int32_t N, const int32_t* begin, const int32_t* node_idx,
                   const int32_t* r_begin, const int32_t* r_node_idx, int has_reverse,
                   int32_t root, int16_t* levels, int32_t* state_trace, int32_t* level_count,
                   int32_t trace_cap) {
    return bfs_levels_impl(N, begin, node_idx, r_begin, r_node_idx, has_reverse, 0, root, levels, state_trace, level_count, trace_cap);
}
int orc_bfs_levels_save_child(int32_t N, const int32_t* begin, const int32_t* node_idx,
                              const int32_t* r_begin, const int32_t* r_node_idx, int has_reverse,
                              int32_t root, int16_t* levels, int32_t* state_trace, int32_t* level_count,
                              int32_t trace_cap) {
    return bfs_levels_impl(N, begin, node_idx, r_begin, r_node_idx, has_reverse, 1, root, levels, state_trace, level_count, trace_cap);
}


/* EDGE_LIST reader: gm_edge_list_graph_reader.cc:68-171 (builtGraph).  Pass 1 marks the ids that occur
 * (:91-110), the lookup table numbers them densely in ascending raw-id order (:112-117), pass 2 adds the edges
 * (:137-160); freeze() lays the rows out by source with the edges in file order, do_semi_sort() (:168) sorts
 * every row by destination.  The reference's fixed LOOKUP_SIZE (80 000 000, :66) is the id bound kept here. */
#define ORC_LOOKUP_SIZE 80000000
static int orc_el_line(char* line, long* a, long* b, int* star) {
    if (strlen(line) == 0 || line[0] == '#') return 0;                  /* :99, :140 */
    char* save = NULL;
    char* p = strtok_r(line, " \t", &save);                             /* :101 */
    if (!p) return 0;
    *a = atol(p);
    p = strtok_r(NULL, " \t\n\r", &save);                               /* :105 */
    if (!p) return -1;
    *star = (*p == '*');
    if (!*star) *b = atol(p);
    return 1;
}
int orc_edge_list_read(const char* path, int32_t* N, int32_t* M, int32_t* begin, int32_t* node_idx) {
    FILE* f = fopen(path, "r");
    if (!f) return -1;
    int32_t* lookup = (int32_t*)malloc(sizeof(int32_t) * (size_t)ORC_LOOKUP_SIZE);
    for (long i = 0; i < ORC_LOOKUP_SIZE; i++) lookup[i] = -1;
    char line[1024];
    long a = 0, b = 0; int star = 0, rc = 0;
    int64_t m = 0;
    while (fgets(line, sizeof line, f)) {
        line[strcspn(line, "\n")] = 0;
        int k = orc_el_line(line, &a, &b, &star);
        if (k < 0 || (k > 0 && (a < 0 || a >= ORC_LOOKUP_SIZE || (!star && (b < 0 || b >= ORC_LOOKUP_SIZE))))) { rc = -2; break; }
        if (k == 0) continue;
        lookup[a] = 1;
        if (!star) { lookup[b] = 1; m++; }
    }
    int32_t n = 0;
    for (long i = 0; i < ORC_LOOKUP_SIZE; i++) lookup[i] = lookup[i] > 0 ? n++ : -1;
    *N = n; *M = (int32_t)m;
    if (rc == 0 && begin) {
        int32_t* deg = (int32_t*)calloc((size_t)n + 1, sizeof(int32_t));
        rewind(f);
        while (fgets(line, sizeof line, f)) {
            line[strcspn(line, "\n")] = 0;
            if (orc_el_line(line, &a, &b, &star) > 0 && !star) deg[lookup[a]]++;
        }
        begin[0] = 0;
        for (int32_t i = 0; i < n; i++) begin[i + 1] = begin[i] + deg[i];
        memset(deg, 0, sizeof(int32_t) * ((size_t)n + 1));
        rewind(f);
        while (fgets(line, sizeof line, f)) {
            line[strcspn(line, "\n")] = 0;
            if (orc_el_line(line, &a, &b, &star) > 0 && !star) { int32_t s = lookup[a]; node_idx[begin[s] + deg[s]++] = lookup[b]; }
        }
        free(deg);
        orc_semi_sort(n, begin, node_idx, NULL);
    }
    free(lookup);
    fclose(f);
    return rc;
}


/* get_edge_idx_for_src_dest, gm_graph.cc:589-633 (rows semi-sorted): the reference's own search, ends compared first,
 * then bisection over the inclusive range; != NIL_EDGE is has_edge_to (:60-66). */
static int orc_has_edge_to(const int32_t* begin, const int32_t* node_idx, int32_t src, int32_t to) {
    int32_t b = begin[src], e = begin[src + 1] - 1;
    if (b > e) return 0;
    if (to == node_idx[b] || to == node_idx[e]) return 1;
    while (b < e) {
        if (to < node_idx[b] || to > node_idx[e]) return 0;
        int32_t mid = (b + e) / 2, mv = node_idx[mid];
        if (to == mv) return 1;
        if (to < mv) { if (e == mid) return 0; e = mid; }
        else { if (b == mid) return 0; b = mid; }
    }
    return 0;
}
/* triangle_counting.gm:5-11 as the C++ back end emits it: parallel over v (dynamic,128), private counter + atomic add */
int64_t orc_triangle_counting(int32_t N, const int32_t* begin, const int32_t* node_idx) {
    int64_t T = 0;
#pragma omp parallel
    {
        int64_t T_prv = 0;
#pragma omp for nowait schedule(dynamic, 128)
        for (int32_t v = 0; v < N; v++)
            for (int32_t ui = begin[v]; ui < begin[v + 1]; ui++) {
                int32_t u = node_idx[ui];
                if (!(u > v)) continue;
                for (int32_t wi = begin[v]; wi < begin[v + 1]; wi++) {
                    int32_t w = node_idx[wi];
                    if (w > u && orc_has_edge_to(begin, node_idx, w, u)) T_prv += 1;
                }
            }
#pragma omp atomic
        T += T_prv;
    }
    return T;
}


/* comp_BC, apps/src/bc.gm:4-29.  Levels from the BFS template restatement above; forward pass level by level
 * (visit_fw), reverse pass from the deepest level back to the root (do_bfs_reverse, gm_bfs_template.h:273-312).
 * Down edges (save_child): iterate_neighbor_small/que/rd mark every edge into level + 1 (:496-509,:603-617,:663-677);
 * the bottom-up levels do not -- check_parent_rrd (:689-708) has no save_down_edge call -- so the out-edges of a level
 * that was expanded in state RRD or RR2Q are NOT down edges (single-thread semantics; with several threads the marking
 * in the que/rd states also races, see tests/test_bc.py). */
static void comp_bc_impl(int32_t N, const int32_t* begin, const int32_t* node_idx, const int32_t* r_begin,
                         const int32_t* r_node_idx, const int32_t* seeds, int32_t n_seeds, float* bc, int first_level) {
    int16_t* lev = (int16_t*)malloc(sizeof(int16_t) * (size_t)(N > 0 ? N : 1));
    float* sigma = (float*)malloc(sizeof(float) * (size_t)(N > 0 ? N : 1));
    float* delta = (float*)malloc(sizeof(float) * (size_t)(N > 0 ? N : 1));
    int32_t* trace = (int32_t*)malloc(sizeof(int32_t) * 32768);                     /* level_t = short */
    for (int32_t t = 0; t < N; t++) bc[t] = 0.0f;                                   /* bc.gm:6 */
    for (int32_t k = 0; k < n_seeds; k++) {
        const int32_t s = seeds[k];
        for (int32_t t = 0; t < N; t++) sigma[t] = 0.0f;                            /* bc.gm:13 */
        sigma[s] = 1.0f;                                                            /* bc.gm:14 */
        const int nl = bfs_levels_impl(N, begin, node_idx, r_begin, r_node_idx, 1, 1, s, lev, trace, NULL, 32768);
        for (int l = first_level; l < nl; l++)                                      /* InBFS, bc.gm:17-20 */
            for (int32_t v = 0; v < N; v++) {
                if (lev[v] != l) continue;
                float S = 0.0f;
                for (int32_t e = r_begin[v]; e < r_begin[v + 1]; e++) {
                    const int32_t w = r_node_idx[e];
                    if (lev[w] != l - 1) continue;
                    S = S + sigma[w];
                }
                sigma[v] = S;
            }
        for (int l = nl - 1; l >= first_level; l--)                                 /* InReverse, bc.gm:21-27 */
            for (int32_t v = 0; v < N; v++) {
                if (lev[v] != l) continue;
                float S = 0.0f;
                const int marked = trace[l] != ST_RRD && trace[l] != ST_RR2Q;
                for (int32_t e = begin[v]; e < begin[v + 1]; e++) {
                    const int32_t w = node_idx[e];
                    if (!marked || lev[w] != l + 1) continue;                       /* is_down_edge */
                    S = S + sigma[v] / sigma[w] * (1 + delta[w]);
                }
                delta[v] = S;
                bc[v] = bc[v] + delta[v];
            }
    }
    free(lev); free(sigma); free(delta); free(trace);
}
void orc_comp_bc(int32_t N, const int32_t* begin, const int32_t* node_idx, const int32_t* r_begin,
                 const int32_t* r_node_idx, const int32_t* seeds, int32_t n_seeds, float* bc) {
    comp_bc_impl(N, begin, node_idx, r_begin, r_node_idx, seeds, n_seeds, bc, 0);
}
/* apps/src/bc_adj.gm:4-29: InBFS(v: G.Nodes From s)(v != s) / InReverse(v != s) -- the root (the only vertex of level 0)
 * is skipped by both passes, so its sigma stays 1 and the path counts are the real ones */
void orc_comp_bc_adj(int32_t N, const int32_t* begin, const int32_t* node_idx, const int32_t* r_begin,
                     const int32_t* r_node_idx, const int32_t* seeds, int32_t n_seeds, float* bc) {
    comp_bc_impl(N, begin, node_idx, r_begin, r_node_idx, seeds, n_seeds, bc, 1);
}
