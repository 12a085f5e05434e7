// Host-side synthetic generators with the reference's signatures and RNG streams
// (apps/output_cpp/gm_graph/src/graph_gen.cc:12-287), so a seeded graph is edge-for-edge the graph the
// reference's graph_gen produces (same CSR after the same counting sort).  These are serial by nature
// (libc rand()/drand48() are one global stream); the large configs use the device generators
// (gmc_graph_generate_*, csrc/gmc_gen.cu) instead.
#include <algorithm>
#include <math.h>
#include <string.h>
#include <sys/time.h>
#include <vector>
#include "graph_gen.h"
#include "gm_rand.h"

namespace {

// Edge list -> frozen CSR of G.  Row u is filled from its LAST slot towards its first in edge order,
// which is the layout the reference's counting sort leaves (graph_gen.cc:38-49: pos = degree[u]--).
void fill_csr(gm_graph& G, node_t N, edge_t M, const std::vector<node_t>& from, const std::vector<node_t>& to) {
    std::vector<edge_t> left((size_t)N, 0);
    for (edge_t k = 0; k < M; k++) left[from[k]]++;
    edge_t run = 0;
    for (node_t v = 0; v < N; v++) {
        G.begin[v] = run;
        run += left[v];
    }
    G.begin[N] = run;
    for (edge_t k = 0; k < M; k++) {
        const node_t u = from[k];
        G.node_idx[G.begin[u] + (--left[u])] = to[k];
    }
}

struct LibcRand {   // srand/rand stream (graph_gen.cc:14-17,29-33)
    explicit LibcRand(long seed) { srand((unsigned)seed); }
    node_t next() { return (node_t)rand(); }
};
struct XorRand {    // gm_rand xorshift stream (gm_rand.cc:10-23)
    gm_rand r;
    explicit XorRand(long seed) : r(seed) {}
    node_t next() { return (node_t)(edge_t)r.rand(); }
};

template <class R>
void uniform_edges(R& rng, node_t N, edge_t M, std::vector<node_t>& from, std::vector<node_t>& to) {
    from.resize(M > 0 ? M : 1);
    to.resize(M > 0 ? M : 1);
    for (edge_t k = 0; k < M; k++) {   // source first, then destination
        from[k] = rng.next() % N;
        to[k] = rng.next() % N;
    }
}

// one quadrant choice of the R-MAT recursion: which halves (row = u, column = v) the edge falls into
inline void pick_quadrant(double p, double pa, double pb, double pc, node_t step, node_t& u, node_t& v) {
    if (p < pa) return;
    if (p < pa + pb) { v += step; return; }
    if (p < pa + pb + pc) { u += step; return; }
    u += step;
    v += step;
}

}  // namespace

void create_uniform_random_graph_new(gm_graph& G, node_t N, edge_t M, long seed, bool use_xorshift_rng) {
    std::vector<node_t> from, to;
    if (use_xorshift_rng) { XorRand r(seed); G.prepare_external_creation(N, M); uniform_edges(r, N, M, from, to); }
    else { LibcRand r(seed); G.prepare_external_creation(N, M); uniform_edges(r, N, M, from, to); }
    fill_csr(G, N, M, from, to);
}

gm_graph* create_uniform_random_graph(node_t N, edge_t M, long seed, bool use_xorshift_rng) {
    gm_graph* g = new gm_graph();
    create_uniform_random_graph_new(*g, N, M, seed, use_xorshift_rng);
    return g;
}

// add_node/add_edge + freeze() route (graph_gen.cc:109-130); freeze() semi-sorts on the device
gm_graph* create_uniform_random_graph2(node_t N, edge_t M, long seed) {
    srand((unsigned)seed);
    gm_graph* g = new gm_graph();
    for (node_t v = 0; v < N; v++) g->add_node();
    for (edge_t k = 0; k < M; k++) {
        node_t a = rand() % N;
        node_t b = rand() % N;
        g->add_edge(a, b);
    }
    struct timeval t0, t1;
    gettimeofday(&t0, NULL);
    g->freeze();
    gettimeofday(&t1, NULL);
    printf("time for freeze (ms) = %lf\n", (t1.tv_sec - t0.tv_sec) * 1000.0 + (t1.tv_usec - t0.tv_usec) * 0.001);
    return g;
}

gm_graph* create_uniform_random_nonmulti_graph(node_t N, edge_t M, long seed) {
    srand((unsigned)seed);
    gm_graph* g = new gm_graph();
    for (node_t v = 0; v < N; v++) g->add_node();
    edge_t made = 0;
    while (made < M) {   // duplicates are re-drawn (graph_gen.cc:142-149)
        node_t a = rand() % N;
        node_t b = rand() % N;
        if (g->has_edge(a, b)) continue;
        g->add_edge(a, b);
        made++;
    }
    g->freeze();
    return g;
}

// R-MAT (graph_gen.cc:159-287): SCALE = floor(log2 N) levels; quadrant probabilities start at (a,b,c,d),
// and before every further level each is scaled by 0.95 + 0.1*drand48() and the four renormalised;
// self edges are drawn again; optional vertex relabelling by a drand48-driven swap sequence.
gm_graph* create_RMAT_graph(node_t N, edge_t M, long rseed, double a, double b, double c, bool permute) {
    assert(a + b + c < 1);
    const double d0 = 1 - (a + b + c);
    srand48(rseed);
    gm_graph* g = new gm_graph();
    g->prepare_external_creation(N, M);
    std::vector<node_t> from(M > 0 ? M : 1), to(M > 0 ? M : 1);
    const node_t levels = (node_t)log2((double)N);
    edge_t k = 0;
    while (k < M) {
        node_t u = 1, v = 1, step = N / 2;
        double pa = a, pb = b, pc = c, pd = d0;
        pick_quadrant(drand48(), pa, pb, pc, step, u, v);
        for (node_t lv = 1; lv < levels; lv++) {
            step /= 2;
            pa *= 0.95 + 0.1 * drand48();
            pb *= 0.95 + 0.1 * drand48();
            pc *= 0.95 + 0.1 * drand48();
            pd *= 0.95 + 0.1 * drand48();
            const double norm = pa + pb + pc + pd;
            pa = pa / norm; pb = pb / norm; pc = pc / norm; pd = pd / norm;
            pick_quadrant(drand48(), pa, pb, pc, step, u, v);
        }
        if (u == v) continue;   // no self edges: draw this edge again
        from[k] = u - 1;
        to[k] = v - 1;
        k++;
    }
    if (permute) {
        std::vector<node_t> label((size_t)N);
        for (node_t i = 0; i < N; i++) label[i] = i;
        for (node_t i = 0; i < N; i++) std::swap(label[i], label[(node_t)(N * drand48())]);
        for (edge_t e = 0; e < M; e++) { from[e] = label[from[e]]; to[e] = label[to[e]]; }
    }
    fill_csr(*g, N, M, from, to);
    return g;
}
