// TEST INFRASTRUCTURE ONLY (oracle/): C entry points over the reference's OWN classes, compiled
// from the sources where they lie under /root/reference (see ../Makefile) into oracle/_ref/.
// Nothing here is on the product path; tests and bench.py's cpu_baseline / --impl reference legs
// load it through ctypes as the checker / the timed CPU arm.
//
// Reference code exercised (unmodified): gm_graph (apps/output_cpp/gm_graph/src/gm_graph.cc),
// load/store_binary (src/gm_graph_binary_loader.cc:42-252), generators (src/graph_gen.cc:58-287),
// gm_bfs_template (inc/gm_bfs_template.h:14-754).  pagerank()/hop_dist() are the restated generated
// procedures in this directory (gm_comp is not buildable here).
#include <stdio.h>
#include <string.h>
#include <sys/time.h>
#include <vector>
#include "gm.h"
#include "graph_gen.h"
#include "pagerank.h"
#include "hop_dist.h"

static double now_ms() {
    struct timeval t;
    gettimeofday(&t, NULL);
    return t.tv_sec * 1000.0 + t.tv_usec * 0.001;
}

// gm_bfs_template<short, multithread, no navigator, forward edges, no save_child>: the shape the
// code generator emits for a plain InBFS (src/backend_cpp/gm_cpp_gen_bfs.cc:88-221, level_t = short :90).
class level_recorder_bfs : public gm_bfs_template<short, true, false, false, false> {
  public:
    level_recorder_bfs(gm_graph& g, short* out) : gm_bfs_template<short, true, false, false, false>(g), lev(out), levels(0) {}
    int levels;
  protected:
    virtual void visit_fw(node_t t) { lev[t] = get_curr_level(); }
    virtual void visit_rv(node_t) {}
    virtual bool check_navigator(node_t, edge_t) { return true; }
    virtual void do_end_of_level_fw() { levels++; }
  private:
    short* lev;
};

extern "C" {

void ref_set_threads(int n) {
    gm_rt_initialize();
    gm_rt_set_num_threads(n);
}
int ref_get_threads(void) { return gm_rt_get_num_threads(); }

// after external creation the reference's load path runs do_semi_sort() then make_reverse_edges()
// (gm_graph_binary_loader.cc:190-197); mirror that so a generated graph equals a stored+loaded one.
static void* finish(gm_graph* g) {
    g->do_semi_sort();
    g->make_reverse_edges();
    return g;
}

void* ref_graph_uniform(int32_t N, int32_t M, long seed, int use_xorshift) {
    return finish(create_uniform_random_graph(N, M, seed, use_xorshift != 0));
}
void* ref_graph_rmat(int32_t N, int32_t M, long seed, double a, double b, double c, int permute) {
    return finish(create_RMAT_graph(N, M, seed, a, b, c, permute != 0));
}
void* ref_graph_from_csr(int32_t N, int32_t M, const int32_t* begin, const int32_t* node_idx) {
    gm_graph* g = new gm_graph();
    g->prepare_external_creation(N, M);
    memcpy(g->begin, begin, sizeof(int32_t) * ((size_t)N + 1));
    memcpy(g->node_idx, node_idx, sizeof(int32_t) * (size_t)M);
    return finish(g);
}
void* ref_graph_load(const char* path) {
    gm_graph* g = new gm_graph();
    if (!g->load_binary((char*)path)) { delete g; return NULL; }
    return g;
}
// the reference's EDGE_LIST reader (gm_edge_list_graph_reader.cc:68-171), no property columns
void* ref_graph_load_edge_list(const char* path) {
    gm_graph* g = new gm_graph();
    std::vector<VALUE_TYPE> vs, es;
    std::vector<void*> vp, ep;
    if (!g->load_edge_list((char*)path, vs, es, vp, ep, false)) { delete g; return NULL; }
    return g;
}
// comp_BC (apps/src/bc.gm:4-29) as the C++ back end emits it, over the reference's REAL gm_bfs_template (save_child)
class ref_bc_bfs : public gm_bfs_template<short, true, false, false, true> {
  public:
    ref_bc_bfs(gm_graph& _G, float* bc, float* sg, float* dl, node_t _s, bool _adj)
        : gm_bfs_template<short, true, false, false, true>(_G), G(_G), G_BC(bc), G_sigma(sg), G_delta(dl), s(_s), adj(_adj) {}
  private:
    gm_graph& G; float* G_BC; float* G_sigma; float* G_delta;
    node_t s; bool adj;       // adj: bc_adj.gm:17,21 -- InBFS(v: G.Nodes From s)(v != s) / InReverse(v != s)
  protected:
    virtual void visit_fw(node_t v) {
        if (adj && v == s) return;
        float S = (float)0;
        for (edge_t w_idx = G.r_begin[v]; w_idx < G.r_begin[v + 1]; w_idx++) {
            node_t w = G.r_node_idx[w_idx];
            if (get_level(w) != (get_curr_level() - 1)) continue;
            S = S + G_sigma[w];
        }
        G_sigma[v] = S;
    }
    virtual void visit_rv(node_t v) {
        if (adj && v == s) return;
        float S = (float)0;
        for (edge_t w_idx = G.begin[v]; w_idx < G.begin[v + 1]; w_idx++) {
            if (!is_down_edge(w_idx)) continue;
            node_t w = G.node_idx[w_idx];
            S = S + G_sigma[v] / G_sigma[w] * (1 + G_delta[w]);
        }
        G_delta[v] = S;
        G_BC[v] = G_BC[v] + G_delta[v];
    }
    virtual bool check_navigator(node_t, edge_t) { return true; }
};
static void ref_comp_bc_impl(void* h, const int32_t* seeds, int32_t n_seeds, float* bc, bool adj) {
    gm_graph& G = *(gm_graph*)h;
    const node_t N = G.num_nodes();
    std::vector<float> sigma(N > 0 ? N : 1), delta(N > 0 ? N : 1);
    for (node_t t = 0; t < N; t++) bc[t] = 0;
    for (int32_t k = 0; k < n_seeds; k++) {
        node_t s = seeds[k];
        for (node_t t = 0; t < N; t++) sigma[t] = 0;
        sigma[s] = 1;
        // deliberately leaked: the reference's destructor does `delete[] down_edge_set` on a `new std::set`
        // (gm_bfs_template.h:31,40), undefined behaviour that crashes under glibc when save_child is on
        ref_bc_bfs* B = new ref_bc_bfs(G, bc, &sigma[0], &delta[0], s, adj);
        B->prepare(s, gm_rt_get_num_threads());
        B->do_bfs_forward();
        B->do_bfs_reverse();
    }
}
void ref_comp_bc(void* h, const int32_t* seeds, int32_t n_seeds, float* bc) { ref_comp_bc_impl(h, seeds, n_seeds, bc, false); }
// apps/src/bc_adj.gm:4-29: the same procedure with the (v != s) filters
void ref_comp_bc_adj(void* h, const int32_t* seeds, int32_t n_seeds, float* bc) { ref_comp_bc_impl(h, seeds, n_seeds, bc, true); }
// triangle_counting.gm:5-11 over the reference's REAL gm_graph::has_edge_to (gm_graph.cc:60-66,589-633)
long long ref_triangle_counting(void* h) {
    gm_graph& G = *(gm_graph*)h;
    int64_t T = 0;
#pragma omp parallel
    {
        int64_t T_prv = 0;
#pragma omp for nowait schedule(dynamic, 128)
        for (node_t v = 0; v < G.num_nodes(); v++)
            for (edge_t u_idx = G.begin[v]; u_idx < G.begin[v + 1]; u_idx++) {
                node_t u = G.node_idx[u_idx];
                if (u > v)
                    for (edge_t w_idx = G.begin[v]; w_idx < G.begin[v + 1]; w_idx++) {
                        node_t w = G.node_idx[w_idx];
                        if (w > u && G.has_edge_to(w, u)) T_prv = T_prv + 1;
                    }
            }
        ATOMIC_ADD<int64_t>(&T, T_prv);
    }
    return T;
}
int ref_graph_store(void* h, const char* path) { return ((gm_graph*)h)->store_binary((char*)path) ? 1 : 0; }
void ref_graph_free(void* h) { delete (gm_graph*)h; }
int32_t ref_graph_num_nodes(void* h) { return ((gm_graph*)h)->num_nodes(); }
int32_t ref_graph_num_edges(void* h) { return ((gm_graph*)h)->num_edges(); }
const int32_t* ref_graph_begin(void* h) { return ((gm_graph*)h)->begin; }
const int32_t* ref_graph_node_idx(void* h) { return ((gm_graph*)h)->node_idx; }
const int32_t* ref_graph_r_begin(void* h) { return ((gm_graph*)h)->r_begin; }
const int32_t* ref_graph_r_node_idx(void* h) { return ((gm_graph*)h)->r_node_idx; }
const int32_t* ref_graph_e_rev2idx(void* h) { return ((gm_graph*)h)->e_rev2idx; }
const int32_t* ref_graph_e_idx2idx(void* h) { return ((gm_graph*)h)->e_idx2idx; }

// returns elapsed ms of the procedure call (what common_main.h:198-201 prints as "running time=")
double ref_pagerank(void* h, double e, double d, int32_t max_iter, double* rank) {
    double t0 = now_ms();
    pagerank(*(gm_graph*)h, e, d, max_iter, rank);
    return now_ms() - t0;
}
double ref_hop_dist(void* h, int32_t root, int32_t* dist) {
    node_t r = root;
    double t0 = now_ms();
    hop_dist(*(gm_graph*)h, dist, r);
    return now_ms() - t0;
}
// levels[t] = BFS level of t from root, -2 (the template's __INVALID_LEVEL, gm_bfs_template.h:725) if unreached.
double ref_bfs_levels(void* h, int32_t root, int16_t* levels, int32_t* n_levels) {
    gm_graph& G = *(gm_graph*)h;
    for (node_t i = 0; i < G.num_nodes(); i++) levels[i] = -2;
    double t0 = now_ms();
    level_recorder_bfs bfs(G, levels);
    bfs.prepare(root, gm_rt_get_num_threads());
    bfs.do_bfs_forward();
    double t = now_ms() - t0;
    if (n_levels) *n_levels = bfs.levels;
    return t;
}

}  // extern "C"
