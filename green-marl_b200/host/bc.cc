// The "generated" comp_BC procedure: what gm_comp emits for apps/src/bc.gm:4-29 -- a subclass of
// gm_bfs_template<short, multithread, no navigator, forward edges, save_child> per InBFS/InReverse pair
// (src/backend_cpp/gm_cpp_gen_bfs.cc:88-221) whose visit_fw sums sigma over the UpNbrs (in-neighbours one level
// up, gm_cpp_gen_foreach.cc:163-175) and whose visit_rv sums over the DownNbrs (is_down_edge).  The traversal itself
// runs on the device (include/gm/gm_bfs_template.h -> gmc_bfs_levels_masked).
//
// comp_BC itself is ONE device call (csrc/gmc_bc.cu: traversal + both callbacks as kernels, per level).  The class below
// is the literal generated form over the facade's gm_bfs_template -- arbitrary user callbacks run as host C++ replayed
// level by level -- kept as comp_BC_via_template (GM_BC_VIA_TEMPLATE=1 selects it) so that the template facade stays
// exercised by the same tests; both give the same bits.
//
// NB (reference behaviour, kept): bc.gm has no `(v != s)` filter on the InBFS body, so the root's sigma = 1 is
// overwritten by its empty UpNbrs sum 0; every sigma is then 0 and delta = 0/0*(1+..) = NaN wherever a vertex has
// a BFS child.  BC comes out as NaN for inner vertices of the BFS DAG and 0 elsewhere, in the reference as here.
#include "bc.h"
#include <vector>

class comp_BC_bfs : public gm_bfs_template<short, true, false, false, true> {
  public:
    comp_BC_bfs(gm_graph& _G, float* _G_BC, node_t& _s, float* _G_sigma, float* _G_delta)
        : gm_bfs_template<short, true, false, false, true>(_G), G(_G), G_BC(_G_BC), s(_s), G_sigma(_G_sigma), G_delta(_G_delta) {}

  private:
    gm_graph& G;
    float* G_BC;
    node_t& s;
    float* G_sigma;
    float* G_delta;

  protected:
    virtual void visit_fw(node_t v) {
        float __S3 = (float)0;
        for (edge_t w_idx = G.r_begin[v]; w_idx < G.r_begin[v + 1]; w_idx++) {
            node_t w = G.r_node_idx[w_idx];
            if (get_level(w) != (get_curr_level() - 1)) continue;
            __S3 = __S3 + G_sigma[w];
        }
        G_sigma[v] = __S3;
    }
    virtual void visit_rv(node_t v) {
        float __S4 = (float)0;
        for (edge_t w_idx = G.begin[v]; w_idx < G.begin[v + 1]; w_idx++) {
            if (!is_down_edge(w_idx)) continue;
            node_t w = G.node_idx[w_idx];
            __S4 = __S4 + G_sigma[v] / G_sigma[w] * (1 + G_delta[w]);
        }
        G_delta[v] = __S4;
        G_BC[v] = G_BC[v] + G_delta[v];
    }
    virtual bool check_navigator(node_t v, edge_t v_idx) { return true; }
};

void comp_BC(gm_graph& G, float* G_BC, gm_node_seq& Seeds) {
    if (getenv("GM_BC_VIA_TEMPLATE")) { comp_BC_via_template(G, G_BC, Seeds); return; }
    gm_rt_initialize();
    G.freeze();
    G.make_reverse_edges();

    shl__start_timer(10);
    gmc_graph* dev = G.device_mirror();
    shl__end_timer("SHOAL_Copyin");

    std::vector<int32_t> seeds;
    gm_node_seq::seq_iter s_I = Seeds.prepare_seq_iteration();
    while (s_I.has_next()) seeds.push_back((int32_t)s_I.get_next());

    gmc_bc_info info;
    shl__start();
    if (G.num_nodes() > 0 && gmc_comp_bc(dev, seeds.data(), (int32_t)seeds.size(), G_BC, &info) != GMC_OK) {
        fprintf(stderr, "comp_BC: %s\n", gmc_last_error());
        exit(EXIT_FAILURE);
    }
    shl__end();
    if (G.num_nodes() > 0)
        printf("comp_BC: seeds=%d launches=%d max_levels=%d device_ms=%.3f\n", info.seeds, info.launches, info.max_levels, info.device_ms);
    gm_rt_cleanup();
}

void comp_BC_via_template(gm_graph& G, float* G_BC, gm_node_seq& Seeds) {
    gm_rt_initialize();
    G.freeze();
    G.make_reverse_edges();

    #pragma omp parallel for
    for (node_t t0 = 0; t0 < G.num_nodes(); t0++) G_BC[t0] = (float)0;

    gm_node_seq::seq_iter s_I = Seeds.prepare_seq_iteration();
    while (s_I.has_next()) {
        node_t s = s_I.get_next();
        float* G_sigma = gm_rt_allocate_float(G.num_nodes(), gm_rt_thread_id());
        float* G_delta = gm_rt_allocate_float(G.num_nodes(), gm_rt_thread_id());

        #pragma omp parallel for
        for (node_t t1 = 0; t1 < G.num_nodes(); t1++) G_sigma[t1] = (float)0;
        G_sigma[s] = (float)1;

        comp_BC_bfs _BFS(G, G_BC, s, G_sigma, G_delta);
        _BFS.prepare(s, gm_rt_get_num_threads());
        _BFS.do_bfs_forward();
        _BFS.do_bfs_reverse();

        gm_rt_deallocate(G_sigma, gm_rt_thread_id());
        gm_rt_deallocate(G_delta, gm_rt_thread_id());
    }
    gm_rt_cleanup();
}
