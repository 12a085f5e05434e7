// TEST INFRASTRUCTURE ONLY (oracle/): hand restatement of the C++ that the reference's gm_comp
// would emit for apps/src/pagerank.gm:1-20.  gm_comp cannot be built here (no flex/bison,
// src/parse/gm_lex.c absent) so the emitted shape is followed from the emitters:
//   procedure prologue        src/backend_cpp/gm_cpp_gen.cc:1307-1334
//   Shoal copy-in / copy-back src/backend_cpp/gm_cpp_gen.cc:685-775, 1457-1507
//   foreach / in-nbr loop     src/backend_cpp/gm_cpp_gen_foreach.cc:214-217, 297-311, 182-186
//   OutDegree()               src/backend_cpp/gm_cpplib_gen.cc:463-471
//   Sum -> temp + plain add   src/opt/gm_syntax_sugar2.cc:242,285; src/opt/gm_opt_nonconf_reduce.cc:18-58
//   diff privatisation        src/backend_cpp/gm_cpp_opt_reduce_scalar.cc:141-257,310-359
//   deferred write + updater  src/backend_cpp/gm_cpp_opt_defer.cc:152-159,221-222
//   schedule(dynamic,128)     src/backend_cpp/gm_cpp_gen.cc:1877-1888
// Array accessors are the non-WR_REP expansion (plain indexing), gm_cpp_gen.cc:390-402.
#include "pagerank.h"

void pagerank(gm_graph& G, double e, double d, int32_t max, double* G_pg_rank) {
    gm_rt_initialize();
    G.freeze();
    G.make_reverse_edges();

    shl__start_timer(10);
    shl_array<edge_t>* a_begin = shl__malloc_array<edge_t>(G.num_nodes() + 1, "shl__G_begin", true, false, true, true, true, true);
    shl_array<edge_t>* a_r_begin = shl__malloc_array<edge_t>(G.num_nodes() + 1, "shl__G_r_begin", true, false, true, true, true, true);
    shl_array<node_t>* a_r_node_idx = shl__malloc_array<node_t>(G.num_edges() + 1, "shl__G_r_node_idx", true, false, true, true, false, true);
    shl_array<double>* a_rank = shl__malloc_array<double>(G.num_nodes(), "shl__G_pg_rank", false, false, true, false, true, true);
    shl_array<double>* a_rank_nxt = shl__malloc_array<double>(G.num_nodes(), "shl__G_pg_rank_nxt", false, true, true, false, true, true);
    a_begin->alloc(); a_r_begin->alloc(); a_r_node_idx->alloc(); a_rank->alloc(); a_rank_nxt->alloc();
    a_begin->copy_from(G.begin);
    a_r_begin->copy_from(G.r_begin);
    memcpy(a_r_node_idx->get_array(), G.r_node_idx, sizeof(node_t) * (size_t)G.num_edges());
    shl__end_timer("SHOAL_Copyin");

    const node_t n_nodes = G.num_nodes();
    double diff = 0.0;
    int32_t cnt = 0;
    double N = 0.0;

    shl__start();
    cnt = 0;
    N = (double)G.num_nodes();
    #pragma omp parallel
    {
        double* shl__G_pg_rank = a_rank->get_array();
        #pragma omp for
        for (node_t t0 = 0; t0 < n_nodes; t0++) shl__G_pg_rank[t0] = 1 / N;
    }
    do {
        diff = 0.000000;
        #pragma omp parallel
        {
            edge_t* shl__begin = a_begin->get_array();
            edge_t* shl__r_begin = a_r_begin->get_array();
            node_t* shl__r_node_idx = a_r_node_idx->get_array();
            double* shl__G_pg_rank = a_rank->get_array();
            double* shl__G_pg_rank_nxt = a_rank_nxt->get_array();
            double diff_prv = 0.0;
            diff_prv = 0.000000;
            #pragma omp for nowait schedule(dynamic,128)
            for (node_t t = 0; t < n_nodes; t++) {
                double val = 0.0;
                double __S1 = 0.0;
                __S1 = 0.000000;
                for (edge_t w_idx = shl__r_begin[t]; w_idx < shl__r_begin[t + 1]; w_idx++) {
                    node_t w = shl__r_node_idx[w_idx];
                    __S1 = __S1 + shl__G_pg_rank[w] / ((double)((shl__begin[w + 1] - shl__begin[w])));
                }
                val = (1 - d) / N + d * __S1;
                diff_prv = diff_prv + std::abs((val - shl__G_pg_rank[t]));
                shl__G_pg_rank_nxt[t] = val;
            }
            ATOMIC_ADD<double>(&diff, diff_prv);
        }
        #pragma omp parallel
        {
            double* shl__G_pg_rank = a_rank->get_array();
            double* shl__G_pg_rank_nxt = a_rank_nxt->get_array();
            #pragma omp for
            for (node_t i3 = 0; i3 < n_nodes; i3++) shl__G_pg_rank[i3] = shl__G_pg_rank_nxt[i3];
        }
        cnt = cnt + 1;
    } while ((diff > e) && (cnt < max));
    shl__end();

    shl__start_timer(5);
    a_rank->copy_back(G_pg_rank);
    shl__end_timer("SHOAL_Copyback");
    a_rank->print_crc();
    delete a_begin; delete a_r_begin; delete a_r_node_idx; delete a_rank; delete a_rank_nxt;
    gm_rt_cleanup();
}
