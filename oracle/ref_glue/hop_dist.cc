// TEST INFRASTRUCTURE ONLY (oracle/): hand restatement of the C++ that the reference's gm_comp
// would emit for apps/src/hop_dist.gm:3-31 (see pagerank.cc in this directory for why).
//   argmin-assign under hashed spinlock   src/backend_cpp/gm_cpp_gen.cc:1563-1741
//   +INF -> INT_MAX                       src/backend_cpp/gm_cpp_gen.cc:1778-1782
//   Exist -> OR reduction, ATOMIC_OR      apps/output_cpp/gm_graph/inc/gm_atomic_operations.h:55-60
//   spinlock table                        apps/output_cpp/gm_graph/src/gm_lock.cc:48-58
//   temp properties                       apps/output_cpp/gm_graph/src/gm_mem_helper.cc:80-159
#include "hop_dist.h"

void hop_dist(gm_graph& G, int32_t* G_dist, node_t& root) {
    gm_rt_initialize();
    G.freeze();

    const node_t n_nodes = G.num_nodes();
    bool fin = false;
    bool* G_updated = gm_rt_allocate_bool(G.num_nodes(), gm_rt_thread_id());
    bool* G_updated_nxt = gm_rt_allocate_bool(G.num_nodes(), gm_rt_thread_id());
    int32_t* G_dist_nxt = gm_rt_allocate_int(G.num_nodes(), gm_rt_thread_id());
    edge_t* begin = G.begin;
    node_t* node_idx = G.node_idx;

    shl__start();
    fin = false;
    #pragma omp parallel for
    for (node_t t0 = 0; t0 < n_nodes; t0++) {
        G_dist[t0] = (t0 == root) ? 0 : INT_MAX;
        G_updated[t0] = (t0 == root) ? true : false;
        G_dist_nxt[t0] = G_dist[t0];
        G_updated_nxt[t0] = G_updated[t0];
    }
    while (!fin) {
        bool __E8 = false;
        fin = true;
        __E8 = false;
        #pragma omp parallel for schedule(dynamic,128)
        for (node_t n = 0; n < n_nodes; n++) {
            if (G_updated[n]) {
                for (edge_t s_idx = begin[n]; s_idx < begin[n + 1]; s_idx++) {
                    node_t s = node_idx[s_idx];
                    {   // argmin(argmax) - test and test-and-set
                        int32_t G_dist_nxt_new = G_dist[n] + 1;
                        if (G_dist_nxt[s] > G_dist_nxt_new) {
                            bool G_updated_nxt_arg = true;
                            gm_spinlock_acquire_for_node(s);
                            if (G_dist_nxt[s] > G_dist_nxt_new) {
                                G_dist_nxt[s] = G_dist_nxt_new;
                                G_updated_nxt[s] = G_updated_nxt_arg;
                            }
                            gm_spinlock_release_for_node(s);
                        }
                    }
                }
            }
        }
        #pragma omp parallel
        {
            bool __E8_prv = false;
            __E8_prv = false;
            #pragma omp for nowait
            for (node_t t4 = 0; t4 < n_nodes; t4++) {
                G_dist[t4] = G_dist_nxt[t4];
                G_updated[t4] = G_updated_nxt[t4];
                G_updated_nxt[t4] = false;
                __E8_prv = __E8_prv || G_updated[t4];
            }
            ATOMIC_OR(&__E8, __E8_prv);
        }
        fin = !__E8;
    }
    shl__end();
    gm_rt_cleanup();
}
