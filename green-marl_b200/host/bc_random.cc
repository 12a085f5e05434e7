// The "generated" bc_random procedure (apps/src/bc_random.gm:4-33): comp_BC's loop body (bc.gm, no (v != s) filter) over K
// seeds drawn with G.PickRandom() -- gm_graph::pick_random_node, rand() % num_nodes (gm_graph.h:389-391), one draw at
// the top of each iteration.  Nothing else in the loop calls rand(), so drawing the K seeds first yields the same
// sequence; the K traversals and both callbacks are then one device call (csrc/gmc_bc.cu).
#include "bc_random.h"
#include <vector>

void bc_random(gm_graph& G, float* G_BC, int32_t K) {
    gm_rt_initialize();
    G.freeze();
    G.make_reverse_edges();
    if (G.num_nodes() == 0) { gm_rt_cleanup(); return; }

    shl__start_timer(10);
    gmc_graph* dev = G.device_mirror();
    shl__end_timer("SHOAL_Copyin");

    std::vector<int32_t> seeds;
    for (int32_t k = 0; k < K; k++) seeds.push_back((int32_t)G.pick_random_node());      // bc_random.gm:9-10

    gmc_bc_info info;
    shl__start();
    if (gmc_comp_bc(dev, seeds.data(), (int32_t)seeds.size(), G_BC, &info) != GMC_OK) {
        fprintf(stderr, "bc_random: %s\n", gmc_last_error());
        exit(EXIT_FAILURE);
    }
    shl__end();
    printf("bc_random: seeds=%d launches=%d max_levels=%d device_ms=%.3f\n", info.seeds, info.launches, info.max_levels, info.device_ms);
    gm_rt_cleanup();
}
