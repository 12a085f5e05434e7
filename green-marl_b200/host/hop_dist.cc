// The "generated" hop_dist procedure: what gm_comp emits for apps/src/hop_dist.gm:3-31 (argmin-assign
// under a hashed spinlock, src/backend_cpp/gm_cpp_gen.cc:1563-1741), with the per-level relaxation and
// swap sweeps replaced by one device call.  dist[v] = hop distance from root, INT_MAX when unreachable
// (hop_dist.gm:9: +INF).  apps/output_cpp/src/hop_dist_main.cc links and runs unchanged.
#include "hop_dist.h"

void hop_dist(gm_graph& G, int32_t* G_dist, node_t& root) {
    gm_rt_initialize();
    G.freeze();

    if (gmc_group* grp = G.group_mirror()) {      // GM_CUDA_DEVICES > 1: vertex-range shards over the device group
        shl__start_timer(10);
        shl__print_array_conf("shl__G_begin", (size_t)G.num_nodes() + 1, sizeof(edge_t), 1, 1, 0);
        shl__print_array_conf("shl__G_node_idx", (size_t)G.num_edges(), sizeof(node_t), 1, 1, 0);
        shl__print_array_conf("shl__G_dist", (size_t)G.num_nodes(), sizeof(int32_t), 1, 0, 1);
        shl__end_timer("SHOAL_Copyin");
        gmc_bfs_info ginfo;
        shl__start();
        if (G.num_nodes() > 0 && gmc_group_hop_dist(grp, root, G_dist, &ginfo) != GMC_OK) {
            fprintf(stderr, "hop_dist: %s\n", gmc_last_error());
            exit(EXIT_FAILURE);
        }
        shl__end();
        if (G.num_nodes() > 0) {
            int ndev = 0;
            gmc_group_size(grp, &ndev);
            printf("hop_dist: levels=%d reached=%d edges=%lld bottom_up_levels=%d device_ms=%.3f launches=%d devices=%d\n", ginfo.levels,
                   ginfo.reached, (long long)ginfo.edges_traversed, ginfo.bottom_up_levels, ginfo.kernel_ms, ginfo.launches, ndev);
        }
        printf("CRC shl__G_dist 0x%04x\n", (unsigned)shl__crc16(G_dist, sizeof(int32_t) * (size_t)G.num_nodes()));
        gm_rt_cleanup();
        return;
    }
    shl__start_timer(10);
    gmc_graph* dev = G.device_mirror();
    shl_array<int32_t>* a_dist = shl__malloc_array<int32_t>(G.num_nodes(), "shl__G_dist", false, false, true, false, true, true);
    a_dist->alloc();
    shl__end_timer("SHOAL_Copyin");

    gmc_bfs_info info;
    shl__start();
    if (G.num_nodes() > 0) {
        if (gmc_hop_dist_dev(dev, root, a_dist->get_array(), &info, NULL) != GMC_OK) {
            fprintf(stderr, "hop_dist: %s\n", gmc_last_error());
            exit(EXIT_FAILURE);
        }
    }
    shl__end();
    if (G.num_nodes() > 0)
        printf("hop_dist: levels=%d reached=%d edges=%lld bottom_up_levels=%d device_ms=%.3f launches=%d\n", info.levels,
               info.reached, (long long)info.edges_traversed, info.bottom_up_levels, info.kernel_ms, info.launches);

    shl__start_timer(5);
    a_dist->copy_back(G_dist);
    shl__end_timer("SHOAL_Copyback");
    a_dist->print_crc();
    delete a_dist;
    gm_rt_cleanup();
}
