// The "generated" PageRank procedure: what gm_comp emits for apps/src/pagerank.gm:1-20 (prologue
// src/backend_cpp/gm_cpp_gen.cc:1307-1334, Shoal copy-in :685-775, copy-back + CRC :1457-1507), with the
// OpenMP loop nest replaced by one device call.  Signature and stdout contract are the reference's, so
// apps/output_cpp/src/pagerank_main.cc links and runs unchanged against libgm_cuda.
//
//   rank = 1/N; do { sweep (K1) ; cnt++ } while (diff > e && cnt < max)      -> gmc_pagerank_dev
//
// The CSR arrays the reference copies into shl_arrays (shl__G_begin, shl__G_r_begin, shl__G_r_node_idx)
// are already resident in the graph's HBM mirror; only the rank property is a Shoal array here.
#include "pagerank.h"

void pagerank(gm_graph& G, double e, double d, int32_t max, double* G_pg_rank) {
    gm_rt_initialize();
    G.freeze();
    G.make_reverse_edges();

    if (gmc_group* grp = G.group_mirror()) {
        // GM_CUDA_DEVICES > 1: the same procedure over the device group -- copy-in = every device takes 1/N of the CSR over
        // its own PCIe link + NVLink allgather; the loop nest = sharded sweeps with the exchange fused into the epilogue;
        // copy-back = every device writes its slice of rank[] into G_pg_rank
        shl__start_timer(10);
        // the arrays the generated procedure hands to Shoal (gm_cpp_gen.cc:685-775) and where they live in the group
        shl__print_array_conf("shl__G_begin", (size_t)G.num_nodes() + 1, sizeof(edge_t), 1, 1, 0);
        shl__print_array_conf("shl__G_r_begin", (size_t)G.num_nodes() + 1, sizeof(edge_t), 1, 1, 0);
        shl__print_array_conf("shl__G_r_node_idx", (size_t)G.num_edges(), sizeof(node_t), 1, 1, 0);
        shl__print_array_conf("shl__G_pg_rank", (size_t)G.num_nodes(), sizeof(double), 1, 0, 1);
        shl__end_timer("SHOAL_Copyin");
        gmc_pagerank_info ginfo;
        shl__start();
        if (G.num_nodes() > 0 && gmc_group_pagerank(grp, e, d, max, G_pg_rank, &ginfo) != GMC_OK) {
            fprintf(stderr, "pagerank: %s\n", gmc_last_error());
            exit(EXIT_FAILURE);
        }
        shl__end();
        if (G.num_nodes() > 0) {
            int ndev = 0;
            gmc_group_size(grp, &ndev);
            printf("pagerank: iterations=%d diff=%.6e device_ms=%.3f launches=%d devices=%d\n", ginfo.iterations, ginfo.diff, ginfo.kernel_ms, ginfo.launches, ndev);
        }
        printf("CRC shl__G_pg_rank 0x%04x\n", (unsigned)shl__crc16(G_pg_rank, sizeof(double) * (size_t)G.num_nodes()));
        gm_rt_cleanup();
        return;
    }
    shl__start_timer(10);
    gmc_graph* dev = G.device_mirror();   // copy-in of begin / r_begin / r_node_idx (no-op when resident)
    shl_array<double>* a_rank = shl__malloc_array<double>(G.num_nodes(), "shl__G_pg_rank", false, false, true, false, true, true);
    a_rank->alloc();
    shl__end_timer("SHOAL_Copyin");

    gmc_pagerank_info info;
    shl__start();
    if (G.num_nodes() > 0) {
        if (gmc_pagerank_dev(dev, e, d, max, a_rank->get_array(), &info, NULL) != GMC_OK) {
            fprintf(stderr, "pagerank: %s\n", gmc_last_error());
            exit(EXIT_FAILURE);
        }
    }
    shl__end();
    if (G.num_nodes() > 0)
        printf("pagerank: iterations=%d diff=%.6e device_ms=%.3f launches=%d\n", info.iterations, info.diff, info.kernel_ms, info.launches);

    shl__start_timer(5);
    a_rank->copy_back(G_pg_rank);
    shl__end_timer("SHOAL_Copyback");
    a_rank->print_crc();
    delete a_rank;
    gm_rt_cleanup();
}
