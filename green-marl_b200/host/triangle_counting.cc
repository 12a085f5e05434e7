// The "generated" triangle_counting procedure: what gm_comp emits for apps/src/triangle_counting.gm:1-13 (three nested
// loops over the out-edge slots of v with G.has_edge_to(w, u) innermost, a private counter per thread combined with
// ATOMIC_ADD), with the loop nest replaced by one device call (csrc/gmc_tc.cu: one sorted-list intersection per edge
// slot against the reverse CSR).  apps/output_cpp/src/triangle_counting_main.cc links and runs unchanged.
#include "triangle_counting.h"

int64_t triangle_counting(gm_graph& G) {
    gm_rt_initialize();
    G.freeze();
    G.make_reverse_edges();          // the intersection walks the in-neighbours of u; the reference needs only forward rows

    shl__start_timer(10);
    gmc_graph* dev = G.device_mirror();
    shl__end_timer("SHOAL_Copyin");

    int64_t T = 0;
    float ms = 0.f;
    shl__start();
    if (G.num_nodes() > 0 && gmc_triangle_counting(dev, &T, &ms) != GMC_OK) {
        fprintf(stderr, "triangle_counting: %s\n", gmc_last_error());
        exit(EXIT_FAILURE);
    }
    shl__end();
    printf("triangle_counting: T=%lld device_ms=%.3f\n", (long long)T, ms);
    gm_rt_cleanup();
    return T;
}
