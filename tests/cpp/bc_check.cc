// Test driver (tests/test_bc.py): comp_BC of the libgm_cuda facade on a binary graph file with the given seeds,
// the whole BC array written raw (float32) to a file so the test can compare it with the oracle element by element.
//   bc_check <graph.bin> <out.f32> <seed> [<seed> ...]
#include <stdio.h>
#include <stdlib.h>
#include "bc.h"

int main(int argc, char** argv) {
    if (argc < 4) { fprintf(stderr, "usage: %s <graph.bin> <out.f32> <seed>...\n", argv[0]); return 2; }
    gm_rt_initialize();
    gm_graph G;
    if (!G.load_binary(argv[1])) return 1;
    gm_node_seq seeds;
    for (int i = 3; i < argc; i++) seeds.push_back((node_t)atol(argv[i]));
    float* BC = new float[G.num_nodes() > 0 ? G.num_nodes() : 1];
    comp_BC(G, BC, seeds);
    FILE* f = fopen(argv[2], "wb");
    if (!f) return 1;
    fwrite(BC, sizeof(float), (size_t)G.num_nodes(), f);
    fclose(f);
    delete[] BC;
    return 0;
}
