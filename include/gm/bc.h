// Header of the "generated" betweenness-centrality procedure.  Signature is the one gm_comp emits for
// apps/src/bc.gm:4 (Procedure comp_BC(G: Graph, BC: N_P<Float>, Seeds: Node_Sequence); in-args then out-args,
// properties as raw host T*, collections by reference: src/backend_cpp/gm_cpp_gen.cc:520-608), so
// apps/output_cpp/src/bc_main.cc links unchanged.
#ifndef GM_GENERATED_CPP_BC_H
#define GM_GENERATED_CPP_BC_H
#include <stdio.h>
#include <stdlib.h>
#include <stdint.h>
#include <float.h>
#include <limits.h>
#include <cmath>
#include <algorithm>
#include <omp.h>
#include "shl.h"
#include "shl_graph.h"
#include "shl_arrays.hpp"
#include "gm.h"

void comp_BC(gm_graph& G, float* G_BC, gm_node_seq& Seeds);
// the same procedure in its literal generated form: a gm_bfs_template subclass with visit_fw / visit_rv on the host
void comp_BC_via_template(gm_graph& G, float* G_BC, gm_node_seq& Seeds);
#endif
