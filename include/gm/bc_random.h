// Header of the "generated" bc_random procedure.  Signature is the one gm_comp emits for apps/src/bc_random.gm:4
// (Procedure bc_random(G: Graph, BC: N_P<Float>, K: Int); in-args then out-args, properties as raw host T*:
// src/backend_cpp/gm_cpp_gen.cc:520-608), so apps/output_cpp/src/bc_random_main.cc links unchanged.
#ifndef GM_GENERATED_CPP_BC_RANDOM_H
#define GM_GENERATED_CPP_BC_RANDOM_H
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

void bc_random(gm_graph& G, float* G_BC, int32_t K);
#endif
