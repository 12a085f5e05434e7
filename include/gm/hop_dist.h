// Header of the "generated" hop_dist procedure (apps/src/hop_dist.gm; signature per
// src/backend_cpp/gm_cpp_gen.cc:520-608,545-548) so apps/output_cpp/src/hop_dist_main.cc links unchanged.
#ifndef GM_GENERATED_CPP_HOP_DIST_H
#define GM_GENERATED_CPP_HOP_DIST_H
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

void hop_dist(gm_graph& G, int32_t* G_dist, node_t& root);
#endif
