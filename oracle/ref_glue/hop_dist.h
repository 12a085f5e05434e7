// TEST INFRASTRUCTURE ONLY (oracle/): header of the restated generated procedure.
// Signature per src/backend_cpp/gm_cpp_gen.cc:520-608,545-548 (node args by reference).
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
#include "shl_arrays.hpp"
#include "gm.h"

void hop_dist(gm_graph& G, int32_t* G_dist, node_t& root);
#endif
