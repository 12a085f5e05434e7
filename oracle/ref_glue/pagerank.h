// TEST INFRASTRUCTURE ONLY (oracle/): header of the restated generated procedure.
// Signature per the reference emitter src/backend_cpp/gm_cpp_gen.cc:520-608
// (in-args, then out-args; properties as raw T*); guard name per :131-139.
#ifndef GM_GENERATED_CPP_PAGERANK_H
#define GM_GENERATED_CPP_PAGERANK_H
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

void pagerank(gm_graph& G, double e, double d, int32_t max, double* G_pg_rank);
#endif
