// Header of the "generated" PageRank procedure.  Signature is the one gm_comp emits for
// apps/src/pagerank.gm (src/backend_cpp/gm_cpp_gen.cc:520-608: in-args, then out-args; node properties as
// raw host T*), so the reference driver apps/output_cpp/src/pagerank_main.cc links unchanged.
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
#include "shl_graph.h"
#include "shl_arrays.hpp"
#include "gm.h"

void pagerank(gm_graph& G, double e, double d, int32_t max, double* G_pg_rank);
#endif
