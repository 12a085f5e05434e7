// Header of the "generated" triangle_counting procedure.  Signature is the one gm_comp emits for
// apps/src/triangle_counting.gm:1 (Procedure triangle_counting(G: Graph): Long -> int64_t return,
// src/backend_cpp/gm_cpp_gen.cc:520-608), so apps/output_cpp/src/triangle_counting_main.cc links unchanged.
#ifndef GM_GENERATED_CPP_TRIANGLE_COUNTING_H
#define GM_GENERATED_CPP_TRIANGLE_COUNTING_H
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

int64_t triangle_counting(gm_graph& G);
#endif
