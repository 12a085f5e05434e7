// Thin runtime wrapper with the reference's names (gm_runtime.h:50-60, gm_runtime.cc:128-162).
// "Threads" only matter for host callbacks and host-side graph preparation; kernels run on the device.
#ifndef GM_RUNTIME_H_
#define GM_RUNTIME_H_
#include <assert.h>
#include <omp.h>
#include <stdlib.h>
#include <stdio.h>
#include <sys/time.h>
#include "gm_mem_helper.h"
void gm_rt_initialize();
bool gm_rt_is_initialized();
int gm_rt_get_num_threads();
void gm_rt_set_num_threads(int n);
int gm_rt_thread_id();
double gm_rt_uniform(int tid = 0);
int gm_rt_rand(int min, int max, int tid = 0);
long gm_rt_rand_long(long max, int tid = 0);
#endif
