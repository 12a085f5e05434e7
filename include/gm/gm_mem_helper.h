// Temporary-property allocation with per-thread ownership lists (reference gm_mem_helper.h:52-68,
// gm_mem_helper.cc:80-159): everything handed out is released by gm_rt_cleanup() at procedure exit.
#ifndef GM_MEM_HELPER_H
#define GM_MEM_HELPER_H
#include <stddef.h>
#include <stdint.h>
#include "gm_internal.h"
int64_t* gm_rt_allocate_long(size_t sz, int thread_id = 0);
int32_t* gm_rt_allocate_int(size_t sz, int thread_id = 0);
float* gm_rt_allocate_float(size_t sz, int thread_id = 0);
bool* gm_rt_allocate_bool(size_t sz, int thread_id = 0);
double* gm_rt_allocate_double(size_t sz, int thread_id = 0);
node_t* gm_rt_allocate_node_t(size_t sz, int thread_id = 0);
edge_t* gm_rt_allocate_edge_t(size_t sz, int thread_id = 0);
void gm_rt_deallocate(void* ptr, int thread_id = 0);
void gm_rt_cleanup();
#endif
