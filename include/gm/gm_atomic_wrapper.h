// Host-side atomic helpers with the names the generated code uses (reference
// gm_atomic_wrapper.h + platform/x86/inc/gm_platform_helpers.h:8-32,95-154).  Implemented with the
// GCC __atomic builtins instead of the reference's inline asm.  They serve host callbacks
// (visit_fw / visit_rv bodies); device kernels use CUDA atomics.
#ifndef GM_ATOMIC_WRAPPER_H
#define GM_ATOMIC_WRAPPER_H
#include <stdint.h>
#include <string.h>

template <typename T>
static inline bool _gm_atomic_compare_and_swap(T* dest, T old_val, T new_val) {
    return __atomic_compare_exchange_n(dest, &old_val, new_val, false, __ATOMIC_SEQ_CST, __ATOMIC_SEQ_CST);
}
static inline bool _gm_atomic_compare_and_swap(double* dest, double old_val, double new_val) {
    int64_t o, n;
    memcpy(&o, &old_val, 8);
    memcpy(&n, &new_val, 8);
    return __atomic_compare_exchange_n((int64_t*)dest, &o, n, false, __ATOMIC_SEQ_CST, __ATOMIC_SEQ_CST);
}
static inline bool _gm_atomic_compare_and_swap(float* dest, float old_val, float new_val) {
    int32_t o, n;
    memcpy(&o, &old_val, 4);
    memcpy(&n, &new_val, 4);
    return __atomic_compare_exchange_n((int32_t*)dest, &o, n, false, __ATOMIC_SEQ_CST, __ATOMIC_SEQ_CST);
}
static inline bool _gm_atomic_compare_and_swap(bool* dest, bool old_val, bool new_val) {
    return __atomic_compare_exchange_n((unsigned char*)dest, (unsigned char*)&old_val, (unsigned char)new_val, false,
                                       __ATOMIC_SEQ_CST, __ATOMIC_SEQ_CST);
}
static inline int32_t _gm_atomic_fetch_and_add_int32(int32_t* p, int32_t v) { return __atomic_fetch_add(p, v, __ATOMIC_SEQ_CST); }
static inline int64_t _gm_atomic_fetch_and_add_int64(int64_t* p, int64_t v) { return __atomic_fetch_add(p, v, __ATOMIC_SEQ_CST); }
static inline unsigned char _gm_atomic_fetch_and_or_char(unsigned char* p, unsigned char v) { return __atomic_fetch_or(p, v, __ATOMIC_SEQ_CST); }
static inline unsigned char _gm_atomic_fetch_and_and_char(unsigned char* p, unsigned char v) { return __atomic_fetch_and(p, v, __ATOMIC_SEQ_CST); }
#include "gm_graph_typedef.h"
static inline node_t _gm_atomic_fetch_and_add_node(node_t* p, node_t v) { return __atomic_fetch_add(p, v, __ATOMIC_SEQ_CST); }
static inline edge_t _gm_atomic_fetch_and_add_edge(edge_t* p, edge_t v) { return __atomic_fetch_add(p, v, __ATOMIC_SEQ_CST); }
#endif
