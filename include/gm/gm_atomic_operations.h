// Reduction helpers the code generator emits calls to (reference gm_atomic_operations.h:4-60):
// ATOMIC_ADD / MULT / MIN / MAX as CAS loops, ATOMIC_AND / OR on bool flags.
#ifndef GM_ATOMIC_OPERATIONS_H_
#define GM_ATOMIC_OPERATIONS_H_
#include <algorithm>
#include "gm_atomic_wrapper.h"

template <typename T, typename F>
static inline void _gm_atomic_update(T* target, F f) {
    for (;;) {
        T seen = *(volatile T*)target;
        T want = f(seen);
        if (want == seen || _gm_atomic_compare_and_swap(target, seen, want)) return;
    }
}
template <typename T> inline void ATOMIC_ADD(T* target, T value) { if (value == 0) return; _gm_atomic_update(target, [=](T o) { return (T)(o + value); }); }
template <typename T> inline void ATOMIC_MULT(T* target, T value) { if (value == 1) return; _gm_atomic_update(target, [=](T o) { return (T)(o * value); }); }
template <typename T> inline void ATOMIC_MIN(T* target, T value) { _gm_atomic_update(target, [=](T o) { return std::min(o, value); }); }
template <typename T> inline void ATOMIC_MAX(T* target, T value) { _gm_atomic_update(target, [=](T o) { return std::max(o, value); }); }
inline void ATOMIC_AND(bool* target, bool v) { if (!v && *target) *target = false; }
inline void ATOMIC_OR(bool* target, bool v) { if (v && !*target) *target = true; }
#endif
