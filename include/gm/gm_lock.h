// Hashed spinlock table (reference gm_lock.h / gm_lock.cc:8-70).  Kept for host-side generated code
// (argmin-assign under a lock); the device hop_dist needs no lock (atomic test-and-set per vertex).
#ifndef GM_LOCK_H_
#define GM_LOCK_H_
#include <stdint.h>
#include "gm_internal.h"
typedef volatile int32_t gm_spinlock_t;
bool gm_spinlock_try_acquire(gm_spinlock_t* ptr);
void gm_spinlock_acquire(gm_spinlock_t* ptr);
void gm_spinlock_release(gm_spinlock_t* ptr);
#define GM_CACHELINE 8
#define GM_SPINLOCK_TAB_ENTRY 10224
#define GM_SPINLOCK_TAB_SZ (GM_CACHELINE * GM_SPINLOCK_TAB_ENTRY)
extern uint32_t gm_spinlock_tab_sz;
extern gm_spinlock_t gm_spinlock_tab[];
void gm_spinlock_table_init();
void gm_spinlock_acquire_for_ptr(void* ptr);
void gm_spinlock_release_for_ptr(void* ptr);
void gm_spinlock_acquire_for_node(node_t ptr);
void gm_spinlock_release_for_node(node_t ptr);
void gm_spinlock_acquire_for_edge(edge_t ptr);
void gm_spinlock_release_for_edge(edge_t ptr);
#endif
