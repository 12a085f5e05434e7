// Runtime helpers: thread bookkeeping, temp-property allocation, spinlocks, bitmap, xorshift RNG,
// shl_graph lookups.  Same entry points as the reference's gm_runtime.cc / gm_mem_helper.cc / gm_lock.cc /
// gm_bitmap.cc / gm_rand.cc / shl_graph.cc; implementations are this library's own.
#include <algorithm>
#include <list>
#include <mutex>
#include <vector>
#include "gm.h"
#include "gm_rand.h"

// ---- runtime ------------------------------------------------------------------------------------
static bool g_rt_init = false;
static int g_rt_threads = 1;
static std::vector<unsigned short> g_seeds;   // 3 x 16-bit erand48 state per thread

static void grow_seeds(int n) {
    size_t old = g_seeds.size() / 3;
    g_seeds.resize((size_t)n * 3);
    for (size_t t = old; t < (size_t)n; t++) {
        g_seeds[t * 3 + 0] = (unsigned short)(0x330E + t);
        g_seeds[t * 3 + 1] = (unsigned short)(0xABCD ^ (t * 7919));
        g_seeds[t * 3 + 2] = (unsigned short)(0x1234 + t * 31);
    }
}
void gm_rt_initialize() {
    if (g_rt_init) return;
    g_rt_init = true;
    if (g_rt_threads < 1) g_rt_threads = 1;
    gm_rt_set_num_threads(omp_get_max_threads());
    gm_spinlock_table_init();
}
bool gm_rt_is_initialized() { return g_rt_init; }
int gm_rt_get_num_threads() { return g_rt_threads; }
void gm_rt_set_num_threads(int n) {
    if (n < 1) n = 1;
    g_rt_threads = n;
    omp_set_num_threads(n);
    if ((int)(g_seeds.size() / 3) < n) grow_seeds(n);
}
int gm_rt_thread_id() { return omp_get_thread_num(); }
double gm_rt_uniform(int tid) {
    if ((int)(g_seeds.size() / 3) <= tid) grow_seeds(tid + 1);
    return erand48(&g_seeds[(size_t)tid * 3]);
}
long gm_rt_rand_long(long max, int tid) { return (long)(gm_rt_uniform(tid) * max); }
int gm_rt_rand(int min, int max, int tid) { return min + (int)(gm_rt_uniform(tid) * (max - min)); }

// ---- temp-property allocation -------------------------------------------------------------------
static std::mutex g_mem_mu;
static std::vector<std::list<void*> > g_owned;   // per thread id

template <typename T>
static T* rt_alloc(size_t sz, int tid) {
    T* p = new T[sz ? sz : 1];
    std::lock_guard<std::mutex> lk(g_mem_mu);
    if ((int)g_owned.size() <= tid) g_owned.resize(tid + 1);
    g_owned[tid].push_back((void*)p);
    return p;
}
int64_t* gm_rt_allocate_long(size_t sz, int tid) { return rt_alloc<int64_t>(sz, tid); }
int32_t* gm_rt_allocate_int(size_t sz, int tid) { return rt_alloc<int32_t>(sz, tid); }
float* gm_rt_allocate_float(size_t sz, int tid) { return rt_alloc<float>(sz, tid); }
bool* gm_rt_allocate_bool(size_t sz, int tid) { return rt_alloc<bool>(sz, tid); }
double* gm_rt_allocate_double(size_t sz, int tid) { return rt_alloc<double>(sz, tid); }
node_t* gm_rt_allocate_node_t(size_t sz, int tid) { return rt_alloc<node_t>(sz, tid); }
edge_t* gm_rt_allocate_edge_t(size_t sz, int tid) { return rt_alloc<edge_t>(sz, tid); }
void gm_rt_deallocate(void* ptr, int tid) {
    std::lock_guard<std::mutex> lk(g_mem_mu);
    if ((int)g_owned.size() > tid) g_owned[tid].remove(ptr);
    delete[] (char*)ptr;
}
void gm_rt_cleanup() {
    std::lock_guard<std::mutex> lk(g_mem_mu);
    for (size_t t = 0; t < g_owned.size(); t++) {
        for (std::list<void*>::iterator it = g_owned[t].begin(); it != g_owned[t].end(); ++it) delete[] (char*)*it;
        g_owned[t].clear();
    }
}

// ---- spinlocks ----------------------------------------------------------------------------------
uint32_t gm_spinlock_tab_sz = GM_SPINLOCK_TAB_SZ;
gm_spinlock_t gm_spinlock_tab[GM_SPINLOCK_TAB_SZ];
void gm_spinlock_table_init() { for (int i = 0; i < GM_SPINLOCK_TAB_SZ; i++) gm_spinlock_tab[i] = 0; }
bool gm_spinlock_try_acquire(gm_spinlock_t* p) { return __sync_lock_test_and_set(p, 1) == 0; }
void gm_spinlock_acquire(gm_spinlock_t* p) {
    while (__sync_lock_test_and_set(p, 1))
        while (*p) __builtin_ia32_pause();
}
void gm_spinlock_release(gm_spinlock_t* p) { __sync_lock_release(p); }
static inline gm_spinlock_t* slot_for(uint64_t key) { return &gm_spinlock_tab[(key % GM_SPINLOCK_TAB_ENTRY) * GM_CACHELINE]; }
void gm_spinlock_acquire_for_ptr(void* p) { gm_spinlock_acquire(slot_for(((uintptr_t)p) >> 4)); }
void gm_spinlock_release_for_ptr(void* p) { gm_spinlock_release(slot_for(((uintptr_t)p) >> 4)); }
void gm_spinlock_acquire_for_node(node_t v) { gm_spinlock_acquire(slot_for((uint64_t)(uint32_t)v)); }
void gm_spinlock_release_for_node(node_t v) { gm_spinlock_release(slot_for((uint64_t)(uint32_t)v)); }
void gm_spinlock_acquire_for_edge(edge_t e) { gm_spinlock_acquire(slot_for((uint64_t)(uint32_t)e)); }
void gm_spinlock_release_for_edge(edge_t e) { gm_spinlock_release(slot_for((uint64_t)(uint32_t)e)); }

// ---- bitmap -------------------------------------------------------------------------------------
unsigned _gm_get_bit(unsigned char* b, int n) { return (b[n >> 3] >> (n & 7)) & 1u; }
void _gm_set_bit(unsigned char* b, int n) { b[n >> 3] = (unsigned char)(b[n >> 3] | (1u << (n & 7))); }
bool _gm_set_bit_atomic(unsigned char* b, int n) {
    unsigned char m = (unsigned char)(1u << (n & 7));
    return (_gm_atomic_fetch_and_or_char(&b[n >> 3], m) & m) == 0;
}
void _gm_clear_bit(unsigned char* b, int n) { b[n >> 3] = (unsigned char)(b[n >> 3] & ~(1u << (n & 7))); }
bool _gm_clear_bit_atomic(unsigned char* b, int n) {
    unsigned char m = (unsigned char)(1u << (n & 7));
    return (_gm_atomic_fetch_and_and_char(&b[n >> 3], (unsigned char)~m) & m) != 0;
}

// ---- xorshift -----------------------------------------------------------------------------------
gm_rand64::gm_rand64() : next((int64_t)0x0139408DCBBF7A44ull) {}
int64_t gm_rand64::rand() {
    uint64_t x = (uint64_t)next;
    x ^= x << 13;
    x = (uint64_t)((int64_t)x ^ ((int64_t)x >> 7));   // arithmetic shift on the signed state, as in the reference
    x ^= x << 17;
    next = (int64_t)x;
    return next;
}
int32_t gm_rand32::rand() {
    next ^= (int32_t)((uint32_t)next << 13);
    next = (next >> 17);                               // reference gm_rand.cc:20: assignment, not xor
    next ^= (int32_t)((uint32_t)next << 5);
    return next;
}

// ---- shl_graph ----------------------------------------------------------------------------------
edge_t shl_graph::get_edge_idx_for_src_dest(node_t src, node_t dest) {
    node_t* lo = node_idx + begin[src];
    node_t* hi = node_idx + begin[src + 1];
    node_t* p = std::lower_bound(lo, hi, dest);
    return (p != hi && *p == dest) ? (edge_t)(p - node_idx) : NIL_EDGE;
}
bool shl_graph::is_neighbor(node_t src, node_t to) { return get_edge_idx_for_src_dest(src, to) != NIL_EDGE; }
