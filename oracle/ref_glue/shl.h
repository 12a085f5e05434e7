// TEST INFRASTRUCTURE ONLY (oracle/): pass-through stand-in for libshoal's shl.h.
//
// libshoal is not part of /root/reference (README.md:7-9 expects a sibling checkout), so the
// reference's unchanged drivers (apps/output_cpp/src/common_main.h:11 `#include "shl.h"`,
// :95-99 `shl__init(nthreads, 0|1)`) cannot link without one.  This stub holds NO arithmetic
// of the path: every array is plain host memory and every timer is a no-op, which is what
// Shoal degenerates to with replication/partitioning/distribution all switched off.
// Call sites it satisfies: src/backend_cpp/gm_cpp_gen.cc:181-227,685-775,1457-1507 and
// src/inc/gm_backend_cpp.h:671-704 of the reference.
#ifndef ORACLE_REF_GLUE_SHL_H
#define ORACLE_REF_GLUE_SHL_H
#include <stddef.h>
#include <stdio.h>
#include <omp.h>
#include <sys/time.h>

static inline int shl__init(int nthreads, int /*partitioned*/) {
    if (nthreads <= 0) nthreads = omp_get_max_threads();
    return nthreads;
}
static inline void shl__start(void) {}
static inline void shl__end(void) {}
static inline void shl__thread_init(void) {}
static inline void shl__start_timer(int) {}
static inline void shl__step_timer(const char*) {}
static inline void shl__end_timer(const char*) {}
static inline int shl__get_rep_id(void) { return 0; }

class Timer {
  public:
    void start() { gettimeofday(&t0, NULL); }
    double stop() {
        struct timeval t1;
        gettimeofday(&t1, NULL);
        return (t1.tv_sec - t0.tv_sec) * 1000.0 + (t1.tv_usec - t0.tv_usec) * 0.001;
    }
  private:
    struct timeval t0;
};
#endif
