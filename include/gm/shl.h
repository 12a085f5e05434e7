// Shoal entry points used by the Green-Marl drivers and generated code, re-targeted from NUMA nodes to
// B200 HBM.  libshoal is not part of the reference tree; the surface below is what its call sites need:
//   shl__init(nthreads, partitioned)           apps/output_cpp/src/common_main.h:95-99
//   shl__start / shl__end / shl__thread_init    src/backend_cpp/shl_extensions.cc:30-40, gm_cpp_gen.cc:1457
//   shl__start_timer / step_timer / end_timer   src/backend_cpp/gm_cpp_gen.cc:685,760,769,775,1468-1486
//   shl__get_rep_id                             src/backend_cpp/gm_cpp_gen.cc:715-716
//   class Timer                                 src/backend_cpp/gm_cpp_gen.cc:1488-1503
#ifndef SHL_H_
#define SHL_H_
#include <stddef.h>
#include <sys/time.h>
#ifdef __cplusplus
extern "C" {
#endif
// Selects the device (SHL_DEVICE / GM_CUDA_DEVICE env, default 0), initialises libgm_cuda and returns
// the number of host threads to use for driver-side work.  Exits if no CUDA device is usable: there is
// no CPU fallback.
int shl__init(int nthreads, int partitioned);
void shl__start(void);            // marks the start of the compute phase (stdout timing contract)
void shl__end(void);
void shl__thread_init(void);
void shl__start_timer(int n_steps);
double shl__step_timer(const char* label);
double shl__end_timer(const char* label);
int shl__get_rep_id(void);
void* shl__alloc_memory(size_t bytes);
// libgm_cuda extension: the device group selected at shl__init (GM_CUDA_DEVICES=2|4|8), NULL when one device is used.
// The reference chooses NUMA placement at the same point (common_main.h:95-99, scripts/run.sh -r / -p).
struct gmc_group;
struct gmc_group* shl__device_group(void);
void shl__print_array_conf(const char* name, size_t elements, size_t elem_bytes, int used, int replicated, int partitioned);
// The Shoal settings file scripts/generate_settings.py:53-75 writes (scripts/run.sh:256-267: shl__settings.lua): a `global = { ... }`
// block of `name = value,` lines (distribution, replication, partitioning, hugepage, trim, stride, crc, static) and a `dma = { enable = .. }`
// block.  shl__init reads the file named by SHL_SETTINGS (or ./shl__settings.lua when present); the SHL_* environment variables
// run.sh exports win over it.  Returns 0, or -1 when the file cannot be read; fields a file does not name keep their defaults.
typedef struct shl_settings {
    int distribution, replication, partitioning, hugepage, trim, stride, crc, static_schedule, dma_enable;
} shl_settings;
int shl__load_settings(const char* path, shl_settings* out);
#ifdef __cplusplus
}
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
#endif
