// Shoal entry points (include/gm/shl.h) re-targeted from NUMA nodes to B200 HBM.  libshoal is not in the
// reference tree; behaviour follows its call sites (common_main.h:95-99; gm_cpp_gen.cc:685-775,1457-1507)
// and the stdout contract the reference's scripts parse (scripts/extract_result.py:162-207).
#include <stdarg.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <omp.h>
#include "shl.h"
#include "shl_arrays.hpp"
#include "gm_cuda.h"

static int g_device = -1;
static gmc_group* g_group = NULL;      // GM_CUDA_DEVICES > 1: every generated procedure runs over this device group
static struct timeval g_t_compute, g_t_timer;
static double g_compute_ms = 0.0;

static double ms_since(const struct timeval& t0) {
    struct timeval t1;
    gettimeofday(&t1, NULL);
    return (t1.tv_sec - t0.tv_sec) * 1000.0 + (t1.tv_usec - t0.tv_usec) * 0.001;
}

extern "C" {

// `global = {` / `dma = {` blocks of `name = value,` lines, `--` comments (scripts/generate_settings.py:53-75)
int shl__load_settings(const char* path, shl_settings* out) {
    if (!out) return -1;
    *out = shl_settings{0, 0, 0, 0, 0, 4096, 1, 0, 0};        // the generator's defaults (generate_settings.py:53-62)
    FILE* f = path ? fopen(path, "r") : NULL;
    if (!f) return -1;
    char line[512], block[64] = "";
    while (fgets(line, sizeof line, f)) {
        if (char* c = strstr(line, "--")) *c = 0;
        char name[64];
        long value = 0;
        if (sscanf(line, " %63[A-Za-z_0-9] = {", name) == 1 && strchr(line, '{')) { snprintf(block, sizeof block, "%s", name); continue; }
        if (strchr(line, '}')) { block[0] = 0; continue; }
        if (sscanf(line, " %63[A-Za-z_0-9] = %li", name, &value) != 2) continue;
        if (!strcmp(block, "global")) {
            if (!strcmp(name, "distribution")) out->distribution = (int)value;
            else if (!strcmp(name, "replication")) out->replication = (int)value;
            else if (!strcmp(name, "partitioning")) out->partitioning = (int)value;
            else if (!strcmp(name, "hugepage")) out->hugepage = (int)value;
            else if (!strcmp(name, "trim")) out->trim = (int)value;
            else if (!strcmp(name, "stride")) out->stride = (int)value;
            else if (!strcmp(name, "crc")) out->crc = (int)value;
            else if (!strcmp(name, "static")) out->static_schedule = (int)value;
        } else if (!strcmp(block, "dma") && !strcmp(name, "enable")) out->dma_enable = (int)value;
    }
    fclose(f);
    return 0;
}

int shl__init(int nthreads, int partitioned) {
    const char* env = getenv("SHL_DEVICE");
    if (!env) env = getenv("GM_CUDA_DEVICE");
    int dev = env ? atoi(env) : 0;
    if (gmc_init(dev) != GMC_OK) {
        fprintf(stderr, "shl__init: %s\n", gmc_last_error());
        exit(EXIT_FAILURE);   // libgm_cuda has no CPU path
    }
    g_device = dev;
    // Placement switch of the reference (shl__init(nthreads, partitioned); run.sh -p / -r) re-targeted: GM_CUDA_DEVICES=N
    // runs the generated procedures over N B200s of this box -- CSR replicated, rows partitioned (csrc/gmc_group.cu)
    // The reference's own switches (scripts/run.sh:70-97,248-251 exports them; scripts/generate_settings.py:53-60 writes them
    // into the Shoal settings): -r SHL_REPLICATION (replicate read-only arrays over the NUMA nodes), -p SHL_PARTITION (partition
    // indexed arrays), -d SHL_DISTRIBUTION, -h SHL_HUGEPAGE.  Replication / partitioning over NUMA nodes = the device group over
    // the GPUs of the box (CSR replicated, rows partitioned): either switch selects every visible device (1, 2, 4 or 8) unless
    // GM_CUDA_DEVICES says otherwise; distribution and huge pages have no HBM equivalent and are reported as ignored.
    // The settings file run.sh generates beside them (shl__settings.lua) gives the defaults, the environment wins.
    shl_settings file_set;
    const char* spath = getenv("SHL_SETTINGS");
    const bool have_file = shl__load_settings(spath ? spath : "shl__settings.lua", &file_set) == 0;
    if (spath && !have_file) fprintf(stderr, "shl__init: cannot read the settings file %s (SHL_SETTINGS): using the environment only\n", spath);
    const int s_rep = getenv("SHL_REPLICATION") ? atoi(getenv("SHL_REPLICATION")) : file_set.replication;
    const int s_part = getenv("SHL_PARTITION") ? atoi(getenv("SHL_PARTITION")) : file_set.partitioning;
    const int s_dist = getenv("SHL_DISTRIBUTION") ? atoi(getenv("SHL_DISTRIBUTION")) : file_set.distribution;
    const int s_huge = getenv("SHL_HUGEPAGE") ? atoi(getenv("SHL_HUGEPAGE")) : file_set.hugepage;
    if (have_file)
        printf("SHOAL settings file %s: distribution=%d replication=%d partitioning=%d hugepage=%d trim=%d stride=%d crc=%d static=%d dma=%d "
               "(trim / stride / static / dma: NUMA and OpenMP knobs without an HBM equivalent, ignored)\n", spath ? spath : "shl__settings.lua",
               file_set.distribution, file_set.replication, file_set.partitioning, file_set.hugepage, file_set.trim, file_set.stride, file_set.crc,
               file_set.static_schedule, file_set.dma_enable);
    int ndev = 1;
    if (const char* nd = getenv("GM_CUDA_DEVICES")) ndev = atoi(nd);
    else if (s_rep || s_part) {       // (the second argument of shl__init is the SHL_STATIC build flag, common_main.h:95-99: not a placement switch)
        int have = 1;
        if (gmc_device_count(&have) == GMC_OK) for (ndev = 1; ndev * 2 <= have && ndev * 2 <= 8; ndev *= 2) {}
    }
    printf("SHOAL settings: replication=%d partition=%d distribution=%d hugepage=%d -> devices=%d%s\n", s_rep, s_part, s_dist, s_huge, ndev,
           (s_dist || s_huge) ? " (distribution / hugepage: no HBM equivalent, ignored)" : "");
    if (ndev > 1 && !g_group) {
        if (gmc_group_create(&g_group, ndev, NULL) != GMC_OK) {
            fprintf(stderr, "shl__init: GM_CUDA_DEVICES=%d: %s\n", ndev, gmc_last_error());
            exit(EXIT_FAILURE);
        }
        printf("SHOAL (libgm_cuda) device group: %d devices, CSR replicated, rows partitioned by destination range\n", ndev);
    }
    int sms = 0;
    size_t total = 0, avail = 0;
    gmc_device_info(&sms, &total, &avail);
    if (nthreads <= 0) nthreads = omp_get_max_threads();
    printf("SHOAL (libgm_cuda) device=%d SMs=%d HBM=%.1f GB free=%.1f GB host-threads=%d partitioned=%d\n", dev, sms,
           total / 1e9, avail / 1e9, nthreads, partitioned);
    return nthreads;
}

void shl__start(void) { gettimeofday(&g_t_compute, NULL); }
void shl__end(void) {
    g_compute_ms = ms_since(g_t_compute);
    printf("SHOAL_Compute: %.3lf\n", g_compute_ms);
}
void shl__thread_init(void) {}
void shl__start_timer(int) { gettimeofday(&g_t_timer, NULL); }
double shl__step_timer(const char* label) {
    double ms = ms_since(g_t_timer);
    printf("%s: %.3lf\n", label, ms);
    gettimeofday(&g_t_timer, NULL);
    return ms;
}
double shl__end_timer(const char* label) { return shl__step_timer(label); }
int shl__get_rep_id(void) { return g_device < 0 ? 0 : g_device; }
struct gmc_group* shl__device_group(void) { return g_group; }
void* shl__alloc_memory(size_t bytes) { return malloc(bytes); }

// "Array[<name>]: elements= size= -- hugepage=[ ] -- used=[X] replication=[X] ..." : the array configuration line Shoal prints per
// array and scripts/extract_result.py:171-190 (ArrayConfChecker) parses
void shl__print_array_conf(const char* name, size_t elements, size_t elem_bytes, int used, int replicated, int partitioned) {
    printf("Array[%20s]: elements=%10zu size=%12zu -- hugepage=[ ] -- used=[%c] replication=[%c] partition=[%c] hbm=[X]\n", name, elements,
           elements * elem_bytes, used ? 'X' : ' ', replicated ? 'X' : ' ', partitioned ? 'X' : ' ');
}

// CRC-16/CCITT-FALSE (poly 0x1021, init 0xFFFF).  The polynomial libshoal's print_crc uses cannot be
// recovered from the reference tree; only the "CRC <name> 0x%04x" line format is (extract_result.py:162).
uint16_t shl__crc16(const void* data, size_t bytes) {
    const unsigned char* p = (const unsigned char*)data;
    uint16_t crc = 0xFFFF;
    for (size_t i = 0; i < bytes; i++) {
        crc ^= (uint16_t)p[i] << 8;
        for (int b = 0; b < 8; b++) crc = (crc & 0x8000) ? (uint16_t)((crc << 1) ^ 0x1021) : (uint16_t)(crc << 1);
    }
    return crc;
}

}  // extern "C"

// COST macro of the generated code (gm_cpp_gen.cc:181-227): n sizes (size_t) -> working-set line
void shl__estimate_working_set_size(int n, ...) {
    va_list ap;
    va_start(ap, n);
    size_t total = 0;
    for (int i = 0; i < n; i++) total += va_arg(ap, size_t);
    va_end(ap);
    printf("Working set size: %zu bytes (%.3f GB of HBM)\n", total, total / 1e9);
}
