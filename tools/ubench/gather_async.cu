// Microbenchmark 5 (round 2): does any load path beat one L1TEX miss per clock per SM for random 8-byte gathers from an
// L2-resident table, and does it survive a large shared-memory carve-out?
//   mode 0  ld.global.nc.f64                       (round-1 baseline)
//   mode 1  ld.global.cg.f64                       (cache at L2 only)
//   mode 2  ld.global.nc.L1::no_allocate.f64
//   mode 3  cp.async.ca 8 B -> private shared slot, wait, LDS           (LDGSTS through L1)
//   mode 4  cp.async.cg 16 B (aligned down) -> private slot, wait, LDS  (LDGSTS.BYPASS: no L1 line)
//   mode 5  ld.global.nc.v2.f64 of the aligned 16 bytes (same sector)
//   mode 6  ld.relaxed.gpu.global.f64             (coherent load: bypasses L1)
// every mode: T threads per CTA, B CTAs per SM, `ballast` KB of extra dynamic shared memory (what a hot block would take),
// D index vectors (int4) in flight per lane => 4*D gathers in flight per lane.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o gather_async gather_async.cu && ./gather_async
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <vector>
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("%s:%d %s\n", __FILE__, __LINE__, cudaGetErrorString(e)); exit(1); } } while (0)

__device__ __forceinline__ double ld_mode(int mode, const double* p) {
    double v;
    if (mode == 0) asm volatile("ld.global.nc.f64 %0, [%1];" : "=d"(v) : "l"(p));
    else if (mode == 1) asm volatile("ld.global.cg.f64 %0, [%1];" : "=d"(v) : "l"(p));
    else if (mode == 2) asm volatile("ld.global.nc.L1::no_allocate.f64 %0, [%1];" : "=d"(v) : "l"(p));
    else asm volatile("ld.relaxed.gpu.global.f64 %0, [%1];" : "=d"(v) : "l"(p));
    return v;
}

template <int MODE, int D>
__global__ void k_reg(const int4* __restrict__ idx, size_t n4, const double* __restrict__ tab, double* __restrict__ out) {
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    double acc = 0;
    for (; i + (D - 1) * stride < n4; i += D * stride) {
        int4 w[D];
#pragma unroll
        for (int u = 0; u < D; u++) w[u] = __ldcs(idx + i + u * stride);
        double v[4 * D];
#pragma unroll
        for (int u = 0; u < D; u++) {
            const int ww[4] = {w[u].x, w[u].y, w[u].z, w[u].w};
#pragma unroll
            for (int q = 0; q < 4; q++) {
                if (MODE == 5) {
                    double a, b;
                    const double* p = tab + (ww[q] & ~1);
                    asm volatile("ld.global.nc.v2.f64 {%0,%1}, [%2];" : "=d"(a), "=d"(b) : "l"(p));
                    v[4 * u + q] = (ww[q] & 1) ? b : a;
                } else v[4 * u + q] = ld_mode(MODE, tab + ww[q]);
            }
        }
#pragma unroll
        for (int u = 0; u < 4 * D; u++) acc += v[u];
    }
    if (acc == 123.456) out[0] = acc;
}

// LDGSTS variants: every thread owns 4*D landing slots (8 or 16 bytes each) in shared memory, laid out slot-major so that
// the 32 lanes of a warp hit 32 different banks groups
template <int BYTES, int D>
__global__ void k_async(const int4* __restrict__ idx, size_t n4, const double* __restrict__ tab, double* __restrict__ out, int ballast_bytes) {
    extern __shared__ __align__(16) unsigned char smem[];
    unsigned char* land = smem + ballast_bytes;
    const uint32_t my = (uint32_t)__cvta_generic_to_shared(land) + threadIdx.x * BYTES;
    const uint32_t slot_stride = blockDim.x * BYTES;
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    double acc = 0;
    for (; i + (D - 1) * stride < n4; i += D * stride) {
        int4 w[D];
#pragma unroll
        for (int u = 0; u < D; u++) w[u] = __ldcs(idx + i + u * stride);
#pragma unroll
        for (int u = 0; u < D; u++) {
            const int ww[4] = {w[u].x, w[u].y, w[u].z, w[u].w};
#pragma unroll
            for (int q = 0; q < 4; q++) {
                const uint32_t dst = my + (4 * u + q) * slot_stride;
                if (BYTES == 8) asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(dst), "l"(tab + ww[q]) : "memory");
                else asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst), "l"(tab + (ww[q] & ~1)) : "memory");
            }
        }
        asm volatile("cp.async.commit_group;" ::: "memory");
        asm volatile("cp.async.wait_group 0;" ::: "memory");
#pragma unroll
        for (int u = 0; u < D; u++) {
            const int ww[4] = {w[u].x, w[u].y, w[u].z, w[u].w};
#pragma unroll
            for (int q = 0; q < 4; q++) {
                const uint32_t src = my + (4 * u + q) * slot_stride + (BYTES == 16 ? 8 * (ww[q] & 1) : 0);
                double v;
                asm volatile("ld.shared.f64 %0, [%1];" : "=d"(v) : "r"(src));
                acc += v;
            }
        }
    }
    if (acc == 123.456) out[0] = acc;
}

template <typename F>
static float best_of(F launch) {
    cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
    launch();
    float best = 1e9;
    for (int r = 0; r < 3; r++) {
        cudaEventRecord(a);
        launch();
        cudaEventRecord(b); cudaEventSynchronize(b);
        float ms; cudaEventElapsedTime(&ms, a, b); if (ms < best) best = ms;
    }
    CK(cudaGetLastError());
    cudaEventDestroy(a); cudaEventDestroy(b);
    return best;
}

struct Cfg { int threads, per_sm, ballast_kb; };

template <int MODE, int D>
static void run_reg(const char* name, const int4* idx, size_t n4, const double* tab, double* out, Cfg c) {
    CK(cudaFuncSetAttribute(k_reg<MODE, D>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
    const size_t smem = (size_t)c.ballast_kb * 1024;
    int occ = 0;
    CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, k_reg<MODE, D>, c.threads, smem));
    if (occ < c.per_sm) { printf("%-22s T=%4d B=%d ballast=%3dKB D=%d: occupancy %d < %d\n", name, c.threads, c.per_sm, c.ballast_kb, D, occ, c.per_sm); return; }
    float t = best_of([&] { k_reg<MODE, D><<<148 * c.per_sm, c.threads, smem>>>(idx, n4, tab, out); });
    const double g = n4 * 4 / t / 1e6;
    printf("%-22s T=%4d B=%d ballast=%3dKB D=%d: %7.1f G/s  %.2f /clk/SM\n", name, c.threads, c.per_sm, c.ballast_kb, D, g, g / 1.965 / 148);
}
template <int BYTES, int D>
static void run_async(const char* name, const int4* idx, size_t n4, const double* tab, double* out, Cfg c) {
    CK(cudaFuncSetAttribute(k_async<BYTES, D>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
    const size_t smem = (size_t)c.ballast_kb * 1024 + (size_t)c.threads * BYTES * 4 * D;
    if (smem > 227 * 1024) { printf("%-22s T=%4d B=%d ballast=%3dKB D=%d: smem %zu too large\n", name, c.threads, c.per_sm, c.ballast_kb, D, smem); return; }
    int occ = 0;
    CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, k_async<BYTES, D>, c.threads, smem));
    if (occ < c.per_sm) { printf("%-22s T=%4d B=%d ballast=%3dKB D=%d: occupancy %d < %d\n", name, c.threads, c.per_sm, c.ballast_kb, D, occ, c.per_sm); return; }
    float t = best_of([&] { k_async<BYTES, D><<<148 * c.per_sm, c.threads, smem>>>(idx, n4, tab, out, c.ballast_kb * 1024); });
    const double g = n4 * 4 / t / 1e6;
    printf("%-22s T=%4d B=%d ballast=%3dKB D=%d: %7.1f G/s  %.2f /clk/SM  (landing %zu KB)\n", name, c.threads, c.per_sm, c.ballast_kb, D, g,
           g / 1.965 / 148, (size_t)c.threads * BYTES * 4 * D / 1024);
}

// L2-resident streaming read (coalesced 16-byte loads over a 64 MB buffer, repeated): the LTS throughput ceiling on this part
__global__ void k_stream(const int4* __restrict__ p, size_t n4, int reps, int* out) {
    int acc = 0;
    for (int r = 0; r < reps; r++)
        for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (size_t)gridDim.x * blockDim.x) {
            int4 v;
            asm volatile("ld.global.cg.v4.s32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p + i));
            acc += v.x ^ v.y ^ v.z ^ v.w;
        }
    if (acc == 0x12345) out[0] = acc;
}

int main() {
    const size_t n = 64u << 20;
    const size_t TAB = 4850000;
    int32_t* idx; double* tab; double* out;
    CK(cudaMalloc(&idx, n * 4)); CK(cudaMalloc(&tab, TAB * 8 + 64)); CK(cudaMalloc(&out, 64));
    CK(cudaMemset(tab, 0, TAB * 8));
    std::vector<int32_t> h(n);
    uint64_t s = 88172645463325252ull;
    for (size_t i = 0; i < n; i++) { s ^= s << 13; s ^= s >> 7; s ^= s << 17; h[i] = (int32_t)(s % TAB); }
    CK(cudaMemcpy(idx, h.data(), n * 4, cudaMemcpyHostToDevice));
    const int4* i4 = (const int4*)idx;
    const size_t n4 = n / 4;

    {   // L2 streaming ceiling
        const size_t bytes = 64u << 20;
        int4* buf; CK(cudaMalloc(&buf, bytes)); CK(cudaMemset(buf, 1, bytes));
        const int reps = 8;
        float t = best_of([&] { k_stream<<<148 * 2, 1024>>>(buf, bytes / 16, reps, (int*)out); });
        printf("L2-resident stream 64 MB x %d: %.1f GB/s = %.0f B/clk @1.965 GHz\n", reps, bytes * (double)reps / t / 1e6, bytes * (double)reps / t / 1e6 / 1.965);
        cudaFree(buf);
    }
    const Cfg cfgs[] = {{1024, 1, 0}, {1024, 1, 130}, {512, 1, 130}, {512, 2, 65}, {1024, 2, 0}, {512, 4, 0}, {1024, 1, 64}};
    for (const Cfg& c : cfgs) {
        run_reg<0, 2>("ld.nc", i4, n4, tab, out, c);
        run_reg<0, 4>("ld.nc", i4, n4, tab, out, c);
        run_reg<1, 2>("ld.cg", i4, n4, tab, out, c);
        run_reg<2, 2>("ld.nc.no_allocate", i4, n4, tab, out, c);
        run_reg<5, 2>("ld.nc.v2 (16B)", i4, n4, tab, out, c);
        run_reg<6, 2>("ld.relaxed.gpu", i4, n4, tab, out, c);
        run_async<8, 2>("cp.async.ca 8B", i4, n4, tab, out, c);
        run_async<16, 2>("cp.async.cg 16B", i4, n4, tab, out, c);
        run_async<8, 4>("cp.async.ca 8B", i4, n4, tab, out, c);
        run_async<16, 4>("cp.async.cg 16B", i4, n4, tab, out, c);
        printf("\n");
    }
    // the big-carve-out question: 192 KB of ballast (24.5 K doubles of hot block), which path still runs?
    const Cfg big[] = {{512, 1, 192}, {1024, 1, 176}};
    for (const Cfg& c : big) {
        run_reg<0, 2>("ld.nc", i4, n4, tab, out, c);
        run_reg<1, 2>("ld.cg", i4, n4, tab, out, c);
        run_reg<6, 2>("ld.relaxed.gpu", i4, n4, tab, out, c);
        run_async<8, 2>("cp.async.ca 8B", i4, n4, tab, out, c);
        run_async<16, 2>("cp.async.cg 16B", i4, n4, tab, out, c);
        printf("\n");
    }
    return 0;
}
