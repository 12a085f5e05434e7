// Microbenchmark: throughput of random 8-byte gathers from a table of W bytes (L1 / L2 / HBM resident),
// as a function of load flavour, unroll (loads in flight per thread) and CTA shape.  B200, sm_100a.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o gather_bw gather_bw.cu && ./gather_bw
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <vector>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("%s:%d %s\n", __FILE__, __LINE__, cudaGetErrorString(e)); exit(1); } } while (0)

template <int MODE> __device__ __forceinline__ double ld(const double* p) {
    if (MODE == 0) return __ldg(p);
    if (MODE == 1) return __ldcg(p);
    if (MODE == 2) { double v; asm volatile("ld.global.nc.L1::no_allocate.f64 %0, [%1];" : "=d"(v) : "l"(p)); return v; }
    if (MODE == 3) return *p;
    return 0;
}

// idx stream read with coalesced int4; U int4 per thread per iteration => 4U gathers in flight
template <int MODE, int U>
__global__ void __launch_bounds__(1024) k_gather(const int4* __restrict__ idx, size_t n4, const double* __restrict__ tab, double* __restrict__ out) {
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    size_t stride = (size_t)gridDim.x * blockDim.x;
    double acc = 0;
    for (; i + (U - 1) * stride < n4; i += U * stride) {
        int4 w[U];
#pragma unroll
        for (int u = 0; u < U; u++) w[u] = __ldcs(idx + i + u * stride);
        double v[4 * U];
#pragma unroll
        for (int u = 0; u < U; u++) { v[4*u] = ld<MODE>(tab + w[u].x); v[4*u+1] = ld<MODE>(tab + w[u].y); v[4*u+2] = ld<MODE>(tab + w[u].z); v[4*u+3] = ld<MODE>(tab + w[u].w); }
#pragma unroll
        for (int u = 0; u < 4 * U; u++) acc += v[u];
    }
    if (acc == 123.456) out[0] = acc;
}

template <int MODE, int U>
float run(const int4* idx, size_t n4, const double* tab, double* out, int grid, int block) {
    cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
    k_gather<MODE, U><<<grid, block>>>(idx, n4, tab, out);
    float best = 1e9;
    for (int r = 0; r < 3; r++) {
        cudaEventRecord(a);
        k_gather<MODE, U><<<grid, block>>>(idx, n4, tab, out);
        cudaEventRecord(b); cudaEventSynchronize(b);
        float ms; cudaEventElapsedTime(&ms, a, b); if (ms < best) best = ms;
    }
    CK(cudaGetLastError());
    return best;
}

int main() {
    const size_t n = 64u << 20;   // gathers per launch
    int32_t* idx; double* tab; double* out;
    CK(cudaMalloc(&idx, n * 4)); CK(cudaMalloc(&tab, 512u << 20)); CK(cudaMalloc(&out, 64));
    CK(cudaMemset(tab, 0, 512u << 20));
    std::vector<int32_t> h(n);
    const char* names[] = {"ldg.nc", "ld.cg", "nc.noalloc", "ld.ca"};
    size_t Ws[] = {128u << 10, 1u << 20, 8u << 20, 38u << 20, 100u << 20, 400u << 20};
    for (size_t W : Ws) {
        uint64_t s = 88172645463325252ull;
        size_t ne = W / 8;
        for (size_t i = 0; i < n; i++) { s ^= s << 13; s ^= s >> 7; s ^= s << 17; h[i] = (int32_t)(s % ne); }
        CK(cudaMemcpy(idx, h.data(), n * 4, cudaMemcpyHostToDevice));
        printf("table %7.1f MB:", W / 1048576.0);
        float t;
        t = run<0, 2>((int4*)idx, n / 4, tab, out, 148, 1024); printf("  %s U2 1024x148: %6.1f G/s", names[0], n / t / 1e6);
        t = run<0, 4>((int4*)idx, n / 4, tab, out, 148, 1024); printf("  U4: %6.1f", n / t / 1e6);
        t = run<0, 2>((int4*)idx, n / 4, tab, out, 148 * 2, 1024); printf("  U2 2cta: %6.1f", n / t / 1e6);
        t = run<0, 4>((int4*)idx, n / 4, tab, out, 148 * 8, 256); printf("  U4 256x8: %6.1f", n / t / 1e6);
        t = run<1, 4>((int4*)idx, n / 4, tab, out, 148 * 8, 256); printf("  %s: %6.1f", names[1], n / t / 1e6);
        t = run<2, 4>((int4*)idx, n / 4, tab, out, 148 * 8, 256); printf("  %s: %6.1f", names[2], n / t / 1e6);
        t = run<3, 4>((int4*)idx, n / 4, tab, out, 148 * 8, 256); printf("  %s: %6.1f", names[3], n / t / 1e6);
        t = run<0, 8>((int4*)idx, n / 4, tab, out, 148 * 4, 256); printf("  U8 256x4: %6.1f\n", n / t / 1e6);
    }
    return 0;
}
