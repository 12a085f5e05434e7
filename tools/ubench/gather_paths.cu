// Microbenchmark 4: alternative data paths for random 8-byte gathers on B200 -- which one can serve the "warm"
// contributions (the ids after the shared-memory hot block) faster than one L1TEX wavefront per gather?
//   ldg      plain __ldg from an L2-resident 38.8 MB table (baseline)
//   ldg-l1   __ldg from a 32 KB range (L1 hits): the wavefront limit itself
//   tex      tex1Dfetch<int2> on a linear texture object over the same table
//   lds      local shared memory only (range = H doubles)
//   dsmem C  cluster of C CTAs, H doubles per CTA, indices uniform over C*H: ld.shared::cluster.f64 via mapa
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o gather_paths gather_paths.cu && ./gather_paths
#include <cuda_runtime.h>
#include <cooperative_groups.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <vector>
namespace cg = cooperative_groups;
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("%s:%d %s\n", __FILE__, __LINE__, cudaGetErrorString(e)); exit(1); } } while (0)

constexpr int U = 2;   // int4 index loads per thread per iteration => 8 gathers in flight per lane

template <int MODE>
__global__ void __launch_bounds__(1024, 1) k_glob(const int4* __restrict__ idx, size_t n4, const double* __restrict__ tab,
                                                  cudaTextureObject_t tex, double* __restrict__ out) {
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    double acc = 0;
    for (; i + (U - 1) * stride < n4; i += U * stride) {
        int4 w[U];
#pragma unroll
        for (int u = 0; u < U; u++) w[u] = __ldcs(idx + i + u * stride);
        double v[4 * U];
#pragma unroll
        for (int u = 0; u < U; u++) {
            const int ww[4] = {w[u].x, w[u].y, w[u].z, w[u].w};
#pragma unroll
            for (int q = 0; q < 4; q++) {
                if (MODE == 0) v[4 * u + q] = __ldg(tab + ww[q]);
                else { int2 t = tex1Dfetch<int2>(tex, ww[q]); v[4 * u + q] = __hiloint2double(t.y, t.x); }
            }
        }
#pragma unroll
        for (int u = 0; u < 4 * U; u++) acc += v[u];
    }
    if (acc == 123.456) out[0] = acc;
}

__device__ __forceinline__ double ld_dsmem(uint32_t local_addr, uint32_t rank) {
    uint32_t remote;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(remote) : "r"(local_addr), "r"(rank));
    double v;
    asm volatile("ld.shared::cluster.f64 %0, [%1];" : "=d"(v) : "r"(remote));
    return v;
}

// H doubles per CTA; index j -> CTA rank j / H (block layout), offset j % H.  C = cluster size (1 = plain LDS).
// LOCAL_FAST: ranks equal to the own rank use a plain ld.shared (the split a real kernel would make)
template <bool LOCAL_FAST>
__global__ void __launch_bounds__(1024, 1) k_dsmem(const int4* __restrict__ idx, size_t n4, const double* __restrict__ tab, int H,
                                                   int C, double* __restrict__ out) {
    extern __shared__ double hot[];
    cg::cluster_group cl = cg::this_cluster();
    const unsigned my = C > 1 ? cl.block_rank() : 0;
    for (int i = threadIdx.x; i < H; i += blockDim.x) hot[i] = tab[my * H + i];
    if (C > 1) cl.sync(); else __syncthreads();
    const uint32_t base = (uint32_t)__cvta_generic_to_shared(hot);
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    double acc = 0;
    for (; i + (U - 1) * stride < n4; i += U * stride) {
        int4 w[U];
#pragma unroll
        for (int u = 0; u < U; u++) w[u] = __ldcs(idx + i + u * stride);
        double v[4 * U];
#pragma unroll
        for (int u = 0; u < U; u++) {
            const int ww[4] = {w[u].x, w[u].y, w[u].z, w[u].w};
#pragma unroll
            for (int q = 0; q < 4; q++) {
                const unsigned r = (unsigned)ww[q] / (unsigned)H, o = (unsigned)ww[q] - r * (unsigned)H;
                if (C == 1) v[4 * u + q] = hot[o];
                else if (LOCAL_FAST && r == my) v[4 * u + q] = hot[o];
                else v[4 * u + q] = ld_dsmem(base + 8u * o, r);
            }
        }
#pragma unroll
        for (int u = 0; u < 4 * U; u++) acc += v[u];
    }
    if (acc == 123.456) out[0] = acc;
    if (C > 1) cl.sync();     // nobody leaves while a peer may still read its shared memory
}

static float time_launch(void (*launch)(void*), void* ctx) {
    cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
    launch(ctx);
    float best = 1e9;
    for (int r = 0; r < 3; r++) {
        cudaEventRecord(a);
        launch(ctx);
        cudaEventRecord(b); cudaEventSynchronize(b);
        float ms; cudaEventElapsedTime(&ms, a, b); if (ms < best) best = ms;
    }
    CK(cudaGetLastError());
    cudaEventDestroy(a); cudaEventDestroy(b);
    return best;
}

struct Ctx { const int4* idx; size_t n4; const double* tab; cudaTextureObject_t tex; double* out; int H, C, grid, mode; bool fast; };

static void launch_glob(void* p) {
    Ctx* c = (Ctx*)p;
    if (c->mode == 0) k_glob<0><<<c->grid, 1024>>>(c->idx, c->n4, c->tab, c->tex, c->out);
    else k_glob<1><<<c->grid, 1024>>>(c->idx, c->n4, c->tab, c->tex, c->out);
}
static void launch_dsmem(void* p) {
    Ctx* c = (Ctx*)p;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(c->grid); cfg.blockDim = dim3(1024); cfg.dynamicSmemBytes = (size_t)c->H * 8;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeClusterDimension; at[0].val.clusterDim.x = c->C; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
    cfg.attrs = at; cfg.numAttrs = 1;
    if (c->fast) CK(cudaLaunchKernelEx(&cfg, k_dsmem<true>, c->idx, c->n4, c->tab, c->H, c->C, c->out));
    else CK(cudaLaunchKernelEx(&cfg, k_dsmem<false>, c->idx, c->n4, c->tab, c->H, c->C, c->out));
}

int main() {
    const size_t n = 64u << 20;
    const size_t TAB = 4850000;
    int32_t* idx; double* tab; double* out;
    CK(cudaMalloc(&idx, n * 4)); CK(cudaMalloc(&tab, TAB * 8)); CK(cudaMalloc(&out, 64));
    CK(cudaMemset(tab, 0, TAB * 8));
    cudaResourceDesc rd = {}; rd.resType = cudaResourceTypeLinear; rd.res.linear.devPtr = tab;
    rd.res.linear.desc = cudaCreateChannelDesc<int2>(); rd.res.linear.sizeInBytes = TAB * 8;
    cudaTextureDesc td = {}; td.readMode = cudaReadModeElementType;
    cudaTextureObject_t tex; CK(cudaCreateTextureObject(&tex, &rd, &td, nullptr));
    std::vector<int32_t> h(n);
    auto fill = [&](size_t range) {
        uint64_t s = 88172645463325252ull;
        for (size_t i = 0; i < n; i++) { s ^= s << 13; s ^= s >> 7; s ^= s << 17; h[i] = (int32_t)(s % range); }
        CK(cudaMemcpy(idx, h.data(), n * 4, cudaMemcpyHostToDevice));
    };
    Ctx c = {(const int4*)idx, n / 4, tab, tex, out, 0, 1, 148, 0, false};
    fill(TAB);
    c.mode = 0; printf("ldg    range 38.8MB        : %7.1f G gathers/s\n", n / time_launch(launch_glob, &c) / 1e6);
    c.mode = 1; printf("tex    range 38.8MB        : %7.1f G gathers/s\n", n / time_launch(launch_glob, &c) / 1e6);
    fill(4096);
    c.mode = 0; printf("ldg-l1 range 32KB          : %7.1f G gathers/s\n", n / time_launch(launch_glob, &c) / 1e6);
    c.mode = 1; printf("tex-l1 range 32KB          : %7.1f G gathers/s\n", n / time_launch(launch_glob, &c) / 1e6);

    const int H = 20480;
    CK(cudaFuncSetAttribute(k_dsmem<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, H * 8));
    CK(cudaFuncSetAttribute(k_dsmem<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, H * 8));
    CK(cudaFuncSetAttribute(k_dsmem<true>, cudaFuncAttributeNonPortableClusterSizeAllowed, 1));
    CK(cudaFuncSetAttribute(k_dsmem<false>, cudaFuncAttributeNonPortableClusterSizeAllowed, 1));
    for (int C : {1, 2, 4, 8, 16}) {
        c.H = H; c.C = C;
        int nclusters = 148;
        if (C > 1) {
            cudaLaunchConfig_t cfg = {};
            cfg.gridDim = dim3(C * 148); cfg.blockDim = dim3(1024); cfg.dynamicSmemBytes = (size_t)H * 8;
            cudaLaunchAttribute at[1];
            at[0].id = cudaLaunchAttributeClusterDimension; at[0].val.clusterDim.x = C; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
            cfg.attrs = at; cfg.numAttrs = 1;
            cudaError_t e = cudaOccupancyMaxActiveClusters(&nclusters, k_dsmem<true>, &cfg);
            if (e != cudaSuccess || nclusters < 1) { printf("dsmem C=%d: not launchable (%s)\n", C, cudaGetErrorString(e)); cudaGetLastError(); continue; }
        }
        c.grid = nclusters * C;
        fill((size_t)H * C);
        for (int fast = 0; fast < (C > 1 ? 2 : 1); fast++) {
            c.fast = fast;
            float t = time_launch(launch_dsmem, &c);
            printf("dsmem  C=%2d H=%d (%4d CTAs, local %s): %7.1f G gathers/s  (%.2f per clk per active SM @1.965GHz)\n", C, H, c.grid,
                   fast ? "lds " : "mapa", n / t / 1e6, n / t / 1e6 / 1.965 / c.grid);
        }
    }
    return 0;
}
