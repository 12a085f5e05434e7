// libgm_cuda core: error text, device setup, raw HBM arrays, a hand-written exclusive scan.
#include "gmc_common.cuh"
#include <mutex>

namespace gmc {

static thread_local char g_err[512] = "";

void set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

int sm_count() {
    static int cached = 0;
    if (!cached) {
        int dev = 0;
        if (cudaGetDevice(&dev) != cudaSuccess) return 148;
        cudaDeviceGetAttribute(&cached, cudaDevAttrMultiProcessorCount, dev);
        if (cached <= 0) cached = 148;
    }
    return cached;
}

// ---------------------------------------------------------------------------------------------
// Exclusive scan: tile of SCAN_TILE elements per CTA; (1) per-tile sums, (2) recursive scan of the
// sums, (3) per-tile exclusive scan seeded with its offset.  2 reads + 1 write of the array.
// ---------------------------------------------------------------------------------------------
constexpr int SCAN_THREADS = 256;
constexpr int SCAN_ITEMS = 8;
constexpr int SCAN_TILE = SCAN_THREADS * SCAN_ITEMS;

__device__ __forceinline__ int32_t block_exclusive_scan(int32_t v, int32_t* total, int32_t* smem /*[32]*/) {
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    int32_t inc = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        int32_t t = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= o) inc += t;
    }
    if (lane == 31) smem[wid] = inc;
    __syncthreads();
    if (wid == 0) {
        int32_t w = (lane < (blockDim.x >> 5)) ? smem[lane] : 0;
        int32_t winc = w;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            int32_t t = __shfl_up_sync(0xffffffffu, winc, o);
            if (lane >= o) winc += t;
        }
        smem[lane] = winc - w;  // exclusive warp offsets
        if (lane == 31) smem[32] = winc;
    }
    __syncthreads();
    int32_t excl = inc - v + smem[wid];
    if (total) *total = smem[32];
    return excl;
}

__global__ void __launch_bounds__(SCAN_THREADS) scan_tile_sums(const int32_t* __restrict__ in, int32_t* __restrict__ sums, size_t n) {
    __shared__ int32_t sm[33];
    size_t base = (size_t)blockIdx.x * SCAN_TILE;
    int32_t s = 0;
#pragma unroll
    for (int k = 0; k < SCAN_ITEMS; k++) {
        size_t i = base + (size_t)k * SCAN_THREADS + threadIdx.x;
        if (i < n) s += in[i];
    }
    s = warp_sum(s);
    if ((threadIdx.x & 31) == 0) sm[threadIdx.x >> 5] = s;
    __syncthreads();
    if (threadIdx.x < 32) {
        int32_t v = threadIdx.x < (SCAN_THREADS / 32) ? sm[threadIdx.x] : 0;
        v = warp_sum(v);
        if (threadIdx.x == 0) sums[blockIdx.x] = v;
    }
}

__global__ void __launch_bounds__(SCAN_THREADS) scan_tiles(const int32_t* __restrict__ in, int32_t* __restrict__ out,
                                                           const int32_t* __restrict__ offsets, size_t n,
                                                           int32_t* __restrict__ total_out) {
    __shared__ int32_t sm[33];
    size_t base = (size_t)blockIdx.x * SCAN_TILE + (size_t)threadIdx.x * SCAN_ITEMS;
    int32_t v[SCAN_ITEMS];
    int32_t s = 0;
#pragma unroll
    for (int k = 0; k < SCAN_ITEMS; k++) {
        v[k] = (base + k < n) ? in[base + k] : 0;
        s += v[k];
    }
    int32_t tile_total;
    int32_t excl = block_exclusive_scan(s, &tile_total, sm) + (offsets ? offsets[blockIdx.x] : 0);
#pragma unroll
    for (int k = 0; k < SCAN_ITEMS; k++) {
        if (base + k < n) out[base + k] = excl;
        excl += v[k];
    }
    if (total_out && blockIdx.x == gridDim.x - 1 && threadIdx.x == SCAN_THREADS - 1) *total_out = excl;
}

int exclusive_scan_i32(const int32_t* in, int32_t* out, size_t n, int32_t* total_dev, cudaStream_t st) {
    if (n == 0) {
        if (total_dev) GMC_CUDA(cudaMemsetAsync(total_dev, 0, sizeof(int32_t), st));
        return GMC_OK;
    }
    size_t tiles = div_up(n, (size_t)SCAN_TILE);
    if (tiles == 1) {
        scan_tiles<<<1, SCAN_THREADS, 0, st>>>(in, out, nullptr, n, total_dev);
        GMC_CUDA(cudaGetLastError());
        return GMC_OK;
    }
    int32_t* sums = nullptr;
    GMC_CUDA(cudaMallocAsync((void**)&sums, tiles * sizeof(int32_t), st));
    scan_tile_sums<<<(unsigned)tiles, SCAN_THREADS, 0, st>>>(in, sums, n);
    GMC_CUDA(cudaGetLastError());
    int rc = exclusive_scan_i32(sums, sums, tiles, nullptr, st);
    if (rc != GMC_OK) return rc;
    scan_tiles<<<(unsigned)tiles, SCAN_THREADS, 0, st>>>(in, out, sums, n, total_dev);
    GMC_CUDA(cudaGetLastError());
    GMC_CUDA(cudaFreeAsync(sums, st));
    return GMC_OK;
}

}  // namespace gmc

using namespace gmc;

extern "C" {

const char* gmc_last_error(void) { return gmc::g_err; }
int gmc_version(void) { return 100; }

int gmc_device_count(int* count) {
    GMC_REQUIRE(count, GMC_ERR_ARG, "gmc_device_count: NULL");
    GMC_CUDA(cudaGetDeviceCount(count));
    return GMC_OK;
}

int gmc_init(int device) {
    int n = 0;
    GMC_CUDA(cudaGetDeviceCount(&n));
    GMC_REQUIRE(n > 0, GMC_ERR_CUDA, "gmc_init: no CUDA device (libgm_cuda has no CPU fallback)");
    GMC_REQUIRE(device >= 0 && device < n, GMC_ERR_ARG, "gmc_init: device %d out of range [0,%d)", device, n);
    GMC_CUDA(cudaSetDevice(device));
    GMC_CUDA(cudaFree(0));
    return GMC_OK;
}

int gmc_device_info(int* sms, size_t* total_bytes, size_t* free_bytes) {
    int dev = 0;
    GMC_CUDA(cudaGetDevice(&dev));
    if (sms) GMC_CUDA(cudaDeviceGetAttribute(sms, cudaDevAttrMultiProcessorCount, dev));
    size_t f = 0, t = 0;
    GMC_CUDA(cudaMemGetInfo(&f, &t));
    if (total_bytes) *total_bytes = t;
    if (free_bytes) *free_bytes = f;
    return GMC_OK;
}

int gmc_sync(void* stream) {
    GMC_CUDA(cudaStreamSynchronize(as_stream(stream)));
    return GMC_OK;
}

// Device arrays come from the device's default memory pool with an unbounded release threshold: graph / plan
// objects are created and destroyed per call on the reference-facing path (a fresh gm_graph mirror and plan per
// pagerank() call), and cudaMalloc/cudaFree of hundreds of MB cost milliseconds each and serialise on the driver.
// The allocation is made on the legacy default stream and completed before returning, so the pointer may be used
// on any stream; cudaFree() accepts pool allocations (memory goes back to the pool).
static thread_local int pool_ready_for_device = -1;     // per thread: a device group runs one host thread per device
int gmc_malloc(void** p, size_t bytes) {
    GMC_REQUIRE(p, GMC_ERR_ARG, "gmc_malloc: NULL");
    *p = nullptr;
    int dev = 0;
    GMC_CUDA(cudaGetDevice(&dev));
    if (pool_ready_for_device != dev) {
        cudaMemPool_t pool;
        GMC_CUDA(cudaDeviceGetDefaultMemPool(&pool, dev));
        unsigned long long keep = ~0ull;
        GMC_CUDA(cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &keep));
        pool_ready_for_device = dev;
    }
    cudaError_t e = cudaMallocAsync(p, bytes ? bytes : 1, (cudaStream_t)0);
    if (e == cudaSuccess) e = cudaStreamSynchronize((cudaStream_t)0);
    if (e == cudaErrorMemoryAllocation) {
        cudaGetLastError();
        set_error("gmc_malloc: out of device memory (%zu bytes)", bytes);
        return GMC_ERR_NOMEM;
    }
    GMC_CUDA(e);
    return GMC_OK;
}

// plain cudaMalloc: memory that is exported to peer processes with CUDA IPC (pool memory cannot be)
int gmc_malloc_ipc(void** p, size_t bytes) {
    GMC_REQUIRE(p, GMC_ERR_ARG, "gmc_malloc_ipc: NULL");
    *p = nullptr;
    cudaError_t e = cudaMalloc(p, bytes ? bytes : 1);
    if (e == cudaErrorMemoryAllocation) {
        cudaGetLastError();
        set_error("gmc_malloc_ipc: out of device memory (%zu bytes)", bytes);
        return GMC_ERR_NOMEM;
    }
    GMC_CUDA(e);
    return GMC_OK;
}

int gmc_free(void* p) {
    if (p) GMC_CUDA(cudaFree(p));
    return GMC_OK;
}

int gmc_h2d(void* dst, const void* src, size_t bytes, void* stream) {
    GMC_REQUIRE((dst && src) || bytes == 0, GMC_ERR_ARG, "gmc_h2d: NULL");
    if (bytes) GMC_CUDA(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyHostToDevice, as_stream(stream)));
    return GMC_OK;
}

int gmc_d2h(void* dst, const void* src, size_t bytes, void* stream) {
    GMC_REQUIRE((dst && src) || bytes == 0, GMC_ERR_ARG, "gmc_d2h: NULL");
    if (bytes) {
        GMC_CUDA(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToHost, as_stream(stream)));
        GMC_CUDA(cudaStreamSynchronize(as_stream(stream)));
    }
    return GMC_OK;
}

int gmc_host_register(void* p, size_t bytes) {
    GMC_REQUIRE(p, GMC_ERR_ARG, "gmc_host_register: NULL");
    GMC_CUDA(cudaHostRegister(p, bytes, cudaHostRegisterDefault));
    return GMC_OK;
}

int gmc_host_unregister(void* p) {
    if (p) GMC_CUDA(cudaHostUnregister(p));
    return GMC_OK;
}

}  // extern "C"
