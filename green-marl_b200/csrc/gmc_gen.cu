// Device generators (K9) for the synthetic inputs of every config.
// Distributions follow the reference (apps/output_cpp/gm_graph/src/graph_gen.cc): uniform multigraph
// :58-104 (src = r % N, dst = r % N), R-MAT :159-287 (SCALE = floor(log2 N) levels, step = N/2 halving,
// a/b/c/d jittered by 0.95 + 0.1*U and renormalised at every level after the first, self edges redrawn,
// optional vertex permutation).  The reference draws from libc's serial rand()/drand48(); here each edge
// has its own counter-based stream (splitmix64 of seed, edge id, attempt), so graphs are not
// edge-for-edge those of the reference -- parity runs always give both sides the same CSR.
#include "gmc_common.cuh"
#include <cub/device/device_radix_sort.cuh>
#include <math.h>

using namespace gmc;

namespace {

__host__ __device__ __forceinline__ uint64_t splitmix64(uint64_t& s) {
    uint64_t z = (s += 0x9E3779B97F4A7C15ull);
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}
__device__ __forceinline__ double u01(uint64_t& s) { return (double)(splitmix64(s) >> 11) * (1.0 / 9007199254740992.0); }

__global__ void gen_uniform_edges(int32_t n, int32_t m, uint64_t seed, uint64_t* __restrict__ keys) {
    int32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= m) return;
    uint64_t s = seed * 0xD1B54A32D192ED03ull + (uint64_t)i * 0x2545F4914F6CDD1Dull;
    uint32_t src = (uint32_t)((splitmix64(s) >> 33) % (uint64_t)n);
    uint32_t dst = (uint32_t)((splitmix64(s) >> 33) % (uint64_t)n);
    keys[i] = ((uint64_t)src << 32) | dst;
}

__global__ void gen_rmat_edges(int32_t n, int32_t m, int32_t scale, uint64_t seed, double a, double b, double c,
                               const int32_t* __restrict__ perm, uint64_t* __restrict__ keys) {
    int32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= m) return;
    const double d = 1 - (a + b + c);
    int32_t u, v;
    for (uint32_t attempt = 0;; attempt++) {
        uint64_t s = seed * 0xD1B54A32D192ED03ull + (uint64_t)i * 0x2545F4914F6CDD1Dull + (uint64_t)attempt * 0x9E3779B97F4A7C15ull;
        u = 1; v = 1;
        int32_t step = n / 2;
        double av = a, bv = b, cv = c, dv = d;
        double p = u01(s);
        if (p < av) { } else if (p < av + bv) { v += step; } else if (p < av + bv + cv) { u += step; } else { v += step; u += step; }
        for (int32_t j = 1; j < scale; j++) {
            step = step / 2;
            av *= 0.95 + 0.1 * u01(s);
            bv *= 0.95 + 0.1 * u01(s);
            cv *= 0.95 + 0.1 * u01(s);
            dv *= 0.95 + 0.1 * u01(s);
            double S = av + bv + cv + dv;
            av /= S; bv /= S; cv /= S; dv /= S;
            p = u01(s);
            if (p < av) { } else if (p < av + bv) { v += step; } else if (p < av + bv + cv) { u += step; } else { v += step; u += step; }
        }
        if (u != v) break;   // "avoid self edges": redraw (graph_gen.cc:232-235)
    }
    uint32_t src = (uint32_t)(u - 1), dst = (uint32_t)(v - 1);
    if (perm) { src = (uint32_t)perm[src]; dst = (uint32_t)perm[dst]; }
    keys[i] = ((uint64_t)src << 32) | dst;
}

__global__ void perm_keys(int32_t n, uint64_t seed, uint64_t* __restrict__ keys, int32_t* __restrict__ ids) {
    int32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    uint64_t s = seed ^ 0xA5A5A5A55A5A5A5Aull;
    s += (uint64_t)i * 0x2545F4914F6CDD1Dull;
    keys[i] = splitmix64(s);
    ids[i] = i;
}

__global__ void split_sorted_edges(const uint64_t* __restrict__ keys, int32_t m, int32_t* __restrict__ node_idx, int32_t* __restrict__ degree) {
    int32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= m) return;
    uint64_t k = keys[i];
    node_idx[i] = (int32_t)(uint32_t)k;
    atomicAdd(&degree[(int32_t)(k >> 32)], 1);
}

int bits_for(int32_t n) { int b = 1; while (b < 31 && (1 << b) < n) b++; return b; }

// edge keys (src<<32|dst) -> frozen, semi-sorted CSR with reverse edges
int keys_to_graph(gmc_graph** out, int32_t n, int32_t m, uint64_t* keys, uint64_t* keys_alt) {
    gmc_graph* g = nullptr;
    GMC_TRY(gmc_graph_create(&g, n, m));
    int rc = GMC_OK;
    void* tmp = nullptr;
    size_t tmp_bytes = 0;
    cub::DoubleBuffer<uint64_t> db(keys, keys_alt);
    const int end_bit = 32 + bits_for(n);
    if (m > 0) {
        cub::DeviceRadixSort::SortKeys(nullptr, tmp_bytes, db, (int64_t)m, 0, end_bit);
        rc = gmc_malloc(&tmp, tmp_bytes);
        if (rc == GMC_OK) {
            cudaError_t e = cub::DeviceRadixSort::SortKeys(tmp, tmp_bytes, db, (int64_t)m, 0, end_bit);
            if (e != cudaSuccess) { set_error("edge sort: %s", cudaGetErrorString(e)); rc = GMC_ERR_CUDA; }
        }
    }
    if (rc == GMC_OK) {
        cudaMemset(g->begin, 0, sizeof(int32_t) * ((size_t)n + 1));
        if (m > 0) split_sorted_edges<<<div_up(m, 256), 256>>>(db.Current(), m, g->node_idx, g->begin);
        rc = exclusive_scan_i32(g->begin, g->begin, (size_t)n + 1, nullptr, nullptr);
    }
    if (rc == GMC_OK && cudaDeviceSynchronize() != cudaSuccess) { set_error("generator: %s", cudaGetErrorString(cudaGetLastError())); rc = GMC_ERR_CUDA; }
    cudaFree(tmp);
    if (rc != GMC_OK) { gmc_graph_destroy(g); return rc; }
    g->semi_sorted = true;
    *out = g;
    return GMC_OK;
}

}  // namespace

extern "C" {

// internal entry of the EDGE_LIST ingest (gmc_file.cc): host edge keys (src << 32 | dst) -> device graph.  One radix sort
// of the keys is the reference's add_edge + freeze + do_semi_sort (gm_edge_list_graph_reader.cc:137-168).
int gmc_graph_from_edge_keys(gmc_graph** out, int32_t n, int32_t m, const uint64_t* keys_host, int make_reverse) {
    GMC_REQUIRE(out && n >= 0 && m >= 0 && (keys_host || m == 0), GMC_ERR_ARG, "gmc_graph_from_edge_keys: bad args");
    uint64_t *k0 = nullptr, *k1 = nullptr;
    GMC_TRY(gmc_malloc((void**)&k0, sizeof(uint64_t) * (size_t)(m ? m : 1)));
    int rc = gmc_malloc((void**)&k1, sizeof(uint64_t) * (size_t)(m ? m : 1));
    if (rc == GMC_OK && m && cudaMemcpy(k0, keys_host, sizeof(uint64_t) * (size_t)m, cudaMemcpyHostToDevice) != cudaSuccess) {
        set_error("edge keys copy-in: %s", cudaGetErrorString(cudaGetLastError())); rc = GMC_ERR_CUDA;
    }
    if (rc == GMC_OK) rc = keys_to_graph(out, n, m, k0, k1);
    cudaFree(k0); cudaFree(k1);
    if (rc == GMC_OK && make_reverse) {
        rc = gmc_graph_make_reverse(*out, 0, nullptr);
        if (rc != GMC_OK) { gmc_graph_destroy(*out); *out = nullptr; }
    }
    return rc;
}

int gmc_graph_generate_uniform(gmc_graph** out, int32_t n, int32_t m, uint64_t seed) {
    GMC_REQUIRE(out && n > 0 && m >= 0, GMC_ERR_ARG, "gmc_graph_generate_uniform: bad args");
    uint64_t *k0 = nullptr, *k1 = nullptr;
    GMC_TRY(gmc_malloc((void**)&k0, sizeof(uint64_t) * (size_t)(m ? m : 1)));
    int rc = gmc_malloc((void**)&k1, sizeof(uint64_t) * (size_t)(m ? m : 1));
    if (rc == GMC_OK) {
        if (m) gen_uniform_edges<<<div_up(m, 256), 256>>>(n, m, seed, k0);
        rc = keys_to_graph(out, n, m, k0, k1);
    }
    cudaFree(k0); cudaFree(k1);
    if (rc == GMC_OK) rc = gmc_graph_make_reverse(*out, 0, nullptr);
    return rc;
}

int gmc_graph_generate_rmat(gmc_graph** out, int32_t n, int32_t m, uint64_t seed, double a, double b, double c, int permute) {
    GMC_REQUIRE(out && n > 1 && m >= 0, GMC_ERR_ARG, "gmc_graph_generate_rmat: bad args");
    GMC_REQUIRE(a > 0 && b >= 0 && c >= 0 && a + b + c < 1, GMC_ERR_ARG, "gmc_graph_generate_rmat: need a+b+c < 1");
    const int32_t scale = (int32_t)log2((double)n);    // SCALE = (node_t) log2(N), graph_gen.cc:185
    int32_t* perm = nullptr;
    int rc = GMC_OK;
    if (permute) {
        uint64_t *pk = nullptr, *pk2 = nullptr; int32_t *ids = nullptr; void* tmp = nullptr; size_t tb = 0;
        rc = gmc_malloc((void**)&pk, sizeof(uint64_t) * (size_t)n);
        if (rc == GMC_OK) rc = gmc_malloc((void**)&pk2, sizeof(uint64_t) * (size_t)n);
        if (rc == GMC_OK) rc = gmc_malloc((void**)&ids, sizeof(int32_t) * (size_t)n);
        if (rc == GMC_OK) rc = gmc_malloc((void**)&perm, sizeof(int32_t) * (size_t)n);
        if (rc == GMC_OK) {
            perm_keys<<<div_up(n, 256), 256>>>(n, seed, pk, ids);
            cub::DeviceRadixSort::SortPairs(nullptr, tb, pk, pk2, ids, perm, (int64_t)n);
            rc = gmc_malloc(&tmp, tb);
            if (rc == GMC_OK && cub::DeviceRadixSort::SortPairs(tmp, tb, pk, pk2, ids, perm, (int64_t)n) != cudaSuccess) rc = GMC_ERR_CUDA;
        }
        cudaDeviceSynchronize();
        cudaFree(pk); cudaFree(pk2); cudaFree(ids); cudaFree(tmp);
        if (rc != GMC_OK) { cudaFree(perm); return rc; }
    }
    uint64_t *k0 = nullptr, *k1 = nullptr;
    rc = gmc_malloc((void**)&k0, sizeof(uint64_t) * (size_t)(m ? m : 1));
    if (rc == GMC_OK) rc = gmc_malloc((void**)&k1, sizeof(uint64_t) * (size_t)(m ? m : 1));
    if (rc == GMC_OK) {
        if (m) gen_rmat_edges<<<div_up(m, 256), 256>>>(n, m, scale, seed, a, b, c, perm, k0);
        rc = keys_to_graph(out, n, m, k0, k1);
    }
    cudaFree(k0); cudaFree(k1); cudaFree(perm);
    if (rc == GMC_OK) rc = gmc_graph_make_reverse(*out, 0, nullptr);
    return rc;
}

}  // extern "C"
