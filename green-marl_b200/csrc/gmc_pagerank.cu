// PageRank pull sweep (K1) with the updater (K2) and init (K3) fused away.
//
// Reference loop being replaced (generated from apps/src/pagerank.gm:8-19, shape in SURVEY.md 3.1):
//   foreach t: S = sum_{w in InNbrs(t), r_node_idx order} rank[w] / (double)outdeg(w)
//              val = (1-d)/N + d*S;  diff += |val - rank[t]|;  rank_nxt[t] = val
//   rank = rank_nxt; cnt++           while (diff > e && cnt < max)
//
// Device formulation.  contrib[w] = rank[w]/(double)outdeg(w) is produced once per vertex by the
// epilogue of the previous sweep (the same IEEE division the reference performs once per EDGE), so a
// sweep gathers one double per edge.  rank/contrib are ping-ponged, which removes the copy-back loop.
//
// Load balance ("CSR-stream" over a merge-path split): CTA b owns the rows whose coordinate
// r_begin[t]+t falls in [b*W,(b+1)*W), i.e. about W rows+edges whatever the degree skew.  The CTA
// streams its contiguous slice of r_node_idx with coalesced read-once loads, gathers contrib[] for every
// edge into shared memory (8 independent gathers in flight per thread), then one thread per row adds
// the row's terms from shared memory strictly left to right -- the reference's summation order, so
// every row of in-degree <= PR_LONG is bit-identical to the CPU result.  Rows longer than PR_LONG
// (power-law hubs) are cut into PR_CHUNK-edge chunks, one CTA each, and combined in chunk order by a
// finalize kernel: deterministic, but a different association than left-to-right (<= few ulp).
//
// HBM traffic per sweep: 4 B/edge (index stream) + 40 B/vertex (row pointer, old rank, new rank, new
// contrib, out-degree, first touch of contrib); the per-edge gathers are served by L2/L1 when
// contrib[] (8N bytes) fits the 126 MB L2.
#include "gmc_common.cuh"
#include <vector>

using namespace gmc;

constexpr int PR_THREADS = 256;
constexpr int PR_W = 2048;        // rows + edges per CTA (merge-path window)
constexpr int PR_LONG = 2048;     // rows with more in-edges than this take the chunked path
constexpr int PR_STAGE = PR_W + PR_LONG;   // max edges staged in shared memory by one CTA
constexpr int PR_CHUNK = 8192;    // edges per CTA on the long-row path
constexpr int PR_GATHER = 8;      // independent gathers in flight per thread

struct gmc_pr_ctrl {   // device-side loop control: lets the host enqueue sweeps without a sync per iteration
    double diff;
    int32_t cnt;
    int32_t done;
};

struct gmc_pr_shard {
    int32_t n_global = 0, lo = 0, hi = 0, nrows = 0, nedges = 0;
    const int32_t* r_begin = nullptr;     // [nrows+1], local edge offsets
    const int32_t* r_node_idx = nullptr;  // [nedges], global source ids
    const int32_t* outdeg = nullptr;      // [nrows]
    bool owns_csr = false;
    // plan
    int32_t n_blocks = 0;
    int32_t* block_first_row = nullptr;   // [n_blocks+1]
    int32_t n_long = 0, n_long_chunks = 0;
    int32_t* long_rows = nullptr;         // [n_long] local row ids, ascending
    int32_t* long_chunk_begin = nullptr;  // [n_long+1]
    double* long_partials = nullptr;      // [n_long_chunks]
    double* diff_parts = nullptr;         // [n_blocks + n_long]
    // peers (multi-GPU direct exchange): device arrays of pointers, one per parity
    int n_peers = 0;
    double** peers[2] = {nullptr, nullptr};
};

namespace {

__global__ void compute_outdeg(const int32_t* __restrict__ begin, int32_t n, int32_t* __restrict__ outdeg) {
    int32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) outdeg[i] = begin[i + 1] - begin[i];
}

// block_first_row[b] = first local row t with r_begin[t] + t >= b*W   (b = 0..n_blocks; last = nrows)
__global__ void plan_blocks(const int32_t* __restrict__ r_begin, int32_t nrows, int32_t n_blocks, int32_t* __restrict__ first_row) {
    int32_t b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b > n_blocks) return;
    if (b == n_blocks) { first_row[b] = nrows; return; }
    int64_t target = (int64_t)b * PR_W;
    int32_t lo = 0, hi = nrows;  // answer in [lo, hi]
    while (lo < hi) {
        int32_t mid = lo + ((hi - lo) >> 1);
        if ((int64_t)r_begin[mid] + mid >= target) hi = mid; else lo = mid + 1;
    }
    first_row[b] = lo;
}

__global__ void flag_long_rows(const int32_t* __restrict__ r_begin, int32_t nrows, int32_t* __restrict__ flag, int32_t* __restrict__ nchunks) {
    int32_t t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= nrows) return;
    int32_t len = r_begin[t + 1] - r_begin[t];
    int32_t is_long = len > PR_LONG;
    flag[t] = is_long;
    nchunks[t] = is_long ? (len + PR_CHUNK - 1) / PR_CHUNK : 0;
}

__global__ void compact_long_rows(const int32_t* __restrict__ r_begin, int32_t nrows, const int32_t* __restrict__ pos,
                                  const int32_t* __restrict__ chunk_pos, int32_t* __restrict__ long_rows,
                                  int32_t* __restrict__ long_chunk_begin, int32_t n_long, int32_t n_chunks) {
    int32_t t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t == 0) long_chunk_begin[n_long] = n_chunks;
    if (t >= nrows) return;
    if (r_begin[t + 1] - r_begin[t] > PR_LONG) {
        long_rows[pos[t]] = t;
        long_chunk_begin[pos[t]] = chunk_pos[t];
    }
}

__global__ void pr_init(int32_t n_global, int32_t lo, int32_t nrows, const int32_t* __restrict__ outdeg,
                        double* __restrict__ contrib_full, double* __restrict__ rank_slice) {
    int32_t r = blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= nrows) return;
    const double v = 1 / (double)n_global;                 // G.pg_rank = 1 / N  (pagerank.gm:6-7)
    rank_slice[r] = v;
    int32_t od = outdeg[r];
    contrib_full[(size_t)lo + r] = od > 0 ? v / (double)od : 0.0;
}

struct RowOut {
    double d, base;
    const double* rank_in;
    double* rank_out;
    double* contrib_out;       // full-length, indexed by global id
    const int32_t* outdeg;
    double* const* peers;      // n_peers full-length contrib buffers on other GPUs (may be null)
    int n_peers;
    int32_t lo;
};

// per-row epilogue; mul and add kept separate (no FMA contraction) to match the CPU build's rounding
__device__ __forceinline__ double finish_row(const RowOut& o, int32_t r, double S) {
    double val = __dadd_rn(o.base, __dmul_rn(o.d, S));
    double old = o.rank_in[r];
    o.rank_out[r] = val;
    int32_t od = o.outdeg[r];
    double c = od > 0 ? val / (double)od : 0.0;
    size_t gi = (size_t)o.lo + r;
    o.contrib_out[gi] = c;
    for (int p = 0; p < o.n_peers; p++) o.peers[p][gi] = c;   // NVLink stores into peer HBM
    return fabs(__dsub_rn(val, old));
}

__device__ __forceinline__ double block_sum(double v, double* sm /*[PR_THREADS/32]*/) {
    v = warp_sum(v);
    if ((threadIdx.x & 31) == 0) sm[threadIdx.x >> 5] = v;
    __syncthreads();
    double t = 0.0;
    if (threadIdx.x < 32) {
        t = threadIdx.x < (PR_THREADS / 32) ? sm[threadIdx.x] : 0.0;
        t = warp_sum(t);
    }
    return t;  // valid in thread 0
}

__global__ void __launch_bounds__(PR_THREADS)
pr_sweep_stream(const int32_t* __restrict__ r_begin, const int32_t* __restrict__ r_node_idx,
                const int32_t* __restrict__ block_first_row, const double* __restrict__ contrib_in,
                RowOut o, double* __restrict__ diff_parts, const gmc_pr_ctrl* __restrict__ ctrl) {
    if (ctrl && ctrl->done) return;
    __shared__ double s_val[PR_STAGE];
    __shared__ int32_t s_off[PR_W + 2];
    __shared__ double s_red[PR_THREADS / 32];

    const int32_t first = block_first_row[blockIdx.x];
    const int32_t last = block_first_row[blockIdx.x + 1];
    const int32_t nrows = last - first;
    for (int32_t i = threadIdx.x; i <= nrows; i += PR_THREADS) s_off[i] = r_begin[first + i];
    __syncthreads();
    const int32_t e0 = s_off[0];
    int32_t nstage_rows = nrows;   // rows reduced from shared memory
    if (nrows > 0 && s_off[nrows] - s_off[nrows - 1] > PR_LONG) nstage_rows = nrows - 1;   // hub row: chunked path
    const int32_t ne = s_off[nstage_rows] - e0;

    // stream the index slice, gather contributions into shared memory
    const int32_t* __restrict__ idx = r_node_idx + e0;
    for (int32_t base = 0; base < ne; base += PR_THREADS * PR_GATHER) {
        int32_t w[PR_GATHER];
        double c[PR_GATHER];
#pragma unroll
        for (int k = 0; k < PR_GATHER; k++) {
            int32_t i = base + k * PR_THREADS + threadIdx.x;
            w[k] = (i < ne) ? ld_stream(idx + i) : -1;
        }
#pragma unroll
        for (int k = 0; k < PR_GATHER; k++) c[k] = (w[k] >= 0) ? contrib_in[w[k]] : 0.0;
#pragma unroll
        for (int k = 0; k < PR_GATHER; k++) {
            int32_t i = base + k * PR_THREADS + threadIdx.x;
            if (i < ne) s_val[i] = c[k];
        }
    }
    __syncthreads();

    // one thread per row, left-to-right (the reference's order)
    double diff = 0.0;
    for (int32_t r = threadIdx.x; r < nstage_rows; r += PR_THREADS) {
        int32_t b = s_off[r] - e0, e = s_off[r + 1] - e0;
        double S = 0.0;
        for (int32_t j = b; j < e; j++) S = S + s_val[j];
        diff += finish_row(o, first + r, S);
    }
    double tot = block_sum(diff, s_red);
    if (threadIdx.x == 0) diff_parts[blockIdx.x] = tot;
}

// hub rows: CTA c sums one PR_CHUNK-edge chunk of one long row
__global__ void __launch_bounds__(PR_THREADS)
pr_long_chunks(const int32_t* __restrict__ r_begin, const int32_t* __restrict__ r_node_idx,
               const int32_t* __restrict__ long_rows, const int32_t* __restrict__ long_chunk_begin, int32_t n_long,
               const double* __restrict__ contrib_in, double* __restrict__ partials, const gmc_pr_ctrl* __restrict__ ctrl) {
    if (ctrl && ctrl->done) return;
    __shared__ double s_red[PR_THREADS / 32];
    const int32_t c = blockIdx.x;
    int32_t lo = 0, hi = n_long;   // long_chunk_begin[lo] <= c < long_chunk_begin[hi]
    while (hi - lo > 1) {
        int32_t mid = (lo + hi) >> 1;
        if (long_chunk_begin[mid] <= c) lo = mid; else hi = mid;
    }
    const int32_t row = long_rows[lo];
    const int32_t k = c - long_chunk_begin[lo];
    const int32_t rb = r_begin[row], re = r_begin[row + 1];
    const int32_t cb = rb + k * PR_CHUNK;
    const int32_t ce = min(cb + PR_CHUNK, re);
    double acc[4] = {0.0, 0.0, 0.0, 0.0};
    int32_t i = cb + threadIdx.x;
    for (; i + 3 * PR_THREADS < ce; i += 4 * PR_THREADS) {
        int32_t w0 = ld_stream(r_node_idx + i), w1 = ld_stream(r_node_idx + i + PR_THREADS);
        int32_t w2 = ld_stream(r_node_idx + i + 2 * PR_THREADS), w3 = ld_stream(r_node_idx + i + 3 * PR_THREADS);
        acc[0] += contrib_in[w0]; acc[1] += contrib_in[w1]; acc[2] += contrib_in[w2]; acc[3] += contrib_in[w3];
    }
    for (; i < ce; i += PR_THREADS) acc[0] += contrib_in[ld_stream(r_node_idx + i)];
    double tot = block_sum((acc[0] + acc[1]) + (acc[2] + acc[3]), s_red);
    if (threadIdx.x == 0) partials[c] = tot;
}

__global__ void pr_long_finalize(const int32_t* __restrict__ long_rows, const int32_t* __restrict__ long_chunk_begin,
                                 int32_t n_long, const double* __restrict__ partials, RowOut o,
                                 double* __restrict__ diff_parts, const gmc_pr_ctrl* __restrict__ ctrl) {
    if (ctrl && ctrl->done) return;
    int32_t j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= n_long) return;
    double S = 0.0;
    for (int32_t c = long_chunk_begin[j]; c < long_chunk_begin[j + 1]; c++) S += partials[c];
    diff_parts[j] = finish_row(o, long_rows[j], S);
}

// diff = sum(diff_parts) in a fixed order; then either loop control (ctrl) or plain output (diff_out)
__global__ void __launch_bounds__(1024)
pr_reduce_diff(const double* __restrict__ parts, int32_t n_parts, gmc_pr_ctrl* ctrl, double e, int32_t max_iter,
               double* __restrict__ diff_out) {
    if (ctrl && ctrl->done) return;
    __shared__ double sm[32];
    double v = 0.0;
    for (int32_t i = threadIdx.x; i < n_parts; i += 1024) v += parts[i];
    v = warp_sum(v);
    if ((threadIdx.x & 31) == 0) sm[threadIdx.x >> 5] = v;
    __syncthreads();
    if (threadIdx.x < 32) {
        double t = sm[threadIdx.x];
        t = warp_sum(t);
        if (threadIdx.x == 0) {
            if (diff_out) *diff_out = t;
            if (ctrl) {
                int32_t cnt = ctrl->cnt + 1;
                ctrl->diff = t;
                ctrl->cnt = cnt;
                ctrl->done = !((t > e) && (cnt < max_iter));   // } While ((diff > e) && (cnt < max))
            }
        }
    }
}

int build_plan(gmc_pr_shard* s, cudaStream_t st) {
    const int32_t nrows = s->nrows;
    int64_t work = (int64_t)s->nedges + nrows;
    s->n_blocks = (int32_t)div_up(work, (int64_t)PR_W);
    if (s->n_blocks == 0) s->n_blocks = nrows > 0 ? 1 : 0;
    GMC_TRY(gmc_malloc((void**)&s->block_first_row, sizeof(int32_t) * ((size_t)s->n_blocks + 1)));
    plan_blocks<<<div_up(s->n_blocks + 1, 256), 256, 0, st>>>(s->r_begin, nrows, s->n_blocks, s->block_first_row);
    GMC_CUDA(cudaGetLastError());
    // long rows
    s->n_long = 0; s->n_long_chunks = 0;
    if (nrows > 0) {
        int32_t *flag = nullptr, *nchunks = nullptr, *tot = nullptr;
        GMC_TRY(gmc_malloc((void**)&flag, sizeof(int32_t) * (size_t)nrows));
        GMC_TRY(gmc_malloc((void**)&nchunks, sizeof(int32_t) * (size_t)nrows));
        GMC_TRY(gmc_malloc((void**)&tot, sizeof(int32_t) * 2));
        flag_long_rows<<<div_up(nrows, 256), 256, 0, st>>>(s->r_begin, nrows, flag, nchunks);
        GMC_TRY(exclusive_scan_i32(flag, flag, nrows, tot, st));
        GMC_TRY(exclusive_scan_i32(nchunks, nchunks, nrows, tot + 1, st));
        int32_t h[2];
        GMC_CUDA(cudaMemcpyAsync(h, tot, sizeof(h), cudaMemcpyDeviceToHost, st));
        GMC_CUDA(cudaStreamSynchronize(st));
        s->n_long = h[0]; s->n_long_chunks = h[1];
        if (s->n_long > 0) {
            GMC_TRY(gmc_malloc((void**)&s->long_rows, sizeof(int32_t) * (size_t)s->n_long));
            GMC_TRY(gmc_malloc((void**)&s->long_chunk_begin, sizeof(int32_t) * ((size_t)s->n_long + 1)));
            GMC_TRY(gmc_malloc((void**)&s->long_partials, sizeof(double) * (size_t)s->n_long_chunks));
            compact_long_rows<<<div_up(nrows, 256), 256, 0, st>>>(s->r_begin, nrows, flag, nchunks, s->long_rows, s->long_chunk_begin, s->n_long, s->n_long_chunks);
            GMC_CUDA(cudaGetLastError());
        }
        GMC_CUDA(cudaStreamSynchronize(st));
        cudaFree(flag); cudaFree(nchunks); cudaFree(tot);
    }
    GMC_TRY(gmc_malloc((void**)&s->diff_parts, sizeof(double) * ((size_t)s->n_blocks + s->n_long + 1)));
    return GMC_OK;
}

// one sweep = up to 3 launches (+1 reduce); returns the number of kernels launched
int launch_sweep(gmc_pr_shard* s, double d, const double* contrib_in, double* contrib_out, const double* rank_in,
                 double* rank_out, int parity, const gmc_pr_ctrl* ctrl, cudaStream_t st, int* launches) {
    RowOut o;
    o.d = d;
    o.base = (1 - d) / (double)s->n_global;     // (1-d) / N, pagerank.gm:11
    o.rank_in = rank_in; o.rank_out = rank_out; o.contrib_out = contrib_out;
    o.outdeg = s->outdeg;
    o.peers = s->n_peers ? s->peers[parity & 1] : nullptr;
    o.n_peers = s->n_peers;
    o.lo = s->lo;
    int n = 0;
    if (s->n_blocks > 0) {
        pr_sweep_stream<<<s->n_blocks, PR_THREADS, 0, st>>>(s->r_begin, s->r_node_idx, s->block_first_row, contrib_in, o, s->diff_parts + s->n_long, ctrl);
        n++;
    }
    if (s->n_long > 0) {
        pr_long_chunks<<<s->n_long_chunks, PR_THREADS, 0, st>>>(s->r_begin, s->r_node_idx, s->long_rows, s->long_chunk_begin, s->n_long, contrib_in, s->long_partials, ctrl);
        pr_long_finalize<<<div_up(s->n_long, 128), 128, 0, st>>>(s->long_rows, s->long_chunk_begin, s->n_long, s->long_partials, o, s->diff_parts, ctrl);
        n += 2;
    }
    GMC_CUDA(cudaGetLastError());
    if (launches) *launches += n;
    return GMC_OK;
}

void free_shard(gmc_pr_shard* s) {
    if (!s) return;
    if (s->owns_csr) { cudaFree((void*)s->r_begin); cudaFree((void*)s->r_node_idx); }
    cudaFree((void*)s->outdeg);
    cudaFree(s->block_first_row); cudaFree(s->long_rows); cudaFree(s->long_chunk_begin);
    cudaFree(s->long_partials); cudaFree(s->diff_parts);
    cudaFree(s->peers[0]); cudaFree(s->peers[1]);
    delete s;
}

}  // namespace

void gmc_pr_plan_free(gmc_graph* g) {
    if (g->pr) { free_shard(g->pr); g->pr = nullptr; }
    if (g->pr_buf) { cudaFree(g->pr_buf); g->pr_buf = nullptr; }
}

static int graph_plan(gmc_graph* g, cudaStream_t st) {
    if (g->pr) return GMC_OK;
    GMC_REQUIRE(g->has_reverse, GMC_ERR_STATE, "pagerank needs the reverse CSR: call gmc_graph_make_reverse first");
    gmc_pr_shard* s = new gmc_pr_shard();
    s->n_global = g->N; s->lo = 0; s->hi = g->N; s->nrows = g->N; s->nedges = g->M;
    s->r_begin = g->r_begin; s->r_node_idx = g->r_node_idx; s->owns_csr = false;
    int32_t* od = nullptr;
    int rc = gmc_malloc((void**)&od, sizeof(int32_t) * (size_t)(g->N ? g->N : 1));
    if (rc != GMC_OK) { delete s; return rc; }
    if (g->N) compute_outdeg<<<div_up(g->N, 256), 256, 0, st>>>(g->begin, g->N, od);
    s->outdeg = od;
    rc = build_plan(s, st);
    if (rc == GMC_OK) rc = gmc_malloc((void**)&g->pr_buf, sizeof(double) * 4 * (size_t)(g->N ? g->N : 1) + sizeof(gmc_pr_ctrl));
    if (rc != GMC_OK) { free_shard(s); return rc; }
    g->pr = s;
    return GMC_OK;
}

extern "C" {

int gmc_pagerank_dev(gmc_graph* g, double e, double d, int32_t max_iter, double* rank_dev, gmc_pagerank_info* info, void* stream) {
    GMC_REQUIRE(g && rank_dev, GMC_ERR_ARG, "gmc_pagerank_dev: NULL");
    cudaStream_t st = as_stream(stream);
    GMC_TRY(graph_plan(g, st));
    gmc_pr_shard* s = g->pr;
    const size_t n = (size_t)g->N;
    const size_t nn = n ? n : 1;
    double* contrib[2] = {g->pr_buf, g->pr_buf + nn};
    double* rank[2] = {g->pr_buf + 2 * nn, g->pr_buf + 3 * nn};
    gmc_pr_ctrl* ctrl = (gmc_pr_ctrl*)(g->pr_buf + 4 * nn);
    GMC_CUDA(cudaMemsetAsync(ctrl, 0, sizeof(gmc_pr_ctrl), st));
    cudaEvent_t ev0, ev1;
    GMC_CUDA(cudaEventCreate(&ev0));
    GMC_CUDA(cudaEventCreate(&ev1));
    int launches = 0;
    GMC_CUDA(cudaEventRecord(ev0, st));
    if (n) { pr_init<<<div_up((int32_t)n, 256), 256, 0, st>>>(g->N, 0, g->N, s->outdeg, contrib[0], rank[0]); launches++; }
    // do { sweep; cnt++ } while (diff > e && cnt < max): sweeps are enqueued in batches; kernels of a
    // finished loop exit at once on ctrl->done, so the host only synchronises once per batch.
    gmc_pr_ctrl h = {0.0, 0, 0};
    int32_t issued = 0;
    const int32_t BATCH = 16;
    do {
        int32_t todo = max_iter - issued;
        if (todo > BATCH) todo = BATCH;
        if (todo < 1) todo = 1;   // a do-while body runs at least once (max <= 0 still sweeps once)
        for (int32_t k = 0; k < todo; k++, issued++) {
            int p = issued & 1;
            GMC_TRY(launch_sweep(s, d, contrib[p], contrib[p ^ 1], rank[p], rank[p ^ 1], p, ctrl, st, &launches));
            pr_reduce_diff<<<1, 1024, 0, st>>>(s->diff_parts, s->n_blocks + s->n_long, ctrl, e, max_iter, nullptr);
            launches++;
        }
        GMC_CUDA(cudaMemcpyAsync(&h, ctrl, sizeof(h), cudaMemcpyDeviceToHost, st));
        GMC_CUDA(cudaStreamSynchronize(st));
    } while (!h.done);
    GMC_CUDA(cudaEventRecord(ev1, st));
    if (n) GMC_CUDA(cudaMemcpyAsync(rank_dev, rank[h.cnt & 1], sizeof(double) * n, cudaMemcpyDeviceToDevice, st));
    GMC_CUDA(cudaStreamSynchronize(st));
    if (info) {
        info->iterations = h.cnt;
        info->diff = h.diff;
        info->launches = launches;
        cudaEventElapsedTime(&info->kernel_ms, ev0, ev1);
    }
    cudaEventDestroy(ev0); cudaEventDestroy(ev1);
    return GMC_OK;
}

// measurement aid for the roofline figure: every sweep bracketed by its own pair of CUDA events
int gmc_pagerank_time_sweeps(gmc_graph* g, double d, int32_t warmup, int32_t n, float* ms_out, int32_t* launches_per_sweep, void* stream) {
    GMC_REQUIRE(g && ms_out && n > 0 && warmup >= 0, GMC_ERR_ARG, "gmc_pagerank_time_sweeps: bad args");
    GMC_REQUIRE(g->N > 0, GMC_ERR_ARG, "gmc_pagerank_time_sweeps: empty graph");
    cudaStream_t st = as_stream(stream);
    GMC_TRY(graph_plan(g, st));
    gmc_pr_shard* s = g->pr;
    const size_t nn = (size_t)g->N;
    double* contrib[2] = {g->pr_buf, g->pr_buf + nn};
    double* rank[2] = {g->pr_buf + 2 * nn, g->pr_buf + 3 * nn};
    pr_init<<<div_up(g->N, 256), 256, 0, st>>>(g->N, 0, g->N, s->outdeg, contrib[0], rank[0]);
    std::vector<cudaEvent_t> ev(2 * (size_t)n);
    for (auto& e : ev) GMC_CUDA(cudaEventCreate(&e));
    int launches = 0;
    for (int32_t k = 0; k < warmup + n; k++) {
        int p = k & 1;
        launches = 0;
        if (k >= warmup) GMC_CUDA(cudaEventRecord(ev[2 * (k - warmup)], st));
        GMC_TRY(launch_sweep(s, d, contrib[p], contrib[p ^ 1], rank[p], rank[p ^ 1], p, nullptr, st, &launches));
        if (k >= warmup) GMC_CUDA(cudaEventRecord(ev[2 * (k - warmup) + 1], st));
    }
    GMC_CUDA(cudaStreamSynchronize(st));
    for (int32_t k = 0; k < n; k++) GMC_CUDA(cudaEventElapsedTime(&ms_out[k], ev[2 * k], ev[2 * k + 1]));
    for (auto& e : ev) cudaEventDestroy(e);
    if (launches_per_sweep) *launches_per_sweep = launches;
    return GMC_OK;
}

int gmc_pagerank(gmc_graph* g, double e, double d, int32_t max_iter, double* rank_host, gmc_pagerank_info* info) {
    GMC_REQUIRE(g && (rank_host || g->N == 0), GMC_ERR_ARG, "gmc_pagerank: NULL");
    double* tmp = nullptr;
    GMC_TRY(gmc_malloc((void**)&tmp, sizeof(double) * (size_t)(g->N ? g->N : 1)));
    int rc = gmc_pagerank_dev(g, e, d, max_iter, tmp, info, nullptr);
    if (rc == GMC_OK) rc = gmc_d2h(rank_host, tmp, sizeof(double) * (size_t)g->N, nullptr);
    cudaFree(tmp);
    return rc;
}

// ---- destination-range shards (multi-GPU) -----------------------------------------------------
int gmc_pr_shard_create(gmc_pr_shard** out, int32_t n_global, int32_t lo, int32_t hi, const int32_t* r_begin_slice,
                        const int32_t* r_node_idx_slice, const int32_t* outdeg_slice) {
    GMC_REQUIRE(out && r_begin_slice && outdeg_slice, GMC_ERR_ARG, "gmc_pr_shard_create: NULL");
    GMC_REQUIRE(0 <= lo && lo <= hi && hi <= n_global, GMC_ERR_ARG, "gmc_pr_shard_create: bad range [%d,%d) of %d", lo, hi, n_global);
    const int32_t nrows = hi - lo;
    GMC_REQUIRE(r_begin_slice[0] == 0, GMC_ERR_ARG, "gmc_pr_shard_create: r_begin slice must be rebased to 0");
    const int32_t ne = r_begin_slice[nrows];
    GMC_REQUIRE(ne == 0 || r_node_idx_slice, GMC_ERR_ARG, "gmc_pr_shard_create: NULL r_node_idx");
    gmc_pr_shard* s = new gmc_pr_shard();
    s->n_global = n_global; s->lo = lo; s->hi = hi; s->nrows = nrows; s->nedges = ne; s->owns_csr = true;
    int32_t *rb = nullptr, *ri = nullptr, *od = nullptr;
    int rc = gmc_malloc((void**)&rb, sizeof(int32_t) * ((size_t)nrows + 1));
    if (rc == GMC_OK) rc = gmc_malloc((void**)&ri, sizeof(int32_t) * (size_t)(ne ? ne : 1));
    if (rc == GMC_OK) rc = gmc_malloc((void**)&od, sizeof(int32_t) * (size_t)(nrows ? nrows : 1));
    s->r_begin = rb; s->r_node_idx = ri; s->outdeg = od;
    if (rc == GMC_OK) rc = gmc_h2d(rb, r_begin_slice, sizeof(int32_t) * ((size_t)nrows + 1), nullptr);
    if (rc == GMC_OK && ne) rc = gmc_h2d(ri, r_node_idx_slice, sizeof(int32_t) * (size_t)ne, nullptr);
    if (rc == GMC_OK && nrows) rc = gmc_h2d(od, outdeg_slice, sizeof(int32_t) * (size_t)nrows, nullptr);
    if (rc == GMC_OK && cudaDeviceSynchronize() != cudaSuccess) rc = GMC_ERR_CUDA;
    if (rc == GMC_OK) rc = build_plan(s, nullptr);
    if (rc != GMC_OK) { free_shard(s); return rc; }
    *out = s;
    return GMC_OK;
}

int gmc_pr_shard_destroy(gmc_pr_shard* s) { free_shard(s); return GMC_OK; }

int gmc_pr_shard_set_peers(gmc_pr_shard* s, int n_peers, double* const* even, double* const* odd) {
    GMC_REQUIRE(s && n_peers >= 0 && (n_peers == 0 || (even && odd)), GMC_ERR_ARG, "gmc_pr_shard_set_peers: bad args");
    cudaFree(s->peers[0]); cudaFree(s->peers[1]);
    s->peers[0] = s->peers[1] = nullptr;
    s->n_peers = 0;
    if (n_peers == 0) return GMC_OK;
    GMC_TRY(gmc_malloc((void**)&s->peers[0], sizeof(double*) * n_peers));
    GMC_TRY(gmc_malloc((void**)&s->peers[1], sizeof(double*) * n_peers));
    GMC_CUDA(cudaMemcpy(s->peers[0], even, sizeof(double*) * n_peers, cudaMemcpyHostToDevice));
    GMC_CUDA(cudaMemcpy(s->peers[1], odd, sizeof(double*) * n_peers, cudaMemcpyHostToDevice));
    s->n_peers = n_peers;
    return GMC_OK;
}

int gmc_pr_shard_init(gmc_pr_shard* s, double* contrib_full, double* rank_slice, void* stream) {
    GMC_REQUIRE(s && contrib_full && rank_slice, GMC_ERR_ARG, "gmc_pr_shard_init: NULL");
    if (s->nrows) pr_init<<<div_up(s->nrows, 256), 256, 0, as_stream(stream)>>>(s->n_global, s->lo, s->nrows, s->outdeg, contrib_full, rank_slice);
    GMC_CUDA(cudaGetLastError());
    return GMC_OK;
}

// parity selects which registered peer buffer set receives the slice (the one matching contrib_out)
int gmc_pr_shard_sweep(gmc_pr_shard* s, double d, const double* contrib_in, double* contrib_out, const double* rank_in,
                       double* rank_out, double* diff_part_dev, int parity, void* stream) {
    GMC_REQUIRE(s && contrib_in && contrib_out && rank_in && rank_out, GMC_ERR_ARG, "gmc_pr_shard_sweep: NULL");
    cudaStream_t st = as_stream(stream);
    GMC_TRY(launch_sweep(s, d, contrib_in, contrib_out, rank_in, rank_out, parity, nullptr, st, nullptr));
    if (diff_part_dev) {
        pr_reduce_diff<<<1, 1024, 0, st>>>(s->diff_parts, s->n_blocks + s->n_long, nullptr, 0.0, 0, diff_part_dev);
        GMC_CUDA(cudaGetLastError());
    }
    return GMC_OK;
}

int gmc_pr_shard_launches_per_sweep(const gmc_pr_shard* s, int32_t* n) {
    GMC_REQUIRE(s && n, GMC_ERR_ARG, "NULL");
    *n = (s->n_blocks > 0 ? 1 : 0) + (s->n_long > 0 ? 2 : 0) + 1;
    return GMC_OK;
}

}  // extern "C"
