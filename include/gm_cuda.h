/* gm_cuda.h -- the thin extern "C" boundary of libgm_cuda (B200 / sm_100a).
 *
 * Everything the C++ facade (the headers under include/gm/: gm_graph, shl_array, gm_bfs_template, the "generated"
 * pagerank()/hop_dist()) and the Python harness need from the device goes through these entry points:
 * plain pointers and sizes, no C++ or torch types.  Ids are the reference's default build,
 * node_t = edge_t = int32_t (reference setup.mk:55-56; apps/output_cpp/gm_graph/inc/gm_graph_typedef.h:8-20).
 * Citations are relative to the reference tree (/root/reference).
 *
 * Conventions: every function returns GMC_OK (0) or a negative gmc_status; gmc_last_error() gives the
 * text (thread-local).  `stream` arguments are a cudaStream_t passed as void* (NULL = default stream).
 * There is NO CPU fallback: without a usable CUDA device every compute entry fails with GMC_ERR_CUDA.
 */
#ifndef GM_CUDA_H
#define GM_CUDA_H
#include <stddef.h>
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

typedef enum {
    GMC_OK = 0,
    GMC_ERR_CUDA = -1,      /* a CUDA runtime call failed / no device */
    GMC_ERR_ARG = -2,       /* bad argument (NULL, negative size, root out of range ...) */
    GMC_ERR_STATE = -3,     /* graph not in the state the call needs (e.g. no reverse CSR) */
    GMC_ERR_IO = -4,        /* file could not be read / written / wrong format */
    GMC_ERR_NOMEM = -5
} gmc_status;

const char* gmc_last_error(void);
int gmc_version(void);

/* --- device ------------------------------------------------------------------------------------
 * replaces shl__init(nthreads, partitioned) (reference apps/output_cpp/src/common_main.h:95-99):
 * "sockets" become devices. */
int gmc_device_count(int* count);
int gmc_init(int device);                          /* cudaSetDevice + context warm-up */
int gmc_device_info(int* sm_count, size_t* total_bytes, size_t* free_bytes);
int gmc_sync(void* stream);

/* --- raw device arrays: what shl_array<T>::alloc/copy_from/copy_back become in HBM
 * (call sites: reference src/backend_cpp/gm_cpp_gen.cc:687-775, 1468-1486). */
int gmc_malloc(void** dev_ptr, size_t bytes);
int gmc_free(void* dev_ptr);
int gmc_h2d(void* dev_dst, const void* host_src, size_t bytes, void* stream);
int gmc_d2h(void* host_dst, const void* dev_src, size_t bytes, void* stream);
int gmc_host_register(void* host_ptr, size_t bytes);   /* pin caller-owned host arrays for fast copies */
int gmc_host_unregister(void* host_ptr);

/* --- device-resident CSR mirror of gm_graph (reference apps/output_cpp/gm_graph/inc/gm_graph.h:119-445) */
typedef struct gmc_graph gmc_graph;
enum { GMC_ARR_BEGIN = 0, GMC_ARR_NODE_IDX = 1, GMC_ARR_R_BEGIN = 2, GMC_ARR_R_NODE_IDX = 3,
       GMC_ARR_E_REV2IDX = 4, GMC_ARR_E_IDX2IDX = 5 };

/* allocate begin[N+1], node_idx[M] in HBM: gm_graph::prepare_external_creation (gm_graph.cc:506-514) */
int gmc_graph_create(gmc_graph** out, int32_t num_nodes, int32_t num_edges);
int gmc_graph_destroy(gmc_graph* g);
int gmc_graph_num_nodes(const gmc_graph* g, int32_t* n);
int gmc_graph_num_edges(const gmc_graph* g, int32_t* m);
/* host -> HBM copy of the frozen CSR (the copy_from of shl__G_begin / shl__G_node_idx).  node_idx may be NULL: a
 * pull-only graph, exactly the arrays the generated pagerank() copies in (G_begin for the out-degrees, G_r_begin,
 * G_r_node_idx: reference src/backend_cpp/gm_cpp_gen.cc:685-775 emits one shl array per array the procedure reads);
 * calls that follow out-edges (BFS, semi-sort, reverse build) then return GMC_ERR_STATE. */
int gmc_graph_upload(gmc_graph* g, const int32_t* begin, const int32_t* node_idx, void* stream);
/* host-built reverse CSR -> HBM (when the caller already holds r_begin / r_node_idx) */
int gmc_graph_upload_reverse(gmc_graph* g, const int32_t* r_begin, const int32_t* r_node_idx, void* stream);
/* gm_graph::do_semi_sort on device (gm_graph.cc:468-503, semi_sort_main :380-423): rows ascending by
 * neighbour id; with want_idx2idx the permutation e_idx2idx is kept (ties: ascending original index). */
int gmc_graph_semi_sort(gmc_graph* g, int want_idx2idx, void* stream);
/* gm_graph::make_reverse_edges on device (gm_graph.cc:205-304 incl. do_semi_sort_reverse :293):
 * builds r_begin, r_node_idx (rows ascending by source id) and, if asked, e_rev2idx (ties: ascending
 * edge index). */
int gmc_graph_make_reverse(gmc_graph* g, int want_rev2idx, void* stream);
int gmc_graph_has_reverse(const gmc_graph* g, int* yes);
/* HBM -> host copy of one array (`which` = GMC_ARR_*); count is checked against the array length */
int gmc_graph_download(const gmc_graph* g, int which, int32_t* host_dst, size_t count, void* stream);
int gmc_graph_device_ptr(const gmc_graph* g, int which, const int32_t** dev_ptr);

/* --- synthetic generators on device (reference apps/output_cpp/gm_graph/src/graph_gen.cc:58-104,159-287).
 * Same edge distributions as the reference generators but a counter-based RNG (the reference's libc
 * rand()/drand48() streams are serial), so graphs are NOT edge-for-edge those of the reference; parity
 * runs feed the same CSR to both sides.  The result is frozen, semi-sorted, with reverse edges. */
int gmc_graph_generate_uniform(gmc_graph** out, int32_t num_nodes, int32_t num_edges, uint64_t seed);
int gmc_graph_generate_rmat(gmc_graph** out, int32_t num_nodes, int32_t num_edges, uint64_t seed,
                            double a, double b, double c, int permute);

/* --- binary CSR file, basic format (reference gm_graph_binary_loader.cc:19-26): bulk read + byte swap */
int gmc_file_read_header(const char* path, int32_t* num_nodes, int32_t* num_edges);
int gmc_file_read(const char* path, int32_t* begin, int32_t* node_idx);
int gmc_file_write(const char* path, int32_t num_nodes, int32_t num_edges, const int32_t* begin,
                   const int32_t* node_idx);

/* --- EDGE_LIST text ingest (reference gm_edge_list_graph_reader.cc:68-171, gm_graph::load_edge_list without
 * property columns; the format gm_format_converter reads, README.md:24-35): lines "src dst" or "src *", '#'
 * comments, ids densified in ascending order.  Parsed by all host threads, CSR + semi-sort (+ reverse CSR)
 * built on the device; the graph stays resident (download with gmc_graph_download). */
int gmc_edge_list_read(const char* path, gmc_graph** out, int make_reverse);

/* --- PageRank: the generated pagerank() (reference apps/src/pagerank.gm:1-20; emitted loop shape
 * reconstructed in SURVEY.md 3.1).  rank = 1/N; do { sweep; cnt++ } while (diff > e && cnt < max_iter).
 * Needs the reverse CSR.  fp64 throughout; per-edge term rank[w]/(double)outdeg(w) as in the reference. */
typedef struct {
    int32_t iterations;      /* cnt at exit */
    double  diff;            /* last diff = sum_t |val_t - rank_t| */
    float   kernel_ms;       /* device time of the iteration loop (CUDA events, sweep kernels only) */
    int32_t launches;        /* kernels launched inside that region */
} gmc_pagerank_info;
/* result copied to caller-owned HOST memory (what shl_array::copy_back does, gm_cpp_gen.cc:1480) */
int gmc_pagerank(gmc_graph* g, double e, double d, int32_t max_iter, double* rank_host,
                 gmc_pagerank_info* info);
/* result left in caller-provided DEVICE memory (N doubles) */
int gmc_pagerank_dev(gmc_graph* g, double e, double d, int32_t max_iter, double* rank_dev,
                     gmc_pagerank_info* info, void* stream);

/* The whole generated procedure on HOST arrays in one call: copy-in of what pagerank() reads (begin for the
 * out-degrees, r_begin, r_node_idx: the Shoal copy-in of gm_cpp_gen.cc:685-775), the loop nest, copy_back of
 * the rank array (:1480).  Same result as gmc_graph_upload + gmc_graph_upload_reverse + gmc_pagerank; the
 * difference is overlap: the copy-in runs on its own stream and the sweep plan is built behind it.  Pin the
 * arrays (gmc_host_register) for full PCIe rate.  Nothing stays on the device afterwards. */
int gmc_pagerank_host_csr(int32_t num_nodes, int32_t num_edges, const int32_t* begin, const int32_t* r_begin,
                          const int32_t* r_node_idx, double e, double d, int32_t max_iter, double* rank_host,
                          gmc_pagerank_info* info);

/* measurement aid (bench.py's roofline figure): rank = 1/N, `warmup` untimed sweeps, then `n` sweeps each
 * bracketed by its own pair of CUDA events on `stream`; ms_out[i] = device time of timed sweep i. */
int gmc_pagerank_time_sweeps(gmc_graph* g, double d, int32_t warmup, int32_t n, float* ms_out,
                             int32_t* launches_per_sweep, void* stream);

/* layout facts of the PageRank plan (relabelled entry array, tasks), for tests and reports */
typedef struct {
    int32_t n_long;         /* rows with in-degree > 512 (row-major, padded to 4 entries) */
    int32_t hot_n;          /* contributions staged in shared memory per CTA (dummy vertex included) */
    int32_t n_slices;       /* sliced-ELL slices (32 rows each) of the rows with in-degree <= 512 */
    int32_t long_entries;   /* entries of the long-row section (edges + padding) */
    int32_t total_entries;  /* entries of the whole array */
    int32_t n_long_tasks, n_short_tasks;
    int32_t grid, threads, smem_bytes;
    /* tiered layout (green-marl_b200/csrc/gmc_pr_plan.h) */
    int32_t n_pos;          /* rows with in-edges: the rows a steady-state sweep processes */
    int32_t n_src;          /* vertices with out-edges: the contributions a sweep can gather */
    int32_t tiers;          /* warm source tiers (entry streams beside the home stream) */
    int32_t segments;       /* (row, tier) segments = part slots beyond the home parts */
    int32_t split_rows;     /* rows with at least one segment (finished by pr_combine) */
    int32_t part_slots;
    int32_t tier_ctas;      /* CTAs that start in a tier stream */
    int32_t home_entries;   /* entry slots of the home stream */
    int64_t hot0_entries;   /* gathers of the home stream served by tier 0 (shared memory) */
    int64_t cold_entries;   /* gathers of the home stream that go to L2 */
    int64_t segment_entries;/* entry slots of the tier streams (shared-memory gathers) */
    int64_t cold_stream_entries; /* entry slots of the cold stream (L2 gathers of split rows, taken by any CTA) */
} gmc_pagerank_plan_info_t;
int gmc_pagerank_plan_info(gmc_graph* g, gmc_pagerank_plan_info_t* out);

/* --- BFS.  One direction-optimizing, frontier-compacted engine behind both reference entry points:
 *  - generated hop_dist() (reference apps/src/hop_dist.gm:3-31): dist int32, INT_MAX = unreachable;
 *  - gm_bfs_template<short,...>::do_bfs_forward (reference inc/gm_bfs_template.h:69-271): level int16,
 *    -2 = unvisited (:725), per-level vertex counts and the state used for each level (:713-719).
 * The level-to-level state machine follows the reference's transition rules (get_next_state :352-421). */
typedef struct {
    int32_t levels;          /* number of BFS levels (root level included) */
    int32_t reached;         /* vertices reached (root included) */
    int64_t edges_traversed; /* sum of out-degrees of reached vertices (M_r of SURVEY.md 8d) */
    float   kernel_ms;       /* device time from first to last kernel */
    int32_t launches;
    int32_t bottom_up_levels;
} gmc_bfs_info;
int gmc_hop_dist(gmc_graph* g, int32_t root, int32_t* dist_host, gmc_bfs_info* info);
int gmc_hop_dist_dev(gmc_graph* g, int32_t root, int32_t* dist_dev, gmc_bfs_info* info, void* stream);
/* use_reverse_edge: the template parameters of the reference class that change the traversal, as flags:
 * GMC_BFS_REVERSE_EDGE (= 1, so a plain boolean keeps working) follows in-edges instead of out-edges;
 * GMC_BFS_SAVE_CHILD selects the state machine of a save_child instantiation, which never switches QUE / Q2R -> RRD
 * directly (gm_bfs_template.h:374,:389), only RD -> RRD (:398-400).  Levels do not depend on it; the states do, and the
 * reference marks down edges only on the levels it expands top-down (states below GMC_BFS_ST_RRD).
 * state_trace / level_count: optional host arrays of capacity trace_cap; states are the reference's (:713-719). */
enum { GMC_BFS_REVERSE_EDGE = 1, GMC_BFS_SAVE_CHILD = 2 };
enum { GMC_BFS_ST_SMALL = 0, GMC_BFS_ST_QUE = 1, GMC_BFS_ST_Q2R = 2, GMC_BFS_ST_RD = 3, GMC_BFS_ST_R2Q = 4,
       GMC_BFS_ST_RRD = 5, GMC_BFS_ST_RR2Q = 6 };
int gmc_bfs_levels(gmc_graph* g, int32_t root, int use_reverse_edge, int16_t* level_host,
                   int32_t* state_trace, int32_t* level_count, int32_t trace_cap, gmc_bfs_info* info);
/* same traversal restricted to the edges with edge_mask_host[e] != 0 (e indexes the arrays being followed:
 * node_idx, or r_node_idx with use_reverse_edge): the has_navigator form of the reference template
 * (check_navigator, gm_bfs_template.h:488-492,596-598).  Stays top-down.  NULL mask = no filter. */
int gmc_bfs_levels_masked(gmc_graph* g, int32_t root, int use_reverse_edge, const uint8_t* edge_mask_host,
                          int16_t* level_host, int32_t* state_trace, int32_t* level_count, int32_t trace_cap,
                          gmc_bfs_info* info);
/* vertices of every level, concatenated in level order (the reference's global_vector /
 * level_queue_begin, gm_bfs_template.h:423-447), for do_bfs_reverse and host callback replay.
 * order_host[N]; level_begin_host[levels+1]. Valid after gmc_bfs_levels on the same graph. */
int gmc_bfs_level_order(gmc_graph* g, int32_t* order_host, int32_t* level_begin_host, int32_t cap_levels);

/* --- comp_BC: the generated comp_BC() (reference apps/src/bc.gm:4-29) with the two traversal
 * callbacks on the device: per seed one BFS, visit_fw (sigma over the UpNbrs, gm_cpp_gen_foreach.cc:163-175) per level in
 * ascending order, visit_rv (delta over the down edges, gm_bfs_template.h:273-319,496-524; BC += delta) per level in
 * descending order.  float arithmetic, every vertex's sum in edge-slot order as in the reference's loops, so BC is
 * bit-identical to the CPU result of a single-thread run (NaN where the reference yields NaN; with several threads the
 * reference's down-edge marking races).  Kept reference behaviour: a level expanded bottom-up (state RRD / RR2Q) has no
 * down edges -- check_parent_rrd (:689-708) does not mark them -- so its vertices get delta = 0.  Seeds are processed in the given order;
 * bc_host[N] is overwritten.  Needs the forward and the reverse CSR. */
typedef struct gmc_bc_info {
    int32_t seeds;        /* traversals done */
    int32_t launches;     /* kernels launched (traversal + bucketing + 2 per level) */
    int32_t max_levels;   /* deepest traversal */
    float device_ms;      /* device time of the whole call up to the copy-out */
} gmc_bc_info;
int gmc_comp_bc(gmc_graph* g, const int32_t* seeds, int32_t n_seeds, float* bc_host, gmc_bc_info* info);
/* the same procedure as apps/src/bc_adj.gm:4-29 writes it: InBFS(v: G.Nodes From s)(v != s) / InReverse(v != s) -- both
 * passes skip the root, whose sigma therefore stays 1 (bc.gm overwrites it with an empty sum: all-zero sigma, NaN delta) */
int gmc_comp_bc_adj(gmc_graph* g, const int32_t* seeds, int32_t n_seeds, float* bc_host, gmc_bc_info* info);

/* --- triangle_counting: the generated triangle_counting() (reference
 * apps/src/triangle_counting.gm:1-13): T = #{(v, slot i, slot j) : u = nbr_i(v) > v, w = nbr_j(v) > u, edge w -> u},
 * the inner has_edge_to loop (gm_graph.cc:60-66,589-633) evaluated as one sorted-list intersection per edge slot.
 * Needs ascending rows and the reverse CSR.  Exact (integer) result.  kernel_ms (nullable) = device time.
 * _range counts the edge slots [e_lo, e_hi) only: ranks of a multi-GPU job take slices and add their counts. */
int gmc_triangle_counting(gmc_graph* g, int64_t* count, float* kernel_ms);
int gmc_triangle_counting_range(gmc_graph* g, int64_t e_lo, int64_t e_hi, int64_t* count, float* kernel_ms, void* stream);
/* No counterpart in the reference: added for callers that cannot call cudaMalloc (code
 * under CUDA graph capture, tight serving loops, harnesses that forbid it).  The same count as gmc_triangle_counting, but the
 * call allocates and frees nothing and does not synchronise: the kernels' scratch is a fixed file-scope device array (about
 * 0.9 MB, held by every GPU that loads the library until exit), so graphs of more than GMC_TC_NOALLOC_MAX_EDGES edge slots are
 * refused with GMC_ERR_ARG.  The graph (inputs) is the caller's resident gmc_graph; the output is ONE int64 in DEVICE memory
 * the caller owns (count_dev), written on `stream`: the count, or -1 when a row of either CSR is not ascending (where
 * gmc_triangle_counting returns GMC_ERR_STATE).  All calls on a device share the scratch array, which keeps the last call's
 * contents (each call clears what must start at zero): calls that overlap on the GPU race -- serialise them. */
#define GMC_TC_NOALLOC_MAX_EDGES (1 << 26)
int gmc_triangle_counting_noalloc(gmc_graph* g, int64_t* count_dev, void* stream);

/* --- multi-GPU: one process per GPU, destination-vertex ranges per rank.  Shoal's
 * replicate / partition placement (reference src/backend_cpp/gm_cpp_gen.cc:181-227: shl__malloc_array with
 * is_ro / is_indexed) becomes: CSR and bitmaps replicated, the streamed entry array / owned rows partitioned.
 * The communicator wraps NCCL (bound at run time); rank 0 makes the id, the host side ships its 128 bytes to
 * the other ranks (bench.py uses torch.distributed for that), every rank calls gmc_comm_create. */
#define GMC_COMM_ID_BYTES 128
typedef struct gmc_comm gmc_comm;
int gmc_comm_unique_id(void* id128);
int gmc_comm_create(gmc_comm** out, int world, int rank, const void* id128);   /* world == 1: no NCCL needed */
int gmc_comm_destroy(gmc_comm* c);
int gmc_comm_info(const gmc_comm* c, int* world, int* rank, int* nccl_version);
int gmc_comm_barrier(gmc_comm* c, void* stream);     /* device-side barrier (1-element allreduce) + stream sync */
/* host-side balanced partition of a prefix-sum array into `world` contiguous ranges (cuts[world+1]) */
int gmc_shard_cuts(const int32_t* prefix, int32_t n, int world, int32_t align, int32_t* cuts);

/* sharded copy-in: each rank copies its 1/world slice of every CSR array host->device, the slices are allgathered
 * over NVLink (the job moves the graph across PCIe once, not `world` times).  r_begin may be NULL (no reverse CSR). */
int gmc_graph_upload_sharded(gmc_graph* g, gmc_comm* comm, const int32_t* begin, const int32_t* node_idx,
                             const int32_t* r_begin, const int32_t* r_node_idx, void* stream);

/* PageRank over destination-range shards: rank r sweeps only its share of the entry array, then the new
 * contributions are allgathered (NCCL over NVLink) before the next sweep; diff is combined in rank order, so
 * every rank leaves the loop on the same iteration.  rank_dev receives the full vector on every rank. */
int gmc_pagerank_sharded_dev(gmc_graph* g, gmc_comm* comm, double e, double d, int32_t max_iter, double* rank_dev,
                             gmc_pagerank_info* info, void* stream);
/* test aid: the `world` shard sweeps run one after the other on this device (no NCCL): same plan code,
 * same kernels, same result as a `world`-rank run */
int gmc_pagerank_emulated_shards_dev(gmc_graph* g, int world, double e, double d, int32_t max_iter, double* rank_dev,
                                     gmc_pagerank_info* info, void* stream);
/* how the last sharded call exchanged contributions: 0 = single rank, 1 = NCCL grouped broadcasts after each sweep,
 * 2 = stores into the peers' memory (CUDA IPC mapping, NVLink) from the sweep's row epilogue + sequence flags.
 * Path 2 is taken when every rank could map every peer (world <= 8); GMC_NO_P2P=1 forces path 1. */
int gmc_pagerank_exchange_path(gmc_graph* g, int* path);
/* per-iteration device times of a sharded run: sweep_ms / exchange_ms [n], measured with CUDA events */
int gmc_pagerank_sharded_time_sweeps(gmc_graph* g, gmc_comm* comm, double d, int32_t warmup, int32_t n, float* sweep_ms,
                                     float* exchange_ms, void* stream);

/* --- device group: 1/2/4/8 GPUs behind ONE call of ONE process (one host thread per device inside the library).  This is
 * what the reference's placement switch maps to -- shl__init(nthreads, partitioned), apps/output_cpp/src/common_main.h:95-99;
 * scripts/run.sh:86-92 (-r replicate / -p partition) -- and how the unchanged drivers reach more than one B200
 * (GM_CUDA_DEVICES=N, host/shl.cc).  The CSR is replicated (every device copies 1/N of it over its own PCIe link, the slices
 * are allgathered over NVLink), rows are partitioned into destination ranges, the peers' state blocks are reached with plain
 * device pointers (cudaDeviceEnablePeerAccess); results land in the caller's ONE host array, each device copying its slice.
 * devices == NULL: devices 0..ndev-1.  A group holds one graph at a time. */
typedef struct gmc_group gmc_group;
int gmc_group_create(gmc_group** out, int ndev, const int* devices);
int gmc_group_destroy(gmc_group* grp);
int gmc_group_size(const gmc_group* grp, int* ndev);
int gmc_group_graph_upload(gmc_group* grp, int32_t num_nodes, int32_t num_edges, const int32_t* begin, const int32_t* node_idx,
                           const int32_t* r_begin, const int32_t* r_node_idx);     /* node_idx may be NULL (pull-only), r_* too */
int gmc_group_graph_free(gmc_group* grp);
int gmc_group_pagerank(gmc_group* grp, double e, double d, int32_t max_iter, double* rank_host, gmc_pagerank_info* info);
int gmc_group_hop_dist(gmc_group* grp, int32_t root, int32_t* dist_host, gmc_bfs_info* info);
int gmc_group_exchange_path(gmc_group* grp, int* path);     /* as gmc_pagerank_exchange_path, of the last gmc_group_pagerank */

/* BFS over vertex-range shards: per level each rank expands / scans its owned vertices, then the candidate
 * bitmaps are exchanged with ncclAllGather.  dist_dev / level_host receive the full array on every rank. */
int gmc_hop_dist_sharded_dev(gmc_graph* g, gmc_comm* comm, int32_t root, int32_t* dist_dev, gmc_bfs_info* info, void* stream);
int gmc_bfs_levels_sharded(gmc_graph* g, gmc_comm* comm, int32_t root, int use_reverse_edge, const uint8_t* edge_mask_host,
                           int16_t* level_host, int32_t* state_trace, int32_t* level_count, int32_t trace_cap, gmc_bfs_info* info);

/* triangle_counting on several ranks: rank r counts the edge slots [r*M/world, (r+1)*M/world) of the replicated graph, one
 * 8-byte allreduce adds the partial counts (the procedure has no other exchange step); every rank returns the total. */
int gmc_triangle_counting_sharded(gmc_graph* g, gmc_comm* comm, int64_t* count, float* kernel_ms, void* stream);

#ifdef __cplusplus
}
#endif
#endif
