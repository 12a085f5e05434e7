/* TEST INFRASTRUCTURE ONLY -- CPU restatement (plain C) of the Green-Marl hot path.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
 * load this library, and only as the checker (or the timed CPU arm) -- never on the product path.
 *
 * Parity pin: the reference holds NO golden vector for this path on synthetic graphs (its only
 * numeric goldens are CRCs / first elements on soc-LiveJournal1 and twitter_rv, datasets that are
 * not in the tree: scripts/extract_result.py:79-116,219-250).  The restatement is therefore pinned
 * against outputs of the reference itself run in the build container: oracle/_ref (the reference's
 * own gm_graph / generators / gm_bfs_template compiled from /root/reference, see Makefile) and the
 * fixtures in tests/golden/ produced from it by tests/golden/make_golden.py.
 *
 * All ids are the reference's default build: node_t = edge_t = int32_t (setup.mk:55-56).
 * Citations are relative to /root/reference/.
 */
#ifndef GM_ORACLE_H
#define GM_ORACLE_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

void orc_set_threads(int n);

/* --- synthetic generators (apps/output_cpp/gm_graph/src/graph_gen.cc) --- */
/* create_uniform_random_graph(N,M,seed,false): graph_gen.cc:58-104 (glibc srand/rand). */
int orc_uniform_random(int32_t N, int32_t M, long seed, int32_t* begin, int32_t* node_idx);
/* create_RMAT_graph(N,M,seed,a,b,c,permute): graph_gen.cc:159-287 (srand48/drand48). */
int orc_rmat(int32_t N, int32_t M, long seed, double a, double b, double c, int permute,
             int32_t* begin, int32_t* node_idx);

/* --- CSR preparation (apps/output_cpp/gm_graph/src/gm_graph.cc) --- */
/* do_semi_sort: gm_graph.cc:468-503 + semi_sort_main :380-423.  aux (e_idx2idx) may be NULL. */
void orc_semi_sort(int32_t N, const int32_t* begin, int32_t* node_idx, int32_t* aux);
/* make_reverse_edges (+ do_semi_sort_reverse): gm_graph.cc:205-304,461-466.  e_rev2idx may be NULL.
 * Among parallel edges the reference's order of e_rev2idx is unspecified (arrival order of an
 * atomic fetch-add, then an unstable std::sort); here ties are in ascending edge index. */
void orc_make_reverse(int32_t N, int32_t M, const int32_t* begin, const int32_t* node_idx,
                      int32_t* r_begin, int32_t* r_node_idx, int32_t* e_rev2idx);

/* --- binary CSR file, basic format (gm_graph_binary_loader.cc:19-26,57-205,222-252) --- */
int orc_store_binary(const char* path, int32_t N, int32_t M, const int32_t* begin, const int32_t* node_idx);
int orc_load_binary_header(const char* path, int32_t* N, int32_t* M);
int orc_load_binary(const char* path, int32_t* begin, int32_t* node_idx);

/* --- EDGE_LIST ingest: gm_edge_list_graph_reader::builtGraph (gm_edge_list_graph_reader.cc:68-171).
 * Lines "src dst" or "src *" (a vertex without an edge on this line), blank lines and lines starting with
 * '#' skipped, tokens separated by blanks/tabs; every id seen (source, destination, "*"-line vertex) gets a
 * vertex, numbered densely in ascending order of the raw id (:104-124); edges keep file order per source
 * (freeze), then do_semi_sort.  Two calls: counts first (begin == NULL), then the fill. */
int orc_edge_list_read(const char* path, int32_t* N, int32_t* M, int32_t* begin, int32_t* node_idx);

/* --- generated triangle_counting(): apps/src/triangle_counting.gm:1-13; three nested loops over the out-edge slots of v
 * (filters u > v, w > u) with G.has_edge_to(w, u) innermost = the binary search of gm_graph.cc:589-633. */
int64_t orc_triangle_counting(int32_t N, const int32_t* begin, const int32_t* node_idx);

/* --- generated comp_BC(): apps/src/bc.gm:4-29 over gm_bfs_template<short,...,save_child> (visit_fw: sigma over the
 * UpNbrs, gm_cpp_gen_foreach.cc:163-175; visit_rv: delta over the down edges, gm_bfs_template.h:273-319,496-524).
 * float arithmetic, sums in edge order.  The shipped bc.gm overwrites the root's sigma with 0 (no v != s filter), so
 * NaNs are the expected output wherever a vertex has a BFS child; bc is compared NaN-aware. */
void orc_comp_bc(int32_t N, const int32_t* begin, const int32_t* node_idx, const int32_t* r_begin,
                 const int32_t* r_node_idx, const int32_t* seeds, int32_t n_seeds, float* bc);
void orc_comp_bc_adj(int32_t N, const int32_t* begin, const int32_t* node_idx, const int32_t* r_begin,
                     const int32_t* r_node_idx, const int32_t* seeds, int32_t n_seeds, float* bc);

/* --- generated pagerank(): apps/src/pagerank.gm:1-20 (emitters: see ref_glue/pagerank.cc) ---
 * returns the number of iterations executed; *diff_out = last diff. */
int orc_pagerank(int32_t N, const int32_t* begin, const int32_t* r_begin, const int32_t* r_node_idx,
                 double e, double d, int32_t max_iter, double* rank, double* diff_out);

/* --- generated hop_dist(): apps/src/hop_dist.gm:3-31; returns the number of sweeps.
 * dist = BFS distance from root over out-edges, INT_MAX when unreachable. */
int orc_hop_dist(int32_t N, const int32_t* begin, const int32_t* node_idx, int32_t root, int32_t* dist);

/* --- gm_bfs_template<short,...,use_reverse_edge=false,save_child=false>::do_bfs_forward
 * (inc/gm_bfs_template.h:69-271, state machine :352-421).  levels[t] = level or -2 (:725).
 * state_trace (nullable, capacity trace_cap) receives the state (0..6, :713-719) used to expand
 * each level; level_count (nullable, same capacity) the vertex count of each level.
 * has_reverse = G.has_reverse_edge() (enables the bottom-up states).  Returns number of levels. */
int orc_bfs_levels(int32_t N, const int32_t* begin, const int32_t* node_idx,
                   const int32_t* r_begin, const int32_t* r_node_idx, int has_reverse,
                   int32_t root, int16_t* levels, int32_t* state_trace, int32_t* level_count,
                   int32_t trace_cap);
/* the same traversal under the state machine of a save_child instantiation (gm_bfs_template.h:374,:389) */
int orc_bfs_levels_save_child(int32_t N, const int32_t* begin, const int32_t* node_idx,
                   const int32_t* r_begin, const int32_t* r_node_idx, int has_reverse,
                   int32_t root, int16_t* levels, int32_t* state_trace, int32_t* level_count,
                   int32_t trace_cap);

#ifdef __cplusplus
}
#endif
#endif
