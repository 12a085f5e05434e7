// gm_graph: the CSR graph container of the Green-Marl runtime, API-compatible with the reference class
// (apps/output_cpp/gm_graph/inc/gm_graph.h:119-445) for everything the PageRank / BFS path and its
// drivers use, re-implemented over a device-resident mirror.
//
// Contract kept from the reference: `begin/node_idx/r_begin/r_node_idx/e_rev2idx/e_idx2idx` are PUBLIC
// HOST arrays (generators write through them, generated code and drivers read them), owned by the graph.
// New: the frozen CSR also lives in HBM (a gmc_graph, include/gm_cuda.h).  The two K4/K5 builders --
// make_reverse_edges() and do_semi_sort() -- run on the device and copy their results back into the
// host arrays.  The mirror is created on first use and dropped by thaw()/clear_graph()/
// prepare_external_creation(); code that rewrites begin[]/node_idx[] in place afterwards must call
// invalidate_device_mirror().
#ifndef GM_GRAPH_H_
#define GM_GRAPH_H_
#include <assert.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <omp.h>
#include <map>
#include <string>
#include <unordered_map>
#include <vector>
#include "gm_internal.h"

typedef node_t node_id;
typedef edge_t edge_id;

struct edge_dest_t {   // flexible (thawed) representation: neighbour + edge id
    node_id dest;
    edge_id edge;
};

struct gmc_graph;      // device mirror (gm_cuda.h)

#define MAGIC_WORD_BIN 0x03939999
#define MAGIC_WORD_EBIN 0x99191191

class gm_graph {
  public:
    gm_graph();
    ~gm_graph();

    // ---- direct access (valid while frozen) ----
    edge_t* begin;            // [N+1]
    node_t* node_idx;         // [M] destination of each edge
    node_t* node_idx_src;     // [M] source of each edge           (prepare_edge_source)
    edge_t* r_begin;          // [N+1]                             (make_reverse_edges)
    node_t* r_node_idx;       // [M] source of each reverse edge's original edge
    node_t* r_node_idx_src;   // [M]
    edge_t* e_idx2idx;        // [M] sorted edge idx -> original edge idx   (do_semi_sort)
    edge_t* e_rev2idx;        // [M] reverse edge idx -> forward edge idx   (make_reverse_edges)

    static const node_t NIL_NODE = (node_t)-1;
    static const edge_t NIL_EDGE = (edge_t)-1;

    bool is_neighbor(node_t src, node_t to);
    bool has_edge_to(node_t source, node_t to);
    edge_t get_edge_idx_for_src_dest(node_t src, node_t dest);

    node_t num_nodes() { return _numNodes; }
    edge_t num_edges() { return _numEdges; }
    bool has_reverse_edge() { return _reverse_edge; }
    bool is_frozen() { return _frozen; }
    bool is_directed() { return _directed; }
    bool is_semi_sorted() { return _semi_sorted; }
    bool has_separate_edge_idx() { return e_id2idx != NULL; }
    bool is_edge_source_ready() { return node_idx_src != NULL; }

    // ---- structure changes ----
    void thaw();
    void freeze();
    void make_reverse_edges();     // K4, on device
    void do_semi_sort();           // K5, on device
    void prepare_edge_source();

    node_id add_node();
    edge_id add_edge(node_id n, node_id m);
    bool is_node(node_id n) { return n < _numNodes; }
    bool is_edge(edge_id n) { return n < _numEdges; }
    bool has_edge(node_id from, node_id to);
    edge_t get_num_edges(node_id from, node_id to);
    edge_t get_num_edges(node_id from) { return begin[from + 1] - begin[from]; }

    // ---- binary file, basic format ----
    void prepare_external_creation(node_t n, edge_t m);
    void prepare_external_creation(node_t n, edge_t m, bool clean_key_id_mappings);
    bool store_binary(char* filename);
    bool load_binary(char* filename);
    // ---- EDGE_LIST text (reference gm_edge_list_graph_reader.cc:5-18,68-171); property columns are not read:
    // the schemas must be empty, as they are for gm_format_converter's plain edge lists ----
    bool load_edge_list(char* filename, std::vector<VALUE_TYPE>& vprop_schema, std::vector<VALUE_TYPE>& eprop_schema,
                        std::vector<void*>& vertex_props, std::vector<void*>& edge_props, bool use_hdfs = false);

    inline edge_t get_edge_idx(edge_id e) { return e_id2idx == NULL ? e : e_id2idx[e]; }
    inline node_t get_node_idx(node_id n) { return n; }
    inline edge_id get_edge_id(edge_t e) { return e_idx2id == NULL ? e : e_idx2id[e]; }
    inline node_id get_node_id(node_t n) { return n; }
    inline edge_t get_org_edge_idx(edge_id e) { return e_idx2idx == NULL ? e : e_idx2idx[e]; }

    void clear_graph(bool clean_key_id_mappings);
    void clear_graph();

    node_t pick_random_node() { return rand() % num_nodes(); }
    inline node_t nodekey_to_nodeid(node_t key) { return key; }
    inline node_t nodeid_to_nodekey(node_t nodeid) { return nodeid; }

    // ---- device mirror (new) ----
    gmc_graph* device_mirror();          // creates / refreshes the HBM copy of the frozen CSR (+ reverse CSR)
    void invalidate_device_mirror();
    // GM_CUDA_DEVICES > 1: the frozen CSR (+ reverse CSR when built) replicated on every device of the group
    // (shl__device_group()); returns NULL when no group is active.  The group holds one graph at a time.
    struct gmc_group* group_mirror();

  private:
    void delete_frozen_graph();
    void allocate_memory_for_frozen_graph(node_t n, edge_t m);
    void prepare_edge_source_reverse();

    node_t _numNodes;
    edge_t _numEdges;
    bool _reverse_edge;
    bool _frozen;
    bool _directed;
    bool _semi_sorted;
    std::unordered_map<node_t, std::vector<edge_dest_t> > flexible_graph;
    edge_t* e_id2idx;
    edge_t* e_idx2id;
    gmc_graph* _dev;
    bool _dev_has_reverse;
};
#endif
