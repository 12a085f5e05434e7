// Non-owning four-pointer CSR view the Shoal-enabled generated code builds inside parallel regions
// (reference shl_graph.h:111-128, shl_graph.cc:3-62).
#ifndef SHL_GRAPH_H_
#define SHL_GRAPH_H_
#include "gm_graph_typedef.h"
class shl_graph {
  public:
    shl_graph(edge_t* b, edge_t* r_b, node_t* n, node_t* r_n) : begin(b), node_idx(n), r_begin(r_b), r_node_idx(r_n) {}
    ~shl_graph() {}
    edge_t* begin;
    node_t* node_idx;
    edge_t* r_begin;
    node_t* r_node_idx;
    static const node_t NIL_NODE = (node_t)-1;
    static const edge_t NIL_EDGE = (edge_t)-1;
    bool is_neighbor(node_t src, node_t to);                      // rows must be semi-sorted
    edge_t get_edge_idx_for_src_dest(node_t src, node_t dest);
};
#endif
