// Node / edge id types of the Green-Marl runtime API (drop-in for the reference header
// apps/output_cpp/gm_graph/inc/gm_graph_typedef.h:8-49).  libgm_cuda is built for the reference's
// default configuration only: node_t = edge_t = int32_t (setup.mk:55-56).  An application compiled
// with -DGM_NODE64 / -DGM_EDGE64 fails at link time through the size-check symbol, exactly as it
// would against a mismatching libgmgraph.a.
#ifndef GM_GRAPH_TYPEDEF_H
#define GM_GRAPH_TYPEDEF_H
#include <stdint.h>
#include <vector>

#if defined(GM_NODE64)
typedef int64_t edge_t;
typedef int64_t node_t;
#define GM_SIZE_CHECK_VAR link_error_becuase_gm_graph_lib_is_configured_as_node64_edge64_but_the_application_is_not
#elif defined(GM_EDGE64)
typedef int64_t edge_t;
typedef int32_t node_t;
#define GM_SIZE_CHECK_VAR link_error_becuase_gm_graph_lib_is_configured_as_node32_edge64_but_the_application_is_not
#else
typedef int32_t edge_t;
typedef int32_t node_t;
#define GM_SIZE_CHECK_VAR link_error_becuase_gm_graph_lib_is_configured_as_node32_edge32_but_the_application_is_not
#endif

enum VALUE_TYPE { GMTYPE_BOOL = 0, GMTYPE_INT, GMTYPE_LONG, GMTYPE_FLOAT, GMTYPE_DOUBLE, GMTYPE_NODE, GMTYPE_EDGE, GMTYPE_END };

typedef std::vector<double> GM_DVECT;
typedef std::vector<float> GM_FVECT;
typedef std::vector<bool> GM_BVECT;
typedef std::vector<int64_t> GM_LVECT;
typedef std::vector<int32_t> GM_IVECT;
typedef std::vector<node_t> GM_NVECT;
typedef std::vector<edge_t> GM_EVECT;

extern int GM_SIZE_CHECK_VAR;   // defined once inside libgm_cuda under the 32/32 name
static inline void gm_graph_check_node_edge_size_at_link_time() { GM_SIZE_CHECK_VAR = 0; }
#endif
