// Root header for applications (counterpart of the reference's gm.h:14-38), restricted to the modules
// the PageRank / BFS hot path and its direct callers touch (gm_seq for bc's Node_Sequence).  Sets, maps,
// DFS and the auto-main support are out of scope of this library.
#ifndef GM_H_
#define GM_H_
#include <math.h>
#include <stdlib.h>
#include <string.h>
#ifndef DEFAULT_GM_TOP
#define DEFAULT_GM_TOP "."
#endif
#include "gm_internal.h"
#include "gm_graph_typedef.h"
#include "gm_graph.h"
#include "shl_graph.h"
#include "gm_runtime.h"
#include "gm_atomic_wrapper.h"
#include "gm_atomic_operations.h"
#include "gm_lock.h"
#include "gm_bitmap.h"
#include "gm_mem_helper.h"
#include "gm_bfs_template.h"
#include "gm_seq.h"
#endif
