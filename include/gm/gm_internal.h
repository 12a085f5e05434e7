// Shared leaf header of the runtime (counterpart of the reference's gm_internal.h).
#ifndef GM_INTERNAL_H
#define GM_INTERNAL_H
#include "gm_graph_typedef.h"
#include "gm_atomic_wrapper.h"
#endif
