// gm_graph implementation over the device mirror.  Behaviour follows the reference class
// (apps/output_cpp/gm_graph/src/gm_graph.cc, gm_graph_binary_loader.cc) -- see include/gm/gm_graph.h.
#include <algorithm>
#include <string.h>
#include "gm_graph.h"
#include "gm_cuda.h"
#include "shl.h"

int GM_SIZE_CHECK_VAR;   // the 32/32 link-time size-check symbol (reference gm_graph.cc:652)

static void dev_check(int rc, const char* what) {
    if (rc != GMC_OK) {
        fprintf(stderr, "gm_graph::%s: device error %d: %s\n", what, rc, gmc_last_error());
        exit(EXIT_FAILURE);   // no CPU fallback by design
    }
}

gm_graph::gm_graph()
    : begin(NULL), node_idx(NULL), node_idx_src(NULL), r_begin(NULL), r_node_idx(NULL), r_node_idx_src(NULL),
      e_idx2idx(NULL), e_rev2idx(NULL), _numNodes(0), _numEdges(0), _reverse_edge(false), _frozen(false),
      _directed(true), _semi_sorted(false), e_id2idx(NULL), e_idx2id(NULL), _dev(NULL), _dev_has_reverse(false) {}

gm_graph::~gm_graph() {
    invalidate_device_mirror();
    delete_frozen_graph();
}

// the graph whose CSR the device group currently holds (and whether its reverse CSR is there)
static const gm_graph* g_group_owner = NULL;
static bool g_group_has_reverse = false;

void gm_graph::invalidate_device_mirror() {
    if (_dev) gmc_graph_destroy(_dev);
    _dev = NULL;
    _dev_has_reverse = false;
    if (g_group_owner == this) {
        if (shl__device_group()) gmc_group_graph_free(shl__device_group());
        g_group_owner = NULL;
    }
}

struct gmc_group* gm_graph::group_mirror() {
    gmc_group* grp = shl__device_group();
    if (!grp) return NULL;
    if (!_frozen) freeze();
    if (g_group_owner != this || (_reverse_edge && !g_group_has_reverse)) {
        dev_check(gmc_group_graph_upload(grp, _numNodes, _numEdges, begin, node_idx, _reverse_edge ? r_begin : NULL,
                                         _reverse_edge ? r_node_idx : NULL), "group_mirror(upload)");
        g_group_owner = this;
        g_group_has_reverse = _reverse_edge;
    }
    return grp;
}

gmc_graph* gm_graph::device_mirror() {
    if (!_frozen) freeze();
    if (!_dev) {
        dev_check(gmc_graph_create(&_dev, _numNodes, _numEdges), "device_mirror(create)");
        dev_check(gmc_graph_upload(_dev, begin, node_idx, NULL), "device_mirror(upload)");
        _dev_has_reverse = false;
    }
    if (_reverse_edge && !_dev_has_reverse) {
        dev_check(gmc_graph_upload_reverse(_dev, r_begin, r_node_idx, NULL), "device_mirror(upload reverse)");
        _dev_has_reverse = true;
    }
    return _dev;
}

void gm_graph::allocate_memory_for_frozen_graph(node_t n, edge_t m) {
    begin = new edge_t[(size_t)n + 1];
    node_idx = new node_t[m > 0 ? m : 1];
    _numNodes = n;
    _numEdges = m;
}

void gm_graph::delete_frozen_graph() {
    delete[] begin; begin = NULL;
    delete[] node_idx; node_idx = NULL;
    delete[] node_idx_src; node_idx_src = NULL;
    delete[] r_begin; r_begin = NULL;
    delete[] r_node_idx; r_node_idx = NULL;
    delete[] r_node_idx_src; r_node_idx_src = NULL;
    delete[] e_rev2idx; e_rev2idx = NULL;
    delete[] e_idx2idx; e_idx2idx = NULL;
    delete[] e_id2idx; e_id2idx = NULL;
    delete[] e_idx2id; e_idx2id = NULL;
}

void gm_graph::clear_graph() { clear_graph(true); }

void gm_graph::clear_graph(bool) {
    invalidate_device_mirror();
    flexible_graph.clear();
    delete_frozen_graph();
    _numNodes = 0;
    _numEdges = 0;
    _reverse_edge = _frozen = _semi_sorted = false;
}

void gm_graph::prepare_external_creation(node_t n, edge_t m) { prepare_external_creation(n, m, true); }

void gm_graph::prepare_external_creation(node_t n, edge_t m, bool clean) {
    clear_graph(clean);
    allocate_memory_for_frozen_graph(n, m);
    _frozen = true;
}

node_t gm_graph::add_node() {
    if (_frozen) thaw();
    flexible_graph[_numNodes] = std::vector<edge_dest_t>();
    return _numNodes++;
}

edge_t gm_graph::add_edge(node_t n, node_t m) {
    assert(is_node(n) && is_node(m));
    if (_frozen) thaw();
    edge_dest_t t;
    t.dest = m;
    t.edge = _numEdges;
    flexible_graph[n].push_back(t);
    return _numEdges++;
}

// flexible -> CSR; edge ids are kept through e_idx2id / e_id2idx; always semi-sorts (reference :68-113)
void gm_graph::freeze() {
    if (_frozen) return;
    const node_t n = _numNodes;
    const edge_t m = _numEdges;
    allocate_memory_for_frozen_graph(n, m);
    e_idx2id = new edge_t[m > 0 ? m : 1];
    e_id2idx = new edge_t[m > 0 ? m : 1];
    edge_t next = 0;
    for (node_t i = 0; i < n; i++) {
        begin[i] = next;
        std::vector<edge_dest_t>& nb = flexible_graph[i];
        for (size_t k = 0; k < nb.size(); k++) {
            node_idx[next] = nb[k].dest;
            e_idx2id[next] = nb[k].edge;
            e_id2idx[nb[k].edge] = next;
            next++;
        }
    }
    begin[n] = next;
    flexible_graph.clear();
    _frozen = true;
    _semi_sorted = false;
    _reverse_edge = false;
    do_semi_sort();
}

void gm_graph::thaw() {
    if (!_frozen) return;
    invalidate_device_mirror();
    flexible_graph.clear();
    for (node_t i = 0; i < _numNodes; i++) {
        std::vector<edge_dest_t>& nb = flexible_graph[i];
        for (edge_t e = begin[i]; e < begin[i + 1]; e++) {
            edge_dest_t t;
            t.dest = node_idx[e];
            t.edge = get_edge_id(e);
            nb.push_back(t);
        }
    }
    delete_frozen_graph();
    _frozen = false;
    _semi_sorted = false;
    _reverse_edge = false;
}

// K5 on device, results copied into the public host arrays
void gm_graph::do_semi_sort() {
    if (!_frozen) freeze();
    if (_semi_sorted) return;
    const size_t m = (size_t)_numEdges;
    gmc_graph* d = device_mirror();
    dev_check(gmc_graph_semi_sort(d, 1, NULL), "do_semi_sort");
    if (m) dev_check(gmc_graph_download(d, GMC_ARR_NODE_IDX, node_idx, m, NULL), "do_semi_sort(download)");
    if (!e_idx2idx) e_idx2idx = new edge_t[m ? m : 1];
    if (m) dev_check(gmc_graph_download(d, GMC_ARR_E_IDX2IDX, e_idx2idx, m, NULL), "do_semi_sort(download map)");
    if (e_idx2id) {   // edge ids travel with their edges (reference semi_sort_main aux2, :476-493)
        edge_t* moved = new edge_t[m ? m : 1];
        for (size_t j = 0; j < m; j++) moved[j] = e_idx2id[e_idx2idx[j]];
        delete[] e_idx2id;
        e_idx2id = moved;
        for (size_t j = 0; j < m; j++) e_id2idx[e_idx2id[j]] = (edge_t)j;
    }
    _semi_sorted = true;
    if (_reverse_edge) {   // the reverse CSR refers to the old edge order: rebuild it
        _reverse_edge = false;
        _dev_has_reverse = false;
        make_reverse_edges();
    }
}

// K4 on device (+ the row sort the reference applies when the graph is semi-sorted, :293)
void gm_graph::make_reverse_edges() {
    if (_reverse_edge) return;
    if (!_frozen) freeze();
    const size_t n = (size_t)_numNodes, m = (size_t)_numEdges;
    gmc_graph* d = device_mirror();
    dev_check(gmc_graph_make_reverse(d, 1, NULL), "make_reverse_edges");
    _dev_has_reverse = true;
    if (!r_begin) r_begin = new edge_t[n + 1];
    if (!r_node_idx) r_node_idx = new node_t[m ? m : 1];
    if (!e_rev2idx) e_rev2idx = new edge_t[m ? m : 1];
    dev_check(gmc_graph_download(d, GMC_ARR_R_BEGIN, r_begin, n + 1, NULL), "make_reverse_edges(download)");
    if (m) {
        dev_check(gmc_graph_download(d, GMC_ARR_R_NODE_IDX, r_node_idx, m, NULL), "make_reverse_edges(download)");
        dev_check(gmc_graph_download(d, GMC_ARR_E_REV2IDX, e_rev2idx, m, NULL), "make_reverse_edges(download)");
    }
    _reverse_edge = true;
    if (is_edge_source_ready()) prepare_edge_source_reverse();
}

void gm_graph::prepare_edge_source() {
    assert(node_idx_src == NULL);
    node_idx_src = new node_t[_numEdges > 0 ? _numEdges : 1];
    #pragma omp parallel for schedule(static, 128)
    for (node_t i = 0; i < _numNodes; i++)
        for (edge_t j = begin[i]; j < begin[i + 1]; j++) node_idx_src[j] = i;
    if (has_reverse_edge()) prepare_edge_source_reverse();
}

void gm_graph::prepare_edge_source_reverse() {
    delete[] r_node_idx_src;
    r_node_idx_src = new node_t[_numEdges > 0 ? _numEdges : 1];
    #pragma omp parallel for schedule(static, 128)
    for (node_t i = 0; i < _numNodes; i++)
        for (edge_t j = r_begin[i]; j < r_begin[i + 1]; j++) r_node_idx_src[j] = i;
}

// first edge src->dest, NIL_EDGE if none: binary search on a semi-sorted row, linear scan otherwise
edge_t gm_graph::get_edge_idx_for_src_dest(node_t src, node_t dest) {
    edge_t lo = begin[src], hi = begin[src + 1];
    if (_semi_sorted) {
        node_t* p = std::lower_bound(node_idx + lo, node_idx + hi, dest);
        return (p != node_idx + hi && *p == dest) ? (edge_t)(p - node_idx) : NIL_EDGE;
    }
    for (edge_t e = lo; e < hi; e++)
        if (node_idx[e] == dest) return e;
    return NIL_EDGE;
}
bool gm_graph::is_neighbor(node_t src, node_t dest) { return get_edge_idx_for_src_dest(src, dest) != NIL_EDGE; }
bool gm_graph::has_edge_to(node_t src, node_t to) { return is_neighbor(src, to); }

bool gm_graph::has_edge(node_id from, node_id to) {
    if (_frozen) return is_neighbor(from, to);
    std::vector<edge_dest_t>& nb = flexible_graph[from];
    for (size_t k = 0; k < nb.size(); k++)
        if (nb[k].dest == to) return true;
    return false;
}

edge_t gm_graph::get_num_edges(node_id from, node_id to) {
    assert(_frozen);
    edge_t c = 0;
    for (edge_t e = begin[from]; e < begin[from + 1]; e++) c += (node_idx[e] == to);
    return c;
}

// load: bulk read, then what the reference does after reading (loader :188-197): semi-sort, reverse edges
bool gm_graph::load_binary(char* filename) {
    clear_graph();
    int32_t n = 0, m = 0;
    if (gmc_file_read_header(filename, &n, &m) != GMC_OK) {
        fprintf(stderr, "%s\n", gmc_last_error());
        return false;
    }
    printf("N = %ld, M = %ld\n", (long)n, (long)m);
    allocate_memory_for_frozen_graph(n, m);
    if (gmc_file_read(filename, begin, node_idx) != GMC_OK) {
        fprintf(stderr, "%s\n", gmc_last_error());
        clear_graph();
        return false;
    }
    _frozen = true;
    do_semi_sort();
    make_reverse_edges();
    return true;
}

// EDGE_LIST ingest: parse + CSR build + semi-sort + reverse CSR in one device-side call, then the host view
// the generated code reads (public begin/node_idx/r_begin/r_node_idx pointers) is downloaded from the mirror.
bool gm_graph::load_edge_list(char* filename, std::vector<VALUE_TYPE>& vprop_schema, std::vector<VALUE_TYPE>& eprop_schema,
                              std::vector<void*>& vertex_props, std::vector<void*>& edge_props, bool use_hdfs) {
    assert(!use_hdfs);
    (void)vertex_props; (void)edge_props;
    if (!vprop_schema.empty() || !eprop_schema.empty()) {
        fprintf(stderr, "load_edge_list: property columns are not supported by libgm_cuda\n");
        return false;
    }
    clear_graph();
    gmc_graph* d = NULL;
    if (gmc_edge_list_read(filename, &d, 0) != GMC_OK) {
        fprintf(stderr, "%s\n", gmc_last_error());
        return false;
    }
    int32_t n = 0, m = 0;
    gmc_graph_num_nodes(d, &n); gmc_graph_num_edges(d, &m);
    allocate_memory_for_frozen_graph(n, m);
    dev_check(gmc_graph_download(d, GMC_ARR_BEGIN, begin, (size_t)n + 1, NULL), "load_edge_list(download)");
    if (m) dev_check(gmc_graph_download(d, GMC_ARR_NODE_IDX, node_idx, (size_t)m, NULL), "load_edge_list(download)");
    _dev = d;                     // the parsed graph IS the device mirror
    _dev_has_reverse = false;
    _frozen = _semi_sorted = true;
    make_reverse_edges();         // as the reference reader ends (:168-169); fills r_begin / r_node_idx / e_rev2idx
    return true;
}

bool gm_graph::store_binary(char* filename) {
    if (!_frozen) freeze();
    if (gmc_file_write(filename, _numNodes, _numEdges, begin, node_idx) != GMC_OK) {
        fprintf(stderr, "%s\n", gmc_last_error());
        return false;
    }
    return true;
}
