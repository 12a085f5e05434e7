"""ctypes bindings for the CHECKERS: oracle/liboracle.so (plain-C restatement) and, when it has been
built, oracle/_ref/libgmref.so (the reference's own gm_graph runtime).  Test infrastructure only."""
import ctypes as C
import os
import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_SO = os.path.join(ROOT, "oracle", "liboracle.so")
REF_SO = os.path.join(ROOT, "oracle", "_ref", "libgmref.so")

_i32p = np.ctypeslib.ndpointer(np.int32, flags="C_CONTIGUOUS")
_i16p = np.ctypeslib.ndpointer(np.int16, flags="C_CONTIGUOUS")
_f64p = np.ctypeslib.ndpointer(np.float64, flags="C_CONTIGUOUS")


def _opt(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


class Oracle:
    def __init__(self, path=ORACLE_SO):
        if not os.path.exists(path):
            raise FileNotFoundError(f"{path} missing: run `make -C oracle` (or __graft_entry__.build())")
        L = self.L = C.CDLL(path)
        L.orc_set_threads.argtypes = [C.c_int]
        L.orc_uniform_random.argtypes = [C.c_int32, C.c_int32, C.c_long, _i32p, _i32p]
        L.orc_rmat.argtypes = [C.c_int32, C.c_int32, C.c_long, C.c_double, C.c_double, C.c_double, C.c_int, _i32p, _i32p]
        L.orc_semi_sort.argtypes = [C.c_int32, _i32p, _i32p, C.c_void_p]
        L.orc_make_reverse.argtypes = [C.c_int32, C.c_int32, _i32p, _i32p, _i32p, _i32p, C.c_void_p]
        L.orc_store_binary.argtypes = [C.c_char_p, C.c_int32, C.c_int32, _i32p, _i32p]
        L.orc_load_binary_header.argtypes = [C.c_char_p, C.POINTER(C.c_int32), C.POINTER(C.c_int32)]
        L.orc_load_binary.argtypes = [C.c_char_p, _i32p, _i32p]
        L.orc_pagerank.argtypes = [C.c_int32, _i32p, _i32p, _i32p, C.c_double, C.c_double, C.c_int32, _f64p, C.POINTER(C.c_double)]
        L.orc_hop_dist.argtypes = [C.c_int32, _i32p, _i32p, C.c_int32, _i32p]
        L.orc_bfs_levels.argtypes = [C.c_int32, _i32p, _i32p, _i32p, _i32p, C.c_int, C.c_int32, _i16p, C.c_void_p, C.c_void_p, C.c_int32]

    def set_threads(self, n):
        self.L.orc_set_threads(n)

    def uniform_random(self, N, M, seed=1997):
        begin = np.zeros(N + 1, np.int32); idx = np.zeros(max(M, 1), np.int32)
        assert self.L.orc_uniform_random(N, M, seed, begin, idx) == 0
        return begin, idx[:M]

    def rmat(self, N, M, seed=1997, a=0.57, b=0.19, c=0.19, permute=True):
        begin = np.zeros(N + 1, np.int32); idx = np.zeros(max(M, 1), np.int32)
        assert self.L.orc_rmat(N, M, seed, a, b, c, int(permute), begin, idx) == 0
        return begin, idx[:M]

    def semi_sort(self, begin, node_idx, aux=None):
        self.L.orc_semi_sort(len(begin) - 1, begin, node_idx, _opt(aux))

    def make_reverse(self, begin, node_idx, want_rev2idx=False):
        N = len(begin) - 1; M = len(node_idx)
        rb = np.zeros(N + 1, np.int32); ri = np.zeros(max(M, 1), np.int32)
        r2 = np.zeros(max(M, 1), np.int32) if want_rev2idx else None
        self.L.orc_make_reverse(N, M, begin, np.ascontiguousarray(node_idx), rb, ri, _opt(r2))
        return (rb, ri[:M], r2[:M]) if want_rev2idx else (rb, ri[:M])

    def store_binary(self, path, begin, node_idx):
        return self.L.orc_store_binary(path.encode(), len(begin) - 1, len(node_idx), begin, np.ascontiguousarray(node_idx)) == 1

    def load_binary(self, path):
        N = C.c_int32(); M = C.c_int32()
        if not self.L.orc_load_binary_header(path.encode(), C.byref(N), C.byref(M)):
            return None
        begin = np.zeros(N.value + 1, np.int32); idx = np.zeros(max(M.value, 1), np.int32)
        if not self.L.orc_load_binary(path.encode(), begin, idx):
            return None
        return begin, idx[:M.value]

    def triangle_counting(self, begin, node_idx):
        self.L.orc_triangle_counting.restype = C.c_int64
        self.L.orc_triangle_counting.argtypes = [C.c_int32, _i32p, _i32p]
        return int(self.L.orc_triangle_counting(len(begin) - 1, begin, np.ascontiguousarray(node_idx)))

    def comp_bc(self, begin, node_idx, r_begin, r_node_idx, seeds, adj=False):
        """bc.gm; adj=True: bc_adj.gm (the (v != s) filters)"""
        N = len(begin) - 1
        fn = self.L.orc_comp_bc_adj if adj else self.L.orc_comp_bc
        fn.restype = None
        fn.argtypes = [C.c_int32, _i32p, _i32p, _i32p, _i32p, _i32p, C.c_int32, C.c_void_p]
        bc = np.zeros(max(N, 1), np.float32)
        sd = np.ascontiguousarray(seeds, np.int32)
        fn(N, begin, np.ascontiguousarray(node_idx), r_begin, np.ascontiguousarray(r_node_idx), sd, len(sd), bc.ctypes.data)
        return bc[:N]

    def edge_list_read(self, path):
        """EDGE_LIST ingest (ids densified, rows semi-sorted): returns begin, node_idx or None"""
        self.L.orc_edge_list_read.argtypes = [C.c_char_p, C.POINTER(C.c_int32), C.POINTER(C.c_int32), C.c_void_p, C.c_void_p]
        N = C.c_int32(); M = C.c_int32()
        if self.L.orc_edge_list_read(path.encode(), C.byref(N), C.byref(M), None, None) != 0:
            return None
        begin = np.zeros(N.value + 1, np.int32); idx = np.zeros(max(M.value, 1), np.int32)
        if self.L.orc_edge_list_read(path.encode(), C.byref(N), C.byref(M), begin.ctypes.data, idx.ctypes.data) != 0:
            return None
        return begin, idx[:M.value]

    def pagerank(self, begin, r_begin, r_node_idx, e=1e-30, d=0.85, max_iter=20):
        N = len(begin) - 1
        rank = np.zeros(max(N, 1), np.float64); diff = C.c_double()
        it = self.L.orc_pagerank(N, begin, r_begin, np.ascontiguousarray(r_node_idx), e, d, max_iter, rank, C.byref(diff))
        return rank[:N], it, diff.value

    def hop_dist(self, begin, node_idx, root=0):
        N = len(begin) - 1
        dist = np.zeros(max(N, 1), np.int32)
        sweeps = self.L.orc_hop_dist(N, begin, np.ascontiguousarray(node_idx), root, dist)
        return dist[:N], sweeps

    def bfs_levels(self, begin, node_idx, r_begin, r_node_idx, root=0, has_reverse=True, cap=256, save_child=False):
        N = len(begin) - 1
        lev = np.zeros(max(N, 1), np.int16)
        trace = np.full(cap, -1, np.int32); cnt = np.zeros(cap, np.int32)
        fn = self.L.orc_bfs_levels_save_child if save_child else self.L.orc_bfs_levels
        fn.argtypes = self.L.orc_bfs_levels.argtypes
        fn.restype = self.L.orc_bfs_levels.restype
        n = fn(N, begin, np.ascontiguousarray(node_idx), r_begin, np.ascontiguousarray(r_node_idx),
                                  int(has_reverse), root, lev, _opt(trace), _opt(cnt), cap)
        return lev[:N], n, trace[:min(n, cap)], cnt[:min(n, cap)]


class Ref:
    """The reference's own code (oracle/_ref/libgmref.so); built only where /root/reference exists."""

    def __init__(self, path=REF_SO):
        if not os.path.exists(path):
            raise FileNotFoundError(path)
        L = self.L = C.CDLL(path)
        vp = C.c_void_p
        L.ref_set_threads.argtypes = [C.c_int]
        L.ref_get_threads.restype = C.c_int
        L.ref_graph_uniform.restype = vp; L.ref_graph_uniform.argtypes = [C.c_int32, C.c_int32, C.c_long, C.c_int]
        L.ref_graph_rmat.restype = vp; L.ref_graph_rmat.argtypes = [C.c_int32, C.c_int32, C.c_long, C.c_double, C.c_double, C.c_double, C.c_int]
        L.ref_graph_from_csr.restype = vp; L.ref_graph_from_csr.argtypes = [C.c_int32, C.c_int32, _i32p, _i32p]
        L.ref_graph_load.restype = vp; L.ref_graph_load.argtypes = [C.c_char_p]
        L.ref_graph_store.argtypes = [vp, C.c_char_p]
        if hasattr(L, "ref_graph_load_edge_list"):
            L.ref_graph_load_edge_list.restype = vp; L.ref_graph_load_edge_list.argtypes = [C.c_char_p]
        L.ref_graph_free.argtypes = [vp]
        for f in ("num_nodes", "num_edges"):
            getattr(L, "ref_graph_" + f).argtypes = [vp]; getattr(L, "ref_graph_" + f).restype = C.c_int32
        for f in ("begin", "node_idx", "r_begin", "r_node_idx", "e_rev2idx", "e_idx2idx"):
            getattr(L, "ref_graph_" + f).argtypes = [vp]; getattr(L, "ref_graph_" + f).restype = C.POINTER(C.c_int32)
        L.ref_pagerank.restype = C.c_double; L.ref_pagerank.argtypes = [vp, C.c_double, C.c_double, C.c_int32, _f64p]
        L.ref_hop_dist.restype = C.c_double; L.ref_hop_dist.argtypes = [vp, C.c_int32, _i32p]
        L.ref_bfs_levels.restype = C.c_double; L.ref_bfs_levels.argtypes = [vp, C.c_int32, _i16p, C.POINTER(C.c_int32)]

    def set_threads(self, n):
        self.L.ref_set_threads(n)

    def uniform(self, N, M, seed=1997):
        return self.L.ref_graph_uniform(N, M, seed, 0)

    def rmat(self, N, M, seed=1997, a=0.57, b=0.19, c=0.19, permute=True):
        return self.L.ref_graph_rmat(N, M, seed, a, b, c, int(permute))

    def from_csr(self, begin, node_idx):
        return self.L.ref_graph_from_csr(len(begin) - 1, len(node_idx), begin, np.ascontiguousarray(node_idx))

    def load(self, path):
        return self.L.ref_graph_load(path.encode())

    def triangle_counting(self, g):
        self.L.ref_triangle_counting.restype = C.c_longlong; self.L.ref_triangle_counting.argtypes = [C.c_void_p]
        return int(self.L.ref_triangle_counting(g))

    def comp_bc(self, g, seeds, adj=False):
        N = self.L.ref_graph_num_nodes(g)
        fn = self.L.ref_comp_bc_adj if adj else self.L.ref_comp_bc
        fn.restype = None
        fn.argtypes = [C.c_void_p, _i32p, C.c_int32, C.c_void_p]
        bc = np.zeros(max(N, 1), np.float32)
        sd = np.ascontiguousarray(seeds, np.int32)
        fn(g, sd, len(sd), bc.ctypes.data)
        return bc[:N]

    def load_edge_list(self, path):
        return self.L.ref_graph_load_edge_list(path.encode())

    def store(self, g, path):
        return self.L.ref_graph_store(g, path.encode()) == 1

    def free(self, g):
        self.L.ref_graph_free(g)

    def arrays(self, g):
        N = self.L.ref_graph_num_nodes(g); M = self.L.ref_graph_num_edges(g)
        def arr(name, n):
            p = getattr(self.L, "ref_graph_" + name)(g)
            return np.ctypeslib.as_array(p, shape=(n,)).copy() if n else np.zeros(0, np.int32)
        return dict(N=N, M=M, begin=arr("begin", N + 1), node_idx=arr("node_idx", M),
                    r_begin=arr("r_begin", N + 1), r_node_idx=arr("r_node_idx", M), e_rev2idx=arr("e_rev2idx", M))

    def pagerank(self, g, e=1e-30, d=0.85, max_iter=20):
        N = self.L.ref_graph_num_nodes(g)
        rank = np.zeros(max(N, 1), np.float64)
        ms = self.L.ref_pagerank(g, e, d, max_iter, rank)
        return rank[:N], ms

    def hop_dist(self, g, root=0):
        N = self.L.ref_graph_num_nodes(g)
        dist = np.zeros(max(N, 1), np.int32)
        ms = self.L.ref_hop_dist(g, root, dist)
        return dist[:N], ms

    def bfs_levels(self, g, root=0):
        N = self.L.ref_graph_num_nodes(g)
        lev = np.zeros(max(N, 1), np.int16); nl = C.c_int32()
        ms = self.L.ref_bfs_levels(g, root, lev, C.byref(nl))
        return lev[:N], nl.value, ms


def have_ref():
    return os.path.exists(REF_SO)
