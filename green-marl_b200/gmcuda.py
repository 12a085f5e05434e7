"""ctypes binding of include/gm_cuda.h.  No compute happens in Python and there is no fallback:
if libgm_cuda.so is missing, or no CUDA device is usable, calls raise."""
import ctypes as C
import os
import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))


def lib_path():
    return os.path.join(_HERE, "lib", "libgm_cuda.so")


class GmcError(RuntimeError):
    pass


class PagerankInfo(C.Structure):
    _fields_ = [("iterations", C.c_int32), ("diff", C.c_double), ("kernel_ms", C.c_float), ("launches", C.c_int32)]


class BfsInfo(C.Structure):
    _fields_ = [("levels", C.c_int32), ("reached", C.c_int32), ("edges_traversed", C.c_int64),
                ("kernel_ms", C.c_float), ("launches", C.c_int32), ("bottom_up_levels", C.c_int32)]


class BcInfo(C.Structure):
    _fields_ = [("seeds", C.c_int32), ("launches", C.c_int32), ("max_levels", C.c_int32), ("device_ms", C.c_float)]


class PlanInfo(C.Structure):
    _fields_ = [(k, C.c_int32) for k in ("n_long", "hot_n", "n_slices", "long_entries", "total_entries", "n_long_tasks",
                                         "n_short_tasks", "grid", "threads", "smem_bytes", "n_pos", "n_src", "tiers", "segments",
                                         "split_rows", "part_slots", "tier_ctas", "home_entries")] + \
               [(k, C.c_int64) for k in ("hot0_entries", "cold_entries", "segment_entries", "cold_stream_entries")]


ARR = dict(begin=0, node_idx=1, r_begin=2, r_node_idx=3, e_rev2idx=4, e_idx2idx=5)
_vp = C.c_void_p
_i32 = C.c_int32

# name -> (restype, argtypes); every symbol include/gm_cuda.h declares
SIGNATURES = {
    "gmc_last_error": (C.c_char_p, []),
    "gmc_version": (C.c_int, []),
    "gmc_device_count": (C.c_int, [C.POINTER(C.c_int)]),
    "gmc_init": (C.c_int, [C.c_int]),
    "gmc_device_info": (C.c_int, [C.POINTER(C.c_int), C.POINTER(C.c_size_t), C.POINTER(C.c_size_t)]),
    "gmc_sync": (C.c_int, [_vp]),
    "gmc_malloc": (C.c_int, [C.POINTER(_vp), C.c_size_t]),
    "gmc_free": (C.c_int, [_vp]),
    "gmc_h2d": (C.c_int, [_vp, _vp, C.c_size_t, _vp]),
    "gmc_d2h": (C.c_int, [_vp, _vp, C.c_size_t, _vp]),
    "gmc_host_register": (C.c_int, [_vp, C.c_size_t]),
    "gmc_host_unregister": (C.c_int, [_vp]),
    "gmc_graph_create": (C.c_int, [C.POINTER(_vp), _i32, _i32]),
    "gmc_graph_destroy": (C.c_int, [_vp]),
    "gmc_graph_num_nodes": (C.c_int, [_vp, C.POINTER(_i32)]),
    "gmc_graph_num_edges": (C.c_int, [_vp, C.POINTER(_i32)]),
    "gmc_graph_upload": (C.c_int, [_vp, _vp, _vp, _vp]),
    "gmc_graph_upload_reverse": (C.c_int, [_vp, _vp, _vp, _vp]),
    "gmc_graph_semi_sort": (C.c_int, [_vp, C.c_int, _vp]),
    "gmc_graph_make_reverse": (C.c_int, [_vp, C.c_int, _vp]),
    "gmc_graph_has_reverse": (C.c_int, [_vp, C.POINTER(C.c_int)]),
    "gmc_graph_download": (C.c_int, [_vp, C.c_int, _vp, C.c_size_t, _vp]),
    "gmc_graph_device_ptr": (C.c_int, [_vp, C.c_int, C.POINTER(_vp)]),
    "gmc_graph_generate_uniform": (C.c_int, [C.POINTER(_vp), _i32, _i32, C.c_uint64]),
    "gmc_graph_generate_rmat": (C.c_int, [C.POINTER(_vp), _i32, _i32, C.c_uint64, C.c_double, C.c_double, C.c_double, C.c_int]),
    "gmc_file_read_header": (C.c_int, [C.c_char_p, C.POINTER(_i32), C.POINTER(_i32)]),
    "gmc_file_read": (C.c_int, [C.c_char_p, _vp, _vp]),
    "gmc_file_write": (C.c_int, [C.c_char_p, _i32, _i32, _vp, _vp]),
    "gmc_edge_list_read": (C.c_int, [C.c_char_p, C.POINTER(_vp), C.c_int]),
    "gmc_pagerank": (C.c_int, [_vp, C.c_double, C.c_double, _i32, _vp, C.POINTER(PagerankInfo)]),
    "gmc_pagerank_dev": (C.c_int, [_vp, C.c_double, C.c_double, _i32, _vp, C.POINTER(PagerankInfo), _vp]),
    "gmc_pagerank_host_csr": (C.c_int, [_i32, _i32, _vp, _vp, _vp, C.c_double, C.c_double, _i32, _vp, C.POINTER(PagerankInfo)]),
    "gmc_pagerank_time_sweeps": (C.c_int, [_vp, C.c_double, _i32, _i32, _vp, C.POINTER(_i32), _vp]),
    "gmc_pagerank_plan_info": (C.c_int, [_vp, C.POINTER(PlanInfo)]),
    "gmc_triangle_counting": (C.c_int, [_vp, C.POINTER(C.c_int64), C.POINTER(C.c_float)]),
    "gmc_triangle_counting_range": (C.c_int, [_vp, C.c_int64, C.c_int64, C.POINTER(C.c_int64), C.POINTER(C.c_float), _vp]),
    "gmc_triangle_counting_noalloc": (C.c_int, [_vp, _vp, _vp]),
    "gmc_triangle_counting_sharded": (C.c_int, [_vp, _vp, C.POINTER(C.c_int64), C.POINTER(C.c_float), _vp]),
    "gmc_comp_bc": (C.c_int, [_vp, _vp, _i32, _vp, C.POINTER(BcInfo)]),
    "gmc_comp_bc_adj": (C.c_int, [_vp, _vp, _i32, _vp, C.POINTER(BcInfo)]),
    "gmc_hop_dist": (C.c_int, [_vp, _i32, _vp, C.POINTER(BfsInfo)]),
    "gmc_hop_dist_dev": (C.c_int, [_vp, _i32, _vp, C.POINTER(BfsInfo), _vp]),
    "gmc_bfs_levels": (C.c_int, [_vp, _i32, C.c_int, _vp, _vp, _vp, _i32, C.POINTER(BfsInfo)]),
    "gmc_bfs_levels_masked": (C.c_int, [_vp, _i32, C.c_int, _vp, _vp, _vp, _vp, _i32, C.POINTER(BfsInfo)]),
    "gmc_bfs_level_order": (C.c_int, [_vp, _vp, _vp, _i32]),
    "gmc_comm_unique_id": (C.c_int, [_vp]),
    "gmc_comm_create": (C.c_int, [C.POINTER(_vp), C.c_int, C.c_int, _vp]),
    "gmc_comm_destroy": (C.c_int, [_vp]),
    "gmc_comm_info": (C.c_int, [_vp, C.POINTER(C.c_int), C.POINTER(C.c_int), C.POINTER(C.c_int)]),
    "gmc_comm_barrier": (C.c_int, [_vp, _vp]),
    "gmc_shard_cuts": (C.c_int, [_vp, _i32, C.c_int, _i32, _vp]),
    "gmc_pagerank_sharded_dev": (C.c_int, [_vp, _vp, C.c_double, C.c_double, _i32, _vp, C.POINTER(PagerankInfo), _vp]),
    "gmc_pagerank_emulated_shards_dev": (C.c_int, [_vp, C.c_int, C.c_double, C.c_double, _i32, _vp, C.POINTER(PagerankInfo), _vp]),
    "gmc_graph_upload_sharded": (C.c_int, [_vp, _vp, _vp, _vp, _vp, _vp, _vp]),
    "gmc_pagerank_exchange_path": (C.c_int, [_vp, C.POINTER(C.c_int)]),
    "gmc_pagerank_sharded_time_sweeps": (C.c_int, [_vp, _vp, C.c_double, _i32, _i32, _vp, _vp, _vp]),
    "gmc_hop_dist_sharded_dev": (C.c_int, [_vp, _vp, _i32, _vp, C.POINTER(BfsInfo), _vp]),
    "gmc_group_create": (C.c_int, [C.POINTER(_vp), C.c_int, _vp]),
    "gmc_group_destroy": (C.c_int, [_vp]),
    "gmc_group_size": (C.c_int, [_vp, C.POINTER(C.c_int)]),
    "gmc_group_graph_upload": (C.c_int, [_vp, _i32, _i32, _vp, _vp, _vp, _vp]),
    "gmc_group_graph_free": (C.c_int, [_vp]),
    "gmc_group_pagerank": (C.c_int, [_vp, C.c_double, C.c_double, _i32, _vp, C.POINTER(PagerankInfo)]),
    "gmc_group_hop_dist": (C.c_int, [_vp, _i32, _vp, C.POINTER(BfsInfo)]),
    "gmc_group_exchange_path": (C.c_int, [_vp, C.POINTER(C.c_int)]),
    "gmc_bfs_levels_sharded": (C.c_int, [_vp, _vp, _i32, C.c_int, _vp, _vp, _vp, _vp, _i32, C.POINTER(BfsInfo)]),
}


def _ptr(a):
    if a is None:
        return None
    if isinstance(a, np.ndarray):
        assert a.flags["C_CONTIGUOUS"]
        return a.ctypes.data_as(_vp)
    return _vp(int(a))


def _i32arr(a):
    return np.ascontiguousarray(a, dtype=np.int32)


class GmCuda:
    """Loaded libgm_cuda.so.  Constructing it does not touch the GPU; `init()` does."""

    def __init__(self, path=None):
        path = path or lib_path()
        if not os.path.exists(path):
            raise GmcError(f"{path} not found: build it with `make -C green-marl_b200` "
                           f"(or __graft_entry__.build()); there is no CPU fallback")
        self.path = path
        self.L = C.CDLL(path)
        for name, (res, args) in SIGNATURES.items():
            f = getattr(self.L, name)     # AttributeError if a declared symbol is missing
            f.restype = res
            f.argtypes = args

    def check(self, rc):
        if rc != 0:
            raise GmcError(f"gmc status {rc}: {self.L.gmc_last_error().decode(errors='replace')}")

    def init(self, device=0):
        self.check(self.L.gmc_init(device))

    def device_count(self):
        n = C.c_int()
        self.check(self.L.gmc_device_count(C.byref(n)))
        return n.value

    def device_info(self):
        sm = C.c_int(); tot = C.c_size_t(); free = C.c_size_t()
        self.check(self.L.gmc_device_info(C.byref(sm), C.byref(tot), C.byref(free)))
        return sm.value, tot.value, free.value

    # ---- graphs --------------------------------------------------------------------------
    def graph_from_csr(self, begin, node_idx, semi_sort=False, make_reverse=True, want_maps=False):
        begin = _i32arr(begin); node_idx = _i32arr(node_idx)
        h = _vp()
        self.check(self.L.gmc_graph_create(C.byref(h), len(begin) - 1, len(node_idx)))
        g = Graph(self, h)
        self.check(self.L.gmc_graph_upload(h, _ptr(begin), _ptr(node_idx), None))
        if semi_sort:
            self.check(self.L.gmc_graph_semi_sort(h, int(want_maps), None))
        if make_reverse:
            self.check(self.L.gmc_graph_make_reverse(h, int(want_maps), None))
        return g

    def graph_pull_only(self, begin, r_begin, r_node_idx, comm=None):
        """what the generated pagerank() copies in: begin (out-degrees) + the reverse CSR; no node_idx (no BFS on it)"""
        h = _vp()
        self.check(self.L.gmc_graph_create(C.byref(h), len(begin) - 1, len(r_node_idx)))
        g = Graph(self, h)
        self.check(self.L.gmc_graph_upload_sharded(h, comm.h if comm else None, _ptr(begin), None, _ptr(r_begin), _ptr(r_node_idx), None))
        return g

    def pagerank_host_csr(self, begin, r_begin, r_node_idx, rank_host, e=1e-30, d=0.85, max_iter=20):
        """the generated pagerank() on host arrays in one call: copy-in (overlapped with the plan build), sweeps, copy-back"""
        n, m = len(begin) - 1, len(r_node_idx)
        assert rank_host.dtype == np.float64 and len(rank_host) >= n
        for a in (begin, r_begin, r_node_idx):
            assert a.dtype == np.int32 and a.flags["C_CONTIGUOUS"]
        info = PagerankInfo()
        self.check(self.L.gmc_pagerank_host_csr(n, m, _ptr(begin), _ptr(r_begin), _ptr(r_node_idx), e, d, max_iter, _ptr(rank_host),
                                                C.byref(info)))
        return rank_host, info

    def graph_from_csr_sharded(self, comm, begin, node_idx, r_begin, r_node_idx):
        """multi-GPU copy-in: this rank uploads its slice of each array, the slices are allgathered over NVLink"""
        h = _vp()
        self.check(self.L.gmc_graph_create(C.byref(h), len(begin) - 1, len(node_idx)))
        g = Graph(self, h)
        self.check(self.L.gmc_graph_upload_sharded(h, comm.h if comm else None, _ptr(begin), _ptr(node_idx), _ptr(r_begin),
                                                   _ptr(r_node_idx), None))
        return g

    def edge_list_read(self, path, make_reverse=True):
        """EDGE_LIST text file -> resident graph (ids densified, rows semi-sorted, reverse CSR on request)"""
        h = _vp()
        self.check(self.L.gmc_edge_list_read(path.encode(), C.byref(h), int(make_reverse)))
        return Graph(self, h)

    def generate_rmat(self, n, m, seed=1997, a=0.57, b=0.19, c=0.19, permute=True):
        h = _vp()
        self.check(self.L.gmc_graph_generate_rmat(C.byref(h), n, m, seed, a, b, c, int(permute)))
        return Graph(self, h)

    def generate_uniform(self, n, m, seed=1997):
        h = _vp()
        self.check(self.L.gmc_graph_generate_uniform(C.byref(h), n, m, seed))
        return Graph(self, h)

    # ---- multi-GPU -----------------------------------------------------------------------
    def group(self, ndev, devices=None):
        """device group: `ndev` GPUs driven by this one process (gmc_group_*)"""
        h = _vp()
        dv = None
        if devices is not None:
            dv = (C.c_int * ndev)(*devices)
        self.check(self.L.gmc_group_create(C.byref(h), ndev, dv))
        return Group(self, h)

    def comm_unique_id(self):
        """128 opaque bytes (ncclUniqueId); made on rank 0 and shipped to the other ranks by the host side"""
        buf = np.zeros(128, np.uint8)
        self.check(self.L.gmc_comm_unique_id(_ptr(buf)))
        return buf

    def comm_create(self, world, rank, unique_id=None):
        h = _vp()
        uid = None if unique_id is None else np.ascontiguousarray(unique_id, dtype=np.uint8)
        self.check(self.L.gmc_comm_create(C.byref(h), world, rank, _ptr(uid)))
        return Comm(self, h, world, rank)

    def shard_cuts(self, prefix, world, align=1):
        prefix = _i32arr(prefix)
        cuts = np.zeros(world + 1, np.int32)
        self.check(self.L.gmc_shard_cuts(_ptr(prefix), len(prefix) - 1, world, align, _ptr(cuts)))
        return cuts

    # ---- files ---------------------------------------------------------------------------
    def file_read(self, path):
        n = _i32(); m = _i32()
        self.check(self.L.gmc_file_read_header(path.encode(), C.byref(n), C.byref(m)))
        begin = np.zeros(n.value + 1, np.int32); idx = np.zeros(max(m.value, 1), np.int32)
        self.check(self.L.gmc_file_read(path.encode(), _ptr(begin), _ptr(idx)))
        return begin, idx[:m.value]

    def file_write(self, path, begin, node_idx):
        begin = _i32arr(begin); node_idx = _i32arr(node_idx)
        self.check(self.L.gmc_file_write(path.encode(), len(begin) - 1, len(node_idx), _ptr(begin), _ptr(node_idx)))

    # ---- raw memory ----------------------------------------------------------------------
    def malloc(self, nbytes):
        p = _vp()
        self.check(self.L.gmc_malloc(C.byref(p), nbytes))
        return p.value

    def free(self, p):
        self.check(self.L.gmc_free(_vp(p)))

    def h2d(self, dev, host, stream=None):
        self.check(self.L.gmc_h2d(_vp(dev), _ptr(host), host.nbytes, _vp(stream) if stream else None))

    def d2h(self, host, dev, stream=None):
        self.check(self.L.gmc_d2h(_ptr(host), _vp(dev), host.nbytes, _vp(stream) if stream else None))

    def sync(self, stream=None):
        self.check(self.L.gmc_sync(_vp(stream) if stream else None))


class Group:
    """1/2/4/8 GPUs behind one process: CSR replicated, rows partitioned, one host thread per device inside the library"""

    def __init__(self, lib, handle):
        self.lib, self.h = lib, handle

    def upload(self, begin, node_idx, r_begin=None, r_node_idx=None):
        self._keep = [np.ascontiguousarray(a, dtype=np.int32) if a is not None else None for a in (begin, node_idx, r_begin, r_node_idx)]
        b, i, rb, ri = self._keep
        self.n, self.m = len(b) - 1, (len(i) if i is not None else len(ri))
        self.lib.check(self.lib.L.gmc_group_graph_upload(self.h, self.n, self.m, _ptr(b), _ptr(i) if i is not None else None,
                                                         _ptr(rb) if rb is not None else None, _ptr(ri) if ri is not None else None))

    def pagerank(self, e=1e-30, d=0.85, max_iter=20, out=None):
        rank = out if out is not None else np.zeros(max(self.n, 1), np.float64)
        info = PagerankInfo()
        self.lib.check(self.lib.L.gmc_group_pagerank(self.h, e, d, max_iter, _ptr(rank), C.byref(info)))
        return rank[:self.n], info

    def hop_dist(self, root=0, out=None):
        dist = out if out is not None else np.zeros(max(self.n, 1), np.int32)
        info = BfsInfo()
        self.lib.check(self.lib.L.gmc_group_hop_dist(self.h, root, _ptr(dist), C.byref(info)))
        return dist[:self.n], info

    def exchange_path(self):
        p = C.c_int()
        self.lib.check(self.lib.L.gmc_group_exchange_path(self.h, C.byref(p)))
        return {0: "single", 1: "nccl", 2: "peer-memory"}.get(p.value, str(p.value))

    def size(self):
        k = C.c_int()
        self.lib.check(self.lib.L.gmc_group_size(self.h, C.byref(k)))
        return k.value

    def close(self):
        if self.h:
            self.lib.L.gmc_group_destroy(self.h)
            self.h = None


class Comm:
    """gmc_comm: this process's rank in a one-process-per-GPU NCCL communicator."""

    def __init__(self, lib, handle, world, rank):
        self.lib, self.h, self.world, self.rank = lib, handle, world, rank

    def barrier(self, stream=None):
        self.lib.check(self.lib.L.gmc_comm_barrier(self.h, _vp(stream) if stream else None))

    def nccl_version(self):
        v = C.c_int()
        self.lib.check(self.lib.L.gmc_comm_info(self.h, None, None, C.byref(v)))
        return v.value

    def close(self):
        if self.h:
            self.lib.L.gmc_comm_destroy(self.h)
            self.h = None


class Graph:
    """Device-resident CSR (gmc_graph)."""

    def __init__(self, lib, handle):
        self.lib = lib
        self.h = handle

    def close(self):
        if self.h:
            self.lib.L.gmc_graph_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    @property
    def num_nodes(self):
        n = _i32(); self.lib.check(self.lib.L.gmc_graph_num_nodes(self.h, C.byref(n))); return n.value

    @property
    def num_edges(self):
        m = _i32(); self.lib.check(self.lib.L.gmc_graph_num_edges(self.h, C.byref(m))); return m.value

    def array(self, name):
        n, m = self.num_nodes, self.num_edges
        ln = n + 1 if name in ("begin", "r_begin") else m
        out = np.zeros(max(ln, 1), np.int32)
        self.lib.check(self.lib.L.gmc_graph_download(self.h, ARR[name], _ptr(out), ln, None))
        return out[:ln]

    def device_ptr(self, name):
        p = _vp()
        self.lib.check(self.lib.L.gmc_graph_device_ptr(self.h, ARR[name], C.byref(p)))
        return p.value

    def semi_sort(self, want_idx2idx=False):
        self.lib.check(self.lib.L.gmc_graph_semi_sort(self.h, int(want_idx2idx), None))

    def make_reverse(self, want_rev2idx=False):
        self.lib.check(self.lib.L.gmc_graph_make_reverse(self.h, int(want_rev2idx), None))

    def upload_reverse(self, r_begin, r_node_idx):
        r_begin = _i32arr(r_begin); r_node_idx = _i32arr(r_node_idx)
        self.lib.check(self.lib.L.gmc_graph_upload_reverse(self.h, _ptr(r_begin), _ptr(r_node_idx), None))

    def pagerank(self, e=1e-30, d=0.85, max_iter=20):
        rank = np.zeros(max(self.num_nodes, 1), np.float64)
        info = PagerankInfo()
        self.lib.check(self.lib.L.gmc_pagerank(self.h, e, d, max_iter, _ptr(rank), C.byref(info)))
        return rank[:self.num_nodes], info

    def triangle_counting(self, e_lo=None, e_hi=None):
        """returns (count, device ms); with a slot range: the partial count of edge slots [e_lo, e_hi)"""
        cnt = C.c_int64(); ms = C.c_float()
        if e_lo is None:
            self.lib.check(self.lib.L.gmc_triangle_counting(self.h, C.byref(cnt), C.byref(ms)))
        else:
            self.lib.check(self.lib.L.gmc_triangle_counting_range(self.h, e_lo, e_hi, C.byref(cnt), C.byref(ms), None))
        return cnt.value, ms.value

    def triangle_counting_noalloc(self, count_dev, stream=None):
        """gmc_triangle_counting_noalloc: allocates nothing, does not synchronise; the count (or -1: rows not ascending) is
        written into the caller's device int64 `count_dev` (an address from GmCuda.malloc) on `stream`"""
        self.lib.check(self.lib.L.gmc_triangle_counting_noalloc(self.h, _vp(count_dev), _vp(stream) if stream else None))

    def comp_bc(self, seeds, adj=False):
        """comp_BC (bc.gm; adj=True: bc_adj.gm) over the given seeds in order: returns (BC float32[N], BcInfo)"""
        seeds = np.ascontiguousarray(seeds, dtype=np.int32)
        bc = np.empty(max(self.num_nodes, 1), np.float32)
        info = BcInfo()
        self.lib.check((self.lib.L.gmc_comp_bc_adj if adj else self.lib.L.gmc_comp_bc)(self.h, _ptr(seeds), len(seeds), _ptr(bc), C.byref(info)))
        return bc[:self.num_nodes], info

    def triangle_counting_sharded(self, comm):
        cnt = C.c_int64(); ms = C.c_float()
        self.lib.check(self.lib.L.gmc_triangle_counting_sharded(self.h, comm.h if comm else None, C.byref(cnt), C.byref(ms), None))
        return cnt.value, ms.value

    def pagerank_into(self, rank_host, e=1e-30, d=0.85, max_iter=20):
        """result into a caller-owned (ideally pinned) host array of N doubles"""
        assert rank_host.dtype == np.float64 and len(rank_host) >= self.num_nodes
        info = PagerankInfo()
        self.lib.check(self.lib.L.gmc_pagerank(self.h, e, d, max_iter, _ptr(rank_host), C.byref(info)))
        return rank_host, info

    def pagerank_dev(self, rank_dev_ptr, e=1e-30, d=0.85, max_iter=20, stream=None):
        info = PagerankInfo()
        self.lib.check(self.lib.L.gmc_pagerank_dev(self.h, e, d, max_iter, _vp(rank_dev_ptr), C.byref(info),
                                                   _vp(stream) if stream else None))
        return info

    def pagerank_sharded_dev(self, comm, rank_dev_ptr, e=1e-30, d=0.85, max_iter=20, stream=None):
        info = PagerankInfo()
        self.lib.check(self.lib.L.gmc_pagerank_sharded_dev(self.h, comm.h if comm else None, e, d, max_iter, _vp(rank_dev_ptr),
                                                           C.byref(info), _vp(stream) if stream else None))
        return info

    def pagerank_emulated_shards(self, world, e=1e-30, d=0.85, max_iter=20):
        """test aid: `world` shard plans swept one after the other on this device; returns host ranks"""
        n = self.num_nodes
        dev = self.lib.malloc(8 * max(n, 1))
        try:
            info = PagerankInfo()
            self.lib.check(self.lib.L.gmc_pagerank_emulated_shards_dev(self.h, world, e, d, max_iter, _vp(dev), C.byref(info), None))
            rank = np.zeros(max(n, 1), np.float64)
            self.lib.d2h(rank, dev)
        finally:
            self.lib.free(dev)
        return rank[:n], info

    def pagerank_exchange_path(self):
        v = C.c_int()
        self.lib.check(self.lib.L.gmc_pagerank_exchange_path(self.h, C.byref(v)))
        return {0: "single", 1: "nccl", 2: "peer-memory"}[v.value]

    def pagerank_sharded_time_sweeps(self, comm, d=0.85, warmup=3, n=20, stream=None):
        sw = np.zeros(n, np.float32); ex = np.zeros(n, np.float32)
        self.lib.check(self.lib.L.gmc_pagerank_sharded_time_sweeps(self.h, comm.h if comm else None, d, warmup, n, _ptr(sw), _ptr(ex),
                                                                   _vp(stream) if stream else None))
        return sw, ex

    def hop_dist_sharded_dev(self, comm, dist_dev_ptr, root=0, stream=None):
        info = BfsInfo()
        self.lib.check(self.lib.L.gmc_hop_dist_sharded_dev(self.h, comm.h if comm else None, root, _vp(dist_dev_ptr), C.byref(info),
                                                           _vp(stream) if stream else None))
        return info

    def bfs_levels_sharded(self, comm, root=0, use_reverse_edge=False, cap=256):
        lev = np.zeros(max(self.num_nodes, 1), np.int16)
        trace = np.full(cap, -1, np.int32); cnt = np.zeros(cap, np.int32)
        info = BfsInfo()
        self.lib.check(self.lib.L.gmc_bfs_levels_sharded(self.h, comm.h if comm else None, root, int(use_reverse_edge), None, _ptr(lev),
                                                         _ptr(trace), _ptr(cnt), cap, C.byref(info)))
        k = min(info.levels, cap)
        return lev[:self.num_nodes], info, trace[:k], cnt[:k]

    def pagerank_plan_info(self):
        pi = PlanInfo()
        self.lib.check(self.lib.L.gmc_pagerank_plan_info(self.h, C.byref(pi)))
        return {k: getattr(pi, k) for k, _ in PlanInfo._fields_}

    def pagerank_time_sweeps(self, d=0.85, warmup=3, n=20, stream=None):
        ms = np.zeros(n, np.float32); lps = _i32()
        self.lib.check(self.lib.L.gmc_pagerank_time_sweeps(self.h, d, warmup, n, _ptr(ms), C.byref(lps),
                                                           _vp(stream) if stream else None))
        return ms, lps.value

    def hop_dist(self, root=0):
        dist = np.zeros(max(self.num_nodes, 1), np.int32)
        info = BfsInfo()
        self.lib.check(self.lib.L.gmc_hop_dist(self.h, root, _ptr(dist), C.byref(info)))
        return dist[:self.num_nodes], info

    def hop_dist_dev(self, dist_dev_ptr, root=0, stream=None):
        info = BfsInfo()
        self.lib.check(self.lib.L.gmc_hop_dist_dev(self.h, root, _vp(dist_dev_ptr), C.byref(info),
                                                   _vp(stream) if stream else None))
        return info

    def bfs_levels(self, root=0, use_reverse_edge=False, cap=256):
        lev = np.zeros(max(self.num_nodes, 1), np.int16)
        trace = np.full(cap, -1, np.int32); cnt = np.zeros(cap, np.int32)
        info = BfsInfo()
        self.lib.check(self.lib.L.gmc_bfs_levels(self.h, root, int(use_reverse_edge), _ptr(lev), _ptr(trace), _ptr(cnt), cap, C.byref(info)))
        k = min(info.levels, cap)
        return lev[:self.num_nodes], info, trace[:k], cnt[:k]

    def bfs_level_order(self, n_levels):
        order = np.zeros(max(self.num_nodes, 1), np.int32)
        lb = np.zeros(n_levels + 1, np.int32)
        self.lib.check(self.lib.L.gmc_bfs_level_order(self.h, _ptr(order), _ptr(lb), n_levels + 1))
        return order[:lb[-1]], lb
