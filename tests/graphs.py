"""Seeded synthetic inputs shared by the parity tests (built by the ORACLE's generators, i.e. the
reference's generator algorithms; see oracle/gm_oracle.c)."""
import numpy as np


def prepared(oracle, begin, node_idx):
    """semi-sort + reverse edges, the state gm_graph::load_binary leaves a graph in."""
    node_idx = np.ascontiguousarray(node_idx).copy()
    oracle.semi_sort(begin, node_idx)
    r_begin, r_node_idx = oracle.make_reverse(begin, node_idx)
    return dict(begin=begin, node_idx=node_idx, r_begin=r_begin, r_node_idx=r_node_idx,
                N=len(begin) - 1, M=len(node_idx))


def c1(oracle):
    """BASELINE config 1: graph_gen 65536 1048576 <f> 0 (uniform multigraph, seed 1997)."""
    b, i = oracle.uniform_random(65536, 1048576, 1997)
    return prepared(oracle, b, i)


def rmat(oracle, n, m, seed=1997, permute=True):
    b, i = oracle.rmat(n, m, seed, 0.57, 0.19, 0.19, permute)
    return prepared(oracle, b, i)


def from_edges(oracle, n, edges):
    """tiny hand-made graphs: list of (src, dst)"""
    edges = sorted(edges)
    begin = np.zeros(n + 1, np.int32)
    for s, _ in edges:
        begin[s + 1] += 1
    begin = np.cumsum(begin).astype(np.int32)
    idx = np.array([d for _, d in edges], np.int32) if edges else np.zeros(0, np.int32)
    return prepared(oracle, begin, idx)


def two_blobs(begin, node_idx, path_len=40):
    """CSR of: the given graph (blob A) | a directed path of `path_len` vertices | a copy of the graph (blob B); the
    largest-out-degree vertex of A feeds the path, the end of the path feeds the largest-out-degree vertex of B.  A
    traversal from inside A grows, shrinks to ONE vertex for `path_len` levels, then grows again: every transition
    between the small-level, the grid and the bottom-up forms of a BFS level happens twice."""
    begin = np.asarray(begin, np.int64); node_idx = np.asarray(node_idx, np.int64)
    n, m = len(begin) - 1, len(node_idx)
    deg = np.diff(begin)
    hub = int(deg.argmax())
    src = np.repeat(np.arange(n, dtype=np.int64), deg)
    p0, b0 = n, n + path_len
    s = np.concatenate([src, [hub], np.arange(p0, p0 + path_len - 1), [p0 + path_len - 1], src + b0])
    d = np.concatenate([node_idx, [p0], np.arange(p0 + 1, p0 + path_len), [b0 + hub], node_idx + b0])
    order = np.lexsort((d, s))
    s, d = s[order], d[order]
    nn = 2 * n + path_len
    nb = np.zeros(nn + 1, np.int64)
    np.add.at(nb, s + 1, 1)
    return np.cumsum(nb).astype(np.int32), d.astype(np.int32), hub


def first_root_with_edges(g):
    deg = np.diff(g["begin"])
    nz = np.nonzero(deg)[0]
    return int(nz[0]) if len(nz) else 0


def max_ulp(a, b):
    a = np.ascontiguousarray(a, np.float64).view(np.int64)
    b = np.ascontiguousarray(b, np.float64).view(np.int64)
    return int(np.abs(a - b).max()) if len(a) else 0
