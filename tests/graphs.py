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


def first_root_with_edges(g):
    deg = np.diff(g["begin"])
    nz = np.nonzero(deg)[0]
    return int(nz[0]) if len(nz) else 0


def max_ulp(a, b):
    a = np.ascontiguousarray(a, np.float64).view(np.int64)
    b = np.ascontiguousarray(b, np.float64).view(np.int64)
    return int(np.abs(a - b).max()) if len(a) else 0
