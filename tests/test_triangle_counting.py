"""triangle_counting.  CPU: the oracle's restatement of the generated loop nest against the same
loops over the reference's own gm_graph::has_edge_to (oracle/_ref) and against counts the reference produced for the
committed fixtures.  GPU: gmc_triangle_counting against the oracle (exact), and at a size the quadratic oracle cannot
reach against the identity T = trace(A^3)/6 on a symmetric simple graph."""
import json
import os
import numpy as np
import pytest

import graphs
from oracle_lib import have_ref

HERE = os.path.dirname(os.path.abspath(__file__))
GOLD = json.load(open(os.path.join(HERE, "golden", "triangle_counts.json")))


def fixture_graph(name):
    z = np.load(os.path.join(HERE, "golden", name + ".npz"))
    return z["begin"], z["node_idx"]


def tiny_cases(oracle):
    # 0->1, 0->2, 2->1: v=0, u=1, w=2, edge 2->1: one count.  Parallel edges of row v count per slot.
    yield graphs.from_edges(oracle, 3, [(0, 1), (0, 2), (2, 1)]), 1
    yield graphs.from_edges(oracle, 3, [(0, 1), (0, 2), (1, 2)]), 0              # the closing edge must go w -> u
    yield graphs.from_edges(oracle, 3, [(0, 1), (0, 1), (0, 2), (0, 2), (2, 1), (2, 1)]), 4
    yield graphs.from_edges(oracle, 4, [(1, 1), (0, 0), (3, 2)]), 0
    yield graphs.from_edges(oracle, 5, []), 0


def test_oracle_tiny_and_golden(oracle):
    for g, want in tiny_cases(oracle):
        assert oracle.triangle_counting(g["begin"], g["node_idx"]) == want
    for name, want in GOLD.items():
        b, i = fixture_graph(name)
        assert oracle.triangle_counting(b, i) == want, name


@pytest.mark.skipif(not have_ref(), reason="oracle/_ref not built (no /root/reference on this box)")
def test_oracle_matches_reference_has_edge_to(oracle, ref):
    for seed, (n, m) in enumerate([(300, 3000), (5000, 100000), (2000, 200000)]):
        g = ref.rmat(n, m, 1997 + seed)
        a = ref.arrays(g)
        assert oracle.triangle_counting(a["begin"], a["node_idx"]) == ref.triangle_counting(g)
        ref.free(g)
    g = ref.uniform(4096, 200000)
    a = ref.arrays(g)
    assert oracle.triangle_counting(a["begin"], a["node_idx"]) == ref.triangle_counting(g)
    ref.free(g)


@pytest.mark.gpu
def test_triangle_counting_parity(gm, oracle):
    cases = [g for g, _ in tiny_cases(oracle) if g["N"] > 0]
    cases += [graphs.c1(oracle), graphs.rmat(oracle, 20000, 400000), graphs.rmat(oracle, 3000, 300000, seed=7)]
    for g in cases:
        dg = gm.graph_from_csr(g["begin"], g["node_idx"], semi_sort=False, make_reverse=False)
        dg.upload_reverse(g["r_begin"], g["r_node_idx"])
        want = oracle.triangle_counting(g["begin"], g["node_idx"])
        got, _ = dg.triangle_counting()
        assert got == want
        # slot ranges add up (how a multi-GPU job splits the work)
        m = g["M"]
        cuts = [0, m // 3, m // 3, (2 * m) // 3 + 1 if m else 0, m]
        cuts = [min(c, m) for c in cuts]
        assert sum(dg.triangle_counting(cuts[k], cuts[k + 1])[0] for k in range(4)) == want
    for name, want in GOLD.items():
        b, i = fixture_graph(name)
        dg = gm.graph_from_csr(b, i, semi_sort=False, make_reverse=True)
        assert dg.triangle_counting()[0] == want


@pytest.mark.gpu
def test_triangle_counting_symmetric_identity(gm):
    """size-independent property: on a symmetric simple graph the procedure counts every triangle once = trace(A^3)/6"""
    import scipy.sparse as sp
    dg = gm.generate_rmat(200000, 3000000, 11)
    b, i = dg.array("begin"), dg.array("node_idx")
    n = len(b) - 1
    A = sp.csr_matrix((np.ones(len(i), np.int64), i, b), shape=(n, n))
    A = ((A + A.T) > 0).astype(np.int64)
    A.setdiag(0); A.eliminate_zeros(); A.sort_indices()
    want = int((A @ A).multiply(A).sum() // 6)
    sg = gm.graph_from_csr(A.indptr.astype(np.int32), A.indices.astype(np.int32), semi_sort=False, make_reverse=True)
    got, ms = sg.triangle_counting()
    assert got == want and want > 0


@pytest.mark.gpu
def test_triangle_counting_needs_reverse(gm, oracle):
    g = graphs.c1(oracle)
    dg = gm.graph_from_csr(g["begin"], g["node_idx"], semi_sort=False, make_reverse=False)
    with pytest.raises(Exception):
        dg.triangle_counting()


@pytest.mark.gpu
def test_triangle_counting_rejects_unsorted_rows(gm, oracle):
    """a CSR filled through prepare_external_creation need not have ascending rows: the count would be silently wrong, so the
    device entry point checks the rows once per graph and fails instead"""
    from gmb200_loader import load_package
    b, i = oracle.rmat(5000, 80000, 7, 0.57, 0.19, 0.19, True)          # generator output: rows NOT sorted
    g = graphs.prepared(oracle, b, i)                                      # the sorted graph, for its reverse CSR
    dg = gm.graph_from_csr(b, i, semi_sort=False, make_reverse=False)
    dg.upload_reverse(g["r_begin"], g["r_node_idx"])
    with pytest.raises(load_package().GmcError):
        dg.triangle_counting()
    dg.semi_sort()
    got = dg.triangle_counting()
    assert (got[0] if isinstance(got, tuple) else got) == oracle.triangle_counting(g["begin"], g["node_idx"])
