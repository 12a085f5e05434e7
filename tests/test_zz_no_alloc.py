"""gmc_triangle_counting_noalloc: the entry point for callers that cannot call cudaMalloc (no counterpart in the reference).
Its scratch is a fixed file-scope device array shared by all calls, so the test makes
two calls in a row on each graph -- the second one starts from the first one's leftovers -- and a call on a DIFFERENT graph in
between, and compares every result with the allocating entry point gmc_triangle_counting through the same C-ABI.  The count is
an integer: the tolerance is zero, as in tests/test_triangle_counting.py.

Collected last on purpose: a failure here hides nothing after it."""
import numpy as np
import pytest

import graphs

MAX_EDGES = 1 << 26        # GMC_TC_NOALLOC_MAX_EDGES, include/gm_cuda.h


def _noalloc_count(gm, dg, count_dev):
    dg.triangle_counting_noalloc(count_dev)
    gm.sync()
    out = np.full(1, -7, np.int64)
    gm.d2h(out, count_dev)
    return int(out[0])


@pytest.mark.gpu
def test_noalloc_matches_allocating_entry_point_twice_in_a_row(gm, oracle):
    count_dev = gm.malloc(8)          # the caller's output buffer: the entry point under test allocates nothing itself
    try:
        small = graphs.rmat(oracle, 2000, 30000, seed=11)
        tiny = graphs.from_edges(oracle, 3, [(0, 1), (0, 1), (0, 2), (0, 2), (2, 1), (2, 1)])       # 4, parallel edges count per slot
        empty = graphs.from_edges(oracle, 5, [])
        cases = []
        for g in (small, tiny, empty):
            dg = gm.graph_from_csr(g["begin"], g["node_idx"], semi_sort=True, make_reverse=True)
            cases.append((dg, dg.triangle_counting()[0]))
        big = gm.generate_rmat(1 << 17, 1 << 21, 1997)      # heavy intersections: the segment list of the scratch is in use
        cases.append((big, big.triangle_counting()[0]))
        assert cases[0][1] == oracle.triangle_counting(small["begin"], small["node_idx"]) and cases[1][1] == 4 and cases[2][1] == 0
        assert cases[3][1] > 0
        for dg, want in cases + cases[::-1]:
            first = _noalloc_count(gm, dg, count_dev)
            second = _noalloc_count(gm, dg, count_dev)        # same array, the first call's contents still in it
            assert first == want and second == want, (first, second, want)
    finally:
        gm.free(count_dev)


@pytest.mark.gpu
def test_noalloc_reports_unsorted_rows_and_bad_arguments(gm, oracle):
    count_dev = gm.malloc(8)
    try:
        # rows not ascending: the allocating entry point fails with an error, this one delivers -1 with the result
        b, i = oracle.rmat(5000, 80000, 7, 0.57, 0.19, 0.19, True)          # generator output: rows NOT sorted
        g = graphs.prepared(oracle, b, i)                                      # the sorted graph, for its reverse CSR
        dg = gm.graph_from_csr(b, i, semi_sort=False, make_reverse=False)
        dg.upload_reverse(g["r_begin"], g["r_node_idx"])
        assert _noalloc_count(gm, dg, count_dev) == -1
        dg.semi_sort()
        assert _noalloc_count(gm, dg, count_dev) == oracle.triangle_counting(g["begin"], g["node_idx"])
        ok = graphs.from_edges(oracle, 3, [(0, 1), (0, 2), (2, 1)])
        dg2 = gm.graph_from_csr(ok["begin"], ok["node_idx"], semi_sort=True, make_reverse=True)
        assert _noalloc_count(gm, dg2, count_dev) == 1          # and the flag does not stick to the next call
        with pytest.raises(Exception, match="NULL"):
            dg2.triangle_counting_noalloc(0)
        no_rev = gm.graph_from_csr(ok["begin"], ok["node_idx"], semi_sort=True, make_reverse=False)
        with pytest.raises(Exception, match="reverse"):
            no_rev.triangle_counting_noalloc(count_dev)
    finally:
        gm.free(count_dev)


def test_noalloc_limit_is_declared_in_the_header():
    import os, re
    hdr = open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "include", "gm_cuda.h")).read()
    m = re.search(r"#define\s+GMC_TC_NOALLOC_MAX_EDGES\s+\(1\s*<<\s*(\d+)\)", hdr)
    assert m and (1 << int(m.group(1))) == MAX_EDGES
    assert "gmc_triangle_counting_noalloc" in hdr
