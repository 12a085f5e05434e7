"""GPU parity tests proper: the CUDA path, called through the C ABI (include/gm_cuda.h), against the
oracle on the same seeded inputs.  BFS distances/levels and CSR builds: bit-exact.  PageRank: the
north-star tolerance max|gpu - ref| <= 1e-9 at a fixed iteration count (and in practice far tighter:
rows of in-degree <= 2048 are summed in the reference's order and must be bit-identical)."""
import numpy as np
import pytest

import graphs

pytestmark = pytest.mark.gpu
PR_TOL = 1e-9          # BASELINE.json north_star: max abs diff <= 1e-9 in double at fixed iteration count
INT_MAX = 2**31 - 1


def dev_graph(gm, g):
    dg = gm.graph_from_csr(g["begin"], g["node_idx"], semi_sort=False, make_reverse=False)
    dg.upload_reverse(g["r_begin"], g["r_node_idx"])
    return dg


@pytest.fixture(scope="module")
def g_c1(oracle):
    return graphs.c1(oracle)


@pytest.fixture(scope="module")
def g_rmat(oracle):
    return graphs.rmat(oracle, 200000, 3000000)


def test_reverse_build_c1(gm, oracle, g_c1):
    dg = gm.graph_from_csr(g_c1["begin"], g_c1["node_idx"], semi_sort=True, make_reverse=True, want_maps=True)
    assert np.array_equal(dg.array("node_idx"), g_c1["node_idx"])
    assert np.array_equal(dg.array("r_begin"), g_c1["r_begin"])
    assert np.array_equal(dg.array("r_node_idx"), g_c1["r_node_idx"])
    _, _, r2 = oracle.make_reverse(g_c1["begin"], g_c1["node_idx"], want_rev2idx=True)
    assert np.array_equal(dg.array("e_rev2idx"), r2)


def test_semi_sort_unsorted_input(gm, oracle):
    b, i = oracle.rmat(50000, 600000, 7, 0.57, 0.19, 0.19, True)     # generator output: rows NOT sorted
    dg = gm.graph_from_csr(b, i, semi_sort=True, make_reverse=True, want_maps=True)
    g = graphs.prepared(oracle, b, i)
    assert np.array_equal(dg.array("node_idx"), g["node_idx"])
    assert np.array_equal(dg.array("r_begin"), g["r_begin"])
    assert np.array_equal(dg.array("r_node_idx"), g["r_node_idx"])
    idx2idx = dg.array("e_idx2idx")
    assert np.array_equal(np.ascontiguousarray(i)[idx2idx], g["node_idx"])          # it IS the permutation applied
    rev2idx = dg.array("e_rev2idx")
    assert np.array_equal(np.sort(rev2idx), np.arange(len(i), dtype=np.int32))
    src_of_edge = np.repeat(np.arange(g["N"], dtype=np.int32), np.diff(g["begin"]))
    assert np.array_equal(src_of_edge[rev2idx], g["r_node_idx"])
    assert np.array_equal(g["node_idx"][rev2idx], np.repeat(np.arange(g["N"], dtype=np.int32), np.diff(g["r_begin"])))


def test_pagerank_c1_fixed_20(gm, oracle, g_c1):
    dg = dev_graph(gm, g_c1)
    rank, info = dg.pagerank(e=1e-30, d=0.85, max_iter=20)
    want, it, diff = oracle.pagerank(g_c1["begin"], g_c1["r_begin"], g_c1["r_node_idx"], 1e-30, 0.85, 20)
    assert info.iterations == it == 20
    assert np.abs(rank - want).max() <= PR_TOL
    # no row of C1 is longer than 2048 in-edges: summation order is the reference's => bit-identical
    assert np.array_equal(rank, want), f"max ulp {graphs.max_ulp(rank, want)}"
    assert abs(info.diff - diff) <= 1e-12 * max(1.0, abs(diff))
    assert info.launches > 0


def test_pagerank_convergence_exit(gm, oracle, g_c1):
    """default arguments of the reference driver (pagerank_main.cc:11-16): e=0.001, d=0.85, max=100"""
    dg = dev_graph(gm, g_c1)
    rank, info = dg.pagerank(e=0.001, d=0.85, max_iter=100)
    want, it, _ = oracle.pagerank(g_c1["begin"], g_c1["r_begin"], g_c1["r_node_idx"], 0.001, 0.85, 100)
    assert info.iterations == it
    assert np.array_equal(rank, want)


@pytest.mark.parametrize("max_iter", [1, 2, 17, 33])
def test_pagerank_iteration_counts(gm, oracle, g_c1, max_iter):
    dg = dev_graph(gm, g_c1)
    rank, info = dg.pagerank(e=1e-30, d=0.85, max_iter=max_iter)
    want, it, _ = oracle.pagerank(g_c1["begin"], g_c1["r_begin"], g_c1["r_node_idx"], 1e-30, 0.85, max_iter)
    assert info.iterations == it == max_iter
    assert np.array_equal(rank, want)


def test_pagerank_rmat_hubs(gm, oracle, g_rmat):
    """power-law rows: hubs above 2048 in-edges take the chunked path (different association)"""
    assert np.diff(g_rmat["r_begin"]).max() > 2048
    dg = dev_graph(gm, g_rmat)
    rank, info = dg.pagerank(e=1e-30, d=0.85, max_iter=20)
    want, it, _ = oracle.pagerank(g_rmat["begin"], g_rmat["r_begin"], g_rmat["r_node_idx"], 1e-30, 0.85, 20)
    assert info.iterations == it
    err = np.abs(rank - want).max()
    assert err <= PR_TOL
    assert err <= 1e-15, f"max abs err {err}, max ulp {graphs.max_ulp(rank, want)}"


@pytest.mark.parametrize("big", ["0", "1"])
def test_pagerank_task_sizes(gm, oracle, g_rmat, g_c1, big, monkeypatch):
    """both task-size sets of the sweep plan (default 2048/2048; large shards 8192/4096, normally chosen from the
    shard's entry count) must give the same result on the same graph: rows cut by task boundaries, <= 30 slices per task"""
    monkeypatch.setenv("GMC_PR_BIGTASKS", big)
    for g in (g_rmat, g_c1):
        dg = dev_graph(gm, g)
        rank, info = dg.pagerank(e=1e-30, d=0.85, max_iter=20)
        want, it, _ = oracle.pagerank(g["begin"], g["r_begin"], g["r_node_idx"], 1e-30, 0.85, 20)
        assert info.iterations == it
        assert np.abs(rank - want).max() <= 1e-15


def test_pagerank_host_csr_one_call(gm, oracle, g_rmat, g_c1):
    """gmc_pagerank_host_csr (copy-in overlapped with the plan build, chunked entry fill) == upload + gmc_pagerank, bit for bit;
    5 M edges: above the chunking threshold, so the entry array is filled in 8 chunks behind the copies"""
    big = graphs.rmat(oracle, 300000, 5000000)
    for g in (g_c1, g_rmat, big):
        rank = np.zeros(g["N"], np.float64)
        _, info = gm.pagerank_host_csr(g["begin"], g["r_begin"], g["r_node_idx"], rank, 1e-30, 0.85, 20)
        want, it, _ = oracle.pagerank(g["begin"], g["r_begin"], g["r_node_idx"], 1e-30, 0.85, 20)
        assert info.iterations == it
        assert np.abs(rank - want).max() <= 1e-15
        ref_rank, _ = dev_graph(gm, g).pagerank(e=1e-30, d=0.85, max_iter=20)
        if len(g["r_node_idx"]) < (4 << 20):
            assert np.array_equal(rank, ref_rank)
        else:       # the resident graph has tier streams from 4 M edges on, a one-shot call of 20 sweeps builds the plan without them
            assert np.abs(rank - ref_rank).max() <= 1e-15
    # a one-shot call long enough for the tiered plan (>= 100 sweeps): pass 1 behind the copy-in, pass 2 after it
    rank = np.zeros(big["N"], np.float64)
    _, info = gm.pagerank_host_csr(big["begin"], big["r_begin"], big["r_node_idx"], rank, 1e-30, 0.85, 120)
    want, it, _ = oracle.pagerank(big["begin"], big["r_begin"], big["r_node_idx"], 1e-30, 0.85, 120)
    assert info.iterations == it == 120 and np.abs(rank - want).max() <= 1e-15
    # degenerate inputs: no edges / no vertices
    b = np.zeros(6, np.int32); ri = np.zeros(0, np.int32); rank = np.zeros(5, np.float64)
    gm.pagerank_host_csr(b, b, ri, rank, 1e-30, 0.85, 3)
    assert np.allclose(rank, 0.15 / 5)
    gm.pagerank_host_csr(np.zeros(1, np.int32), np.zeros(1, np.int32), ri, np.zeros(0, np.float64), 1e-30, 0.85, 3)


def test_pagerank_device_built_reverse(gm, oracle, g_rmat):
    dg = gm.graph_from_csr(g_rmat["begin"], g_rmat["node_idx"], semi_sort=True, make_reverse=True)
    rank, _ = dg.pagerank(max_iter=5)
    want, _, _ = oracle.pagerank(g_rmat["begin"], g_rmat["r_begin"], g_rmat["r_node_idx"], 1e-30, 0.85, 5)
    assert np.abs(rank - want).max() <= 1e-15


def test_pagerank_needs_reverse(gm, g_c1):
    from gmb200_loader import load_package
    dg = gm.graph_from_csr(g_c1["begin"], g_c1["node_idx"], make_reverse=False)
    with pytest.raises(load_package().GmcError):
        dg.pagerank()


def test_hop_dist_c1(gm, oracle, g_c1):
    dg = dev_graph(gm, g_c1)
    dist, info = dg.hop_dist(0)
    want, _ = oracle.hop_dist(g_c1["begin"], g_c1["node_idx"], 0)
    assert np.array_equal(dist, want)
    assert list(dist[:10]) == [0, 4, 4, 4, 4, 4, 4, 4, 5, 5]      # SURVEY.md F10, from the reference binary
    assert info.reached == 65536 and info.levels == 7
    assert info.edges_traversed == g_c1["M"]


def test_bfs_levels_c1_state_machine(gm, oracle, g_c1):
    dg = dev_graph(gm, g_c1)
    lev, info, trace, cnt = dg.bfs_levels(0)
    want, nl, wtrace, wcnt = oracle.bfs_levels(g_c1["begin"], g_c1["node_idx"], g_c1["r_begin"], g_c1["r_node_idx"], 0)
    assert np.array_equal(lev, want)
    assert info.levels == nl
    assert list(trace) == list(wtrace)        # same direction switches as gm_bfs_template
    assert list(cnt) == list(wcnt)
    assert info.bottom_up_levels == sum(1 for s in wtrace if s in (5, 6))


@pytest.mark.parametrize("root_kind", ["zero", "first_with_edges", "hub", "random"])
def test_bfs_rmat_roots(gm, oracle, g_rmat, root_kind):
    deg = np.diff(g_rmat["begin"])
    root = {"zero": 0, "first_with_edges": graphs.first_root_with_edges(g_rmat), "hub": int(deg.argmax()),
            "random": int(np.random.RandomState(1997).randint(g_rmat["N"]))}[root_kind]
    dg = dev_graph(gm, g_rmat)
    dist, info = dg.hop_dist(root)
    want, _ = oracle.hop_dist(g_rmat["begin"], g_rmat["node_idx"], root)
    assert np.array_equal(dist, want)
    lev, linfo, trace, cnt = dg.bfs_levels(root)
    wlev, nl, wtrace, wcnt = oracle.bfs_levels(g_rmat["begin"], g_rmat["node_idx"], g_rmat["r_begin"], g_rmat["r_node_idx"], root)
    assert np.array_equal(lev, wlev)
    assert list(trace) == list(wtrace) and list(cnt) == list(wcnt)
    # the state machine of a save_child instantiation (flag 2): same levels, no direct QUE / Q2R -> RRD switch
    slev, _, strace, scnt = dg.bfs_levels(root, use_reverse_edge=2)
    _, _, wstrace, _ = oracle.bfs_levels(g_rmat["begin"], g_rmat["node_idx"], g_rmat["r_begin"], g_rmat["r_node_idx"], root, save_child=True)
    assert np.array_equal(slev, wlev) and list(scnt) == list(wcnt) and list(strace) == list(wstrace)
    # hop_dist and the BFS template agree: INT_MAX <-> -2
    assert np.array_equal(np.where(dist == INT_MAX, -2, dist).astype(np.int16), lev)
    assert info.reached == int((want != INT_MAX).sum())
    assert info.edges_traversed == int(deg[want != INT_MAX].sum())


def test_bfs_reverse_edges_variant(gm, oracle, g_rmat):
    """use_reverse_edge=true (template parameter): traversal over in-edges == BFS on the transposed graph"""
    root = int(np.diff(g_rmat["r_begin"]).argmax())
    dg = dev_graph(gm, g_rmat)
    lev, info, _, _ = dg.bfs_levels(root, use_reverse_edge=True)
    wlev, _, _, _ = oracle.bfs_levels(g_rmat["r_begin"], g_rmat["r_node_idx"], g_rmat["begin"], g_rmat["node_idx"], root)
    assert np.array_equal(lev, wlev)


def test_bfs_level_order(gm, oracle, g_rmat):
    root = int(np.diff(g_rmat["begin"]).argmax())
    dg = dev_graph(gm, g_rmat)
    lev, info, _, cnt = dg.bfs_levels(root)
    order, lb = dg.bfs_level_order(info.levels)
    assert list(np.diff(lb)) == list(cnt)
    for l in range(info.levels):
        assert np.array_equal(np.sort(order[lb[l]:lb[l + 1]]), np.nonzero(lev == l)[0])


def test_bfs_no_reverse_top_down_only(gm, oracle, g_c1):
    dg = gm.graph_from_csr(g_c1["begin"], g_c1["node_idx"], make_reverse=False)
    dist, info = dg.hop_dist(0)
    want, _ = oracle.hop_dist(g_c1["begin"], g_c1["node_idx"], 0)
    assert np.array_equal(dist, want)
    assert info.bottom_up_levels == 0


def test_bfs_two_blobs_and_a_path(gm, oracle):
    """grow -> shrink to one vertex for 40 levels -> grow again: small levels, grid levels and bottom-up levels follow each
    other in every order; hop_dist (edge-based direction policy) and gm_bfs_template levels (reference policy + trace)"""
    base = graphs.rmat(oracle, 60000, 900000)
    b, i, hub = graphs.two_blobs(base["begin"], base["node_idx"])
    g = graphs.prepared(oracle, b, i)
    dg = dev_graph(gm, g)
    for root in (hub, graphs.first_root_with_edges(g), g["N"] - 1):
        dist, info = dg.hop_dist(root)
        want, _ = oracle.hop_dist(g["begin"], g["node_idx"], root)
        assert np.array_equal(dist, want), root
        lev, linfo, trace, cnt = dg.bfs_levels(root)
        wlev, nl, wtrace, wcnt = oracle.bfs_levels(g["begin"], g["node_idx"], g["r_begin"], g["r_node_idx"], root)
        assert np.array_equal(lev, wlev) and list(trace) == list(wtrace) and list(cnt) == list(wcnt), root
    dist, info = dg.hop_dist(hub)
    assert info.levels > 45 and info.bottom_up_levels >= 2       # both blobs reached, each through bottom-up levels


def test_bfs_best_parent_and_reference_policy(gm, oracle, g_rmat, monkeypatch):
    """bottom-up scans with the best-parent table from the first traversal on (normally built for reused graphs only), and
    hop_dist under the reference's direction rule: same distances"""
    root = int(np.diff(g_rmat["begin"]).argmax())
    want, _ = oracle.hop_dist(g_rmat["begin"], g_rmat["node_idx"], root)
    monkeypatch.setenv("GMC_BFS_BEST_PARENT_NOW", "1")
    dg = dev_graph(gm, g_rmat)
    dist, info = dg.hop_dist(root)
    assert np.array_equal(dist, want) and info.bottom_up_levels >= 1
    monkeypatch.delenv("GMC_BFS_BEST_PARENT_NOW")
    monkeypatch.setenv("GMC_BFS_REF_POLICY", "1")
    dist2, info2 = dev_graph(gm, g_rmat).hop_dist(root)
    assert np.array_equal(dist2, want)
    _, _, wtrace, _ = oracle.bfs_levels(g_rmat["begin"], g_rmat["node_idx"], g_rmat["r_begin"], g_rmat["r_node_idx"], root)
    assert info2.bottom_up_levels == sum(1 for s in wtrace if s in (5, 6))      # the reference's rule: the template's switches


def test_micro_graphs(gm, oracle):
    """hand-checkable cases: path, star, 2-cycle + dangling vertex, isolated vertex, self loop, multi-edge"""
    cases = {
        "path": (5, [(0, 1), (1, 2), (2, 3), (3, 4)]),
        "star_out": (6, [(0, i) for i in range(1, 6)]),
        "star_in": (6, [(i, 0) for i in range(1, 6)]),
        "cycle2_dangling": (4, [(0, 1), (1, 0), (1, 2)]),
        "isolated": (3, [(0, 1)]),
        "selfloop_multi": (3, [(0, 0), (0, 1), (0, 1), (1, 2), (2, 2)]),
        "no_edges": (4, []),
        "single": (1, []),
    }
    for name, (n, edges) in cases.items():
        g = graphs.from_edges(oracle, n, edges)
        dg = gm.graph_from_csr(g["begin"], g["node_idx"], semi_sort=True, make_reverse=True)
        assert np.array_equal(dg.array("r_begin"), g["r_begin"]), name
        assert np.array_equal(dg.array("r_node_idx"), g["r_node_idx"]), name
        rank, info = dg.pagerank(e=1e-30, d=0.85, max_iter=10)
        want, it, _ = oracle.pagerank(g["begin"], g["r_begin"], g["r_node_idx"], 1e-30, 0.85, 10)
        assert info.iterations == it, name
        assert np.array_equal(rank, want), name
        for root in range(n):
            dist, _ = dg.hop_dist(root)
            wd, _ = oracle.hop_dist(g["begin"], g["node_idx"], root)
            assert np.array_equal(dist, wd), (name, root)
            lev, _, _, _ = dg.bfs_levels(root)
            wl, _, _, _ = oracle.bfs_levels(g["begin"], g["node_idx"], g["r_begin"], g["r_node_idx"], root)
            assert np.array_equal(lev, wl), (name, root)


def test_bad_arguments(gm, g_c1):
    from gmb200_loader import load_package
    Err = load_package().GmcError
    dg = dev_graph(gm, g_c1)
    with pytest.raises(Err):
        dg.hop_dist(-1)
    with pytest.raises(Err):
        dg.hop_dist(g_c1["N"])


def test_device_generators(gm, oracle):
    """device R-MAT / uniform generators: valid semi-sorted CSR, no self loops (R-MAT), reverse CSR equal
    to the oracle's build, and a degree skew comparable to the reference generator's."""
    n, m = 1 << 16, 1 << 20
    dg = gm.generate_rmat(n, m, seed=1997)
    b, i = dg.array("begin"), dg.array("node_idx")
    assert b[0] == 0 and b[-1] == m and np.all(np.diff(b) >= 0)
    assert i.min() >= 0 and i.max() < n
    src = np.repeat(np.arange(n, dtype=np.int32), np.diff(b))
    assert not np.any(src == i)                                  # self edges are redrawn
    rows_sorted = np.all((np.diff(i) >= 0) | (np.diff(src) > 0))
    assert rows_sorted
    rb, ri = oracle.make_reverse(b, i)
    assert np.array_equal(dg.array("r_begin"), rb) and np.array_equal(dg.array("r_node_idx"), ri)
    rb_ref, _ = oracle.rmat(n, m, 1997, 0.57, 0.19, 0.19, True)
    top_dev = np.sort(np.diff(b))[::-1][:n // 100].sum() / m    # edge share of the top 1% sources
    top_ref = np.sort(np.diff(rb_ref))[::-1][:n // 100].sum() / m
    assert abs(top_dev - top_ref) < 0.05, (top_dev, top_ref)
    du = gm.generate_uniform(n, m, seed=5)
    bu = du.array("begin")
    assert bu[-1] == m and np.diff(bu).max() < 64


def test_bfs_many_low_degree_rows(gm, oracle):
    """top-down chunks that span far more than 32 frontier rows (degree-0/1 vertices): a root fanning out to
    6000 vertices, every third of them a dead end, the others with one edge into a long tail"""
    k = 6000
    edges = [(0, 1 + i) for i in range(k)]
    edges += [(1 + i, 1 + k + i) for i in range(k) if i % 3]                 # degree 0 / 1 rows interleaved
    edges += [(1 + k + i, 1 + 2 * k + (i // 7)) for i in range(k) if i % 3]  # converge 7:1
    n = 1 + 2 * k + k // 7 + 2
    g = graphs.from_edges(oracle, n, edges)
    for has_rev in (True, False):
        dg = dev_graph(gm, g) if has_rev else gm.graph_from_csr(g["begin"], g["node_idx"], semi_sort=False, make_reverse=False)
        dist, info = dg.hop_dist(0)
        want, _ = oracle.hop_dist(g["begin"], g["node_idx"], 0)
        assert np.array_equal(dist, want)
        assert info.levels == 4


@pytest.mark.parametrize("n,m", [(300000, 4000000), (50000, 5000000)])
def test_bfs_invariants_device_graph(gm, n, m):
    """size-independent BFS properties on a device-generated R-MAT graph, several roots, repeated runs:
    no edge (u,v) with dist[v] > dist[u]+1, every reached vertex has a parent at dist-1, runs repeat bit for bit"""
    g = gm.generate_rmat(n, m, 1997)
    b, idx = g.array("begin"), g.array("node_idx")
    deg = np.diff(b)
    src = np.repeat(np.arange(n, dtype=np.int32), deg)
    roots = [int(np.nonzero(deg)[0][0]), int(np.argmax(deg)), int(np.nonzero(deg == 1)[0][0])]
    for root in roots:
        first = None
        for _ in range(3):
            d, _ = g.hop_dist(root)
            if first is None:
                first = d.copy()
            assert np.array_equal(d, first)
        du, dv = first[src].astype(np.int64), first[idx].astype(np.int64)
        assert not np.any((du < INT_MAX) & (dv > du + 1))
        best = np.full(n, INT_MAX, np.int64)
        np.minimum.at(best, idx, np.where(du < INT_MAX, du + 1, INT_MAX))
        reached = (first < INT_MAX) & (np.arange(n) != root)
        assert np.array_equal(best[reached], first[reached].astype(np.int64)) and first[root] == 0
        lev, _, _, cnt = g.bfs_levels(root)
        assert np.array_equal(np.where(lev >= 0, lev.astype(np.int64), INT_MAX), first.astype(np.int64))
        assert cnt.sum() == int((first < INT_MAX).sum())


def test_pull_only_graph(gm, oracle, g_c1):
    """the copy-in of the generated pagerank(): begin + reverse CSR only; out-edge traversals refuse loudly"""
    from gmb200_loader import load_package
    dg = gm.graph_pull_only(g_c1["begin"], g_c1["r_begin"], g_c1["r_node_idx"])
    rank, info = dg.pagerank(e=1e-30, d=0.85, max_iter=20)
    want, _, _ = oracle.pagerank(g_c1["begin"], g_c1["r_begin"], g_c1["r_node_idx"], 1e-30, 0.85, 20)
    assert np.array_equal(rank, want)
    with pytest.raises(load_package().GmcError):
        dg.hop_dist(0)
    with pytest.raises(load_package().GmcError):
        dg.array("node_idx")


@pytest.mark.parametrize("tiers,kmin,big,split", [(16, 3, "0", "0"), (31, 1, "0", "1"), (3, 2, "1", "0"), (16, 6, "1", "1"), (16, 4, "0", "1")])
def test_pagerank_tiered_streams(gm, oracle, g_rmat, tiers, kmin, big, split, monkeypatch):
    """the tiered layout (gmc_pr_plan.h) forced onto the 200 K / 3 M graph: (row, tier) segments in their own entry streams,
    partial sums through the part slots, pr_combine finishing the split rows -- against the oracle, and the plan must
    really have segments.  Production graphs get it from 4 M edges on (LJ- and TW-shaped runs, bench.py parity)."""
    monkeypatch.setenv("GMC_PR_TIER_MIN_EDGES", "0")
    monkeypatch.setenv("GMC_PR_TIERS", str(tiers))
    monkeypatch.setenv("GMC_PR_KMIN", str(kmin))
    monkeypatch.setenv("GMC_PR_BIGTASKS", big)
    monkeypatch.setenv("GMC_PR_SPLIT", split)
    monkeypatch.setenv("GMC_PR_NO_LITE", "1")          # every plan of this test with its tier streams, the first call's too
    monkeypatch.setenv("GMC_PR_COLD_STREAM", "1" if kmin == 2 else "0")       # one case with the cold stream of the split rows
    dg = dev_graph(gm, g_rmat)
    pi = dg.pagerank_plan_info()
    assert pi["tiers"] >= 3 and pi["segments"] > 0 and pi["split_rows"] > 0 and pi["part_slots"] == pi["segments"] + pi["split_rows"]
    assert pi["tier_ctas"] > 0
    rank, info = dg.pagerank(e=1e-30, d=0.85, max_iter=20)
    want, it, diff = oracle.pagerank(g_rmat["begin"], g_rmat["r_begin"], g_rmat["r_node_idx"], 1e-30, 0.85, 20)
    assert info.iterations == it == 20
    err = np.abs(rank - want).max()
    assert err <= 1e-15, f"max abs err {err}, max ulp {graphs.max_ulp(rank, want)}"
    assert abs(info.diff - diff) <= 1e-12 * max(1.0, abs(diff))
    # deterministic: a second plan of the same graph gives the same bits (nothing depends on the CTA schedule)
    rank2, _ = dev_graph(gm, g_rmat).pagerank(e=1e-30, d=0.85, max_iter=20)
    assert np.array_equal(rank, rank2)
    # convergence exit through pr_combine's loop control
    rank3, info3 = dg.pagerank(e=1e-4, d=0.85, max_iter=100)
    want3, it3, _ = oracle.pagerank(g_rmat["begin"], g_rmat["r_begin"], g_rmat["r_node_idx"], 1e-4, 0.85, 100)
    assert info3.iterations == it3 and np.abs(rank3 - want3).max() <= 1e-15


def test_pagerank_tiered_emulated_shards(gm, oracle, g_rmat, monkeypatch):
    """shard plans of the tiered layout (3 shards on one device, shared state): same result as the oracle"""
    monkeypatch.setenv("GMC_PR_TIER_MIN_EDGES", "0")
    dg = dev_graph(gm, g_rmat)
    rank, info = dg.pagerank_emulated_shards(3, e=1e-30, d=0.85, max_iter=20)
    want, it, _ = oracle.pagerank(g_rmat["begin"], g_rmat["r_begin"], g_rmat["r_node_idx"], 1e-30, 0.85, 20)
    assert info.iterations == it
    assert np.abs(rank - want).max() <= 1e-15


def test_bfs_host_loop_and_deep_graph(gm, oracle, g_rmat, monkeypatch):
    """the device-driven traversal (GMC_BFS_DEVICE_LOOP=1: state machine in bfs_finish, kernels of several levels enqueued at
    once) and the host-driven loop (default; multi-GPU runs; graphs deeper than the device trace) give the same levels and trace.  A path graph of 3000 vertices has 3000 levels: more than the device trace holds -> falls back to the host loop."""
    root = graphs.first_root_with_edges(g_rmat)
    lev2, _, trace2, cnt2 = dev_graph(gm, g_rmat).bfs_levels(root)               # host-driven (default)
    monkeypatch.setenv("GMC_BFS_DEVICE_LOOP", "1")
    lev, _, trace, cnt = dev_graph(gm, g_rmat).bfs_levels(root)
    assert np.array_equal(lev, lev2) and list(trace) == list(trace2) and list(cnt) == list(cnt2)
    n = 3000
    begin = np.minimum(np.arange(n + 1), n - 1).astype(np.int32)          # v -> v + 1
    idx = np.arange(1, n, dtype=np.int32)
    dg = gm.graph_from_csr(begin, idx, semi_sort=True, make_reverse=True)
    dist, info = dg.hop_dist(0)
    assert np.array_equal(dist, np.arange(n, dtype=np.int32)) and info.levels == n
    dist, info = dg.hop_dist(n - 200)                                     # 200 levels: inside the device trace
    want = np.full(n, INT_MAX, np.int32); want[n - 200:] = np.arange(200)
    assert np.array_equal(dist, want) and info.levels == 200


@pytest.mark.gpu
@pytest.mark.parametrize("mode", ["GMC_BFS_NO_SMALL", "GMC_BFS_NO_SEEN", "GMC_BFS_NO_BU_MARK", "GMC_BFS_FUSED_BU", "GMC_BFS_PERSISTENT", "GMC_BFS_DEVICE_LOOP"])
def test_bfs_loop_modes_agree(gm, oracle, g_rmat, monkeypatch, mode):
    """the default traversal (runs of small top-down levels in one CTA, bfs_small_levels; the rest as prepare / expand / finish
    launches) against the other ways to drive the same level kernels: no small-level kernel, one cooperative launch per
    traversal, the device-side state machine.  Levels, state trace and level counts must be the oracle's in every mode, from a quiet root, from
    the hub, and on graphs small enough that a small level is followed by a bottom-up one."""
    deg = np.diff(g_rmat["begin"])
    small = graphs.rmat(oracle, 600, 5000, seed=3)
    sdeg = np.diff(small["begin"])
    cases = [(g_rmat, graphs.first_root_with_edges(g_rmat)), (g_rmat, int(np.argmax(deg)))]
    cases += [(small, int(r)) for r in np.nonzero(sdeg)[0][[0, 1, 2, 5, 11]]] + [(small, int(np.argmax(sdeg)))]
    want = [oracle.bfs_levels(g["begin"], g["node_idx"], g["r_begin"], g["r_node_idx"], r) for g, r in cases]
    for env in (None, mode):
        if env: monkeypatch.setenv(env, "1")
        for (g, r), (wlev, nl, wtrace, wcnt) in zip(cases, want):
            lev, info, trace, cnt = dev_graph(gm, g).bfs_levels(r)
            assert info.levels == nl and np.array_equal(lev, wlev), (env, r)
            assert list(trace) == list(wtrace[:nl]) and list(cnt) == list(wcnt[:nl]), (env, r)


def _ks(a, b):
    """Kolmogorov-Smirnov distance of two integer samples (max gap of their CDFs)"""
    hi = int(max(a.max(), b.max())) + 1
    ca = np.cumsum(np.bincount(a, minlength=hi)) / len(a)
    cb = np.cumsum(np.bincount(b, minlength=hi)) / len(b)
    return float(np.abs(ca - cb).max())


def test_device_generator_matches_reference_distribution(gm, oracle):
    """the device R-MAT generator makes every benchmarked graph: its in- and out-degree distributions must be the reference
    generator's (create_RMAT_graph, graph_gen.cc:159-287, restated edge for edge by the oracle) -- KS distance of the degree
    CDFs <= 0.01, share of vertices without in-edges (it sets the roofline bytes) within 1 point, top-0.1 % edge share within
    2 points, maximum degree within 15 % -- at 2^18 vertices / 2^22 edges, three seeds against one reference graph"""
    n, m = 1 << 18, 1 << 22
    rb, ri = oracle.rmat(n, m, 1997, 0.57, 0.19, 0.19, True)
    ref_out = np.diff(rb).astype(np.int64)
    ref_in = np.bincount(ri, minlength=n).astype(np.int64)
    for seed in (1997, 5, 77):
        dg = gm.generate_rmat(n, m, seed=seed)
        b, i = dg.array("begin"), dg.array("node_idx")
        out_d = np.diff(b).astype(np.int64)
        in_d = np.bincount(i, minlength=n).astype(np.int64)
        for name, d, r in (("out", out_d, ref_out), ("in", in_d, ref_in)):
            assert _ks(d, r) <= 0.01, (seed, name, _ks(d, r))
            assert abs(float((d == 0).mean()) - float((r == 0).mean())) <= 0.01, (seed, name)
            top = lambda x: np.sort(x)[::-1][:n // 1000].sum() / m
            assert abs(top(d) - top(r)) <= 0.02, (seed, name, top(d), top(r))
            assert abs(d.max() - r.max()) <= 0.15 * r.max(), (seed, name, d.max(), r.max())


def test_device_generator_golden_digest(gm):
    """bit-for-bit reproducibility of the device generator across rounds: sha-256 of the generated CSR against the committed
    fixture (tests/golden/device_rmat_digest.json, written by tools/gen_digest.py on a B200)"""
    import hashlib, json, os
    p = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "device_rmat_digest.json")
    if not os.path.exists(p):
        pytest.skip("fixture not generated yet (tools/gen_digest.py)")
    for name, want in json.load(open(p)).items():
        dg = gm.generate_rmat(want["n"], want["m"], seed=want["seed"])
        b, i = dg.array("begin"), dg.array("node_idx")
        assert hashlib.sha256(np.ascontiguousarray(b).tobytes()).hexdigest() == want["sha256_begin"], name
        assert hashlib.sha256(np.ascontiguousarray(i).tobytes()).hexdigest() == want["sha256_node_idx"], name
