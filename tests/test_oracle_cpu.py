"""CPU suite (-m "not gpu"): pins the plain-C oracle (oracle/gm_oracle.c) against (a) the committed golden
fixtures produced by the reference itself (tests/golden/make_golden.py) and (b), where oracle/_ref has been
built, the reference's own code run live.  No GPU, no libgm_cuda compute."""
import hashlib
import os

import numpy as np
import pytest

import graphs

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
INT_MAX = 2**31 - 1


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


@pytest.fixture(scope="module")
def g_c1(oracle):
    return graphs.c1(oracle)


def test_golden_c1_generator_and_csr(oracle, g_c1):
    """BASELINE config 1: graph_gen 65536 1048576 <f> 0 -- same glibc rand() stream, same CSR, same reverse CSR"""
    z = np.load(os.path.join(GOLD, "c1_digest.npz"))
    assert g_c1["N"] == int(z["N"]) and g_c1["M"] == int(z["M"])
    assert sha(g_c1["begin"]) == str(z["begin_sha"])
    assert sha(g_c1["node_idx"]) == str(z["node_idx_sha"])
    assert sha(g_c1["r_begin"]) == str(z["r_begin_sha"])
    assert sha(g_c1["r_node_idx"]) == str(z["r_node_idx_sha"])


def test_golden_c1_pagerank_bit_exact(oracle, g_c1):
    z = np.load(os.path.join(GOLD, "c1_digest.npz"))
    rank, it, _ = oracle.pagerank(g_c1["begin"], g_c1["r_begin"], g_c1["r_node_idx"], 1e-30, 0.85, 20)
    assert it == 20
    assert np.array_equal(rank[:64], z["rank_head"])
    assert sha(rank) == str(z["rank_sha"])
    # what the reference binary printed with %0.9lf
    assert ["%0.9f" % v for v in rank[:4]] == ["0.000013077", "0.000011851", "0.000011365", "0.000010712"]


def test_golden_c1_bfs_bit_exact(oracle, g_c1):
    z = np.load(os.path.join(GOLD, "c1_digest.npz"))
    dist, sweeps = oracle.hop_dist(g_c1["begin"], g_c1["node_idx"], 0)
    assert sha(dist) == str(z["dist_sha"])
    assert list(dist[:10]) == [0, 4, 4, 4, 4, 4, 4, 4, 5, 5]
    assert np.array_equal(np.bincount(dist), z["dist_hist"])
    lev, nl, _, _ = oracle.bfs_levels(g_c1["begin"], g_c1["node_idx"], g_c1["r_begin"], g_c1["r_node_idx"], 0)
    assert sha(lev) == str(z["lev_sha"]) and nl == int(z["n_levels"])


@pytest.mark.parametrize("name", ["rmat4096", "uniform2000"])
def test_golden_small_graphs(oracle, name):
    """full arrays from the reference: CSR prep, PageRank (fixed 20 and converged), hop_dist, gm_bfs_template"""
    z = np.load(os.path.join(GOLD, name + ".npz"))
    begin, idx = z["begin"], z["node_idx"]
    i2 = idx.copy()
    oracle.semi_sort(begin, i2)
    assert np.array_equal(i2, idx)                       # already semi-sorted by the reference
    rb, ri = oracle.make_reverse(begin, idx)
    assert np.array_equal(rb, z["r_begin"]) and np.array_equal(ri, z["r_node_idx"])
    rank, it, _ = oracle.pagerank(begin, rb, ri, 1e-30, 0.85, 20)
    assert it == 20 and np.array_equal(rank, z["rank20"])
    if "rank_conv" in z:
        rank, _, _ = oracle.pagerank(begin, rb, ri, 0.001, 0.85, 100)
        assert np.array_equal(rank, z["rank_conv"])
    roots = z["roots"] if "roots" in z else np.array([0])
    dists = z["dist"] if z["dist"].ndim == 2 else z["dist"][None]
    levs = z["levels"] if z["levels"].ndim == 2 else z["levels"][None]
    for k, r in enumerate(roots):
        d, _ = oracle.hop_dist(begin, idx, int(r))
        assert np.array_equal(d, dists[k])
        l, _, _, _ = oracle.bfs_levels(begin, idx, rb, ri, int(r))
        assert np.array_equal(l, levs[k])


def test_rmat_generator_matches_reference_stream(oracle):
    """the oracle's R-MAT draws the reference's drand48 stream: same CSR as the golden graph"""
    z = np.load(os.path.join(GOLD, "rmat4096.npz"))
    g = graphs.rmat(oracle, 4096, 65536)
    assert np.array_equal(g["begin"], z["begin"]) and np.array_equal(g["node_idx"], z["node_idx"])


# ---- live cross-check against the reference's own code (container only: oracle/_ref) --------------------
def test_live_reference_rmat_pagerank_bfs(oracle, ref):
    g = ref.rmat(30000, 400000, 7)
    a = ref.arrays(g)
    og = graphs.rmat(oracle, 30000, 400000, seed=7)
    for k in ("begin", "node_idx", "r_begin", "r_node_idx"):
        assert np.array_equal(a[k], og[k]), k
    want, _ = ref.pagerank(g, 1e-30, 0.85, 10)
    got, it, _ = oracle.pagerank(og["begin"], og["r_begin"], og["r_node_idx"], 1e-30, 0.85, 10)
    assert np.array_equal(got, want)
    for root in (0, graphs.first_root_with_edges(og), int(np.diff(og["begin"]).argmax())):
        wd, _ = ref.hop_dist(g, root)
        gd, _ = oracle.hop_dist(og["begin"], og["node_idx"], root)
        assert np.array_equal(gd, wd)
        wl, nl, _ = ref.bfs_levels(g, root)
        gl, gnl, _, _ = oracle.bfs_levels(og["begin"], og["node_idx"], og["r_begin"], og["r_node_idx"], root)
        assert np.array_equal(gl, wl)
        assert np.array_equal(np.where(gd == INT_MAX, -2, gd).astype(np.int16), gl)
    ref.free(g)


def test_live_reference_thread_count_independent(ref):
    g = ref.uniform(20000, 300000, 3)
    ref.set_threads(1)
    r1, _ = ref.pagerank(g, 1e-30, 0.85, 8)
    d1, _ = ref.hop_dist(g, 0)
    ref.set_threads(min(8, os.cpu_count() or 1))
    r8, _ = ref.pagerank(g, 1e-30, 0.85, 8)
    d8, _ = ref.hop_dist(g, 0)
    assert np.array_equal(r1, r8) and np.array_equal(d1, d8)
    ref.free(g)


def test_binary_file_roundtrip_with_reference(oracle, ref, tmp_path):
    """basic binary format (gm_graph_binary_loader.cc:19-26): oracle writer <-> reference loader and back"""
    og = graphs.rmat(oracle, 5000, 60000, seed=11)
    p = str(tmp_path / "g.bin")
    assert oracle.store_binary(p, og["begin"], og["node_idx"])
    g = ref.load(p)
    assert g
    a = ref.arrays(g)
    assert np.array_equal(a["begin"], og["begin"]) and np.array_equal(a["node_idx"], og["node_idx"])
    p2 = str(tmp_path / "g2.bin")
    assert ref.store(g, p2)
    assert open(p, "rb").read() == open(p2, "rb").read()
    b, i = oracle.load_binary(p2)
    assert np.array_equal(b, og["begin"]) and np.array_equal(i, og["node_idx"])
    ref.free(g)


def test_micro_graphs_hand_checked(oracle):
    """values a reader can verify by hand (d = 0.85, N = 3, one sweep from 1/3):
    edges 0->1, 0->2, 1->2: rank1 = (0.05, 0.05 + .85*(1/3)/2, 0.05 + .85*((1/3)/2 + 1/3))"""
    g = graphs.from_edges(oracle, 3, [(0, 1), (0, 2), (1, 2)])
    rank, it, diff = oracle.pagerank(g["begin"], g["r_begin"], g["r_node_idx"], 1e-30, 0.85, 1)
    base = (1 - 0.85) / 3.0
    third = 1 / 3.0
    want = [base, base + 0.85 * (third / 2.0), base + 0.85 * (third / 2.0 + third / 1.0)]
    assert it == 1 and list(rank) == want
    assert diff == sum(abs(w - third) for w in want)
    d, _ = oracle.hop_dist(g["begin"], g["node_idx"], 1)
    assert list(d) == [INT_MAX, 0, 1]
    l, nl, _, _ = oracle.bfs_levels(g["begin"], g["node_idx"], g["r_begin"], g["r_node_idx"], 1)
    assert list(l) == [-2, 0, 1] and nl == 2


def test_scipy_cross_check(oracle):
    """independent fp64 formulation (sparse mat-vec) within 1e-12 -- sanity, not the parity oracle"""
    import scipy.sparse as sp
    g = graphs.rmat(oracle, 3000, 40000, seed=5)
    n = g["N"]
    src = np.repeat(np.arange(n), np.diff(g["begin"]))
    deg = np.diff(g["begin"]).astype(np.float64)
    A = sp.csr_matrix((1.0 / deg[src], (g["node_idx"], src)), shape=(n, n))
    r = np.full(n, 1.0 / n)
    for _ in range(15):
        r = (1 - 0.85) / n + 0.85 * (A @ r)
    got, _, _ = oracle.pagerank(g["begin"], g["r_begin"], g["r_node_idx"], 1e-30, 0.85, 15)
    assert np.abs(got - r).max() < 1e-12
