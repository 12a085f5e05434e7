"""comp_BC (apps/src/bc.gm, bc_adj.gm): the device procedure (gmc_comp_bc / gmc_comp_bc_adj: traversal
and both callbacks as kernels) and the facade's gm_bfs_template (device traversal, arbitrary visit_fw / visit_rv callbacks
replayed on the host).  The oracle (plain C) is checked against the same generated class over the reference's own
gm_bfs_template (oracle/_ref).  bc.gm as shipped yields NaN wherever a vertex has a BFS child (the root's sigma is
overwritten by an empty sum), so arrays are compared NaN-aware, bit for bit otherwise; bc_adj.gm (the (v != s) filters)
is the variant with real path counts, which is what exercises the float arithmetic."""
import os
import re
import subprocess
import numpy as np
import pytest

import graphs
from oracle_lib import have_ref

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BIN = os.path.join(ROOT, "green-marl_b200", "bin")


def same(a, b):
    return a.shape == b.shape and np.array_equal(np.isnan(a), np.isnan(b)) and np.array_equal(a[~np.isnan(a)].view(np.uint32), b[~np.isnan(b)].view(np.uint32))


@pytest.mark.skipif(not have_ref(), reason="oracle/_ref not built (no /root/reference on this box)")
def test_oracle_matches_reference_bfs_template(oracle, ref):
    # one thread: with several, the reference's down-edge marking races (a thread that finds u's bit already set reads
    # visited_level[u] before the thread that set the bit has written it, gm_bfs_template.h:596-622, and skips the edge),
    # so its output is not deterministic; the single-thread run is the semantics: every edge to level + 1 is a down edge
    ref.set_threads(1)
    for n, m, seed in [(300, 2000, 5), (5000, 60000, 3), (20000, 400000, 7)]:
        g = ref.rmat(n, m, seed)
        a = ref.arrays(g)
        nz = np.nonzero(np.diff(a["begin"]))[0]
        seeds = [int(nz[k]) for k in (0, 5, 9)] + [0]
        want = ref.comp_bc(g, seeds)
        got = oracle.comp_bc(a["begin"], a["node_idx"], a["r_begin"], a["r_node_idx"], seeds)
        assert same(got, want)
        assert np.isnan(want).any() and (want == 0).any()       # the shipped bc.gm's degenerate output, see module docstring
        want = ref.comp_bc(g, seeds, adj=True)
        got = oracle.comp_bc(a["begin"], a["node_idx"], a["r_begin"], a["r_node_idx"], seeds, adj=True)
        assert same(got, want)
        assert not np.isnan(want).any() and (want > 0).any()    # bc_adj.gm: real dependencies
        ref.free(g)
    ref.set_threads(min(8, os.cpu_count() or 1))


def golden_bc_case(oracle, name):
    gold = os.path.join(ROOT, "tests", "golden")
    b = np.load(os.path.join(gold, "bc_golden.npz"))
    if name == "rmat5000":      # not stored: the oracle's generator reproduces the reference's graph, pinned by a digest
        import hashlib
        n, m, seed = (int(x) for x in b["rmat5000_params"])
        z = graphs.rmat(oracle, n, m, seed=seed)
        cat = np.concatenate([z["begin"], z["node_idx"], z["r_begin"], z["r_node_idx"]])
        assert hashlib.sha256(np.ascontiguousarray(cat).tobytes()).hexdigest() == str(b["rmat5000_csr_sha"])
    else:
        z = np.load(os.path.join(gold, name + ".npz"))
    return z, b


@pytest.mark.parametrize("name", ["rmat4096", "uniform2000", "rmat5000"])
def test_oracle_matches_golden(oracle, name):
    """tests/golden/bc_golden.npz (make_golden.py bc): the reference's own template, single thread; travels to the GPU box"""
    z, b = golden_bc_case(oracle, name)
    seeds = b[name + "_seeds"]
    for adj, key in ((False, "_bc"), (True, "_bc_adj")):
        got = oracle.comp_bc(z["begin"], z["node_idx"], z["r_begin"], z["r_node_idx"], seeds, adj=adj)
        assert same(got, b[name + key])
    assert not np.isnan(b[name + "_bc_adj"]).any() and (b[name + "_bc_adj"] > 0).any()
    if name == "rmat5000":
        # this case exercises the unmarked bottom-up levels: the save_child traversal from seeds[0] expands an inner level in RRD
        _, _, trace, cnt = oracle.bfs_levels(z["begin"], z["node_idx"], z["r_begin"], z["r_node_idx"], int(seeds[0]), save_child=True)
        assert 5 in list(trace)[:-1]


def test_oracle_tiny(oracle):
    # path 0 -> 1 -> 2: from seed 0 the inner vertices (0, 1) have a child -> NaN, the leaf 2 stays 0
    g = graphs.from_edges(oracle, 4, [(0, 1), (1, 2)])
    bc = oracle.comp_bc(g["begin"], g["node_idx"], g["r_begin"], g["r_node_idx"], [0])
    assert np.isnan(bc[0]) and np.isnan(bc[1]) and bc[2] == 0 and bc[3] == 0


def test_oracle_adj_tiny(oracle):
    # diamond 0 -> {1, 2} -> 3 -> 4 from seed 0: sigma = 1,1,1,2,2; delta(3) = 1, delta(1) = delta(2) = 1/2 * (1 + 1) = 1
    g = graphs.from_edges(oracle, 5, [(0, 1), (0, 2), (1, 3), (2, 3), (3, 4)])
    bc = oracle.comp_bc(g["begin"], g["node_idx"], g["r_begin"], g["r_node_idx"], [0], adj=True)
    assert bc.tolist() == [0.0, 1.0, 1.0, 1.0, 0.0]


@pytest.mark.gpu
def test_comp_bc_device_parity(gm, oracle):
    """gmc_comp_bc / gmc_comp_bc_adj through the C ABI against the oracle, bit for bit (NaN-aware)"""
    cases = [graphs.from_edges(oracle, 5, [(0, 1), (0, 2), (1, 3), (2, 3), (3, 4)]),
             graphs.rmat(oracle, 3000, 40000, seed=11), graphs.rmat(oracle, 200000, 3000000, seed=12),
             graphs.prepared(oracle, *oracle.uniform_random(50000, 400000, 5))]
    for g in cases:
        G = gm.graph_from_csr(g["begin"], g["node_idx"], semi_sort=False, make_reverse=True)
        deg = np.diff(g["begin"])
        nz = np.nonzero(deg)[0]
        seeds = [int(nz[0]), int(nz[len(nz) // 2]), int(deg.argmax()), int(np.diff(g["r_begin"]).argmax()), 1, int(nz[0])]
        for adj in (False, True):
            got, info = G.comp_bc(seeds, adj=adj)
            want = oracle.comp_bc(g["begin"], g["node_idx"], g["r_begin"], g["r_node_idx"], seeds, adj=adj)
            assert same(got, want), (g["N"], adj)
            assert info.seeds == len(seeds) and info.launches > 0 and info.device_ms > 0
            if adj and g["N"] > 100:
                assert not np.isnan(want).any() and (want > 0).any()
        got, _ = G.comp_bc([], adj=True)
        assert not got.any()
        with pytest.raises(Exception):
            G.comp_bc([g["N"]])
        G.close()


@pytest.mark.gpu
@pytest.mark.parametrize("via_template", [False, True])
def test_comp_bc_facade_parity(gm, oracle, tmp_path, via_template):
    exe = os.path.join(BIN, "bc_check")
    assert os.path.exists(exe), f"{exe} missing: run __graft_entry__.build() where /root/reference is present"
    env = dict(os.environ)
    if via_template:
        env["GM_BC_VIA_TEMPLATE"] = "1"
    for k, (n, m) in enumerate([(3000, 40000), (200000, 3000000)]):
        g = graphs.rmat(oracle, n, m, seed=11 + k)
        f = str(tmp_path / f"g{k}.bin"); out = str(tmp_path / f"bc{k}.f32")
        assert oracle.store_binary(f, g["begin"], g["node_idx"])
        nz = np.nonzero(np.diff(g["begin"]))[0]
        seeds = [int(nz[0]), int(nz[len(nz) // 2]), int(np.diff(g["begin"]).argmax()), 1]
        r = subprocess.run([exe, f, out] + [str(s) for s in seeds], capture_output=True, text=True, timeout=300, env=env)
        assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
        assert ("comp_BC: seeds=" in r.stdout) != via_template      # the device procedure reports itself
        got = np.fromfile(out, np.float32)
        want = oracle.comp_bc(g["begin"], g["node_idx"], g["r_begin"], g["r_node_idx"], seeds)
        assert same(got, want)


@pytest.mark.gpu
def test_bc_driver_unchanged(gm, oracle, tmp_path):
    """apps/output_cpp/src/bc_main.cc, unchanged: 5 seeds from gm_rand32 (the reference's xorshift, quirk included), prints BC[0..3]"""
    g = graphs.rmat(oracle, 20000, 400000)
    f = str(tmp_path / "r.bin")
    assert oracle.store_binary(f, g["begin"], g["node_idx"])
    r = subprocess.run([os.path.join(BIN, "bc"), f, "4", "/dev/null"], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    # gm_rand32 (gm_rand.cc:10-23): x ^= x << 13; x = x >> 17 (assignment, as in the reference); x ^= x << 5
    def i32(x):
        x &= 0xffffffff
        return x - (1 << 32) if x & 0x80000000 else x
    x, seeds = i32(2463534242), []
    while len(seeds) < 5:
        x = i32(x ^ (x << 13)); x = x >> 17; x = i32(x ^ (x << 5))
        if x < g["N"]:
            seeds.append(x)
    want = oracle.comp_bc(g["begin"], g["node_idx"], g["r_begin"], g["r_node_idx"], seeds)
    got = re.findall(r"BC\[\d\] = (\S+)", r.stdout)
    assert len(got) == 4
    for k in range(4):
        assert (np.isnan(want[k]) and "nan" in got[k].lower()) or float(got[k]) == float("%0.9f" % want[k])


@pytest.mark.gpu
def test_bc_random_driver_unchanged(gm, oracle, tmp_path):
    """apps/output_cpp/src/bc_random_main.cc, unchanged: bc_random(G, BC, 10), seeds = rand() % N from libc's default stream"""
    import sys
    g = graphs.rmat(oracle, 20000, 400000)
    f = str(tmp_path / "r.bin")
    assert oracle.store_binary(f, g["begin"], g["node_idx"])
    r = subprocess.run([os.path.join(BIN, "bc_random"), f, "4", "/dev/null"], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    assert "bc_random: seeds=10" in r.stdout
    draws = subprocess.run([sys.executable, "-c", "import ctypes; L = ctypes.CDLL('libc.so.6'); print(*[L.rand() for _ in range(10)])"],
                           capture_output=True, text=True, timeout=60)
    seeds = [int(x) % g["N"] for x in draws.stdout.split()]
    assert len(seeds) == 10
    want = oracle.comp_bc(g["begin"], g["node_idx"], g["r_begin"], g["r_node_idx"], seeds)
    got = re.findall(r"BC\[\d\] = (\S+)", r.stdout)
    assert len(got) == 4
    for k in range(4):
        assert (np.isnan(want[k]) and "nan" in got[k].lower()) or float(got[k]) == float("%0.9f" % want[k])
