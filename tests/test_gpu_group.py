"""Device group (gmc_group_*): 1/2/4/8 GPUs behind ONE process -- what GM_CUDA_DEVICES selects for the reference's unchanged
drivers (host/shl.cc).  Results against the oracle: PageRank <= 1e-9 (in practice 1e-15), hop_dist bit-exact.  Sizes above one
device run only where the box has them (the driver's GPU test box has one; gpurun --gpus 2 runs the rest)."""
import numpy as np
import pytest

import graphs

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def g_rmat(oracle):
    return graphs.rmat(oracle, 200000, 3000000)


@pytest.mark.parametrize("ndev", [1, 2, 4, 8])
def test_group_pagerank_hop_dist(gm, oracle, g_rmat, ndev, monkeypatch):
    if gm.device_count() < ndev:
        pytest.skip(f"needs {ndev} GPUs")
    monkeypatch.setenv("GMC_PR_TIER_MIN_EDGES", "0")      # tier streams on this small graph too
    g = g_rmat
    grp = gm.group(ndev)
    try:
        assert grp.size() == ndev
        grp.upload(g["begin"], g["node_idx"], g["r_begin"], g["r_node_idx"])
        want, it, _ = oracle.pagerank(g["begin"], g["r_begin"], g["r_node_idx"], 1e-30, 0.85, 20)
        for _ in range(2):                                  # second call: plan and peer mappings reused
            rank, info = grp.pagerank(1e-30, 0.85, 20)
            assert info.iterations == it == 20
            assert np.abs(rank - want).max() <= 1e-15
        assert grp.exchange_path() == ("single" if ndev == 1 else "peer-memory")
        root = graphs.first_root_with_edges(g)
        dist, binfo = grp.hop_dist(root)
        wdist, _ = oracle.hop_dist(g["begin"], g["node_idx"], root)
        assert np.array_equal(dist, wdist)
        # a second graph in the same group (pull-only: what the generated pagerank() copies in)
        c1 = graphs.c1(oracle)
        grp.upload(c1["begin"], None, c1["r_begin"], c1["r_node_idx"])
        rank, info = grp.pagerank(1e-30, 0.85, 20)
        want, _, _ = oracle.pagerank(c1["begin"], c1["r_begin"], c1["r_node_idx"], 1e-30, 0.85, 20)
        assert np.abs(rank - want).max() <= 1e-15
    finally:
        grp.close()
