"""Sharded paths.  On one GPU: the shard plans of a `world`-rank PageRank are swept one after
the other on the same device (gmc_pagerank_emulated_shards_dev) -- same plan code, same kernels, same
combination order as a real multi-rank run -- and must reproduce the single-shard result.  With >= 2 GPUs on
the box: a real 2-process NCCL run (tests/multi_gpu_worker.py under torch.distributed.run)."""
import os
import subprocess
import sys

import numpy as np
import pytest

import graphs

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def dev_graph(gm, g):
    dg = gm.graph_from_csr(g["begin"], g["node_idx"], semi_sort=False, make_reverse=False)
    dg.upload_reverse(g["r_begin"], g["r_node_idx"])
    return dg


@pytest.fixture(scope="module")
def g_rmat(oracle):
    return graphs.rmat(oracle, 200000, 3000000)


@pytest.mark.parametrize("world", [2, 3, 8])
def test_pagerank_emulated_shards_match_single(gm, oracle, g_rmat, world):
    dg = dev_graph(gm, g_rmat)
    one, info1 = dg.pagerank(e=1e-30, d=0.85, max_iter=20)
    many, infow = dg.pagerank_emulated_shards(world, e=1e-30, d=0.85, max_iter=20)
    assert infow.iterations == info1.iterations == 20
    want, _, _ = oracle.pagerank(g_rmat["begin"], g_rmat["r_begin"], g_rmat["r_node_idx"], 1e-30, 0.85, 20)
    assert np.abs(many - want).max() <= 1e-9
    # shards cut long rows only at row boundaries, but the task cuts inside a shard differ from the unsharded
    # plan, so rows longer than a task may round differently in the last place (and feed their neighbours)
    assert graphs.max_ulp(many, one) <= 8


def test_pagerank_emulated_shards_c1_bit_exact(gm, oracle):
    g = graphs.c1(oracle)
    dg = dev_graph(gm, g)
    want, it, _ = oracle.pagerank(g["begin"], g["r_begin"], g["r_node_idx"], 1e-30, 0.85, 20)
    for world in (2, 5):
        got, info = dg.pagerank_emulated_shards(world, e=1e-30, d=0.85, max_iter=20)
        assert info.iterations == 20 and np.array_equal(got, want)


def test_pagerank_emulated_shards_convergence_exit(gm, oracle):
    g = graphs.c1(oracle)
    dg = dev_graph(gm, g)
    want, it, _ = oracle.pagerank(g["begin"], g["r_begin"], g["r_node_idx"], 0.001, 0.85, 100)
    got, info = dg.pagerank_emulated_shards(4, e=0.001, d=0.85, max_iter=100)
    assert info.iterations == it and np.array_equal(got, want)


def test_more_shards_than_rows(gm, oracle):
    g = graphs.from_edges(oracle, 5, [(0, 1), (1, 2), (2, 0), (3, 0)])
    dg = dev_graph(gm, g)
    want, _, _ = oracle.pagerank(g["begin"], g["r_begin"], g["r_node_idx"], 1e-30, 0.85, 10)
    got, _ = dg.pagerank_emulated_shards(8, e=1e-30, d=0.85, max_iter=10)
    assert np.array_equal(got, want)


@pytest.mark.parametrize("bfs_peer", ["0", "1"])
def test_two_rank_nccl_run(gm, bfs_peer):
    """real multi-process run: PageRank + hop_dist + bfs levels on 2 GPUs equal the 1-GPU results; bfs_peer = 1: the
    owner-directed exchange of the top-down levels (GMC_BFS_PEER, remote atomics into the owner's candidate slice)"""
    if gm.device_count() < 2:
        pytest.skip("box has one GPU: the 2-rank NCCL run needs two")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr", "127.0.0.1",
           "--master-port", "2954" + bfs_peer, os.path.join(ROOT, "tests", "multi_gpu_worker.py")]
    env = dict(os.environ)
    if bfs_peer == "1":
        env["GMC_BFS_PEER"] = "1"
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600, env=env)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    assert "multi_gpu_worker ok" in r.stdout
