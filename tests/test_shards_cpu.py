"""CPU suite for the host-side logic of the sharded paths: the balanced range partition (gmc_shard_cuts) and
the rank bootstrap / timing reduction of bench_multi.py over a world_size-2 gloo group (no GPU)."""
import os
import subprocess
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def lib():
    from gmb200_loader import load_package
    return load_package().GmCuda()


def test_shard_cuts_cover_and_balance():
    L = lib()
    rs = np.random.RandomState(5)
    w = rs.zipf(1.6, 100000).clip(max=200000).astype(np.int64)
    prefix = np.concatenate([[0], np.cumsum(w)]).astype(np.int32)
    for world in (1, 2, 3, 8):
        cuts = L.shard_cuts(prefix, world)
        assert cuts[0] == 0 and cuts[-1] == len(w) and np.all(np.diff(cuts) >= 0)
        loads = np.diff(prefix[cuts].astype(np.int64))
        assert loads.sum() == prefix[-1]
        assert loads.max() <= prefix[-1] / world + w.max()      # never worse than one row over the ideal share


def test_shard_cuts_alignment_and_degenerate():
    L = lib()
    prefix = np.arange(0, 1001, dtype=np.int32)
    cuts = L.shard_cuts(prefix, 4, align=32)
    assert list(cuts) == [0, 224, 480, 736, 1000] and all(c % 32 == 0 for c in cuts[:-1])
    assert list(L.shard_cuts(np.zeros(1, np.int32), 4)) == [0, 0, 0, 0, 0]          # empty array
    assert list(L.shard_cuts(np.array([0, 0, 0, 7], np.int32), 2)) == [0, 3, 3]      # all weight in the last row: one rank gets it
    cuts = L.shard_cuts(np.array([0, 5, 5, 5, 5], np.int32), 8)                     # more ranks than rows
    assert cuts[0] == 0 and cuts[-1] == 4 and np.all(np.diff(cuts) >= 0)


def test_gloo_world2_bootstrap_and_timing():
    """bench_multi's host plumbing: unique-id broadcast, shard agreement, max-over-ranks time, on CPU"""
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr", "127.0.0.1",
           "--master-port", "29533", os.path.join(ROOT, "tests", "gloo_worker.py")]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=300, env={**os.environ, "CUDA_VISIBLE_DEVICES": ""})
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    assert "gloo_worker ok" in r.stdout
