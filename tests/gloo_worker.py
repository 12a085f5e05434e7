"""World-size-2 CPU worker (gloo): the host-side plumbing of bench_multi.py without a GPU -- id broadcast shape,
identical shard cuts on every rank, max/sum-over-ranks reductions."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench_multi  # noqa: E402
from gmb200_loader import load_package  # noqa: E402


def main():
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    dist.init_process_group("gloo")
    lib = load_package().GmCuda()
    # the id travels as 128 raw bytes from rank 0 (here a stand-in pattern: NCCL itself needs a GPU)
    uid = torch.from_numpy(np.arange(128, dtype=np.uint8) if rank == 0 else np.zeros(128, np.uint8))
    dist.broadcast(uid, 0)
    assert np.array_equal(uid.numpy(), np.arange(128, dtype=np.uint8))
    # world == 1 communicators need no NCCL and no GPU
    c = lib.comm_create(1, 0, None)
    assert (c.world, c.rank) == (1, 0)
    c.close()
    # every rank computes the cuts of every rank from the same prefix array: they must agree and tile the range
    rs = np.random.RandomState(11)
    prefix = np.concatenate([[0], np.cumsum(rs.randint(0, 50, 10000))]).astype(np.int32)
    cuts = lib.shard_cuts(prefix, world)
    t = torch.from_numpy(cuts.astype(np.int64)); dist.broadcast(t, 0)
    assert np.array_equal(t.numpy(), cuts)
    mine = prefix[cuts[rank + 1]] - prefix[cuts[rank]]
    assert bench_multi.sum_over_ranks(dist, torch, mine) == float(prefix[-1])
    assert bench_multi.max_over_ranks(dist, torch, 10.0 + rank) == 10.0 + world - 1
    dist.barrier()
    if rank == 0:
        print("gloo_worker ok", flush=True)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
