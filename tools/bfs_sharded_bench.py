"""Experiment driver (torchrun, N ranks): sharded hop_dist with the small levels run redundantly by every rank (default) against
the round-1 form (GMC_BFS_NO_SMALL_SHARDED=1: every level sharded + exchanged), parity of both against the single-GPU traversal.
usage: python -m torch.distributed.run --nproc-per-node N tools/bfs_sharded_bench.py [nodes edges [roots]]"""
import os, sys
import numpy as np
import torch
import torch.distributed as dist
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from gmb200_loader import load_package  # noqa: E402

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ.get("LOCAL_RANK", "0"))
n, m = (int(sys.argv[1]), int(sys.argv[2])) if len(sys.argv) > 2 else (4850000, 69000000)
nroots = int(sys.argv[3]) if len(sys.argv) > 3 else 8
dist.init_process_group("gloo")
lib = load_package().GmCuda(); lib.init(local); torch.cuda.set_device(local)
uid = torch.from_numpy(lib.comm_unique_id() if rank == 0 else np.zeros(128, np.uint8)); dist.broadcast(uid, 0)
comm = lib.comm_create(world, rank, uid.numpy())
g = lib.generate_rmat(n, m, 1997)
deg = np.diff(g.array("begin")); nzv = np.nonzero(deg)[0]
rs = np.random.RandomState(1997)
roots = [int(nzv[0]), int(np.argmax(deg))] + [int(r) for r in rs.choice(nzv, size=max(0, nroots - 2), replace=False)]
d_dev = torch.empty(n, dtype=torch.int32, device="cuda")
want = {}
for r in roots:
    want[r], _ = g.hop_dist(r)
res = {}
for mode in ("sharded_all_levels", "small_redundant_allgather", "small_redundant_alltoall"):
    # sharded_all_levels = round-1 form; the other two run the small levels in full on every rank and differ in the exchange of
    # the top-down levels that stay sharded: allgather of `world` whole bitmaps / all-to-all of slices + OR + allgather of slices
    if mode == "sharded_all_levels": os.environ["GMC_BFS_NO_SMALL_SHARDED"] = "1"
    else: os.environ.pop("GMC_BFS_NO_SMALL_SHARDED", None)
    os.environ["GMC_BFS_A2A"] = "1" if mode == "small_redundant_alltoall" else "0"
    ms, edges, bad, launches, levels = [], [], 0, [], []
    for r in roots:
        g.hop_dist_sharded_dev(comm, d_dev.data_ptr(), r)
        dist.barrier()
        bi = g.hop_dist_sharded_dev(comm, d_dev.data_ptr(), r)
        t = torch.tensor([bi.kernel_ms], dtype=torch.float64); dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms.append(float(t.item())); edges.append(bi.edges_traversed); launches.append(bi.launches); levels.append(bi.levels)
        bad += int(not np.array_equal(d_dev.cpu().numpy(), want[r]))
        lev, _, tr, cnt = g.bfs_levels_sharded(comm, r)
        lev1, _, tr1, cnt1 = g.bfs_levels(r)
        bad += int(not (np.array_equal(lev, lev1) and list(tr) == list(tr1) and list(cnt) == list(cnt1)))
    res[mode] = (ms, edges, bad, launches, levels)
    if rank == 0:
        print(f"{mode:26s} world {world} n {n} m {m}: {np.mean(ms):.4f} ms per traversal, {sum(edges) / sum(ms) / 1e6:.1f} GTEPS, "
              f"mismatching {bad}, launches {launches}, levels {levels}, ms {[round(x, 3) for x in ms]}", flush=True)
t = torch.tensor([res[k][2] for k in res], dtype=torch.float64); dist.all_reduce(t)
assert float(t.sum().item()) == 0, "sharded BFS differs from the single-GPU traversal"
comm.close(); dist.destroy_process_group()
