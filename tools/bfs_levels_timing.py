"""Profiling aid (torchrun, N ranks): per-level host wall times of sharded hop_dist (GMC_BFS_LEVEL_TIMING=1) for a few roots."""
import os, sys
import numpy as np
import torch
import torch.distributed as dist
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from gmb200_loader import load_package  # noqa: E402
rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ.get("LOCAL_RANK", "0"))
n, m = (int(sys.argv[1]), int(sys.argv[2])) if len(sys.argv) > 2 else (41600000, 1470000000)
dist.init_process_group("gloo")
lib = load_package().GmCuda(); lib.init(local); torch.cuda.set_device(local)
uid = torch.from_numpy(lib.comm_unique_id() if rank == 0 else np.zeros(128, np.uint8)); dist.broadcast(uid, 0)
comm = lib.comm_create(world, rank, uid.numpy()) if world > 1 else None
g = lib.generate_rmat(n, m, 1997)
deg = np.diff(g.array("begin")); nzv = np.nonzero(deg)[0]
rs = np.random.RandomState(1997)
roots = [int(nzv[0]), int(np.argmax(deg))] + [int(r) for r in rs.choice(nzv, size=2, replace=False)]
d_dev = torch.empty(n, dtype=torch.int32, device="cuda")
for r in roots:
    run = (lambda: g.hop_dist_sharded_dev(comm, d_dev.data_ptr(), r)) if world > 1 else (lambda: g.hop_dist_dev(d_dev.data_ptr(), r))
    run()
    dist.barrier()
    os.environ["GMC_BFS_LEVEL_TIMING"] = "1"
    if rank == 0: print(f"--- root {r} (out-degree {deg[r]})", file=sys.stderr, flush=True)
    bi = run()
    os.environ.pop("GMC_BFS_LEVEL_TIMING")
    if rank == 0: print(f"    kernel_ms {bi.kernel_ms:.3f} levels {bi.levels} launches {bi.launches}", file=sys.stderr, flush=True)
if comm: comm.close()
dist.destroy_process_group()
