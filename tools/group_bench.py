"""Experiment driver: the e2e step through the device group (gmc_group_*) on N GPUs of the box, phase by phase.
usage: python tools/group_bench.py <ndev> [nodes edges]      (GMC_PR_CALL_TIMING=1 / GMC_PR_PLAN_TIMING=1 print the inner phases)"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from gmb200_loader import load_package
L = load_package().GmCuda(); L.init(0)
ndev = int(sys.argv[1]) if len(sys.argv) > 1 else 2
n, m = (int(sys.argv[2]), int(sys.argv[3])) if len(sys.argv) > 3 else (10400000, 367500000)
g = L.generate_rmat(n, m, 1997)
pin = {}
for k in ("begin", "r_begin", "r_node_idx"):
    a = g.array(k)
    t = torch.empty(a.shape, dtype=torch.int32).pin_memory(); t.numpy()[:] = a; pin[k] = t.numpy()
one, info1 = g.pagerank(max_iter=20)
g.close()
rank_host = torch.empty(n, dtype=torch.float64).pin_memory().numpy()
grp = L.group(ndev)
for rep in range(4):
    t0 = time.perf_counter()
    grp.upload(pin["begin"], None, pin["r_begin"], pin["r_node_idx"])
    t1 = time.perf_counter()
    _, info = grp.pagerank(1e-30, 0.85, 20, out=rank_host)
    t2 = time.perf_counter()
    print(f"rep {rep}: upload {1e3*(t1-t0):.1f} ms  pagerank {1e3*(t2-t1):.1f} ms (device loop {info.kernel_ms:.2f} ms)  total {1e3*(t2-t0):.1f} ms = {m*20/(t2-t0)/1e9:.1f} GTEPS"
          f"  path {grp.exchange_path()}  max|diff vs 1 GPU| {np.abs(rank_host - one).max():.2e}", flush=True)
grp.close()
