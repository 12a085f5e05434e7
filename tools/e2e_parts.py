"""Experiment: where the e2e step spends its time (upload, plan build, sweeps, copy-back)."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from gmb200_loader import load_package
L = load_package().GmCuda(); L.init(0)
n, m = 4850000, 69000000
g = L.generate_rmat(n, m, 1997)
host = {k: g.array(k) for k in ("begin", "r_begin", "r_node_idx")}
pin = {}
for k, a in host.items():
    t = torch.empty(a.shape, dtype=torch.int32).pin_memory(); t.numpy()[:] = a; pin[k] = t.numpy()
rank_host = torch.empty(n, dtype=torch.float64).pin_memory().numpy()
for rep in range(4):
    t0 = time.perf_counter()
    dg = L.graph_pull_only(pin["begin"], pin["r_begin"], pin["r_node_idx"]); L.sync()
    t1 = time.perf_counter()
    pi = dg.pagerank_plan_info(); L.sync()
    t2 = time.perf_counter()
    _, info = dg.pagerank_into(rank_host, 1e-30, 0.85, 20)
    t3 = time.perf_counter()
    dg.close(); L.sync()
    t4 = time.perf_counter()
    print(f"rep {rep}: upload {1e3*(t1-t0):.1f} ms  plan {1e3*(t2-t1):.1f} ms  pagerank call {1e3*(t3-t2):.1f} ms (device {info.kernel_ms:.2f})  close {1e3*(t4-t3):.1f} ms", flush=True)
