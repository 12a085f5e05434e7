"""Experiment: sweep time vs the long/short row threshold (GMC_PR_SMAX) on an R-MAT graph."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from gmb200_loader import load_package
L = load_package().GmCuda(); L.init(0)
n, m = (4850000, 69000000) if len(sys.argv) < 3 else (int(sys.argv[1]), int(sys.argv[2]))
want = None
for smax in [272, 384, 512, 768, 1024, 2048, 4096]:
    os.environ["GMC_PR_SMAX"] = str(smax)
    g = L.generate_rmat(n, m, 1997)
    pi = g.pagerank_plan_info()
    ms, _ = g.pagerank_time_sweeps(n=20)
    rank, info = g.pagerank(max_iter=5)
    if want is None: want = rank
    print(f"S_MAX {smax:5d}: long rows {pi['n_long']:8d} long entries {pi['long_entries']:10d} total {pi['total_entries']:10d} "
          f"ms/sweep min {ms.min():.4f} mean {ms.mean():.4f}  max|d| vs S_MAX=64 {np.abs(rank - want).max():.2e}", flush=True)
    g.close()
