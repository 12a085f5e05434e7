"""Experiment driver: per-sweep time of pr_sweep under the GMC_PR_MODE switches (LJ-shaped R-MAT)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from gmb200_loader import load_package
pkg = load_package(); L = pkg.GmCuda(); L.init(0)
g = L.generate_rmat(4850000, 69000000, 1997)
modes = [int(x) for x in sys.argv[1].split(",")] if len(sys.argv) > 1 else [0, 8, 16, 32, 24, 40, 0]
for mode in modes:
    os.environ["GMC_PR_MODE"] = str(mode)
    ms, _ = g.pagerank_time_sweeps(n=20)
    print("mode", mode, "ms/sweep min %.4f mean %.4f" % (ms.min(), ms.mean()), flush=True)
