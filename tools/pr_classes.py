"""Experiment: time the sweep per task class (GMC_PR_ONLY) on the LJ-shaped graph."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from gmb200_loader import load_package
pkg = load_package(); L = pkg.GmCuda(); L.init(0)
n, m = (4850000, 69000000) if len(sys.argv) < 3 else (int(sys.argv[1]), int(sys.argv[2]))
g = L.generate_rmat(n, m, 1997)
print(g.pagerank_plan_info())
for only in [None, "long", "short", "zero", "none"]:
    if only: os.environ["GMC_PR_ONLY"] = only
    ms, _ = g.pagerank_time_sweeps(n=10)
    print(only, "ms/sweep min %.4f mean %.4f" % (ms.min(), ms.mean()))
