"""Profiling driver (used under ncu via gpurun): LJ-shaped R-MAT, a few PageRank iterations / one BFS."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from gmb200_loader import load_package
pkg = load_package(); L = pkg.GmCuda(); L.init(0)
what = sys.argv[1] if len(sys.argv) > 1 else "pagerank"
n, m = int(os.environ.get("PR_N", 4850000)), int(os.environ.get("PR_M", 69000000))
g = L.generate_rmat(n, m, 1997)
if what == "pagerank":
    g.pagerank_plan_info()          # the tiered plan of a resident graph (a first short call alone would build the plan without tiers)
    rank, info = g.pagerank(max_iter=int(sys.argv[2]) if len(sys.argv) > 2 else 4)
    print("pagerank ms/it", info.kernel_ms / info.iterations)
else:
    deg = np.diff(g.array("begin"))
    root = int(np.nonzero(deg)[0][0])
    dist, bi = g.hop_dist(root)
    print("bfs ms", bi.kernel_ms, "levels", bi.levels)
