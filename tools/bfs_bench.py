"""BFS timing on the LJ-shaped graph with bench.py's 16 roots: per root the best of 4 traversals (device time), then the mean over
roots and the GTEPS of the sums -- for A/B runs of the GMC_BFS_* switches (set them in the environment)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from gmb200_loader import load_package
pkg = load_package(); L = pkg.GmCuda(); L.init(0)
n, m = int(os.environ.get("PR_N", 4850000)), int(os.environ.get("PR_M", 69000000))
g = L.generate_rmat(n, m, 1997)
deg = np.diff(g.array("begin"))
rs = np.random.RandomState(1997)
roots = [int(np.nonzero(deg)[0][0])] + [int(r) for r in rs.choice(np.nonzero(deg)[0], size=15, replace=False)] + [int(deg.argmax())]
best, edges, lv = [], [], []
for r in roots:
    ms = []
    for rep in range(4):
        dist, bi = g.hop_dist(r)
        ms.append(bi.kernel_ms)
    best.append(min(ms)); edges.append(bi.edges_traversed); lv.append(bi.levels)
b16, e16 = best[:16], edges[:16]
sw = " ".join(k for k in os.environ if k.startswith("GMC_BFS")) or "default"
print(f"[{sw}] 16 roots: mean {np.mean(b16):.4f} ms, {sum(e16) / sum(b16) / 1e6:.1f} GTEPS; root3 {best[0]:.3f} ms ({lv[0]} lv); hub {best[16]:.3f} ms ({lv[16]} lv); "
      f"per root {' '.join('%.3f' % x for x in b16)}")
