"""CPU model (no GPU): how concentrated are pr_sweep's cold gathers in 32-byte sectors of contrib[] under the current
relabelling, and under a relabelling that breaks in-degree ties by out-degree.
Gathers of source w = outdeg(w).  Graph: the oracle's R-MAT generator at the LJ-shaped size (a proxy for the device
generator's graph: same parameters, different RNG)."""
import os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests"))
from oracle_lib import Oracle
n, m = (int(sys.argv[1]), int(sys.argv[2])) if len(sys.argv) > 2 else (4850000, 69000000)
S_MAX, HOT_CAP = 512, 16640
o = Oracle(); o.set_threads(os.cpu_count())
t0 = time.time()
begin, idx = o.rmat(n, m, 1997, 0.57, 0.19, 0.19, True)
outdeg = np.diff(begin).astype(np.int64)
indeg = np.bincount(idx, minlength=n).astype(np.int64)
print("graph in %.1f s; long rows %d" % (time.time() - t0, int((indeg > S_MAX).sum())))
is_long = indeg > S_MAX
long_ids = np.nonzero(is_long)[0]
hot = long_ids[np.argsort(-outdeg[long_ids], kind="stable")][:HOT_CAP - 2]
in_hot = np.zeros(n, bool); in_hot[hot] = True
print("hot block %d ids, %.1f %% of the gathers" % (len(hot), 100.0 * outdeg[hot].sum() / m))
rest = np.nonzero(~in_hot)[0]
cls = np.where(is_long[rest], 1, 2)
for name, keys in (("current (in-degree, ties in id order)", (rest, -indeg[rest], cls)),
                   ("ties by out-degree", (-outdeg[rest], -indeg[rest], cls))):
    order = rest[np.lexsort(keys)]                        # new ids hot_n + position
    new_id = np.empty(n, np.int64); new_id[order] = len(hot) + 1 + np.arange(len(order))
    sector = new_id[order] // 4
    g_per_sector = np.bincount(sector - sector.min(), weights=outdeg[order].astype(np.float64))
    g_sorted = np.sort(g_per_sector)[::-1]
    cold = g_sorted.sum()
    cum = np.cumsum(g_sorted)
    line = ", ".join("top %dK sectors %.1f %%" % (k // 1000, 100.0 * cum[min(k, len(cum)) - 1] / cold) for k in (4000, 16000, 64000, 256000))
    touched = int((g_per_sector > 0).sum())
    print("%-40s cold gathers %.1f M in %d sectors: %s" % (name, cold / 1e6, touched, line))
