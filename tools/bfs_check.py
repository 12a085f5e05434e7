"""Debug aid: repeatability + BFS invariants of hop_dist on a device-generated graph (single GPU)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from gmb200_loader import load_package
L = load_package().GmCuda(); L.init(0)
n, m = (300000, 4000000) if len(sys.argv) < 3 else (int(sys.argv[1]), int(sys.argv[2]))
g = L.generate_rmat(n, m, 1997)
b, idx = g.array("begin"), g.array("node_idx")
src = np.repeat(np.arange(n, dtype=np.int32), np.diff(b))
INF = 2**31 - 1
for root in [int(np.nonzero(np.diff(b))[0][0]), int(np.argmax(np.diff(b)))]:
    res = []
    for rep in range(6):
        d, info = g.hop_dist(root)
        res.append(d.copy())
        du, dv = d[src].astype(np.int64), d[idx].astype(np.int64)
        viol = np.nonzero((du < INF) & (dv > du + 1))[0]
        # every reached vertex (not root) needs a parent at dist-1
        best = np.full(n, INF, np.int64)
        np.minimum.at(best, idx, np.where(du < INF, du + 1, INF))
        orphan = np.nonzero((d < INF) & (np.arange(n) != root) & (best != d))[0]
        print(f"root {root} rep {rep}: levels {info.levels} reached {info.reached} edge-violations {len(viol)} orphans {len(orphan)}"
              + (f" e.g. edge {src[viol[0]]}->{idx[viol[0]]} d {du[viol[0]]},{dv[viol[0]]}" if len(viol) else "")
              + (f" orphan {orphan[:4]} d {d[orphan[:4]]} best {best[orphan[:4]]}" if len(orphan) else ""))
    for r in res[1:]:
        if not np.array_equal(r, res[0]):
            bad = np.nonzero(r != res[0])[0]
            print("  runs differ at", bad[:8], res[0][bad[:8]], r[bad[:8]])
