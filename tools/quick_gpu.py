"""Scratch timing driver for gpurun (not part of the product): LJ-shaped PageRank / BFS timings."""
import sys, time, json
import numpy as np
import os; sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from gmb200_loader import load_package
pkg = load_package(); L = pkg.GmCuda(); L.init(0)
n, m = (4850000, 69000000) if len(sys.argv) < 3 else (int(sys.argv[1]), int(sys.argv[2]))
t = time.time(); g = L.generate_rmat(n, m, 1997); print("gen s", time.time() - t)
rb = g.array("r_begin"); print("max indeg", np.diff(rb).max(), "zero indeg frac", (np.diff(rb) == 0).mean())
for rep in range(3):
    rank, info = g.pagerank(max_iter=20)
    A = 4 * m + 40 * n
    print(f"pagerank: {info.kernel_ms:.3f} ms / {info.iterations} it, {info.kernel_ms/info.iterations:.4f} ms/it, "
          f"GTEPS {m*info.iterations/info.kernel_ms/1e6:.1f}, alg GB/s {A*info.iterations/info.kernel_ms/1e6:.1f}, launches {info.launches}")
b = g.array("begin"); deg = np.diff(b)
for root in [0, int(np.nonzero(deg)[0][0]), int(deg.argmax())]:
    for rep in range(2):
        dist, bi = g.hop_dist(root)
    lev, li, trace, cnt = g.bfs_levels(root)
    print(f"bfs root {root}: {bi.kernel_ms:.3f} ms levels {bi.levels} reached {bi.reached} edges {bi.edges_traversed} "
          f"GTEPS {bi.edges_traversed/max(bi.kernel_ms,1e-6)/1e6:.2f} bu {bi.bottom_up_levels} trace {list(trace)} cnt {list(cnt)}")
