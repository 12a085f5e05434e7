"""Profiling driver (under ncu --profile-from-start off, via gpurun): one hop_dist traversal of the LJ-shaped graph from the
first root with edges ("quiet") or from the vertex of largest out-degree ("hub"), after two warm traversals."""
import ctypes, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from gmb200_loader import load_package
pkg = load_package(); L = pkg.GmCuda(); L.init(0)
n, m = int(os.environ.get("PR_N", 4850000)), int(os.environ.get("PR_M", 69000000))
g = L.generate_rmat(n, m, 1997)
deg = np.diff(g.array("begin"))
root = int(deg.argmax()) if len(sys.argv) > 1 and sys.argv[1] == "hub" else int(np.nonzero(deg)[0][0])
for _ in range(2): g.hop_dist(root)
rt = None
for name in ("libcudart.so.12", "libcudart.so"):
    try: rt = ctypes.CDLL(name); break
    except OSError: pass
if rt: rt.cudaProfilerStart()
dist, bi = g.hop_dist(root)
if rt: rt.cudaProfilerStop()
lev, li, trace, cnt = g.bfs_levels(root)
print("root", root, "ms", bi.kernel_ms, "levels", bi.levels, "launches", bi.launches, "trace", [int(x) for x in trace], "cnt", [int(x) for x in cnt])
