import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from gmb200_loader import load_package
pkg = load_package(); L = pkg.GmCuda(); L.init(0)
g = L.generate_rmat(4850000, 69000000, 1997)
for only in ["long", "short"]:
    os.environ["GMC_PR_ONLY"] = only
    for mode in [0, 1, 2, 4, 5]:
        os.environ["GMC_PR_MODE"] = str(mode)
        ms, _ = g.pagerank_time_sweeps(n=10)
        print(only, "mode", mode, "ms/sweep min %.4f mean %.4f" % (ms.min(), ms.mean()))
