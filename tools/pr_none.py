import os, sys
sys.path.insert(0, "/root/repo")
from gmb200_loader import load_package
pkg = load_package(); L = pkg.GmCuda(); L.init(0)
g = L.generate_rmat(4850000, 69000000, 1997)
os.environ["GMC_PR_ONLY"] = "none"
for mode in ["0", "2", "0", "2"]:
    os.environ["GMC_PR_MODE"] = mode
    ms, _ = g.pagerank_time_sweeps(n=20)
    print("none mode", mode, "min %.4f mean %.4f" % (ms.min(), ms.mean()))
