"""Experiment: sweep time vs the share / size of the small short tasks at the tail (GMC_PR_TAIL="<sixteenths big>,<shift>")."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from gmb200_loader import load_package
L = load_package().GmCuda(); L.init(0)
for tail in ["16,2", "14,2", "12,2", "14,3", "12,3", "10,3", "12,4", "8,2", "14,2"]:
    os.environ["GMC_PR_TAIL"] = tail
    g = L.generate_rmat(4850000, 69000000, 1997)
    ms, _ = g.pagerank_time_sweeps(n=20)
    print("tail", tail, "ms/sweep min %.4f mean %.4f" % (ms.min(), ms.mean()), flush=True)
    g.close()
