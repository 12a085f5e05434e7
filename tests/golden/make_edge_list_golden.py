"""Regenerates tests/golden/edge_list_small.{txt,npz}: a small EDGE_LIST text file and the CSR the REFERENCE's own
reader (gm_graph::load_edge_list, compiled from /root/reference into oracle/_ref/libgmref.so) builds from it.
Run in the build container (needs /root/reference):  python tests/golden/make_edge_list_golden.py"""
import os, sys
import numpy as np
HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE)); sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
from oracle_lib import Ref
from test_edge_list import write_edge_list

txt = os.path.join(HERE, "edge_list_small.txt")
write_edge_list(txt, 400, 1500, seed=1997)
ref = Ref()
g = ref.load_edge_list(txt)
a = ref.arrays(g)
np.savez_compressed(os.path.join(HERE, "edge_list_small.npz"), begin=a["begin"], node_idx=a["node_idx"],
                    r_begin=a["r_begin"], r_node_idx=a["r_node_idx"])
print("N", a["N"], "M", a["M"])
