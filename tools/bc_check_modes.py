"""Quick check (seconds): gmc_comp_bc / _adj against the oracle in the default and in the GMC_BC_OVERLAP=1 mode, then both
modes timed on the LJ-shaped graph."""
import os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
from gmb200_loader import load_package
from oracle_lib import Oracle
import graphs
def same(a, b):
    return np.array_equal(np.isnan(a), np.isnan(b)) and np.array_equal(a[~np.isnan(a)].view(np.uint32), b[~np.isnan(b)].view(np.uint32))
L = load_package().GmCuda(); L.init(0)
o = Oracle(); o.set_threads(os.cpu_count())
g = graphs.rmat(o, 200000, 3000000, seed=12)
deg = np.diff(g["begin"]); nz = np.nonzero(deg)[0]
seeds = [int(nz[0]), int(deg.argmax()), int(np.diff(g["r_begin"]).argmax()), 1]
want = {adj: o.comp_bc(g["begin"], g["node_idx"], g["r_begin"], g["r_node_idx"], seeds, adj=adj) for adj in (False, True)}
G = L.graph_from_csr(g["begin"], g["node_idx"], semi_sort=False, make_reverse=True)
for mode in ("0", "1"):
    os.environ["GMC_BC_OVERLAP"] = mode
    for adj in (False, True):
        got, info = G.comp_bc(seeds, adj=adj)
        print("overlap", mode, "adj", adj, "identical to the oracle:", same(got, want[adj]), "launches", info.launches, flush=True)
G.close()
D = L.generate_rmat(4850000, 69000000, 1997)
nz = np.nonzero(np.diff(D.array("begin")))[0]
seeds = [int(nz[k * (len(nz) // 8)]) for k in range(8)]
res = {}
for mode in ("0", "1", "0", "1"):
    os.environ["GMC_BC_OVERLAP"] = mode
    bc, info = D.comp_bc(seeds, adj=True)
    res.setdefault(mode, []).append((info.device_ms, bc))
    print("LJ-shaped overlap", mode, "ms per seed %.3f" % (info.device_ms / 8), flush=True)
print("LJ-shaped: same bits in both modes:", same(res["0"][0][1], res["1"][0][1]), flush=True)
