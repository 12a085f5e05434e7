"""Profiling driver (under ncu via gpurun, -k regex:bc_): comp_BC (bc_adj.gm) from one seed on the LJ-shaped R-MAT graph."""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from gmb200_loader import load_package
L = load_package().GmCuda(); L.init(0)
g = L.generate_rmat(4850000, 69000000, 1997)
nz = np.nonzero(np.diff(g.array("begin")))[0]
bc, info = g.comp_bc([int(nz[len(nz) // 8])], adj=True)
print(info.device_ms, info.launches, info.max_levels)
