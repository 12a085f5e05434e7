"""Profiling driver (under ncu via gpurun): triangle counting on a 1 M / 16 M R-MAT graph."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from gmb200_loader import load_package
L = load_package().GmCuda(); L.init(0)
g = L.generate_rmat(1000000, 16000000, 1997)
print(g.triangle_counting())
