"""green-marl_b200: B200-native PageRank / BFS hot path of Green-Marl + Shoal.

The product is native: `lib/libgm_cuda.so` (hand-written sm_100a kernels behind the C ABI declared in
`include/gm_cuda.h`, plus the C++ facade of the Green-Marl runtime API in `include/gm/`).  This Python
package is a harness only: a ctypes binding used by tests/, bench.py and __graft_entry__.py.
The directory name contains a hyphen, so load it with `load_package()` from `gmb200_loader.py` at the
repo root (registers it as module `green_marl_b200`).
"""
from .gmcuda import GmCuda, GmcError, lib_path, Graph, Comm  # noqa: F401
