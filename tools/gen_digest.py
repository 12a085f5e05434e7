"""Writes the digest of the device R-MAT generator's output (gmc_graph_generate_rmat(2^16, 2^20, seed 1997): sha-256 of begin[] and of the
semi-sorted node_idx[]) -- the fixture tests/golden/device_rmat_digest.json pins every benchmarked graph to.  Run on a GPU box:
    python tools/gen_digest.py > gpurun_out/device_rmat_digest.json"""
import hashlib, json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from gmb200_loader import load_package
L = load_package().GmCuda(); L.init(0)
out = {}
for name, (n, m, seed) in {"rmat_2e16_2e20_s1997": (1 << 16, 1 << 20, 1997), "rmat_200k_3m_s7": (200000, 3000000, 7)}.items():
    g = L.generate_rmat(n, m, seed)
    b, i = g.array("begin"), g.array("node_idx")
    out[name] = {"n": n, "m": m, "seed": seed, "a": 0.57, "b": 0.19, "c": 0.19, "permute": True,
                 "sha256_begin": hashlib.sha256(np.ascontiguousarray(b).tobytes()).hexdigest(),
                 "sha256_node_idx": hashlib.sha256(np.ascontiguousarray(i).tobytes()).hexdigest(),
                 "max_outdeg": int(np.diff(b).max()), "zero_outdeg": int((np.diff(b) == 0).sum())}
print(json.dumps(out, indent=1))
