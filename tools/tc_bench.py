"""triangle_counting timing: gmc_triangle_counting on R-MAT graphs; the quadratic reference loop nest
(oracle port, all host threads) on a bounded smaller graph of the same shape.  Prints one JSON line."""
import json, os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
from gmb200_loader import load_package
from oracle_lib import Oracle
L = load_package().GmCuda(); L.init(0)
out = {"what": "triangle_counting", "runs": []}
for n, m in [(1 << 16, 1 << 20), (1000000, 16000000), (4850000, 69000000)]:
    g = L.generate_rmat(n, m, 1997)
    t, ms = g.triangle_counting(); t2, ms2 = g.triangle_counting()
    assert t == t2
    out["runs"].append({"nodes": n, "edges": m, "T": t, "device_ms": round(min(ms, ms2), 3), "Medge_slots_per_s": round(m / min(ms, ms2) / 1e3, 1)})
    if n == 1 << 16:
        o = Oracle(); o.set_threads(os.cpu_count())
        b, i = g.array("begin"), g.array("node_idx")
        t0 = time.perf_counter(); want = o.triangle_counting(b, i); dt = time.perf_counter() - t0
        out["cpu_port"] = {"nodes": n, "edges": m, "T": want, "seconds": round(dt, 3), "threads": os.cpu_count(), "identical": want == t}
    g.close()
print(json.dumps(out))
