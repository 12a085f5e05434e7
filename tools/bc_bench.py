"""comp_BC timing: gmc_comp_bc_adj on R-MAT graphs, per seed; the oracle port (one host thread; the
reference's own multi-thread run races on the down-edge marks) on the smaller graph.  Prints one JSON line."""
import json, os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
from gmb200_loader import load_package
from oracle_lib import Oracle
L = load_package().GmCuda(); L.init(0)
out = {"what": "comp_BC (bc_adj.gm)", "runs": []}
for n, m in [(1000000, 16000000), (4850000, 69000000)]:
    g = L.generate_rmat(n, m, 1997)
    b = g.array("begin")
    nz = np.nonzero(np.diff(b))[0]
    seeds = [int(nz[k * (len(nz) // 8)]) for k in range(8)]
    g.comp_bc(seeds[:1], adj=True)
    bc, info = g.comp_bc(seeds, adj=True)
    bc2, info2 = g.comp_bc(seeds, adj=True)
    ms = min(info.device_ms, info2.device_ms)
    out["runs"].append({"nodes": n, "edges": m, "seeds": len(seeds), "device_ms": round(ms, 3), "ms_per_seed": round(ms / len(seeds), 3),
                        "launches": info.launches, "max_levels": info.max_levels, "repeatable": bool(np.array_equal(bc, bc2, equal_nan=True)),
                        "bc_max": float(np.nanmax(bc)), "GTEPS_2M_per_seed": round(2 * m / (ms / len(seeds)) / 1e6, 1)})
    if n == 4850000:      # where the time goes: the traversals alone (save_child state machine), then the knobs of the two callbacks' kernels
        out["bfs_ms_8_seeds"] = round(sum(g.bfs_levels(sd, use_reverse_edge=2)[1].kernel_ms for sd in seeds), 3)
        out["knobs"] = []
        for grp, heavy in [(8, 1024), (32, 1024), (16, 1024), (8, 256), (8, 4096), (32, 256)]:
            os.environ["GMC_BC_GROUP"] = str(grp); os.environ["GMC_BC_HEAVY"] = str(heavy)
            t = min(g.comp_bc(seeds, adj=True)[1].device_ms for _ in range(2))
            got = g.comp_bc(seeds, adj=True)[0]
            out["knobs"].append({"group": grp, "heavy_min": heavy, "device_ms": round(t, 3), "same_bits": bool(np.array_equal(got, bc, equal_nan=True))})
        os.environ.pop("GMC_BC_GROUP"); os.environ.pop("GMC_BC_HEAVY")
    if n == 1000000:
        o = Oracle()
        a = {k: g.array(k) for k in ("begin", "node_idx", "r_begin", "r_node_idx")}
        t0 = time.perf_counter(); want = o.comp_bc(a["begin"], a["node_idx"], a["r_begin"], a["r_node_idx"], seeds[:2], adj=True); dt = time.perf_counter() - t0
        got, _ = g.comp_bc(seeds[:2], adj=True)
        same = bool(np.array_equal(np.isnan(got), np.isnan(want)) and np.array_equal(got[~np.isnan(got)].view(np.uint32), want[~np.isnan(want)].view(np.uint32)))
        out["cpu_port"] = {"nodes": n, "edges": m, "seeds": 2, "seconds_per_seed": round(dt / 2, 3), "threads": 1, "identical": same}
    g.close()
print(json.dumps(out))
