"""Experiment driver: per-sweep time of the LJ-shaped graph for a list of plan settings (env switches of build_plan), one process each.
usage: python tools/pr_tiers_sweep.py ["TIERS,KMIN,SCALE[,COLD_WARPS[,SPLIT]]" ...]"""
import os, sys, subprocess
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if len(sys.argv) > 1 and sys.argv[1] == "one":
    sys.path.insert(0, ROOT)
    from gmb200_loader import load_package
    pkg = load_package(); L = pkg.GmCuda(os.environ.get("GMC_LIB") or None); L.init(0)
    n, m = int(os.environ.get("PR_N", 4850000)), int(os.environ.get("PR_M", 69000000))
    g = L.generate_rmat(n, m, 1997)
    L.sync(); import time; t0 = time.perf_counter()
    pi = g.pagerank_plan_info()
    L.sync(); plan_ms = (time.perf_counter() - t0) * 1e3
    ms, _ = g.pagerank_time_sweeps(n=20)
    rank, info = g.pagerank(max_iter=20)
    ab = 4 * m + 32 * pi["n_pos"] + 8 * pi["n_src"]
    print("%-28s ms/sweep min %.4f mean %.4f (%.0f GB/s alg) | 20 it %.3f ms | T %d segs %d split %d tierCTAs %d cold %.1fM hot0 %.1fM seg %.1fM cstream %.1fM plan %.1f ms" % (
        sys.argv[2], ms.min(), ms.mean(), ab / ms.mean() / 1e6, info.kernel_ms, pi["tiers"], pi["segments"], pi["split_rows"], pi["tier_ctas"],
        pi["cold_entries"] / 1e6, pi["hot0_entries"] / 1e6, pi["segment_entries"] / 1e6, pi["cold_stream_entries"] / 1e6, plan_ms), flush=True)
else:
    cfgs = sys.argv[1:] or ["16,3,1.0,12", "0,3,1.0", "16,4,1.0,12", "16,6,1.0,12", "16,3,1.0,8", "16,3,1.0,16", "16,3,1.0,20", "8,3,1.0,12", "24,3,1.0,12",
                            "16,3,1.3,12", "16,2,1.0,12", "31,4,1.0,12"]
    for c in cfgs:
        f = c.split(",")
        env = dict(os.environ, GMC_PR_TIERS=f[0], GMC_PR_KMIN=f[1], GMC_PR_TIER_SCALE=f[2])
        if len(f) >= 4: env["GMC_PR_COLD_WARPS"] = f[3]
        if len(f) >= 5: env["GMC_PR_SPLIT"] = f[4]
        if len(f) >= 6: env["GMC_PR_COLD_STREAM"] = f[5]
        if len(f) >= 7: env["GMC_PR_KCOLD"] = f[6]
        subprocess.run([sys.executable, __file__, "one", c], env=env)
