"""Experiment driver: per-sweep time of every green-marl_b200/lib/libgm_cuda*.so variant (tools/build_variant.sh)."""
import glob, os, sys, subprocess
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if len(sys.argv) > 1 and sys.argv[1] == "one":
    sys.path.insert(0, ROOT)
    from gmb200_loader import load_package
    pkg = load_package(); L = pkg.GmCuda(sys.argv[2]); L.init(0)
    g = L.generate_rmat(4850000, 69000000, 1997)
    ms, _ = g.pagerank_time_sweeps(n=20)
    rank, info = g.pagerank(max_iter=20)
    print(os.path.basename(sys.argv[2]), "ms/sweep min %.4f mean %.4f | pagerank() 20 it %.3f ms" % (ms.min(), ms.mean(), info.kernel_ms), flush=True)
else:
    for so in sorted(glob.glob(os.path.join(ROOT, "green-marl_b200", "lib", "libgm_cuda*.so"))):
        subprocess.run([sys.executable, __file__, "one", so])
