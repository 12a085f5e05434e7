"""Experiment driver: hop_dist time per root for every green-marl_b200/lib/libgm_cuda*.so present."""
import glob, os, sys, subprocess
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if len(sys.argv) > 2 and sys.argv[1] == "one":
    sys.path.insert(0, ROOT)
    import numpy as np
    from gmb200_loader import load_package
    L = load_package().GmCuda(sys.argv[2]); L.init(0)
    g = L.generate_rmat(4850000, 69000000, 1997)
    deg = np.diff(g.array("begin"))
    rs = np.random.RandomState(1997)
    roots = [int(np.nonzero(deg)[0][0]), int(deg.argmax())] + [int(r) for r in rs.choice(np.nonzero(deg)[0], size=6, replace=False)]
    ms = []
    for r in roots:
        g.hop_dist(r)
        ms.append(min(g.hop_dist(r)[1].kernel_ms for _ in range(3)))
    print(os.path.basename(sys.argv[2]), "ms per root", [round(x, 3) for x in ms], "mean %.4f" % (sum(ms) / len(ms)), flush=True)
else:
    for so in sorted(glob.glob(os.path.join(ROOT, "green-marl_b200", "lib", "libgm_cuda*.so"))):
        subprocess.run([sys.executable, __file__, "one", so])
