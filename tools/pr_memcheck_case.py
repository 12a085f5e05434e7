"""compute-sanitizer target: a small tiered PageRank (tiers forced on a 2^16 / 1 M R-MAT), 3 sweeps, single + 2 emulated shards."""
import os, sys
os.environ.setdefault("GMC_PR_TIER_MIN_EDGES", "0")
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from gmb200_loader import load_package
pkg = load_package(); L = pkg.GmCuda(); L.init(0)
g = L.generate_rmat(1 << 16, 1 << 20, 5)
print(g.pagerank_plan_info())
rank, info = g.pagerank(max_iter=3)
print("single", info.iterations, float(rank.sum()))
rank2, info2 = g.pagerank_emulated_shards(2, max_iter=3)
print("2 shards", info2.iterations, float(abs(rank - rank2).max()))
