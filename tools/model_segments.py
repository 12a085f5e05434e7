"""CPU model (no GPU): (row, source-tier) segments of the tiered sweep.  Sources ranked by out-degree; tier 0 = the H0 hottest
(resident on every home SM), tiers 1..T of HT ids each live on dedicated SMs.  A (row, tier>=1) group of k entries becomes a
segment (partial sum written once, read once) when k >= KMIN, otherwise its entries stay cold gathers of the row's home list.
Prints cold gathers, segment entries and segment counts per (T, HT, KMIN)."""
import os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests"))
from oracle_lib import Oracle
n, m = (int(sys.argv[1]), int(sys.argv[2])) if len(sys.argv) > 2 else (4850000, 69000000)
H0 = 16638
o = Oracle(); o.set_threads(os.cpu_count())
t0 = time.time()
begin, idx = o.rmat(n, m, 1997, 0.57, 0.19, 0.19, True)
print("graph %.1f s" % (time.time() - t0))
outdeg = np.diff(begin).astype(np.int64)
indeg = np.bincount(idx, minlength=n).astype(np.int64)
print("n_pos %d n_src %d" % ((indeg > 0).sum(), (outdeg > 0).sum()))
zin = indeg == 0
print("edges from zero-in-degree sources: %.2f M" % (outdeg[zin].sum() / 1e6))
rank = np.empty(n, np.int64); rank[np.argsort(-outdeg, kind="stable")] = np.arange(n)
src_rank = np.repeat(rank, outdeg).astype(np.int32)       # per edge: rank of source
dst = idx.astype(np.int32)
hot0 = src_rank < H0
print("tier-0 entries %.1f M (%.1f %%)" % (hot0.sum() / 1e6, 100.0 * hot0.sum() / m))
cs = np.sort(outdeg)[::-1].cumsum()
for k in (16638, 50000, 100000, 200000, 400000, 800000):
    print(" top %7d sources: %.1f %% of edges" % (k, 100.0 * cs[k - 1] / m))
rest_rank = src_rank[~hot0] - H0
rest_dst = dst[~hot0]
for HT in (16638, 27000):
    tier = rest_rank // HT
    for T in (4, 8, 12, 16, 24):
        sel = tier < T
        key = rest_dst[sel].astype(np.int64) * 64 + tier[sel]
        u, c = np.unique(key, return_counts=True)
        tot = int(sel.sum())
        line = "HT=%d T=%2d (sources to rank %d): tier entries %.1f M;" % (HT, T, H0 + T * HT, tot / 1e6)
        for kmin in (1, 2, 3, 4, 6, 8):
            ok = c >= kmin
            line += " k>=%d: ent %.1f M seg %.2f M cold %.1f M |" % (kmin, c[ok].sum() / 1e6, ok.sum() / 1e6, ((~hot0).sum() - c[ok].sum()) / 1e6)
        print(line)
