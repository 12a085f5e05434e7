"""CPU model (no GPU) for DESIGN.md 4.1 "CTA specialisation by source tier": with the sources ranked by out-degree and cut
into tiers of H ids (one shared-memory block each), how many of pr_sweep's gathers stay cold if only the entries of rows
with in-degree > S are split by tier (the short rows keep their strict left-to-right order and see tier 0 only).
Graph: the oracle's R-MAT generator at the LJ-shaped size (proxy for the device generator's graph)."""
import os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests"))
from oracle_lib import Oracle
n, m = (int(sys.argv[1]), int(sys.argv[2])) if len(sys.argv) > 2 else (4850000, 69000000)
H = 16638
o = Oracle(); o.set_threads(os.cpu_count())
begin, idx = o.rmat(n, m, 1997, 0.57, 0.19, 0.19, True)
outdeg = np.diff(begin).astype(np.int64)
indeg = np.bincount(idx, minlength=n).astype(np.int64)
rank = np.empty(n, np.int64); rank[np.argsort(-outdeg, kind="stable")] = np.arange(n)     # 0 = largest out-degree
tier_of_src = rank // H
src_tier = np.repeat(tier_of_src, outdeg)                 # per edge slot (CSR order): tier of its source
dst_indeg = indeg[idx]                                    # per edge slot: in-degree of its destination
print("edges %d; sources ranked by out-degree, tier = rank // %d" % (m, H))
for S in (512, 256, 128, 64, 32, 16):
    long_e = dst_indeg > S
    n_long_rows = int((indeg > S).sum())
    le = int(long_e.sum())
    t_long = np.bincount(np.minimum(src_tier[long_e], 63), minlength=64)
    hot0_short = int((src_tier[~long_e] == 0).sum())
    row = ["S=%4d: long rows %7d hold %4.1f %% of the entries" % (S, n_long_rows, 100.0 * le / m)]
    for T in (1, 2, 4, 8, 16):
        cold = m - hot0_short - int(t_long[:T].sum())
        row.append("T=%d cold %.1f M" % (T, cold / 1e6))
    print("; ".join(row))

# ---- SM split for today's row classes (S = 512) and T tiers: group 0 keeps tier 0 and takes the short rows, the long rows' tier-0
# entries and a share of the long rows' cold entries; groups 1..T-1 take their tier's entries and the rest of the cold ones.  Rates per
# SM and clock measured on this B200: 0.95 L1-missing gathers, 3.1 shared-memory gathers (they overlap: the slower side counts).
S, MISS, LDS, SMS, CLK = 512, 0.95, 3.1, 148, 1.965e9
long_e = dst_indeg > S
t_long = np.bincount(np.minimum(src_tier[long_e], 63), minlength=64).astype(np.float64)
short_hot = float((src_tier[~long_e] == 0).sum()); short_cold = float((~long_e).sum()) - short_hot
for T in (1, 2, 4, 8, 16):
    long_cold = float(long_e.sum()) - t_long[:T].sum()
    total_miss = short_cold + long_cold
    # misses are spread over all SMs; an SM of group t also does its tier's hits.  Find the split where every SM finishes together.
    hits = [short_hot + t_long[0]] + [t_long[t] for t in range(1, T)]
    lo, hi = 0.0, 1.0
    for _ in range(60):                                   # bisection on the sweep time (seconds)
        tm = (lo + hi) / 2
        # SMs needed by group t if it runs for tm: its hits at LDS rate, misses fill the rest of the miss path
        need = 0.0
        miss_cap = 0.0
        for h in hits:
            sms = max(h / (LDS * CLK * tm), 1e-9)         # SMs kept busy by the hits alone
            need += sms
        # all SMs serve misses concurrently with hits (different pipes): the miss path needs total_miss / (MISS * CLK * tm) SMs
        need = max(need, total_miss / (MISS * CLK * tm))
        if need > SMS: lo = tm
        else: hi = tm
    groups = [h / (LDS * CLK * hi) for h in hits]
    print("T=%2d: cold %.1f M (short rows %.1f M, long rows %.1f M); ideal sweep %.3f ms; SMs busy with hits per group: %s"
          % (T, total_miss / 1e6, short_cold / 1e6, long_cold / 1e6, hi * 1e3, " ".join("%.1f" % x for x in groups)))
