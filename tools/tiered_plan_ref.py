"""Executable specification (numpy, no GPU) of the tiered pr_sweep plan sketched in DESIGN.md 4.1, for the next round's
CUDA plan builder to be compared against: relabelling with T source tiers of H ids, the long section as T + 1 entry
streams, per-(stream, row) partial sums combined in stream order.  Runs 20 sweeps with that summation order and compares
with the oracle's PageRank (the tolerance of the path is 1e-9).

    python tools/tiered_plan_ref.py [log2_nodes] [log2_edges] [T] [H] [S]
"""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests"))
from oracle_lib import Oracle
import graphs


def build(begin, r_begin, r_node_idx, T, H, S):
    n = len(begin) - 1
    outdeg = np.diff(begin).astype(np.int64); indeg = np.diff(r_begin).astype(np.int64)
    by_out = np.argsort(-outdeg, kind="stable")
    n_tiered = min(T * H, n)
    tiered = by_out[:n_tiered]                                        # new ids 1 .. n_tiered, tier = (id - 1) // H
    is_tiered = np.zeros(n, bool); is_tiered[tiered] = True
    rest = np.nonzero(~is_tiered)[0]
    cls = np.where(indeg[rest] > S, 0, 1)                             # other long rows, then short rows, in-degree descending
    rest = rest[np.lexsort((rest, -indeg[rest], cls))]
    old_of_new = np.concatenate([[-1], tiered, rest]).astype(np.int64)   # id 0 = dummy (contribution 0)
    new_of_old = np.empty(n, np.int64); new_of_old[old_of_new[1:]] = np.arange(1, n + 1)
    n_long = n_tiered + int((indeg[rest] > S).sum())                  # long section = ids [1, n_long]
    # streams of the long section
    rows = old_of_new[1:n_long + 1]
    seg_len = np.zeros((T + 1, n_long), np.int64)
    src_new = new_of_old[r_node_idx]
    stream_of_edge = np.where(src_new <= n_tiered, (src_new - 1) // H, T)
    streams = []
    for s in range(T + 1):
        ent, off = [], [0]
        for j, v in enumerate(rows):
            e = np.arange(r_begin[v], r_begin[v + 1])
            sel = src_new[e][stream_of_edge[e] == s]                  # original source order inside a segment
            pad = (-len(sel)) % 4
            ent.append(np.concatenate([sel, np.zeros(pad, np.int64)]))
            off.append(off[-1] + len(sel) + pad)
            seg_len[s, j] = len(sel)
        streams.append((np.concatenate(ent) if ent else np.zeros(0, np.int64), np.array(off, np.int64)))
    return dict(n=n, n_long=n_long, n_tiered=n_tiered, old_of_new=old_of_new, new_of_old=new_of_old, streams=streams,
                seg_len=seg_len, outdeg=outdeg, indeg=indeg, src_new=src_new)


def sweep20(P, begin, r_begin, r_node_idx, d=0.85, iters=20):
    n, n_long = P["n"], P["n_long"]
    old_of_new = P["old_of_new"]
    outdeg_new = np.concatenate([[0], P["outdeg"][old_of_new[1:]]]).astype(np.float64)
    rank = np.full(n + 1, 1.0 / n); rank[0] = 0.0
    safe = np.where(outdeg_new > 0, outdeg_new, 1.0)
    for _ in range(iters):
        contrib = np.where(outdeg_new > 0, rank / safe, 0.0); contrib[0] = 0.0
        S_row = np.zeros(n + 1)
        for s, (ent, off) in enumerate(P["streams"]):                # partial[s][row], added in stream order
            part = np.add.reduceat(np.concatenate([contrib[ent], [0.0]]), off[:-1])[:n_long] if len(ent) else np.zeros(n_long)
            part[np.diff(off) == 0] = 0.0                             # reduceat returns the next element for empty segments
            S_row[1:n_long + 1] = S_row[1:n_long + 1] + part
        for j in range(n_long + 1, n + 1):                            # short rows: strictly left to right, original order
            v = old_of_new[j]
            acc = 0.0
            for w in P["src_new"][r_begin[v]:r_begin[v + 1]]:
                acc = acc + contrib[w]
            S_row[j] = acc
        rank = (1.0 - d) / n + d * S_row; rank[0] = 0.0
    out = np.empty(n); out[old_of_new[1:]] = rank[1:]
    return out


if __name__ == "__main__":
    a = [int(x) for x in sys.argv[1:]]
    ln, lm, T, H, S = (a + [12, 16, 4, 64, 32][len(a):])[:5]
    o = Oracle()
    g = graphs.rmat(o, 1 << ln, 1 << lm)
    P = build(g["begin"], g["r_begin"], g["r_node_idx"], T, H, S)
    want, it, _ = o.pagerank(g["begin"], g["r_begin"], g["r_node_idx"], 1e-30, 0.85, 20)
    got = sweep20(P, g["begin"], g["r_begin"], g["r_node_idx"])
    m = len(g["r_node_idx"])
    long_entries = int(P["seg_len"].sum())
    print("N=%d M=%d T=%d H=%d S=%d: long section %d rows (%d tiered), %.1f %% of the entries" %
          (P["n"], m, T, H, S, P["n_long"], P["n_tiered"], 100.0 * long_entries / m))
    for s in range(T + 1):
        nz = P["seg_len"][s][P["seg_len"][s] > 0]
        ent, off = P["streams"][s]
        print("  stream %s: %8d entries in %6d segments (mean %.1f), padded to %d" %
              ("cold" if s == T else "t%d" % s, int(nz.sum()), len(nz), nz.mean() if len(nz) else 0.0, len(ent)))
    print("max |tiered order - oracle| after 20 sweeps: %.3e (tolerance 1e-9)" % np.abs(got - want).max())
