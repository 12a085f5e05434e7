"""Regenerates tests/golden/*.npz from the REFERENCE ITSELF (oracle/_ref/libgmref.so = the reference's own
gm_graph runtime, generators and gm_bfs_template compiled from /root/reference by oracle/Makefile, plus the
restated generated procedures in oracle/ref_glue).  Run in the build container (needs /root/reference):

    make -C oracle && python tests/golden/make_golden.py

The fixtures pin the plain-C oracle (oracle/gm_oracle.c) and, through it, the CUDA path.
"""
import hashlib
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
from oracle_lib import Ref  # noqa: E402


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def main():
    ref = Ref()
    ref.set_threads(4)
    # (1) BASELINE config 1: graph_gen 65536 1048576 <f> 0  -> digests + leading values (full arrays are too big to commit)
    g = ref.uniform(65536, 1048576, 1997)
    a = ref.arrays(g)
    rank, _ = ref.pagerank(g, 1e-30, 0.85, 20)
    dist, _ = ref.hop_dist(g, 0)
    lev, nl, _ = ref.bfs_levels(g, 0)
    np.savez_compressed(os.path.join(HERE, "c1_digest.npz"),
                        N=a["N"], M=a["M"], begin_sha=sha(a["begin"]), node_idx_sha=sha(a["node_idx"]),
                        r_begin_sha=sha(a["r_begin"]), r_node_idx_sha=sha(a["r_node_idx"]),
                        rank_sha=sha(rank), rank_head=rank[:64], rank_sum=float(rank.sum()),
                        dist_sha=sha(dist), dist_head=dist[:64], dist_hist=np.bincount(dist),
                        lev_sha=sha(lev), n_levels=nl, begin_head=a["begin"][:64], node_idx_head=a["node_idx"][:64])
    ref.free(g)
    # (2) small R-MAT (reference create_RMAT_graph), full arrays: CSR, reverse CSR, rank after 20 sweeps, dist / levels from 3 roots
    g = ref.rmat(4096, 65536, 1997, 0.57, 0.19, 0.19, True)
    a = ref.arrays(g)
    rank, _ = ref.pagerank(g, 1e-30, 0.85, 20)
    rank_conv, _ = ref.pagerank(g, 0.001, 0.85, 100)
    deg = np.diff(a["begin"])
    roots = np.array([0, int(np.nonzero(deg)[0][0]), int(deg.argmax())], np.int32)
    dists = np.stack([ref.hop_dist(g, int(r))[0] for r in roots])
    levs = np.stack([ref.bfs_levels(g, int(r))[0] for r in roots])
    np.savez_compressed(os.path.join(HERE, "rmat4096.npz"), begin=a["begin"], node_idx=a["node_idx"],
                        r_begin=a["r_begin"], r_node_idx=a["r_node_idx"], rank20=rank, rank_conv=rank_conv,
                        roots=roots, dist=dists, levels=levs)
    ref.free(g)
    # (3) uniform 2000 / 30000 (graph_gen type 0 at small size)
    g = ref.uniform(2000, 30000, 1997)
    a = ref.arrays(g)
    rank, _ = ref.pagerank(g, 1e-30, 0.85, 20)
    dist, _ = ref.hop_dist(g, 0)
    lev, _, _ = ref.bfs_levels(g, 0)
    np.savez_compressed(os.path.join(HERE, "uniform2000.npz"), begin=a["begin"], node_idx=a["node_idx"],
                        r_begin=a["r_begin"], r_node_idx=a["r_node_idx"], rank20=rank, dist=dist, levels=lev)
    ref.free(g)
    print("golden fixtures written to", HERE)


def triangle_counts():
    """triangle_counts.json: the generated triangle_counting() loop nest over the reference's own gm_graph::has_edge_to
    (oracle/ref_glue/ref_capi.cc: ref_triangle_counting) on the two full fixtures written by main()"""
    import json
    ref = Ref()
    tc = {}
    for name in ("rmat4096", "uniform2000"):
        z = np.load(os.path.join(HERE, name + ".npz"))
        g = ref.from_csr(z["begin"], z["node_idx"])
        tc[name] = ref.triangle_counting(g)
        ref.free(g)
    json.dump(tc, open(os.path.join(HERE, "triangle_counts.json"), "w"))
    print("triangle counts", tc)


def bc_golden():
    """bc_golden.npz: comp_BC as generated from apps/src/bc.gm and apps/src/bc_adj.gm (oracle/ref_glue/ref_capi.cc: the
    generated gm_bfs_template<short, true, false, false, true> subclass over the reference's REAL template), ONE host
    thread (the reference's down-edge marking races with several), on the two full fixtures written by main().  Both graphs
    are large enough for the traversal to reach the bottom-up states, whose levels carry no down edges."""
    ref = Ref()
    ref.set_threads(1)
    out = {}
    for name in ("rmat4096", "uniform2000"):
        z = np.load(os.path.join(HERE, name + ".npz"))
        g = ref.from_csr(z["begin"], z["node_idx"])
        deg = np.diff(z["begin"])
        nz = np.nonzero(deg)[0]
        seeds = np.array([int(nz[0]), int(nz[len(nz) // 2]), int(deg.argmax()), 1, int(nz[0])], np.int32)
        out[name + "_seeds"] = seeds
        out[name + "_bc"] = ref.comp_bc(g, seeds)
        out[name + "_bc_adj"] = ref.comp_bc(g, seeds, adj=True)
        ref.free(g)
    # (3) the reference's create_RMAT_graph(5000, 60000, seed 3): from vertex 18 the save_child traversal expands level 4 (332
    # vertices with children) bottom-up.  The graph is not stored: the oracle's generator reproduces it (checked by the digest)
    g = ref.rmat(5000, 60000, 3)
    a = ref.arrays(g)
    seeds = np.array([18, 0, 7], np.int32)
    out["rmat5000_params"] = np.array([5000, 60000, 3], np.int64)
    out["rmat5000_csr_sha"] = sha(np.concatenate([a["begin"], a["node_idx"], a["r_begin"], a["r_node_idx"]]))
    out["rmat5000_seeds"] = seeds
    out["rmat5000_bc"] = ref.comp_bc(g, seeds)
    out["rmat5000_bc_adj"] = ref.comp_bc(g, seeds, adj=True)
    ref.free(g)
    np.savez_compressed(os.path.join(HERE, "bc_golden.npz"), **out)
    print("bc golden:", {k: (v.shape, int(np.isnan(v).sum())) for k, v in out.items() if k.endswith("_bc") or k.endswith("_bc_adj")})


if __name__ == "__main__":
    if len(sys.argv) > 1 and sys.argv[1] == "triangles":
        triangle_counts()
    elif len(sys.argv) > 1 and sys.argv[1] == "bc":
        bc_golden()
    else:
        main()
        triangle_counts()
        bc_golden()
