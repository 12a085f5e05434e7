#!/usr/bin/env python
"""bench.py -- the PageRank / BFS hot path of Green-Marl+Shoal on B200 (BASELINE.json metric).

A "step" is one pass of the hot path over one synthetic graph: one call of the generated procedure
pagerank(G, e=1e-30, d=0.85, max=ITERS) -- ITERS pull sweeps over the reverse CSR with the fused rank
update and convergence diff (reference apps/src/pagerank.gm:1-20) -- on the LJ-shaped R-MAT graph of
BASELINE.json configs[1] (4.85 M vertices, 69 M edges, int32 ids, fp64 ranks).

  value     GTEPS = edges x iterations x steps / device time; CSR already resident in HBM (CUDA events on
            the launching stream, max over ranks).
  e2e       same metric through the reference-facing C-ABI calls with HOST buffers: per step begin[] and the
            reverse CSR (the arrays the generated procedure reads, copied in by its Shoal prologue on every call,
            reference gm_cpp_gen.cc:685-775) go H2D from pinned memory, the plan is built, the sweeps run,
            and rank[] is copied back to host memory (copy_back, :1480).
  roofline  the pull-sweep kernel: algorithmic bytes per sweep (4*M + 40*N) / its average
            launch duration (each sweep bracketed by CUDA events), against MEASURED_PEAKS.json.
  cpu_baseline  the reference's own OpenMP code (oracle/_ref: reference gm_graph + the generated loop nest
            restated in oracle/ref_glue) on this box's host cores, bounded sample, rank 0 at N=1.
  bfs       extra: hop_dist / gm_bfs_template level-synchronous BFS GTEPS on the same graph (configs[2]).
  extras    comp_BC (bc_adj.gm, 8 seeds) and triangle_counting on the same graph, device time.

`--impl reference` times only the CPU reference arm (no GPU code of this repo on that path).
N>1 (torchrun): destination-range shards, one process per GPU, weak scaling (the graph grows with N).
"""
import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import numpy as np  # noqa: E402

WORKLOADS = {
    # name: (N, M) -- BASELINE.json configs
    "c1": (65536, 1048576),
    "lj": (4850000, 69000000),
    "tw": (41600000, 1470000000),
}
RMAT = dict(seed=1997, a=0.57, b=0.19, c=0.19, permute=True)


_REAL_STDOUT = None


def capture_stdout():
    """stdout carries exactly ONE JSON line: native code under us (the reference's print_crc, NCCL banners) writes to
    fd 1, so fd 1 is pointed at stderr for the whole run and the line goes to a duplicate of the original fd 1."""
    global _REAL_STDOUT
    if _REAL_STDOUT is None:
        sys.stdout.flush()
        _REAL_STDOUT = os.dup(1)
        os.dup2(2, 1)


def emit(line):
    data = (json.dumps(line) + "\n").encode()
    if _REAL_STDOUT is None:
        sys.stdout.write(data.decode()); sys.stdout.flush()
    else:
        os.write(_REAL_STDOUT, data)


def alg_bytes_pagerank(n, m, n_pos=None, n_src=None):
    """compulsory HBM traffic of one pull sweep, int32 ids / fp64 ranks: 4 B per edge (index)
    + per vertex 40 B (row pointer 4, first touch of its contribution 8, old rank 8, new rank 8, new contribution 8,
    out-degree 4).  A STEADY-STATE sweep (the third of a call onwards, the ones bench.py times) skips the rows without
    in-edges -- their value is the constant (1-d)/N -- and gathers only from vertices with out-edges, so it is charged
    4*M + 32*n_pos + 8*n_src: the bytes of the work that launch really does (VERDICT r1 / ADVICE r1)."""
    if n_pos is None:
        return 4 * m + 40 * n
    return 4 * m + 32 * n_pos + 8 * n_src


def measured_peak_gbs():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return float(json.load(open(p))["hbm_gbs"]), "measured"
        except Exception:
            pass
    return 6650.0, "fallback"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md recipe)."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index=0):
        self.index = index
        self.rows = []
        self.proc = None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append((time.time(), line.strip()))

    def stop(self, t0=None, t1=None):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, smax, reasons = [], [], set()
        for ts, line in self.rows:
            if t0 is not None and not (t0 - 0.05 <= ts <= t1 + 0.15):
                continue
            f = [x.strip() for x in line.split(",")]
            if len(f) < 7:
                continue
            try:
                sm.append(float(f[0])); smax.append(float(f[1]))
            except ValueError:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        if not sm:    # region shorter than one sample: fall back to every sample taken
            for ts, line in self.rows:
                f = [x.strip() for x in line.split(",")]
                try:
                    sm.append(float(f[0])); smax.append(float(f[1]))
                except Exception:
                    pass
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(smax) if smax else None,
                "reasons": sorted(reasons), "samples": len(sm)}


# --------------------------------------------------------------------------------------------------
# reference arm (CPU): the reference's own gm_graph + generated loop nest, on the host cores
# --------------------------------------------------------------------------------------------------
def cpu_reference(begin, node_idx, iters_sample, steps=1, warmup=0, threads=None):
    """returns dict(value GTEPS, ms_per_step, kind, cores, rank, threads) timing pagerank() of oracle/_ref (or the
    plain-C oracle port when the reference could not be built) for `iters_sample` iterations per step."""
    from oracle_lib import Oracle, Ref, have_ref
    threads = threads or (os.cpu_count() or 1)
    n, m = len(begin) - 1, len(node_idx)
    if have_ref():
        ref = Ref()
        ref.set_threads(threads)
        g = ref.from_csr(begin, node_idx)          # reference do_semi_sort + make_reverse_edges
        times = []
        rank = None
        for k in range(warmup + steps):
            rank, ms = ref.pagerank(g, 1e-30, 0.85, iters_sample)
            if k >= warmup:
                times.append(ms)
        ref.free(g)
        kind = "reference"
    else:
        orc = Oracle()
        orc.set_threads(threads)
        idx = np.ascontiguousarray(node_idx).copy()
        orc.semi_sort(begin, idx)
        rb, ri = orc.make_reverse(begin, idx)
        times = []
        rank = None
        for k in range(warmup + steps):
            t = time.perf_counter()
            rank, _, _ = orc.pagerank(begin, rb, ri, 1e-30, 0.85, iters_sample)
            if k >= warmup:
                times.append((time.perf_counter() - t) * 1e3)
        kind = "port"
    ms = sum(times) / len(times)
    return dict(value=m * iters_sample / ms / 1e6, ms_per_step=ms, kind=kind, cores=threads, rank=rank)


def cpu_reference_bfs(host, roots, gpu_dists, g, n_ref_roots):
    """BFS CPU baseline + LJ-size parity.  Times the reference's hop_dist (apps/src/hop_dist.gm, generated loop nest
    restated in oracle/ref_glue over the reference gm_graph) and its gm_bfs_template<short,true,false,false,false>
    (inc/gm_bfs_template.h, the reference header itself) on `n_ref_roots` of the timed roots, all host threads;
    checks the device dist[] of EVERY timed root bit for bit against the CPU result (reference for the sampled roots,
    the plain-C oracle port for the rest) and the device gm_bfs_template levels against the reference's."""
    from oracle_lib import Oracle, Ref, have_ref
    threads = os.cpu_count() or 1
    begin, node_idx = host["begin"], host["node_idx"]
    deg = np.diff(begin).astype(np.int64)
    orc = Oracle(); orc.set_threads(threads)
    checked, mismatches, lev_checked, lev_mismatch = 0, 0, 0, 0
    hop_ms, hop_edges, tpl_ms, tpl_edges = [], [], [], []
    kind = "port"
    ref = g_ref = None
    if have_ref():
        ref = Ref(); ref.set_threads(threads)
        g_ref = ref.from_csr(begin, node_idx)
        kind = "reference"
    for k, r in enumerate(roots):
        if k < n_ref_roots and ref is not None:
            want, ms = ref.hop_dist(g_ref, r)
            hop_ms.append(ms); hop_edges.append(int(deg[want != np.iinfo(np.int32).max].sum()))
            lev, _, ms2 = ref.bfs_levels(g_ref, r)
            tpl_ms.append(ms2); tpl_edges.append(hop_edges[-1])
            glev, _, _, _ = g.bfs_levels(r)
            lev_checked += 1; lev_mismatch += int(not np.array_equal(glev, lev))
        else:
            t = time.perf_counter()
            want, _ = orc.hop_dist(begin, node_idx, r)
            if k < n_ref_roots:
                hop_ms.append((time.perf_counter() - t) * 1e3); hop_edges.append(int(deg[want != np.iinfo(np.int32).max].sum()))
        checked += 1
        mismatches += int(not np.array_equal(want, gpu_dists[r]))
    if ref is not None:
        ref.free(g_ref)
    cpu = {"value": round(sum(hop_edges) / sum(hop_ms) / 1e6, 4), "unit": "GTEPS", "cores": threads, "kind": kind,
           "sample": f"hop_dist on {len(hop_ms)} of the {len(roots)} timed roots, same CSR, all host threads; {np.mean(hop_ms):.1f} ms per traversal",
           "ms_per_traversal": round(float(np.mean(hop_ms)), 2)}
    if tpl_ms:
        cpu["bfs_template"] = {"value": round(sum(tpl_edges) / sum(tpl_ms) / 1e6, 4), "unit": "GTEPS",
                               "ms_per_traversal": round(float(np.mean(tpl_ms)), 2),
                               "what": "reference gm_bfs_template<short,true,false,false,false>::do_bfs_forward, same roots"}
    parity = {"dist_roots_checked": checked, "dist_roots_mismatching": mismatches, "bit_exact": mismatches == 0 and lev_mismatch == 0,
              "template_levels_checked": lev_checked, "template_levels_mismatching": lev_mismatch,
              "against": "reference hop_dist / gm_bfs_template for the sampled roots, oracle port (orc_hop_dist) for the others"}
    return cpu, parity


def run_reference_arm(args, n, m):
    """bench.py --impl reference: nothing of libgm_cuda on this path.  The graph comes from the
    reference's own R-MAT generator (create_RMAT_graph, graph_gen.cc:159-287); each step is a bounded
    sample (ITERS_SAMPLE sweeps) of the pagerank() call."""
    rank_id = int(os.environ.get("RANK", "0"))
    if rank_id != 0:
        return
    from oracle_lib import Oracle, Ref, have_ref
    threads = os.cpu_count() or 1
    iters = args.cpu_iters
    if m > 100_000_000:          # TW-shaped: ~1.5-3 s per sweep on the host cores; 2 sweeps per step keep K + W steps within minutes
        iters = min(iters, 2)
        args.cpu_bfs_roots = min(args.cpu_bfs_roots, 2)
    full_n, full_m = n, m
    scale = 1
    graph_desc = "reference create_RMAT_graph seed 1997 a=.57 b=.19 c=.19 permuted"
    dev_csr = None
    if m > 100_000_000 and have_ref() and not os.environ.get("GMC_REF_ARM_SCALED"):
        # The TW-shaped graph: the reference's serial generator needs ~0.75 s per million edges (18 minutes here), so the INPUT is
        # made by the device generator -- the very graph our arm runs on (same seed, same code) -- downloaded once and handed to the
        # reference's gm_graph (prepare_external_creation + freeze + make_reverse_edges on the host cores).  Nothing of libgm_cuda is
        # on the timed path: it only produced the input file-equivalent.
        try:
            from gmb200_loader import load_package
            lib = load_package().GmCuda()
            lib.init(int(os.environ.get("LOCAL_RANK", "0")))
            dg = lib.generate_rmat(n, m, RMAT["seed"], RMAT["a"], RMAT["b"], RMAT["c"], RMAT["permute"])
            dev_csr = (dg.array("begin"), dg.array("node_idx"))
            dg.close()
            graph_desc = "device R-MAT generator seed 1997 a=.57 b=.19 c=.19 permuted: the same CSR our arm runs on, loaded into the reference gm_graph"
        except Exception as ex:       # no usable GPU for the generator: fall back to the scaled sample below
            sys.stderr.write(f"reference arm: device generator unavailable ({ex}); scaling the graph instead\n")
            dev_csr = None
    if m > 100_000_000 and dev_csr is None:      # bounded sample: the serial reference generator needs ~0.75 s per million edges
        scale = -(-m // 92_000_000)
        n, m = n // scale, m // scale
    t0 = time.perf_counter()
    if have_ref():
        ref = Ref(); ref.set_threads(threads)
        if dev_csr is not None:
            g = ref.from_csr(dev_csr[0], dev_csr[1])      # reference do_semi_sort + make_reverse_edges
            dev_csr = None
        else:
            g = ref.rmat(n, m, RMAT["seed"], RMAT["a"], RMAT["b"], RMAT["c"], RMAT["permute"])
        gen_s = time.perf_counter() - t0
        times = []
        for k in range(args.warmup + args.steps):
            _, ms = ref.pagerank(g, 1e-30, 0.85, iters)
            if k >= args.warmup:
                times.append(ms)
        # BFS (configs[2]/[4]) on the same graph: the reference's hop_dist and gm_bfs_template, a few roots
        arr = ref.arrays(g)
        deg = np.diff(arr["begin"]).astype(np.int64)
        nz = np.nonzero(deg)[0]
        broots = [int(nz[0])] + [int(r) for r in np.random.RandomState(1997).choice(nz, size=max(0, args.cpu_bfs_roots - 1), replace=False)]
        hop_ms, hop_e, tpl_ms = [], [], []
        for r in broots:
            dist, ms_h = ref.hop_dist(g, r)
            hop_ms.append(ms_h); hop_e.append(int(deg[dist != np.iinfo(np.int32).max].sum()))
            _, _, ms_t = ref.bfs_levels(g, r)
            tpl_ms.append(ms_t)
        bfs_ref = {"metric": "bfs_gteps", "value": round(sum(hop_e) / sum(hop_ms) / 1e6, 4), "unit": "GTEPS", "roots": len(broots),
                   "ms_per_traversal": round(float(np.mean(hop_ms)), 2), "what": "reference hop_dist, all host threads",
                   "bfs_template": {"value": round(sum(hop_e) / sum(tpl_ms) / 1e6, 4), "unit": "GTEPS", "ms_per_traversal": round(float(np.mean(tpl_ms)), 2)}}
        del arr, deg, nz
        ref.free(g)
        kind = "reference"
    else:
        bfs_ref = None
        orc = Oracle(); orc.set_threads(threads)
        b, i = orc.rmat(n, m, RMAT["seed"], RMAT["a"], RMAT["b"], RMAT["c"], RMAT["permute"])
        i = np.ascontiguousarray(i).copy()
        orc.semi_sort(b, i)
        rb, ri = orc.make_reverse(b, i)
        gen_s = time.perf_counter() - t0
        times = []
        for k in range(args.warmup + args.steps):
            t = time.perf_counter()
            orc.pagerank(b, rb, ri, 1e-30, 0.85, iters)
            if k >= args.warmup:
                times.append((time.perf_counter() - t) * 1e3)
        kind = "port"
    ms = sum(times) / len(times)
    v = m * iters / ms / 1e6
    sample = f"{iters} of {args.iters} sweeps per step, same graph shape, all {threads} host threads (OpenMP dynamic,128)"
    if scale > 1:
        sample += f"; graph scaled 1/{scale} (N={n}, M={m}, same a/b/c and average degree) to bound the serial generator's time"
    line = {
        "impl": "reference", "metric": "pagerank_gteps", "value": v, "unit": "GTEPS", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": workload_name(args, full_n, full_m), "nodes": full_n, "edges": full_m, "sample_nodes": n, "sample_edges": m,
                   "iterations_per_step": iters,
                   "graph": graph_desc, "generate_s": round(gen_s, 1), "same_graph_as_our_arm": scale == 1 and full_m > 100_000_000},
        "cpu_baseline": {"value": v, "unit": "GTEPS", "cores": threads, "kind": kind, "sample": sample},
        "e2e": {"value": v, "unit": "GTEPS", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "bfs": bfs_ref,
    }
    emit(line)


def workload_name(args, n, m, workload=None):
    return f"pagerank {args.iters} sweeps, R-MAT {workload or args.workload}-shaped N={n} M={m}, int32 ids, fp64 ranks, d=0.85, e=1e-30"


# --------------------------------------------------------------------------------------------------
# our arm
# --------------------------------------------------------------------------------------------------
def run_single(args, n, m):
    import torch
    from gmb200_loader import load_package
    pkg = load_package()
    lib = pkg.GmCuda()
    dev = int(os.environ.get("LOCAL_RANK", "0"))
    lib.init(dev)
    torch.cuda.set_device(dev)
    iters = args.iters
    t0 = time.perf_counter()
    g = lib.generate_rmat(n, m, RMAT["seed"], RMAT["a"], RMAT["b"], RMAT["c"], RMAT["permute"])
    lib.sync()
    gen_s = time.perf_counter() - t0

    stream = torch.cuda.Stream()
    sp = stream.cuda_stream
    rank_dev = torch.empty(n, dtype=torch.float64, device="cuda")
    launches = 0
    for _ in range(args.warmup):
        g.pagerank_dev(rank_dev.data_ptr(), 1e-30, 0.85, iters, stream=sp)
    torch.cuda.synchronize()
    clocks = ClockSampler(dev)
    clocks.start()
    time.sleep(0.25)
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    w0 = time.time()
    ev0.record(stream)
    for _ in range(args.steps):
        info = g.pagerank_dev(rank_dev.data_ptr(), 1e-30, 0.85, iters, stream=sp)
        launches += info.launches
        assert info.iterations == iters
    ev1.record(stream)
    torch.cuda.synchronize()
    w1 = time.time()
    ms_total = ev0.elapsed_time(ev1)
    # keep the clocks sampler running over the per-sweep roofline pass as well
    sweep_ms, lps = g.pagerank_time_sweeps(0.85, warmup=3, n=max(20, iters), stream=sp)
    w2 = time.time()
    clk = clocks.stop(w0, w2)
    ms_step = ms_total / args.steps
    value = m * iters / ms_step / 1e6
    rank_gpu = rank_dev.cpu().numpy()

    # ---- roofline of the dominant kernel (the pull sweep) --------------------------------------
    peak, peak_kind = measured_peak_gbs()
    pinfo = g.pagerank_plan_info()
    a_bytes = alg_bytes_pagerank(n, m, pinfo["n_pos"], pinfo["n_src"])       # timed sweeps are steady-state ones (warmup=3)
    avg_sweep_ms = float(np.mean(sweep_ms))
    achieved = a_bytes / avg_sweep_ms / 1e6
    traffic = None
    tp = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(tp):
        try:
            tj = json.load(open(tp))
            if tj.get("nodes") == n and tj.get("edges") == m:
                traffic = tj.get("dram_bytes_per_launch")
        except Exception:
            pass
    roofline = {"bound": "hbm", "achieved": round(achieved, 1), "peak": peak, "unit": "GB/s", "frac": round(achieved / peak, 4),
                "traffic": traffic, "peak_kind": peak_kind, "kernel": "pr_sweep (+ pr_combine)", "alg_bytes_per_launch": a_bytes,
                "alg_bytes_formula": "4*M + 32*n_pos + 8*n_src", "n_pos": pinfo["n_pos"], "n_src": pinfo["n_src"],
                "alg_bytes_full_sweep": alg_bytes_pagerank(n, m),
                "plan": {k: int(pinfo[k]) for k in ("tiers", "segments", "split_rows", "tier_ctas", "hot0_entries", "cold_entries", "segment_entries")},
                "avg_launch_ms": round(avg_sweep_ms, 5), "kernels_per_sweep": lps,
                "gteps_kernel_only": round(m / avg_sweep_ms / 1e6, 1)}

    # ---- e2e: host buffers through the C ABI, H2D + D2H inside the timed region ---------------------
    host = {k: g.array(k) for k in ("begin", "node_idx", "r_begin", "r_node_idx")}
    pinned = {}
    for k in ("begin", "r_begin", "r_node_idx"):      # what the generated pagerank() copies in (no node_idx: pull only)
        t = torch.empty(host[k].shape, dtype=torch.int32).pin_memory()
        t.numpy()[:] = host[k]
        pinned[k] = t.numpy()
    rank_host_t = torch.empty(n, dtype=torch.float64).pin_memory()
    rank_host = rank_host_t.numpy()
    h2d = sum(a.nbytes for a in pinned.values())
    d2h = rank_host.nbytes
    parts = {"h2d_ms": [], "plan_and_sweeps_ms": []}

    def e2e_step_parts():          # un-overlapped: copy-in, then plan + sweeps + copy-back (only for the breakdown)
        t_a = time.perf_counter()
        dg = lib.graph_pull_only(pinned["begin"], pinned["r_begin"], pinned["r_node_idx"])
        t_b = time.perf_counter()
        _, inf = dg.pagerank_into(rank_host, 1e-30, 0.85, iters)
        t_c = time.perf_counter()
        dg.close()
        parts["h2d_ms"].append((t_b - t_a) * 1e3); parts["plan_and_sweeps_ms"].append((t_c - t_b) * 1e3)
        return inf

    def e2e_step():                # the call a user makes: gmc_pagerank_host_csr (copy-in overlapped with the plan build)
        _, inf = lib.pagerank_host_csr(pinned["begin"], pinned["r_begin"], pinned["r_node_idx"], rank_host, 1e-30, 0.85, iters)
        return inf

    for _ in range(min(args.warmup, 2)):
        e2e_step()
        e2e_step_parts()
    lib.sync()
    for v in parts.values():
        v.clear()
    for _ in range(2):
        e2e_step_parts()
    lib.sync()
    e2e_steps = max(1, min(args.steps, 5))
    e2e_all = []
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        t_s = time.perf_counter()
        e2e_step()                      # returns with rank[] in host memory: the call is synchronous
        e2e_all.append((time.perf_counter() - t_s) * 1e3)
    lib.sync()
    e2e_ms = (time.perf_counter() - t0) * 1e3 / e2e_steps
    e2e = {"value": round(m * iters / e2e_ms / 1e6, 3), "unit": "GTEPS", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
           "ms_per_step": round(e2e_ms, 3), "steps": e2e_steps,
           "ms_each_step": [round(x, 3) for x in e2e_all],
           "call": "gmc_pagerank_host_csr (pinned host CSR in, host rank[] out; copy-in overlapped with the plan build)",
           "breakdown_ms": {"what": "the same work un-overlapped (upload, then gmc_pagerank)",
                            "h2d_begin_rbegin_rnodeidx": round(float(np.mean(parts["h2d_ms"])), 3),
                            "plan_20_sweeps_d2h": round(float(np.mean(parts["plan_and_sweeps_ms"])), 3)},
           "h2d_gbs": round(h2d / float(np.mean(parts["h2d_ms"])) / 1e6, 1)}
    if os.environ.get("GMC_BENCH_DEBUG"):
        sys.stderr.write("e2e parts: %r\n" % ({k: [round(x, 1) for x in v] for k, v in parts.items()},))
    # the one-shot call builds the plan without tier streams (it serves 20 sweeps once: DESIGN.md 6.1), the resident graph has
    # them: same values to the last bits of the summation order, far inside the 1e-9 tolerance
    e2e["max_abs_vs_resident_result"] = float(np.abs(rank_host - rank_gpu).max())
    assert e2e["max_abs_vs_resident_result"] <= 1e-12, "e2e result differs from the resident-graph result"

    # ---- BFS (configs[2]) on the same graph: hop_dist over 16 roots + the gm_bfs_template traversal ------
    deg = np.diff(host["begin"])
    rs = np.random.RandomState(1997)
    roots = [int(np.nonzero(deg)[0][0])] + [int(r) for r in rs.choice(np.nonzero(deg)[0], size=15, replace=False)]
    dist_dev = torch.empty(n, dtype=torch.int32, device="cuda")
    bfs_ms, bfs_edges, bfs_launches, bfs_abytes, bfs_dists = [], [], 0, [], {}
    for r in roots:
        g.hop_dist_dev(dist_dev.data_ptr(), r, stream=sp)           # warm
        bi = g.hop_dist_dev(dist_dev.data_ptr(), r, stream=sp)
        bfs_ms.append(bi.kernel_ms); bfs_edges.append(bi.edges_traversed); bfs_launches += bi.launches
        # A_BFS = 4*M_r + N_r*(4 + sizeof(level)) + 2*ceil(N/8)*L   (level = int32 dist for hop_dist)
        bfs_abytes.append(4 * bi.edges_traversed + bi.reached * 8 + 2 * ((n + 7) // 8) * bi.levels)
        bfs_dists[r] = dist_dev.cpu().numpy()
    a_bfs = float(np.sum(bfs_abytes)) / float(np.sum(bfs_ms)) / 1e6
    bfs_traffic, bfs_traffic_what = None, None       # dram bytes of one traversal (the first root), from the committed ncu launch list
    try:
        tb = json.load(open(os.path.join(ROOT, "profiles", "traffic.json"))).get("bfs_lj_root3", {})
        if tb.get("nodes") == n and tb.get("edges") == m and tb.get("root") == roots[0]:
            bfs_traffic, bfs_traffic_what = tb.get("dram_bytes_per_traversal"), "one traversal from root %d (%s); A_BFS of that root: %d" % (
                roots[0], tb.get("source"), int(bfs_abytes[0]))
    except Exception:
        pass
    bfs = {"metric": "bfs_gteps", "value": round(sum(bfs_edges) / sum(bfs_ms) / 1e6, 3), "unit": "GTEPS",
           "roots": len(roots), "ms_per_traversal": round(float(np.mean(bfs_ms)), 4),
           "edges_per_traversal": int(np.mean(bfs_edges)), "what": "hop_dist (direction-optimizing: bottom-up when the frontier's out-edges exceed 1/14 of the unexplored edges), device time",
           "gpu_launches": int(bfs_launches),
           "roofline": {"bound": "hbm", "achieved": round(a_bfs, 1), "peak": peak, "unit": "GB/s", "frac": round(a_bfs / peak, 4),
                        "traffic": bfs_traffic, "traffic_what": bfs_traffic_what, "peak_kind": peak_kind, "kernel": "whole traversal (bfs_small_levels for runs of small levels; td_prepare/td_expand/bfs_finish, bu_scan_mark or bu_scan/bfs_finish per wide level)",
                        "alg_bytes_per_traversal": int(np.mean(bfs_abytes)),
                        "formula": "A_BFS = 4*M_r + N_r*(4+4) + 2*ceil(N/8)*L over the traversal's device time"}}
    # gm_bfs_template<short,...> traversal (bc / kosaraju's InBFS) on the first 4 roots
    tl_ms, tl_edges = [], []
    for r in roots[:4]:
        g.bfs_levels(r)
        _, li, _, _ = g.bfs_levels(r)
        tl_ms.append(li.kernel_ms); tl_edges.append(li.edges_traversed)
    bfs["bfs_template"] = {"value": round(sum(tl_edges) / sum(tl_ms) / 1e6, 3), "unit": "GTEPS", "roots": len(tl_ms),
                           "ms_per_traversal": round(float(np.mean(tl_ms)), 4), "what": "gmc_bfs_levels (gm_bfs_template<short> levels), device time"}

    # ---- the other generated procedures on the same graph: comp_BC (bc_adj.gm) and triangle_counting ------
    extras = {}
    try:
        _, bci = g.comp_bc(roots[:1], adj=True)                          # warm
        _, bci = g.comp_bc(roots, adj=True)
        extras["comp_bc"] = {"what": "gmc_comp_bc_adj, device time incl. the traversals", "seeds": len(roots),
                             "ms_per_seed": round(bci.device_ms / len(roots), 3), "launches": int(bci.launches)}
        tcount, tms = g.triangle_counting()
        extras["triangle_counting"] = {"T": int(tcount), "device_ms": round(tms, 3)}
    except Exception as ex:                                              # secondary figures: reported, never fatal to the line
        extras["error"] = repr(ex)

    # ---- CPU baseline: the reference's own code on the host cores, bounded sample -------------------
    cpu = None
    parity = None
    if not args.no_cpu_baseline:
        c = cpu_reference(host["begin"], host["node_idx"], args.cpu_iters, steps=1, warmup=0)
        cpu = {"value": round(c["value"], 4), "unit": "GTEPS", "cores": c["cores"], "kind": c["kind"],
               "sample": f"{args.cpu_iters} of {iters} sweeps of the same graph (identical CSR), one pagerank() call, "
                         f"all host threads; {c['ms_per_step']:.1f} ms"}
        # the CPU result doubles as a full-size parity check (same CSR, same iteration count)
        rk, _ = g.pagerank(1e-30, 0.85, args.cpu_iters)
        parity = {"max_abs_err_vs_cpu_reference": float(np.abs(rk - c["rank"]).max()), "iterations": args.cpu_iters,
                  "tolerance": 1e-9}
        bfs["cpu_baseline"], bfs["parity"] = cpu_reference_bfs(host, roots, bfs_dists, g, args.cpu_bfs_roots)

    line = {
        "metric": "pagerank_gteps", "value": round(value, 3), "unit": "GTEPS", "n_gpus": 1, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": round(ms_step, 4), "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": workload_name(args, n, m), "nodes": n, "edges": m, "iterations_per_step": iters,
                   "graph": "device R-MAT generator seed 1997 a=.57 b=.19 c=.19 permuted (counter-based RNG)",
                   "generate_s": round(gen_s, 2),
                   "l2": "inputs larger than L2 (index stream %d MB per sweep > 126 MB L2); no explicit flush" % (4 * m // 2**20),
                   "parallelism": "1 GPU"},
        "clocks": clk, "e2e": e2e, "gpu_launches": int(launches), "roofline": roofline, "cpu_baseline": cpu,
        "bfs": bfs, "extras": extras, "parity": parity,
    }
    emit(line)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default=None, choices=sorted(WORKLOADS),
                    help="default: lj on 1 GPU (BASELINE configs[1]), tw on 2/4/8 GPUs (configs[3])")
    ap.add_argument("--nodes", type=int, default=None, help="override the workload's vertex count (experiments)")
    ap.add_argument("--edges", type=int, default=None, help="override the workload's edge count (experiments)")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--iters", type=int, default=20, help="pull sweeps per pagerank() call (BASELINE configs: 20)")
    ap.add_argument("--cpu-iters", type=int, default=20, help="sweeps per CPU-reference sample (the config's 20: one full pagerank() call)")
    ap.add_argument("--cpu-bfs-roots", type=int, default=3, help="roots on which the reference's hop_dist / gm_bfs_template are timed")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()
    capture_stdout()
    if args.warmup < 3 and args.impl == "ours":
        args.warmup = 3        # timing rule: at least 3 warm-up steps
    world = int(os.environ.get("WORLD_SIZE", "1"))
    multi = world > 1 or args.gpus > 1
    if args.workload is None:
        args.workload = "tw" if multi else "lj"
    n, m = WORKLOADS[args.workload]
    n, m = args.nodes or n, args.edges or m
    if args.impl == "reference":
        run_reference_arm(args, n, m)
        return
    if multi:
        from bench_multi import run_multi
        run_multi(args, n, m, args.workload)
    else:
        run_single(args, n, m)


if __name__ == "__main__":
    main()
