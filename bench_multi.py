"""bench.py --gpus N (N > 1): one process per GPU (torch.distributed.run), destination-range shards.

Every rank builds the same synthetic graph with the deterministic device generator (counter-based RNG), plans
its own shard of the PageRank entry array, and the ranks exchange the new contributions with NCCL after every
sweep (BASELINE.json configs[3]: the Twitter-shaped R-MAT graph on 2/4/8 B200).  torch.distributed
is used for the bootstrap (shipping the 128-byte NCCL id), the barriers around the timed region and the
max-over-ranks reduction of the device times -- not for the data path, which is libgm_cuda's own communicator.
"""
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))


def bootstrap_comm(lib, dist, torch, rank, world):
    """rank 0 makes the NCCL unique id, everybody receives its 128 bytes, every rank joins (gmc_comm_create)"""
    uid = torch.from_numpy(lib.comm_unique_id() if rank == 0 else np.zeros(128, np.uint8))
    if dist.get_backend() == "nccl":
        t = uid.cuda()
        dist.broadcast(t, 0)
        uid = t.cpu()
    else:
        dist.broadcast(uid, 0)
    return lib.comm_create(world, rank, uid.numpy())


def max_over_ranks(dist, torch, value):
    """the slowest rank's device time is the job's time"""
    t = torch.tensor([float(value)], dtype=torch.float64)
    if dist.get_backend() == "nccl":
        t = t.cuda()
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def sum_over_ranks(dist, torch, value):
    t = torch.tensor([float(value)], dtype=torch.float64)
    if dist.get_backend() == "nccl":
        t = t.cuda()
    dist.all_reduce(t, op=dist.ReduceOp.SUM)
    return float(t.item())


def run_multi(args, n, m, workload):
    import torch
    import torch.distributed as dist
    from bench import RMAT, ClockSampler, alg_bytes_pagerank, measured_peak_gbs, workload_name
    from gmb200_loader import load_package
    rank, world, local = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1")), int(os.environ.get("LOCAL_RANK", "0"))
    if world == 1 and args.gpus > 1:
        raise SystemExit("bench.py --gpus N with N > 1 must be launched with torch.distributed.run (one process per GPU)")
    lib = load_package().GmCuda()
    lib.init(local)
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    cpu_group = dist.new_group(backend="gloo")      # host-side barriers: ranks that wait must not spin a kernel on their GPU
    comm = bootstrap_comm(lib, dist, torch, rank, world)
    iters = args.iters

    t0 = time.perf_counter()
    g = lib.generate_rmat(n, m, RMAT["seed"], RMAT["a"], RMAT["b"], RMAT["c"], RMAT["permute"])
    lib.sync()
    gen_s = time.perf_counter() - t0
    stream = torch.cuda.Stream()
    sp = stream.cuda_stream
    rank_dev = torch.empty(n, dtype=torch.float64, device="cuda")
    t0 = time.perf_counter()
    for _ in range(args.warmup):
        g.pagerank_sharded_dev(comm, rank_dev.data_ptr(), 1e-30, 0.85, iters, stream=sp)
    torch.cuda.synchronize()
    warm_s = time.perf_counter() - t0

    clocks = ClockSampler(local)
    if rank == 0:
        clocks.start()
        time.sleep(0.25)
    launches = 0
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    dist.barrier(); torch.cuda.synchronize()
    w0 = time.time()
    ev0.record(stream)
    for _ in range(args.steps):
        info = g.pagerank_sharded_dev(comm, rank_dev.data_ptr(), 1e-30, 0.85, iters, stream=sp)
        launches += info.launches
        assert info.iterations == iters
    ev1.record(stream)
    torch.cuda.synchronize(); dist.barrier()
    w1 = time.time()
    ms_total = max_over_ranks(dist, torch, ev0.elapsed_time(ev1))
    ms_step = ms_total / args.steps
    value = m * iters / ms_step / 1e6

    # per-sweep device times: the shard sweep kernel alone, and the exchange that follows it
    sweep_ms, exch_ms = g.pagerank_sharded_time_sweeps(comm, 0.85, warmup=3, n=max(20, iters), stream=sp)
    w2 = time.time()
    clk = clocks.stop(w0, w2) if rank == 0 else None
    per_rank = torch.zeros(world, dtype=torch.float64, device="cuda")
    per_rank[rank] = float(np.mean(sweep_ms))
    dist.all_reduce(per_rank)
    pi = g.pagerank_plan_info()
    shard_stats = torch.zeros(world, 4, dtype=torch.float64, device="cuda")
    shard_stats[rank] = torch.tensor([pi["n_long"], pi["n_slices"], pi["long_entries"], pi["total_entries"]], dtype=torch.float64)
    dist.all_reduce(shard_stats)
    avg_sweep = max_over_ranks(dist, torch, float(np.mean(sweep_ms)))
    avg_exch = max_over_ranks(dist, torch, float(np.mean(exch_ms)))
    peak, peak_kind = measured_peak_gbs()
    # steady-state sweeps (warmup=3): rows without in-edges are skipped, so the bytes are those of the rows really processed
    a_bytes = alg_bytes_pagerank(n, m, pi["n_pos"], pi["n_src"]) / world
    roofline = {"bound": "hbm", "achieved": round(a_bytes / avg_sweep / 1e6, 1), "peak": peak, "unit": "GB/s",
                "frac": round(a_bytes / avg_sweep / 1e6 / peak, 4), "traffic": None, "peak_kind": peak_kind, "kernel": "pr_sweep (one shard)",
                "alg_bytes_per_launch": int(a_bytes), "alg_bytes_formula": "(4*M + 32*n_pos + 8*n_src) / world", "n_pos": int(pi["n_pos"]), "n_src": int(pi["n_src"]),
                "avg_launch_ms": round(avg_sweep, 5),
                "avg_launch_ms_per_rank": [round(float(x), 4) for x in per_rank.cpu().numpy()],
                "shards_long_rows_slices_long_entries_entries": [[int(x) for x in row] for row in shard_stats.cpu().numpy()],
                "exchange": {"path": g.pagerank_exchange_path(),
                             "what": "peer-memory: contributions stored into the peers' HBM over NVLink from the sweep epilogue, avg_ms = wait for the peers' flags; "
                                     "nccl: grouped broadcasts of the owned contribution ranges after the sweep",
                             "avg_ms": round(avg_exch, 5), "recv_bytes_per_rank": int(8 * n * (world - 1) / world),
                             "recv_gbs_per_rank": round(8 * n * (world - 1) / world / ((avg_sweep + avg_exch) if g.pagerank_exchange_path() == "peer-memory" else avg_exch) / 1e6, 1),
                             "nvlink_peak_gbs_per_dir": 900.0}}

    # ---- BFS (configs[4]): sharded hop_dist over a few roots, device time = max over ranks ----------------
    deg = np.diff(g.array("begin"))
    nzv = np.nonzero(deg)[0]
    rs = np.random.RandomState(1997)
    roots = [int(nzv[0])] + [int(r) for r in rs.choice(nzv, size=3, replace=False)]
    del deg, nzv
    dist_dev = torch.empty(n, dtype=torch.int32, device="cuda")
    bfs_ms, bfs_edges, bfs_levels = [], [], []
    for r in roots:
        g.hop_dist_sharded_dev(comm, dist_dev.data_ptr(), r, stream=sp)
        dist.barrier()
        bi = g.hop_dist_sharded_dev(comm, dist_dev.data_ptr(), r, stream=sp)
        bfs_ms.append(max_over_ranks(dist, torch, bi.kernel_ms)); bfs_edges.append(bi.edges_traversed); bfs_levels.append(bi.levels)
    # ---- parity of the sharded paths, checkable by the driver: every rank holds the whole graph, so
    # rank 0 repeats both algorithms on its ONE GPU (world = 1 plan, no exchange) and the sharded results must agree:
    # PageRank max |sharded - single| <= 1e-9 (north-star tolerance), hop_dist bit-equal for every timed root; the
    # result vectors of all ranks must be identical bit for bit (they come out of the same exchange).
    sharded_rank = rank_dev.clone()
    ref0 = sharded_rank.clone()
    dist.broadcast(ref0, 0)
    ranks_differ = max_over_ranks(dist, torch, float((sharded_rank - ref0).abs().max().item()))
    del ref0
    bfs_dist_sharded = {}
    for r in roots:
        g.hop_dist_sharded_dev(comm, dist_dev.data_ptr(), r, stream=sp)
        torch.cuda.synchronize()
        bfs_dist_sharded[r] = dist_dev.clone()
    exchange_path = g.pagerank_exchange_path()
    parity = None
    single = None
    traffic_tw = None          # dram bytes of one sweep of this graph on one GPU, from the committed ncu capture
    try:
        tj = json.load(open(os.path.join(ROOT, "profiles", "traffic.json"))).get("tw_1gpu", {})
        if tj.get("nodes") == n and tj.get("edges") == m:
            traffic_tw = tj.get("dram_bytes_per_launch")
    except Exception:
        pass
    if rank == 0:
        t0 = time.perf_counter()
        one_dev = torch.empty(n, dtype=torch.float64, device="cuda")
        g.pagerank_dev(one_dev.data_ptr(), 1e-30, 0.85, iters, stream=sp)          # builds the world = 1 plan of the same graph
        g.pagerank_dev(one_dev.data_ptr(), 1e-30, 0.85, iters, stream=sp)          # (second call: the tiered plan of a reused graph)
        torch.cuda.synchronize()
        plan_s = time.perf_counter() - t0
        ev2, ev3 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        ev2.record(stream)
        info1 = g.pagerank_dev(one_dev.data_ptr(), 1e-30, 0.85, iters, stream=sp)
        ev3.record(stream)
        torch.cuda.synchronize()
        one_ms = ev2.elapsed_time(ev3)
        one_sweeps, _ = g.pagerank_time_sweeps(0.85, warmup=3, n=10, stream=sp)
        pi1 = g.pagerank_plan_info()
        pr_err = float((one_dev - sharded_rank).abs().max().item())
        bad_roots = 0
        one_bfs_ms, one_bfs_edges = [], []
        for r in roots:
            g.hop_dist_dev(dist_dev.data_ptr(), r, stream=sp)
            b1 = g.hop_dist_dev(dist_dev.data_ptr(), r, stream=sp)
            torch.cuda.synchronize()
            one_bfs_ms.append(b1.kernel_ms); one_bfs_edges.append(b1.edges_traversed)
            bad_roots += int(not torch.equal(dist_dev, bfs_dist_sharded[r]))
        parity = {"pagerank_max_abs_vs_single_gpu": pr_err, "pagerank_ok": bool(pr_err <= 1e-9 and info1.iterations == iters),
                  "pagerank_ranks_max_abs_diff": ranks_differ, "hop_dist_roots_checked": len(roots), "hop_dist_roots_mismatching": bad_roots,
                  "exchange_path": exchange_path, "exchange_path_ok": exchange_path == "peer-memory" or world > 8,
                  "ok": bool(pr_err <= 1e-9 and bad_roots == 0 and ranks_differ == 0.0),
                  "against": "the same device-generated graph run on ONE GPU of this job (world = 1 plan, no exchange), same call arguments"}
        a1 = alg_bytes_pagerank(n, m, pi1["n_pos"], pi1["n_src"])
        single = {"what": "same graph on one GPU of this job: the anchor of the strong-scaling curve", "pagerank_gteps": round(m * iters / one_ms / 1e6, 2),
                  "ms_per_step": round(one_ms, 3), "avg_sweep_ms": round(float(np.mean(one_sweeps)), 4),
                  "roofline_frac": round(a1 / float(np.mean(one_sweeps)) / 1e6 / peak, 4), "alg_bytes_per_launch": int(a1),
                  "traffic": traffic_tw,
                  "bfs_gteps": round(sum(one_bfs_edges) / sum(one_bfs_ms) / 1e6, 2), "bfs_ms_per_traversal": round(float(np.mean(one_bfs_ms)), 4),
                  "first_call_s": round(plan_s, 2)}
        del one_dev
    del sharded_rank, bfs_dist_sharded
    torch.cuda.synchronize()
    dist.barrier(group=cpu_group)            # rank 0 alone was busy: the others wait on the host

    bfs = {"metric": "bfs_gteps", "value": round(sum(bfs_edges) / sum(bfs_ms) / 1e6, 3), "unit": "GTEPS", "roots": len(roots),
           "ms_per_traversal": round(float(np.mean(bfs_ms)), 4), "edges_per_traversal": int(np.mean(bfs_edges)),
           "levels": bfs_levels, "what": "sharded hop_dist (direction chosen from the frontier's out-edges; small / medium levels run in full by every rank, the others sharded with a bitmap exchange per level), device time, max over ranks"}

    # ---- e2e: host buffers through the C ABI on every rank (CSR + reverse CSR H2D, plan, sweeps + exchange,
    # rank[] D2H), wall clock between barriers, max over ranks ---------------------------------------------
    e2e = None
    if not args.no_e2e:
        host = {k: g.array(k) for k in ("begin", "r_begin", "r_node_idx")}     # what pagerank() reads: pull only
        pinned, pin_ok = {}, True
        for k, a in host.items():
            try:
                t = torch.empty(a.shape, dtype=torch.int32).pin_memory()
            except Exception:
                t = torch.empty(a.shape, dtype=torch.int32); pin_ok = False
            t.numpy()[:] = a
            pinned[k] = t.numpy()
        del host
        try:
            rank_host = torch.empty(n, dtype=torch.float64).pin_memory().numpy()
        except Exception:
            rank_host = np.empty(n, np.float64); pin_ok = False
        h2d = sum(a.nbytes for a in pinned.values())
        e2e_steps = max(1, min(args.steps, 2))
        times, parts = [], {}
        for k in range(1 + e2e_steps):
            dist.barrier(); lib.sync()
            t0 = time.perf_counter()
            dg = lib.graph_pull_only(pinned["begin"], pinned["r_begin"], pinned["r_node_idx"], comm=comm)
            t1 = time.perf_counter()
            out = torch.empty(n, dtype=torch.float64, device="cuda")
            dg.pagerank_sharded_dev(comm, out.data_ptr(), 1e-30, 0.85, iters)
            t2 = time.perf_counter()
            lib.d2h(rank_host, out.data_ptr())
            t3 = time.perf_counter()
            comm.barrier()          # nobody unmaps / frees its state block while a peer may still store into it
            dg.close()
            lib.sync(); dist.barrier()
            t4 = time.perf_counter()
            if k > 0:
                times.append(max_over_ranks(dist, torch, (t4 - t0) * 1e3))
                for name, dt in (("copy_in_and_allgather", t1 - t0), ("plan_peer_mapping_sweeps", t2 - t1), ("d2h_rank", t3 - t2), ("teardown", t4 - t3)):
                    parts.setdefault(name, []).append(max_over_ranks(dist, torch, dt * 1e3))
        e2e_ms = float(np.mean(times))
        e2e = {"value": round(m * iters / e2e_ms / 1e6, 3), "unit": "GTEPS", "h2d_bytes_per_step": int(h2d),
               "d2h_bytes_per_step": int(rank_host.nbytes * world), "ms_per_step": round(e2e_ms, 3), "steps": e2e_steps,
               "pinned": pin_ok,
               "breakdown_ms_max_over_ranks": {k: round(float(np.mean(v)), 2) for k, v in parts.items()}, "note": "each rank copies its 1/world slice of the CSR host->device, the slices are allgathered over NVLink, every rank plans its shard; all inside the timed region"}

    # ---- the same e2e step through the DEVICE GROUP (gmc_group_*, what GM_CUDA_DEVICES=N selects for the reference's unchanged
    # drivers): ONE process (rank 0 here; the other ranks of the launcher idle) drives all N GPUs -- every device copies 1/N of the
    # host CSR over its own PCIe link, peers through plain pointers, each device copies its slice of rank[] back
    group_e2e = None
    if e2e is not None and not os.environ.get("GMC_BENCH_NO_GROUP"):
        torch.cuda.synchronize()
        dist.barrier(group=cpu_group)        # gloo: the idle ranks wait on the host while rank 0's process uses every GPU
        if rank == 0:
            try:
                sharded_result = rank_host.copy()
                grp = lib.group(world)
                gtimes = []
                for k in range(1 + e2e_steps):
                    t0 = time.perf_counter()
                    grp.upload(pinned["begin"], None, pinned["r_begin"], pinned["r_node_idx"])
                    t1 = time.perf_counter()
                    _, ginfo = grp.pagerank(1e-30, 0.85, iters, out=rank_host)
                    t2 = time.perf_counter()
                    if k > 0:
                        gtimes.append(((t2 - t0) * 1e3, (t1 - t0) * 1e3, (t2 - t1) * 1e3))
                gms = float(np.mean([t[0] for t in gtimes]))
                group_e2e = {"value": round(m * iters / gms / 1e6, 3), "unit": "GTEPS", "ms_per_step": round(gms, 3), "steps": e2e_steps,
                             "copy_in_ms": round(float(np.mean([t[1] for t in gtimes])), 2),
                             "plan_sweeps_copy_out_ms": round(float(np.mean([t[2] for t in gtimes])), 2),
                             "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(rank_host.nbytes), "devices": grp.size(),
                             "exchange_path": grp.exchange_path(), "iterations": int(ginfo.iterations),
                             "max_abs_vs_torchrun_ranks": float(np.abs(rank_host - sharded_result).max()),
                             "what": "gmc_group_graph_upload + gmc_group_pagerank from ONE process over all devices (host CSR in, host rank[] out)"}
                grp.close()
                del sharded_result
            except Exception as ex:
                group_e2e = {"error": str(ex)[:300]}
        dist.barrier(group=cpu_group)

    if rank == 0:
        line = {
            "metric": "pagerank_gteps", "value": round(value, 3), "unit": "GTEPS", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": round(ms_step, 4), "higher_is_better": True, "scaling": "strong",
            "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": workload_name(args, n, m, workload), "nodes": n, "edges": m, "iterations_per_step": iters,
                       "graph": "device R-MAT generator seed 1997 a=.57 b=.19 c=.19 permuted (counter-based RNG), identical on every rank",
                       "generate_s": round(gen_s, 2), "first_calls_s": round(warm_s, 2),
                       "l2": "inputs larger than L2 (index stream %d MB per rank per sweep); no explicit flush" % (4 * m // world // 2**20),
                       "parallelism": f"{world} ranks x 1 GPU, destination-range shards, contributions exchanged per sweep (" + exchange_path + ")"},
            "clocks": clk, "e2e": e2e, "gpu_launches": int(launches), "roofline": roofline, "cpu_baseline": None, "bfs": bfs,
            "parity": parity, "single_gpu_same_graph": single, "e2e_device_group": group_e2e,
            "nccl_version": comm.nccl_version(),
        }
        import __main__ as _main           # bench.py, run as a script: its emit() owns the real stdout
        _main.emit(line) if hasattr(_main, "emit") else print(json.dumps(line), flush=True)
    comm.barrier()
    comm.close()
    dist.destroy_process_group()
