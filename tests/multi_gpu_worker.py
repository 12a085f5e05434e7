"""Worker of test_two_rank_nccl_run / gpurun --gpus N checks: every rank builds the same graph, runs the sharded
PageRank / hop_dist / BFS levels through the C ABI and compares with its own single-GPU run of the same calls."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from gmb200_loader import load_package  # noqa: E402


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ.get("LOCAL_RANK", "0"))
    n, m = (int(sys.argv[1]), int(sys.argv[2])) if len(sys.argv) > 2 else (300000, 4000000)
    dist.init_process_group("gloo")
    lib = load_package().GmCuda()
    lib.init(local)
    torch.cuda.set_device(local)
    uid = torch.from_numpy(lib.comm_unique_id() if rank == 0 else np.zeros(128, np.uint8))
    dist.broadcast(uid, 0)
    comm = lib.comm_create(world, rank, uid.numpy())
    g = lib.generate_rmat(n, m, 1997)
    out = torch.empty(n, dtype=torch.float64, device="cuda")
    info = g.pagerank_sharded_dev(comm, out.data_ptr(), 1e-30, 0.85, 20)
    sharded = out.cpu().numpy()
    path = g.pagerank_exchange_path()
    if os.environ.get("GMC_NO_P2P"):
        assert path == "nccl"
    one, info1 = g.pagerank(1e-30, 0.85, 20)
    assert info.iterations == info1.iterations == 20
    err = float(np.abs(sharded - one).max())
    assert err <= 1e-12, f"rank {rank}: sharded pagerank differs from single-GPU by {err}"
    # every rank must hold the same bits
    ref = torch.from_numpy(sharded.copy()); dist.broadcast(ref, 0)
    assert np.array_equal(ref.numpy(), sharded), "ranks disagree on the PageRank vector"
    deg = np.diff(g.array("begin"))
    for root in [int(np.nonzero(deg)[0][0]), int(np.argmax(deg))]:
        d_dev = torch.empty(n, dtype=torch.int32, device="cuda")
        bi = g.hop_dist_sharded_dev(comm, d_dev.data_ptr(), root)
        d_one, bi1 = g.hop_dist(root)
        d_sh = d_dev.cpu().numpy()
        if not np.array_equal(d_sh, d_one):
            bad = np.nonzero(d_sh != d_one)[0]
            lev_s, _, tr_s, cnt_s = g.bfs_levels_sharded(comm, root)
            lev_1, _, tr_1, cnt_1 = g.bfs_levels(root)
            raise AssertionError(f"rank {rank}: sharded hop_dist differs (root {root}): {len(bad)} vertices, first {bad[:8]}, "
                                 f"sharded {d_sh[bad[:8]]} single {d_one[bad[:8]]}; levels {bi.levels}/{bi1.levels} reached {bi.reached}/{bi1.reached}; "
                                 f"trace sharded {list(tr_s)} {list(cnt_s)} single {list(tr_1)} {list(cnt_1)}")
        assert bi.levels == bi1.levels and bi.reached == bi1.reached
        lev, li, trace, cnt = g.bfs_levels_sharded(comm, root)
        lev1, li1, trace1, cnt1 = g.bfs_levels(root)
        assert np.array_equal(lev, lev1) and list(trace) == list(trace1) and list(cnt) == list(cnt1)
    # two copies of the graph joined by a 40-vertex path: the traversal shrinks to one vertex and grows again, so the sharded
    # loop goes small levels (run in full by every rank) -> sharded levels -> small -> sharded -> small
    import graphs
    b2, i2, hub = graphs.two_blobs(g.array("begin"), g.array("node_idx"))
    g3 = lib.graph_from_csr(b2, i2, semi_sort=True, make_reverse=True)
    for root in (hub, len(b2) - 2):
        d_dev = torch.empty(len(b2) - 1, dtype=torch.int32, device="cuda")
        bi = g3.hop_dist_sharded_dev(comm, d_dev.data_ptr(), root)
        d_one, bi1 = g3.hop_dist(root)
        assert np.array_equal(d_dev.cpu().numpy(), d_one) and bi.levels == bi1.levels, f"rank {rank}: two-blob hop_dist differs (root {root})"
        lev, li, trace, cnt = g3.bfs_levels_sharded(comm, root)
        lev1, li1, trace1, cnt1 = g3.bfs_levels(root)
        assert np.array_equal(lev, lev1) and list(trace) == list(trace1) and list(cnt) == list(cnt1), f"rank {rank}: two-blob levels differ"
    comm.barrier()
    g3.close()
    # triangle counting: slot ranges per rank + one allreduce == the single-GPU count, on every rank
    t_all, _ = g.triangle_counting_sharded(comm)
    t_one, _ = g.triangle_counting()
    assert t_all == t_one and t_one > 0, f"rank {rank}: sharded triangle count {t_all} != {t_one}"
    # sharded copy-in: slices uploaded by their ranks + NVLink allgather reproduce the arrays on every rank
    host = {k: g.array(k) for k in ("begin", "node_idx", "r_begin", "r_node_idx")}
    g2 = lib.graph_from_csr_sharded(comm, host["begin"], host["node_idx"], host["r_begin"], host["r_node_idx"])
    for k, a in host.items():
        assert np.array_equal(g2.array(k), a), f"rank {rank}: sharded upload of {k} differs"
    comm.barrier()
    g2.close()
    comm.barrier()
    if rank == 0:
        print(f"multi_gpu_worker ok: world {world}, exchange {path}, pagerank max|sharded-single| = {err:.3e}, nccl {comm.nccl_version()}", flush=True)
    comm.close()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
