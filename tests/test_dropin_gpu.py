"""The drop-in check: the reference's UNCHANGED drivers -- apps/output_cpp/src/pagerank_main.cc,
hop_dist_main.cc, gm_graph/src/graph_gen_main.cc, compiled from /root/reference against include/gm and linked to
libgm_cuda.so (green-marl_b200/Makefile, target apps; the binaries travel to the GPU box prebuilt) -- run with the
reference's command lines and print the reference's result lines with the oracle's values."""
import os
import re
import subprocess

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BIN = os.path.join(ROOT, "green-marl_b200", "bin")


def run(exe, *args, env=None):
    path = os.path.join(BIN, exe)
    assert os.path.exists(path), f"{path} missing: run __graft_entry__.build() where /root/reference is present"
    r = subprocess.run([path] + [str(a) for a in args], capture_output=True, text=True, timeout=300,
                       env=dict(os.environ, **env) if env else None)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    return r.stdout


def test_reference_drivers_run_unchanged(gm, oracle, tmp_path):
    f = str(tmp_path / "c1.bin")
    # graph_gen <N> <M> <file> <type 0 = uniform multigraph>, seed 1997 (graph_gen_main.cc:20-61)
    run("graph_gen", 65536, 1048576, f, 0)
    b, i = oracle.uniform_random(65536, 1048576, 1997)
    want = str(tmp_path / "want.bin")
    assert oracle.store_binary(want, b, i)                     # the generator stores the rows as filled, unsorted
    assert open(f, "rb").read() == open(want, "rb").read(), "graph_gen output differs from the reference's file"
    i = np.ascontiguousarray(i).copy()
    oracle.semi_sort(b, i)                                     # load_binary ends with do_semi_sort
    rb, ri = oracle.make_reverse(b, i)

    # pagerank <graph> <threads> <nfs> [max_iter] [e] [d]   (pagerank_main.cc:11-16,50-62)
    out = run("pagerank", f, 4, "/dev/null", 20, 1e-30, 0.85)
    rank, it, _ = oracle.pagerank(b, rb, ri, 1e-30, 0.85, 20)
    got = [float(x) for x in re.findall(r"rank\[\d\] = ([0-9.]+)", out)]
    assert len(got) == 4 and it == 20
    assert got == [float("%0.9f" % v) for v in rank[:4]]
    assert "running time=" in out and "reverse edge creation time=" in out and "CRC shl__G_pg_rank" in out

    # hop_dist <graph> <threads> <nfs>: root 0, prints dist[0..9]  (hop_dist_main.cc:21,37)
    out = run("hop_dist", f, 4, "/dev/null")
    dist, _ = oracle.hop_dist(b, i, 0)
    got = [int(x) for x in re.findall(r"dist\[\d\] = (-?\d+)", out)]
    assert got == [int(v) for v in dist[:10]]


def test_triangle_counting_driver_unchanged(gm, oracle, tmp_path):
    """apps/output_cpp/src/triangle_counting_main.cc, unchanged: prints `number of triangles: <T>`"""
    f = str(tmp_path / "r.bin")
    b, i = oracle.rmat(20000, 400000, 1997, 0.57, 0.19, 0.19, True)
    assert oracle.store_binary(f, b, i)
    i = np.ascontiguousarray(i).copy()
    oracle.semi_sort(b, i)
    out = run("triangle_counting", f, 4, "/dev/null")
    got = re.findall(r"number of triangles: (-?\d+)", out)
    assert got and int(got[0]) == oracle.triangle_counting(b, i)


@pytest.mark.parametrize("ndev", [2, 4, 8])
def test_reference_drivers_on_a_device_group(gm, oracle, tmp_path, ndev):
    """GM_CUDA_DEVICES=N: the same unchanged drivers over N GPUs of the box (device group selected in shl__init, the
    reference's placement switch, common_main.h:95-99): same printed values as the oracle"""
    if gm.device_count() < ndev:
        pytest.skip(f"needs {ndev} GPUs")
    f = str(tmp_path / "r.bin")
    b, i = oracle.rmat(200000, 3000000, 1997, 0.57, 0.19, 0.19, True)
    assert oracle.store_binary(f, b, i)
    i = np.ascontiguousarray(i).copy()
    oracle.semi_sort(b, i)
    rb, ri = oracle.make_reverse(b, i)
    env = {"GM_CUDA_DEVICES": str(ndev), "GMC_PR_TIER_MIN_EDGES": "0"}
    out = run("pagerank", f, 4, "/dev/null", 20, 1e-30, 0.85, env=env)
    rank, it, _ = oracle.pagerank(b, rb, ri, 1e-30, 0.85, 20)
    got = [float(x) for x in re.findall(r"rank\[\d\] = ([0-9.]+)", out)]
    assert got == [float("%0.9f" % v) for v in rank[:4]]
    assert f"devices={ndev}" in out and "device group" in out
    out = run("hop_dist", f, 4, "/dev/null", env=env)
    dist, _ = oracle.hop_dist(b, i, 0)
    got = [int(x) for x in re.findall(r"dist\[\d\] = (-?\d+)", out)]
    assert got == [int(v) for v in dist[:10]]
    assert f"devices={ndev}" in out


def test_shoal_placement_switches(gm, oracle, tmp_path):
    """scripts/run.sh -r / -p export SHL_REPLICATION / SHL_PARTITION: shl__init maps them to the device group over every visible GPU
    (one GPU: a no-op), prints the settings line and the Array[...] configuration lines scripts/extract_result.py parses"""
    f = str(tmp_path / "c1.bin")
    run("graph_gen", 65536, 1048576, f, 0)
    have = gm.device_count()
    ndev = 1
    while ndev * 2 <= have and ndev * 2 <= 8:
        ndev *= 2
    out = run("pagerank", f, 4, "/dev/null", 20, 1e-30, 0.85, env={"SHL_REPLICATION": "1", "SHL_PARTITION": "1", "SHL_HUGEPAGE": "1"})
    assert f"SHOAL settings: replication=1 partition=1 distribution=0 hugepage=1 -> devices={ndev}" in out
    assert re.search(r"^Array\[\s*shl__G_pg_rank\]: elements=\s*65536 size=\s*524288 ", out, re.M)
    b, i = oracle.uniform_random(65536, 1048576, 1997)
    i = np.ascontiguousarray(i).copy()
    oracle.semi_sort(b, i)
    rb, ri = oracle.make_reverse(b, i)
    rank, _, _ = oracle.pagerank(b, rb, ri, 1e-30, 0.85, 20)
    got = [float(x) for x in re.findall(r"rank\[\d\] = ([0-9.]+)", out)]
    assert got == [float("%0.9f" % v) for v in rank[:4]]
    if ndev > 1:
        assert "replication=[X]" in out and "partition=[X]" in out
