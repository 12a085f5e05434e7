"""CPU check of the PageRank sweep plan builder (host logic, no GPU): green-marl_b200/csrc/gmc_pr_plan.h is written against a
backend interface, so tests/cpp/pr_plan_check.cc runs its very steps as host loops -- with tiers shrunk to 64 slots so that small
graphs exercise tier streams, segments, part slots and shard cuts -- and checks (1) every reverse edge appears exactly once across the
streams, tasks cover their streams and obey the 32-rows / 31-slices limits, (2) 20 sweeps THROUGH the plan equal the plain pull loop
of apps/src/pagerank.gm:8-19 to 1e-12."""
import os
import subprocess
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def checker(tmp_path_factory):
    out = tmp_path_factory.mktemp("prplan")
    exes = {}
    for name, flags in (("small", ["-DPRP_HT=64", "-DPRP_S_MAX=128"]), ("mid", ["-DPRP_HT=256", "-DPRP_S_MAX=512"])):
        exe = str(out / ("pr_plan_check_" + name))
        subprocess.run(["g++", "-O2", "-fopenmp", "-std=c++17"] + flags + ["-I", os.path.join(ROOT, "green-marl_b200", "csrc"),
                        os.path.join(ROOT, "tests", "cpp", "pr_plan_check.cc"), "-o", exe], check=True)
        exes[name] = exe
    return exes


@pytest.mark.parametrize("args", ["12 65536 3 3 1", "12 65536 8 2 1", "14 400000 16 3 1", "14 400000 31 3 2", "13 200000 5 1 3",
                                  "15 1000000 12 2 4 11", "10 3000 4 3 1", "12 65536 0 3 2", "8 300 31 1 1"])
def test_plan_small_tiers(checker, args):
    for split in ("1", "0"):          # SPLIT mode (tier 0 as a segment stream) and MIXED mode with warm tiers
        r = subprocess.run([checker["small"]] + args.split(), stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=600,
                           env=dict(os.environ, PRP_SPLIT=split))
        assert r.returncode == 0 and r.stdout.strip().endswith("OK"), r.stdout[-3000:]


def test_plan_chunked_edge_arrival(checker):
    """pass 1 runs chunk by chunk behind the copy-in of the edges (gmc_pagerank_host_csr): any chunking gives the same plan"""
    for chunks in ("3", "8"):
        env = dict(os.environ, PRP_CHUNKS=chunks)
        r = subprocess.run([checker["small"]] + "14 400000 16 3 2".split(), stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=600, env=env)
        assert r.returncode == 0 and r.stdout.strip().endswith("OK"), r.stdout[-3000:]


def test_plan_big_tasks(checker):
    env = dict(os.environ, PRP_BIGTASKS="1")
    r = subprocess.run([checker["mid"], "16", "3000000", "16", "3", "1"], stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=900, env=env)
    assert r.returncode == 0 and r.stdout.strip().endswith("OK"), r.stdout[-3000:]
    r = subprocess.run([checker["mid"], "15", "1500000", "8", "3", "8"], stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=900, env=env)
    assert r.returncode == 0 and r.stdout.strip().endswith("OK"), r.stdout[-3000:]
