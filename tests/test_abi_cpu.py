"""CPU suite: the C-ABI library loads and exports every symbol include/*.h declares; the ctypes binding covers
all of them; compute entry points fail loudly (no CPU fallback) when there is no CUDA device."""
import ctypes
import os
import re

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    txt = open(os.path.join(ROOT, "include", "gm_cuda.h")).read()
    txt = re.sub(r"/\*.*?\*/", "", txt, flags=re.S)
    return sorted(set(re.findall(r"\b(gmc_[a-z0-9_]+)\s*\(", txt)))


def test_library_exports_every_declared_symbol():
    from gmb200_loader import load_package
    pkg = load_package()
    lib = ctypes.CDLL(pkg.lib_path())
    syms = declared_symbols()
    assert len(syms) >= 35
    missing = [s for s in syms if not hasattr(lib, s)]
    assert not missing, missing


def test_binding_covers_header():
    from gmb200_loader import load_package
    pkg = load_package()
    from green_marl_b200 import gmcuda
    assert sorted(gmcuda.SIGNATURES) == declared_symbols()
    pkg.GmCuda()          # binds all of them


def test_cpp_facade_symbols_present():
    """the C++ faces of SURVEY.md 8b that the reference's unchanged drivers link against"""
    import subprocess
    from gmb200_loader import load_package
    out = subprocess.run(["nm", "-D", "--defined-only", "-C", load_package().lib_path()], capture_output=True, text=True).stdout
    for needle in ["pagerank(gm_graph&, double, double, int, double*)", "hop_dist(gm_graph&, int*, int&)",
                   "gm_graph::make_reverse_edges()", "gm_graph::load_binary(char*)", "gm_graph::store_binary(char*)",
                   "gm_graph::do_semi_sort()", "gm_graph::prepare_external_creation(int, int)", "shl__init",
                   "gm_rt_initialize", "gm_rt_set_num_threads", "create_uniform_random_graph", "create_RMAT_graph",
                   "_gm_set_bit_atomic"]:
        assert needle in out, needle


@pytest.mark.skipif(os.path.exists("/dev/nvidiactl"), reason="a GPU is present")
def test_no_cpu_fallback():
    from gmb200_loader import load_package
    pkg = load_package()
    lib = pkg.GmCuda()
    with pytest.raises(pkg.GmcError):
        lib.init(0)
    with pytest.raises(pkg.GmcError):
        lib.graph_from_csr(np.array([0, 1, 2], np.int32), np.array([1, 0], np.int32))


def test_binary_file_io_through_abi(oracle, tmp_path):
    """gmc_file_write / gmc_file_read (host-side, no GPU): byte-identical to the reference format as written by the oracle"""
    import graphs
    from gmb200_loader import load_package
    lib = load_package().GmCuda()
    g = graphs.rmat(oracle, 3000, 40000, seed=3)
    p1, p2 = str(tmp_path / "a.bin"), str(tmp_path / "b.bin")
    assert oracle.store_binary(p1, g["begin"], g["node_idx"])
    lib.file_write(p2, g["begin"], g["node_idx"])
    assert open(p1, "rb").read() == open(p2, "rb").read()
    b, i = lib.file_read(p1)
    assert np.array_equal(b, g["begin"]) and np.array_equal(i, g["node_idx"])
    with pytest.raises(load_package().GmcError):
        lib.file_read(str(tmp_path / "missing.bin"))
    open(str(tmp_path / "bad.bin"), "wb").write(b"\x00" * 64)
    with pytest.raises(load_package().GmcError):
        lib.file_read(str(tmp_path / "bad.bin"))


def test_binary_file_rejects_corrupt_csr(oracle, tmp_path):
    """a file whose offsets decrease, do not end at M, or whose neighbour ids leave [0, N) is an I/O error, not a kernel fault"""
    import graphs
    from gmb200_loader import load_package
    pkg = load_package()
    lib = pkg.GmCuda()
    g = graphs.rmat(oracle, 3000, 40000, seed=3)
    for what in ("decreasing", "bad_end", "bad_id"):
        b, i = g["begin"].copy(), g["node_idx"].copy()
        if what == "decreasing":
            b[100], b[101] = b[101] + 5, b[100]
        elif what == "bad_end":
            b[-1] -= 1
        else:
            i[1234] = len(b) + 7
        p = str(tmp_path / (what + ".bin"))
        lib.file_write(p, b, i)
        with pytest.raises(pkg.GmcError):
            lib.file_read(p)


def test_shoal_settings_file(tmp_path):
    """shl__load_settings reads the file scripts/generate_settings.py:53-75 writes (same text, produced here as run.sh:264-267
    would with -R 1 -P 1 -H 0 -T 1 -S 1 -W 4096 -A 0 -C 1); names it does not know and the dma_device blocks are skipped"""
    from gmb200_loader import load_package
    lib = ctypes.CDLL(load_package().lib_path())

    class Settings(ctypes.Structure):
        _fields_ = [(k, ctypes.c_int) for k in ("distribution", "replication", "partitioning", "hugepage", "trim", "stride", "crc",
                                                "static_schedule", "dma_enable")]
    lib.shl__load_settings.argtypes = [ctypes.c_char_p, ctypes.POINTER(Settings)]
    text = ("-- global configuration\nglobal = {\n    distribution = 0,\n    replication = 1,\n    partitioning = 1,\n    hugepage = 0,\n"
            "    trim = 1,\n    stride = 4096,\n    crc = 1,\n    static = 1\n}\n-- dma configuration\ndma = {\n    enable = 1,\n    vendor = 32902,\n"
            "    device = 3616,\n    device_count = 2\n}\ndma_device_0 = {\n    count=8, \n    bus=0, \n    dev=4, \n    fun=0\n}\n")
    path = tmp_path / "shl__settings.lua"
    path.write_text(text)
    s = Settings()
    assert lib.shl__load_settings(str(path).encode(), ctypes.byref(s)) == 0
    assert (s.distribution, s.replication, s.partitioning, s.hugepage, s.trim, s.stride, s.crc, s.static_schedule, s.dma_enable) == \
        (0, 1, 1, 0, 1, 4096, 1, 1, 1)
    assert lib.shl__load_settings(str(tmp_path / "missing.lua").encode(), ctypes.byref(s)) == -1
    assert (s.replication, s.stride, s.crc) == (0, 4096, 1)          # the generator's defaults
