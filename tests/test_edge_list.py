"""EDGE_LIST ingest: the oracle's restatement of gm_edge_list_graph_reader::builtGraph against the
reference reader itself (when oracle/_ref is present) and a committed fixture; the product path (host-parallel parse +
device CSR build) against the oracle on the GPU."""
import os
import numpy as np
import pytest

import graphs
from oracle_lib import Ref, have_ref

HERE = os.path.dirname(os.path.abspath(__file__))
FIXTURE = os.path.join(HERE, "golden", "edge_list_small.txt")


def write_edge_list(path, n_raw, m, seed, sorted_by_src=True):
    """sparse raw ids (gaps), duplicates, self loops, comments, blank lines, tabs, '*' lines, CRLF on some lines"""
    rs = np.random.RandomState(seed)
    ids = np.sort(rs.choice(n_raw, size=max(2, n_raw // 3), replace=False))
    src = ids[rs.randint(0, len(ids), m)]
    dst = ids[rs.randint(0, len(ids), m)]
    if sorted_by_src:
        o = np.argsort(src, kind="stable")
        src, dst = src[o], dst[o]
    lone = ids[rs.randint(0, len(ids), 7)]
    lines = []
    for k, (a, b) in enumerate(zip(src, dst)):
        sep = "\t" if k % 3 == 0 else " "
        end = "\r\n" if k % 11 == 0 else "\n"
        lines.append((int(a), f"{a}{sep}{b}{end}" + ("# a comment line\n" if k % 97 == 0 else "")))
    lines += [(int(a), f"{a} *\n") for a in lone]
    if sorted_by_src:          # the reference reader asserts ascending sources over ALL lines (:145)
        lines.sort(key=lambda t: t[0])
    with open(path, "w") as f:
        f.write("# generated test input\n\n")
        for _, t in lines:
            f.write(t)
    return path


def test_oracle_matches_fixture(oracle):
    z = np.load(os.path.join(HERE, "golden", "edge_list_small.npz"))
    b, i = oracle.edge_list_read(FIXTURE)
    assert np.array_equal(b, z["begin"]) and np.array_equal(i, z["node_idx"])


@pytest.mark.skipif(not have_ref(), reason="oracle/_ref not built (no /root/reference on this box)")
def test_oracle_matches_reference_reader(oracle, tmp_path):
    ref = Ref()
    for seed, (n_raw, m) in enumerate([(50, 200), (3000, 20000), (10, 0)]):
        p = write_edge_list(str(tmp_path / f"el{seed}.txt"), n_raw, m, seed)
        g = ref.load_edge_list(p)
        assert g
        a = ref.arrays(g)
        b, i = oracle.edge_list_read(p)
        assert np.array_equal(b, a["begin"]) and np.array_equal(i, a["node_idx"])
        ref.free(g)


@pytest.mark.gpu
def test_edge_list_device_build(gm, oracle, tmp_path):
    for seed, (n_raw, m, srt) in enumerate([(50, 200, True), (30000, 400000, True), (30000, 400000, False), (200000, 3000000, True)]):
        p = write_edge_list(str(tmp_path / f"el{seed}.txt"), n_raw, m, seed, srt)      # > 1 MB: parsed by all threads
        b, i = oracle.edge_list_read(p)
        g = gm.edge_list_read(p, make_reverse=True)
        assert g.num_nodes == len(b) - 1 and g.num_edges == len(i)
        assert np.array_equal(g.array("begin"), b) and np.array_equal(g.array("node_idx"), i)
        rb, ri = oracle.make_reverse(b, i)
        assert np.array_equal(g.array("r_begin"), rb) and np.array_equal(g.array("r_node_idx"), ri)
        g.close()
    g = gm.edge_list_read(FIXTURE)
    z = np.load(os.path.join(HERE, "golden", "edge_list_small.npz"))
    assert np.array_equal(g.array("begin"), z["begin"]) and np.array_equal(g.array("node_idx"), z["node_idx"])


@pytest.mark.gpu
def test_edge_list_errors(gm, tmp_path):
    GmcError = Exception        # gmcuda.GmcError (RuntimeError): the package is loaded by path, not importable by name
    p = str(tmp_path / "bad.txt")
    open(p, "w").write("1 2\n3\n")
    with pytest.raises(GmcError):
        gm.edge_list_read(p)
    open(p, "w").write("1 2\n3 x\n")
    with pytest.raises(GmcError):
        gm.edge_list_read(p)
    with pytest.raises(GmcError):
        gm.edge_list_read(str(tmp_path / "missing.txt"))
    open(p, "w").write("# nothing\n\n")
    g = gm.edge_list_read(p)
    assert g.num_nodes == 0 and g.num_edges == 0
