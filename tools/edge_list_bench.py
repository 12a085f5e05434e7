"""EDGE_LIST ingest timing: libgm_cuda (host-parallel parse + device CSR build) vs the CPU restatement
of the reference reader (oracle, 3 sequential passes like the reference) on the same file.  Prints one JSON line."""
import json, os, sys, time
import numpy as np, pandas as pd
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
from gmb200_loader import load_package
from oracle_lib import Oracle
n, m = (int(sys.argv[1]), int(sys.argv[2])) if len(sys.argv) > 2 else (1000000, 16000000)
L = load_package().GmCuda(); L.init(0)
g = L.generate_rmat(n, m, 1997)
b, idx = g.array("begin"), g.array("node_idx")
src = np.repeat(np.arange(n, dtype=np.int32), np.diff(b))
path = "/tmp/edge_list_bench.txt"
pd.DataFrame({"s": src, "d": idx}).to_csv(path, sep=" ", header=False, index=False)
size = os.path.getsize(path)
ts = []
for _ in range(3):
    t = time.perf_counter(); h = L.edge_list_read(path, make_reverse=True); ts.append(time.perf_counter() - t)
    nn, mm = h.num_nodes, h.num_edges
    if _ < 2: h.close()
o = Oracle()
t = time.perf_counter(); ob, oi = o.edge_list_read(path); t_cpu = time.perf_counter() - t
ok = bool(np.array_equal(h.array("begin"), ob) and np.array_equal(h.array("node_idx"), oi))
print(json.dumps({"what": "EDGE_LIST ingest", "file_bytes": size, "edges": int(mm), "vertices": int(nn),
                  "gm_cuda_s": round(min(ts), 3), "gm_cuda_Medges_per_s": round(mm / min(ts) / 1e6, 1),
                  "host_threads": os.cpu_count(), "cpu_port_s": round(t_cpu, 3), "cpu_port_Medges_per_s": round(mm / t_cpu / 1e6, 1),
                  "identical_csr": ok}))
