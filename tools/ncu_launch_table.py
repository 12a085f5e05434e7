"""ncu --csv launch list (gpu__time_duration.sum, dram__bytes_read.sum, dram__bytes_write.sum, lts__t_sectors.sum, ...) -> one
markdown row per launch + shares per kernel.  usage: ncu_launch_table.py in.csv out.md "title" """
import collections, csv, sys
src, dst, title = sys.argv[1], sys.argv[2], (sys.argv[3] if len(sys.argv) > 3 else "launch list")
lines = [l for l in open(src) if l.startswith('"')]
L = collections.OrderedDict()
for x in csv.DictReader(lines):
    d = L.setdefault(int(x["ID"]), {"name": x["Kernel Name"].split("(")[0].replace("void ", "").replace("<unnamed>::", ""), "grid": x["Grid Size"], "block": x["Block Size"]})
    d[x["Metric Name"]] = (float(x["Metric Value"].replace(",", "")), x["Metric Unit"])
def nbytes(d, m):
    v, u = d.get(m, (0, "byte"))
    return v * {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}.get(u, 1)
out = ["# " + title, "", "Times are serialised and cold-cache (ncu --clock-control none): read the shares.", "",
       "| # | kernel | grid | block | us | dram MB | L2 sectors |", "|---:|---|---|---|---:|---:|---:|"]
tot = collections.Counter()
for k, d in L.items():
    t = d["gpu__time_duration.sum"]; us = t[0] / 1000 if t[1] == "ns" else (t[0] if t[1] == "us" else t[0] * 1000)
    mb = (nbytes(d, "dram__bytes_read.sum") + nbytes(d, "dram__bytes_write.sum")) / 1e6
    out.append(f"| {k} | `{d['name']}` | {d['grid']} | {d['block']} | {us:.1f} | {mb:.2f} | {d.get('lts__t_sectors.sum', (0, ''))[0]:.0f} |")
    tot[d["name"]] += us
out += ["", "| kernel | total us | share |", "|---|---:|---:|"]
s = sum(tot.values())
for nme, t in tot.most_common(): out.append(f"| `{nme}` | {t:.1f} | {100 * t / s:.1f}% |")
out.append(f"| all | {s:.1f} | |")
open(dst, "w").write("\n".join(out) + "\n")
print("\n".join(out))
