"""Summarise ncu outputs brought back in gpurun_out/ into small text files under profiles/.
  python tools/ncu_summary.py launches <launches.csv> <out.md>     per-kernel totals / shares of a launch list
  python tools/ncu_summary.py full <report.ncu-rep> <out.md>       key metrics of a --set full capture
"""
import csv
import subprocess
import sys
from collections import defaultdict

KEYS = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "lts__t_sector_hit_rate.pct", "lts__t_sectors.sum", "lts__throughput.avg.pct_of_peak_sustained_elapsed",
        "l1tex__t_sector_hit_rate.pct", "l1tex__throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum", "launch__registers_per_thread",
        "launch__grid_size", "launch__block_size", "launch__shared_mem_per_block_dynamic", "launch__occupancy_limit_shared_mem",
        "launch__occupancy_limit_registers", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "sm__cycles_elapsed.max",
        "sm__inst_executed_pipe_fp64.sum", "smsp__average_warp_latency_issue_stalled_long_scoreboard.ratio",
        "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio", "smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_lg_throttle_per_issue_active.ratio", "smsp__average_warps_issue_stalled_mio_throttle_per_issue_active.ratio"]


def launches(path, out):
    rows = list(csv.reader(open(path)))
    h = next(i for i, r in enumerate(rows) if r and r[0] == "ID")
    hdr = rows[h]
    ik, iv = hdr.index("Kernel Name"), hdr.index("Metric Value")
    t, c = defaultdict(float), defaultdict(int)
    for r in rows[h + 1:]:
        if len(r) <= iv:
            continue
        try:
            v = float(r[iv].replace(",", ""))
        except ValueError:
            continue
        name = r[ik].split("(")[0].replace("<unnamed>::", "")
        t[name] += v
        c[name] += 1
    tot = sum(t.values())
    with open(out, "w") as f:
        f.write(f"# ncu launch list summary of {path}\n\n(gpu__time_duration.sum, --clock-control none; cold-cache and serialised: compare shares)\n\n")
        f.write("| kernel | launches | total us | avg us | share |\n|---|---:|---:|---:|---:|\n")
        for k in sorted(t, key=lambda k: -t[k]):
            f.write(f"| `{k[:70]}` | {c[k]} | {t[k]/1e3:.1f} | {t[k]/1e3/c[k]:.2f} | {100*t[k]/tot:.1f}% |\n")
    print(open(out).read())


def full(path, out):
    txt = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(txt.splitlines()))
    hdr, units = rows[0], rows[1]
    ik = hdr.index("Kernel Name")
    with open(out, "w") as f:
        f.write(f"# ncu --set full summary of {path}\n\n")
        for r in rows[2:]:
            f.write(f"## {r[ik].split('(')[0]}  (id {r[0]})\n\n| metric | unit | value |\n|---|---|---:|\n")
            for k in KEYS:
                if k in hdr:
                    i = hdr.index(k)
                    f.write(f"| {k} | {units[i]} | {r[i]} |\n")
            f.write("\n")
    print(open(out).read())


if __name__ == "__main__":
    {"launches": launches, "full": full}[sys.argv[1]](sys.argv[2], sys.argv[3])
