#!/usr/bin/env python
"""Turn ncu output into the small tables committed under profiles/.

  python profiles/summarize.py launches gpurun_out/launches_rX.csv  > profiles/rX_launches.md
  python profiles/summarize.py full gpurun_out/prof_rX.ncu-rep [kernel-substring] > profiles/rX_<kernel>_full.md

`launches` reads the CSV of `ncu --metrics gpu__time_duration.sum --clock-control none --csv`;
`full` reads a `--set full` report through `ncu -i ... --page raw --csv` (ncu must be on PATH)."""
import csv
import io
import re
import subprocess
import sys
from collections import OrderedDict

KEYS = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "lts__t_bytes.sum",
        "dram__throughput.avg.pct_of_peak_sustained_elapsed", "lts__t_sector_hit_rate.pct",
        "l1tex__t_sector_hit_rate.pct", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread", "launch__grid_size",
        "launch__block_size", "launch__shared_mem_per_block_static", "launch__occupancy_limit_registers",
        "launch__waves_per_multiprocessor", "smsp__inst_executed.sum", "sm__inst_executed_pipe_fp64.sum",
        "smsp__average_warp_latency_issue_stalled_long_scoreboard.pct",
        "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio"]


def short(name):
    name = re.sub(r"<unnamed>::", "", name)
    return re.sub(r"\(.*", "", name)


def launches(path):
    rows = [l for l in open(path, newline="") if l.startswith('"')]
    rd = csv.DictReader(io.StringIO("".join(rows)))
    agg = OrderedDict()
    total = 0.0
    for r in rd:
        if r["Metric Name"] != "gpu__time_duration.sum":
            continue
        ns = float(r["Metric Value"].replace(",", ""))
        if r["Metric Unit"] in ("us", "usecond"):
            ns *= 1e3
        elif r["Metric Unit"] in ("ms", "msecond"):
            ns *= 1e6
        k = short(r["Kernel Name"])
        a = agg.setdefault(k, [0, 0.0, 1e30, 0.0, r["Grid Size"], r["Block Size"]])
        a[0] += 1; a[1] += ns; a[2] = min(a[2], ns); a[3] = max(a[3], ns)
        total += ns
    print(f"source: {path} (ncu launch list; per-launch times are cold-cache and serialised — read the SHARE column)\n")
    print("| kernel | launches | total us | share | avg us | min us | max us | last grid | block |")
    print("|---|---|---|---|---|---|---|---|---|")
    for k, a in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        print(f"| {k} | {a[0]} | {a[1] / 1e3:.1f} | {100 * a[1] / total:.1f}% | {a[1] / a[0] / 1e3:.2f} | {a[2] / 1e3:.2f} | {a[3] / 1e3:.2f} | {a[4]} | {a[5]} |")
    print(f"\ntotal kernel time {total / 1e3:.1f} us over {sum(a[0] for a in agg.values())} launches")


def full(path, pick=None):
    out = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    hdr, units = rows[0], rows[1]
    print(f"source: {path} (ncu --set full --clock-control none)\n")
    for r in rows[2:]:
        d = dict(zip(hdr, r))
        if pick and pick not in d.get("Kernel Name", ""):
            continue
        print(f"### {short(d['Kernel Name'])}  grid {d.get('Grid Size')} block {d.get('Block Size')}\n")
        print("| metric | value | unit |\n|---|---|---|")
        for k in KEYS:
            if k in d:
                print(f"| {k} | {d[k]} | {units[hdr.index(k)]} |")
        print()


if __name__ == "__main__":
    if sys.argv[1] == "launches":
        launches(sys.argv[2])
    else:
        full(sys.argv[2], sys.argv[3] if len(sys.argv) > 3 else None)
