#!/usr/bin/env python3
"""Summarise ncu outputs brought back in gpurun_out/ (run here, no GPU needed).
    python scripts/ncu_summary.py launches gpurun_out/launches.csv
    python scripts/ncu_summary.py full gpurun_out/prof_step.ncu-rep
"""
import collections
import csv
import io
import subprocess
import sys

KEYS = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "dram__throughput.avg.pct_of_peak_sustained_elapsed",
        "dram__cycles_active.avg.pct_of_peak_sustained_elapsed",
        "lts__t_sector_hit_rate.pct", "l1tex__t_sector_hit_rate.pct", "lts__t_bytes.sum", "l1tex__t_bytes.sum",
        "launch__registers_per_thread", "launch__waves_per_multiprocessor", "launch__occupancy_limit_registers",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__inst_executed.sum", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active",
        "smsp__inst_executed_pipe_lsu.sum", "sm__cycles_elapsed.max", "gpc__cycles_elapsed.max",
        "smsp__average_warp_latency_issue_stalled_long_scoreboard.ratio",
        "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio"]


def launches(path):
    lines = [l for l in open(path) if l.startswith('"')]
    rows = list(csv.DictReader(io.StringIO("".join(lines))))
    d = collections.OrderedDict()
    for r in rows:
        d.setdefault(r["Kernel Name"].split("(")[0], []).append(float(r["Metric Value"].replace(",", "")))
    tot = sum(sum(v) for v in d.values())
    print(f"{'kernel':60s} {'n':>5s} {'mean us':>10s} {'min us':>10s} {'max us':>10s} {'share':>7s}")
    for k, v in d.items():
        print(f"{k[:60]:60s} {len(v):5d} {sum(v)/len(v)/1e3:10.2f} {min(v)/1e3:10.2f} {max(v)/1e3:10.2f} {sum(v)/tot:7.1%}")


def full(path):
    out = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    lines = [l for l in out.splitlines(True) if l.startswith('"')]
    rows = list(csv.reader(io.StringIO("".join(lines))))
    hdr, units, data = rows[0], rows[1], rows[2:]
    kn = hdr.index("Kernel Name")
    for r in data:
        print("==", r[kn][:100], "grid", r[hdr.index("Grid Size")], "block", r[hdr.index("Block Size")])
        for k in KEYS:
            hits = [i for i, h in enumerate(hdr) if h == k or h.endswith("." + k)]
            for i in hits[:1]:
                print(f"   {k:75s} {r[i]:>18s} {units[i]}")
        rd = [i for i, h in enumerate(hdr) if h.endswith("dram__bytes_read.sum")]
        wr = [i for i, h in enumerate(hdr) if h.endswith("dram__bytes_write.sum")]
        if rd and wr:
            def val(i):
                v = float(r[i].replace(",", ""))
                u = units[i].lower()
                return v * {"byte": 1, "kbyte": 1e3, "mbyte": 1e6, "gbyte": 1e9}.get(u, 1)
            print(f"   traffic (read+write) bytes per launch: {val(rd[0]) + val(wr[0]):.0f}")


if __name__ == "__main__":
    {"launches": launches, "full": full}[sys.argv[1]](sys.argv[2])
