#!/usr/bin/env python
"""Summarise ncu output into the small text files committed under profiles/.

  python profiles/summarize.py launches gpurun_out/launches_X.csv > profiles/rNN_launches_X.txt
  python profiles/summarize.py full gpurun_out/prof_X.ncu-rep     > profiles/rNN_full_X.txt   (runs `ncu -i`)
"""
import csv
import io
import re
import subprocess
import sys
from collections import OrderedDict

KEYS = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "dram__throughput.avg.pct_of_peak_sustained_elapsed",
        "lts__t_bytes.sum", "lts__throughput.avg.pct_of_peak_sustained_elapsed", "l1tex__throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "launch__registers_per_thread", "launch__grid_size", "launch__block_size", "launch__occupancy_limit_registers",
        "smsp__inst_executed.sum", "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
        "smsp__average_warp_latency_issue_stalled_long_scoreboard.ratio",
        "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio",
        "smsp__issue_active.avg.pct_of_peak_sustained_active"]


def launches(path):
    rows = []
    with open(path) as f:
        txt = f.read()
    start = txt.index('"ID"')
    rd = csv.DictReader(io.StringIO(txt[start:]))
    for r in rd:
        if r.get("Metric Name") != "gpu__time_duration.sum":
            continue
        v = float(r["Metric Value"].replace(",", ""))
        unit = r["Metric Unit"]
        us = v / 1000.0 if unit in ("ns", "nsecond") else (v if unit in ("us", "usecond") else v * 1000.0)
        name = re.sub(r"\(.*", "", r["Kernel Name"])
        name = re.sub(r"^void\s+", "", name)
        rows.append((name, us))
    agg = OrderedDict()
    for n, us in rows:
        a = agg.setdefault(n, [0, 0.0])
        a[0] += 1
        a[1] += us
    tot = sum(a[1] for a in agg.values())
    print("# %d launches, %.1f us total (ncu per-launch times are cold-cache and serialised: compare shares)" % (len(rows), tot))
    print("%-70s %8s %12s %10s %7s" % ("kernel", "launches", "total_us", "avg_us", "share"))
    for n, (c, t) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        print("%-70s %8d %12.1f %10.2f %6.1f%%" % (n[:70], c, t, t / c, 100 * t / tot))


def full(path):
    out = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rd = list(csv.reader(io.StringIO(out)))
    hdr, units = rd[0], rd[1]
    idx = {h: i for i, h in enumerate(hdr)}
    for row in rd[2:]:
        print("== %s  grid %s block %s" % (re.sub(r"\(.*", "", row[idx["Kernel Name"]]), row[idx.get("Grid Size", 0)],
                                          row[idx.get("Block Size", 0)]))
        for k in KEYS:
            if k in idx:
                print("   %-80s %s %s" % (k, row[idx[k]], units[idx[k]]))


if __name__ == "__main__":
    {"launches": launches, "full": full}[sys.argv[1]](sys.argv[2])
