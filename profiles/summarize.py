#!/usr/bin/env python
"""Summarise ncu output into the small text files committed under profiles/.

  python profiles/summarize.py launches gpurun_out/launches_X.csv > profiles/rNN_launches_X.txt
  python profiles/summarize.py full gpurun_out/prof_X.ncu-rep     > profiles/rNN_full_X.txt   (runs `ncu -i`)
  python profiles/summarize.py traffic gpurun_out/prof_X.ncu-rep WORKLOAD:NATOMS profiles/rNN_full_X.txt
        merges per-kernel dram bytes / L2 throughput (mean over the captured launches) into profiles/ncu_traffic.json,
        which bench.py reads for roofline.traffic
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


def _num(v):
    return float(v.replace(",", "")) if v not in ("", "n/a") else 0.0


def _bytes(v, unit):
    scale = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "Tbyte": 1e12}
    return _num(v) * scale.get(unit, 1)


FAMILY = [("tn_msg_bwd", "tn_msg_bwd"), ("tn_msg_fwd", "tn_msg_fwd"), ("tn_embed_edge_bwd", "tn_embed_edge_bwd"),
          ("tn_embed_edge_fwd", "tn_embed_edge_fwd"), ("tc_mix_kernel", "node_gemm_tc"), ("sgemm_nt_kernel", "node_gemm"), ("sgemm_small_kernel", "node_gemm_small"), ("tc_gemm_kernel", "node_gemm_tc"),
          ("cell_tile_kernel", "neighbors_csr"), ("et_edge_bwd", "et_edge_bwd"), ("et_edge_fwd", "et_edge_fwd")]


def traffic(path, key, source):
    import json
    import os
    out = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rd = list(csv.reader(io.StringIO(out)))
    hdr, units = rd[0], rd[1]
    idx = {h: i for i, h in enumerate(hdr)}
    acc = {}
    for row in rd[2:]:
        name = row[idx["Kernel Name"]]
        fam = next((f for pat, f in FAMILY if pat in name), None)
        if fam is None:
            continue
        a = acc.setdefault(fam, {"n": 0, "dram": 0.0, "lts": 0.0, "us": 0.0, "issue": 0.0, "tensor": 0.0})
        a["n"] += 1
        a["dram"] += _bytes(row[idx["dram__bytes_read.sum"]], units[idx["dram__bytes_read.sum"]]) + \
            _bytes(row[idx["dram__bytes_write.sum"]], units[idx["dram__bytes_write.sum"]])
        a["lts"] += _num(row[idx["lts__throughput.avg.pct_of_peak_sustained_elapsed"]])
        t, u = row[idx["gpu__time_duration.sum"]], units[idx["gpu__time_duration.sum"]]
        a["us"] += _num(t) * {"ns": 1e-3, "nsecond": 1e-3, "us": 1, "usecond": 1, "ms": 1e3, "msecond": 1e3}.get(u, 1)
        a["issue"] += _num(row[idx["smsp__issue_active.avg.pct_of_peak_sustained_active"]])
        a["tensor"] += _num(row[idx["sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active"]])
    dst = os.path.join(os.path.dirname(os.path.abspath(__file__)), "ncu_traffic.json")
    data = json.load(open(dst)) if os.path.exists(dst) else {}
    data[key] = {f: {"dram_bytes": a["dram"] / a["n"], "lts_throughput_pct": a["lts"] / a["n"],
                     "issue_active_pct": a["issue"] / a["n"], "tensor_pipe_pct": a["tensor"] / a["n"],
                     "ncu_launch_us": a["us"] / a["n"], "launches_captured": a["n"], "source": source}
                 for f, a in acc.items()}
    json.dump(data, open(dst, "w"), indent=1, sort_keys=True)
    print(json.dumps(data[key], indent=1))


if __name__ == "__main__":
    if sys.argv[1] == "traffic":
        traffic(sys.argv[2], sys.argv[3], sys.argv[4])
    else:
        {"launches": launches, "full": full}[sys.argv[1]](sys.argv[2])
