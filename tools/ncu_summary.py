#!/usr/bin/env python
"""Summarise an .ncu-rep (one `ncu --set full` capture) into the handful of numbers DESIGN.md and
bench.py's roofline.traffic quote.  Usage: python tools/ncu_summary.py gpurun_out/prof.ncu-rep [launch_index]"""
import csv
import io
import json
import subprocess
import sys

KEYS = [
    "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
    "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
    "lts__throughput.avg.pct_of_peak_sustained_elapsed",
    "sm__throughput.avg.pct_of_peak_sustained_elapsed",
    "sm__warps_active.avg.pct_of_peak_sustained_active",
    "smsp__issue_active.avg.pct_of_peak_sustained_active",
    "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed",
    "sm__inst_executed_pipe_tensor.sum",
    "launch__registers_per_thread", "launch__grid_size", "launch__block_size",
    "launch__shared_mem_per_block_dynamic", "launch__waves_per_multiprocessor",
    "smsp__inst_executed.sum",
]


def to_bytes(value, unit):
    mult = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "Tbyte": 1e12}
    return float(value) * mult.get(unit, 1)


def main():
    rep = sys.argv[1]
    idx = int(sys.argv[2]) if len(sys.argv) > 2 else 0
    out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    hdr, units, vals = rows[0], rows[1], rows[2 + idx]
    d = {h: (v, u) for h, u, v in zip(hdr, units, vals)}
    res = {"kernel": d.get("Kernel Name", ("?", ""))[0]}
    for k in KEYS:
        if k in d:
            res[k] = f"{d[k][0]} {d[k][1]}".strip()
    stalls = sorted(((float(v.replace(',', '')), h) for h, (v, u) in d.items()
                     if h.startswith("smsp__average_warps_issue_stalled") and h.endswith("per_issue_active.ratio")),
                    reverse=True)[:5]
    res["top_stalls_per_issue"] = {h.split("issue_stalled_")[1].replace("_per_issue_active.ratio", ""): v
                                   for v, h in stalls}
    rd = to_bytes(*d["dram__bytes_read.sum"])
    wr = to_bytes(*d["dram__bytes_write.sum"])
    res["dram_bytes"] = rd + wr
    print(json.dumps(res, indent=1))


if __name__ == "__main__":
    main()
