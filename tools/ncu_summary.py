#!/usr/bin/env python
"""Summarise an `ncu --set full` report: per kernel the counters DESIGN.md / bench.py quote, plus a traffic JSON.

    python tools/ncu_summary.py <report.ncu-rep> <summary.txt> <traffic.json> "<header line>"
"""
import csv
import io
import json
import subprocess
import sys

rep, out_txt, out_json, header = sys.argv[1:5]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
H, U = rows[0], rows[1]
ix = {h: i for i, h in enumerate(H)}
KEYS = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed", "lts__throughput.avg.pct_of_peak_sustained_elapsed",
        "dram__throughput.avg.pct_of_peak_sustained_elapsed",
        "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed",
        "sm__ops_path_tensor_src_fp16_dst_fp32.avg.pct_of_peak_sustained_elapsed",
        "sm__ops_path_tensor_src_tf32_dst_fp32.avg.pct_of_peak_sustained_elapsed",
        "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active",
        "smsp__inst_executed.sum", "launch__registers_per_thread", "launch__grid_size", "launch__block_size",
        "sm__warps_active.avg.pct_of_peak_sustained_active"]
ALIAS = {}     # (r1 aliased enc16_fused_kernel to the bench label of the time)
MULT = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
lines, traffic, seen = [header], {}, set()
for r in rows[2:]:
    name = r[ix["Kernel Name"]]
    short = name.split("(")[0].split("<")[0].split("::")[-1].replace("void ", "").strip()
    if short in seen:
        continue
    seen.add(short)
    lines.append("== " + name[:110])
    for k in KEYS:
        if k in ix:
            lines.append("   %-78s %s %s" % (k, r[ix[k]], U[ix[k]]))
    rd = float(r[ix["dram__bytes_read.sum"]].replace(",", "")) * MULT.get(U[ix["dram__bytes_read.sum"]], 1.0)
    wr = float(r[ix["dram__bytes_write.sum"]].replace(",", "")) * MULT.get(U[ix["dram__bytes_write.sum"]], 1.0)
    traffic[ALIAS.get(short, short)] = {"dram_read_bytes": rd, "dram_write_bytes": wr, "total_bytes": rd + wr}
open(out_txt, "w").write("\n".join(lines) + "\n")
json.dump({"workload": "bench.py default (655360 rows, D=208, K=1024)", "source": out_txt, "kernels": traffic},
          open(out_json, "w"), indent=1)
print("\n".join(lines[:60]))
