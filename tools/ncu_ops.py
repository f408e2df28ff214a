#!/usr/bin/env python3
"""Executed floating-point work per kernel from an ncu --csv metrics log (long format, one row per metric):

  ncu --metrics smsp__sass_thread_inst_executed_op_{dfma,dmul,dadd,ffma,fmul,fadd}_pred_on.sum,... --csv --log-file ops.csv <cmd>
  python tools/ncu_ops.py ops.csv [out.json]

Per kernel name (mean over its captured launches): executed FP64 / FP32 thread-level instructions, executed flops
(FMA = 2, MUL / ADD = 1 - conversions, compares, selects, MUFU seeds and min / max are NOT counted), threads per warp
instruction, FP64-pipe and issue utilisation, DRAM bytes, duration.  bench.py reads the JSON to report
`executed_flops_per_launch` / `frac_executed` next to the frozen flop model."""
import csv
import json
import re
import sys
from collections import defaultdict

SHORT = {"smsp__sass_thread_inst_executed_op_dfma_pred_on.sum": "dfma", "smsp__sass_thread_inst_executed_op_dmul_pred_on.sum": "dmul",
         "smsp__sass_thread_inst_executed_op_dadd_pred_on.sum": "dadd", "smsp__sass_thread_inst_executed_op_ffma_pred_on.sum": "ffma",
         "smsp__sass_thread_inst_executed_op_fmul_pred_on.sum": "fmul", "smsp__sass_thread_inst_executed_op_fadd_pred_on.sum": "fadd",
         "smsp__thread_inst_executed_per_inst_executed.ratio": "threads_per_inst", "smsp__inst_executed.sum": "warp_inst",
         "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active": "fp64_pipe_pct",
         "smsp__issue_active.avg.pct_of_peak_sustained_active": "issue_pct",
         "sm__warps_active.avg.pct_of_peak_sustained_active": "warps_active_pct", "gpu__time_duration.sum": "duration_ns",
         "dram__bytes_read.sum": "dram_read", "dram__bytes_write.sum": "dram_write"}


def parse(path):
    rows = [r for r in csv.reader(open(path, newline="")) if len(r) >= 15 and r[0].isdigit()]
    launches = defaultdict(dict)
    for r in rows:
        name = re.sub(r"^void\s+", "", r[4])
        name = re.sub(r"\(.*\)$", "", name).replace("scb::", "")
        key = SHORT.get(r[12])
        if key:
            launches[(int(r[0]), name)][key] = float(r[14].replace(",", ""))
    per = defaultdict(list)
    for (_, name), m in sorted(launches.items()):
        per[name].append(m)
    out = {}
    for name, ls in per.items():
        mean = {k: sum(m.get(k, 0.0) for m in ls) / len(ls) for k in set().union(*ls)}
        mean["launches"] = len(ls)
        mean["fp64_flops"] = 2 * mean.get("dfma", 0) + mean.get("dmul", 0) + mean.get("dadd", 0)
        mean["fp32_flops"] = 2 * mean.get("ffma", 0) + mean.get("fmul", 0) + mean.get("fadd", 0)
        mean["dram_bytes"] = mean.get("dram_read", 0) + mean.get("dram_write", 0)
        if mean.get("duration_ns"):
            mean["fp64_tflops_under_ncu"] = mean["fp64_flops"] / mean["duration_ns"] / 1e3
        out[name] = mean
    return out


if __name__ == "__main__":
    res = parse(sys.argv[1])
    if len(sys.argv) > 2 and sys.argv[2] == "--merge":  # ncu_ops.py ops.csv --merge profiles/ncu_ops.json TAG "source description"
        path, tag, src = sys.argv[3], sys.argv[4], sys.argv[5]
        try:
            db = json.load(open(path))
        except Exception:
            db = {}
        db[tag] = {"source": src, "kernels": {k: v for k, v in res.items() if not k.startswith("at::")}}
        json.dump(db, open(path, "w"), indent=1, sort_keys=True)
    elif len(sys.argv) > 2:
        json.dump(res, open(sys.argv[2], "w"), indent=1, sort_keys=True)
    for k, m in res.items():
        print("%-46s n=%d  %.3f ms  dfma %.3e dmul %.3e dadd %.3e  -> %.2f GFLOP fp64 (%.1f TF/s)  ffma %.2e  thr/inst %.1f  fp64 pipe %.0f%%  issue %.0f%%  dram %.2f GB"
              % (k[:46], m["launches"], m.get("duration_ns", 0) / 1e6, m.get("dfma", 0), m.get("dmul", 0), m.get("dadd", 0), m["fp64_flops"] / 1e9,
                 m.get("fp64_tflops_under_ncu", 0), m.get("ffma", 0), m.get("threads_per_inst", 0), m.get("fp64_pipe_pct", 0), m.get("issue_pct", 0),
                 m["dram_bytes"] / 1e9))
