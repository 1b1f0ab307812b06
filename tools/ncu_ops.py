#!/usr/bin/env python
"""ncu --metrics ... --csv log of one clip_kernel launch (tools/gpu_r02.sh `ops`) -> profiles/clip_kernel_ops.json:
executed FP64 thread-instruction counts, warp instructions, issue-slot and FP64-pipe utilisation, for bench.py's
`roofline.executed`.   python tools/ncu_ops.py gpurun_out/r02j/clip_kernel_ops.csv profiles/clip_kernel_ops.json --frames 1228800"""
import argparse
import csv
import json

ap = argparse.ArgumentParser()
ap.add_argument("csv")
ap.add_argument("out")
ap.add_argument("--frames", type=int, default=4096 * 300)
a = ap.parse_args()
rows = list(csv.reader(open(a.csv)))
hdr = next(i for i, r in enumerate(rows) if r and r[0] == "ID")
col = {n: i for i, n in enumerate(rows[hdr])}
m = {}
for r in rows[hdr + 1:]:
    if len(r) > col["Metric Value"]:
        m[r[col["Metric Name"]]] = float(r[col["Metric Value"]].replace(",", ""))
        kern = r[col["Kernel Name"]]
g = lambda k: m.get(k, 0.0)
out = {"kernel": kern, "frames": a.frames,
       "dfma": g("smsp__sass_thread_inst_executed_op_dfma_pred_on.sum"), "dmul": g("smsp__sass_thread_inst_executed_op_dmul_pred_on.sum"),
       "dadd": g("smsp__sass_thread_inst_executed_op_dadd_pred_on.sum"), "ffma": g("smsp__sass_thread_inst_executed_op_ffma_pred_on.sum"),
       "inst_executed": g("smsp__inst_executed.sum"), "thread_inst_executed": g("smsp__thread_inst_executed.sum"),
       "issue_active_pct": g("smsp__issue_active.avg.pct_of_peak_sustained_active"),
       "pipe_fp64_pct": g("sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active"),
       "pipe_lsu_pct": g("sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active"),
       "pipe_alu_pct": g("sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active"),
       "warps_active_pct": g("sm__warps_active.avg.pct_of_peak_sustained_active"),
       "ncu_duration_ms": g("gpu__time_duration.sum") / 1e6, "source": a.csv.replace("gpurun_out/", "gpurun_out/ (scratch) ")}
out["flops_per_frame"] = (2 * out["dfma"] + out["dmul"] + out["dadd"]) / a.frames
out["lanes_per_instruction"] = out["thread_inst_executed"] / max(out["inst_executed"], 1)
json.dump(out, open(a.out, "w"), indent=1)
print(json.dumps(out))
