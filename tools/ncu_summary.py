#!/usr/bin/env python
"""Reduce one `ncu --set full` capture to the numbers DESIGN.md / bench.py quote.

  python tools/ncu_summary.py gpurun_out/r01h/clip_kernel.ncu-rep profiles/r01_clip_kernel_ncu_summary.json \
         [--traffic profiles/clip_kernel_traffic.json]

Reads `ncu -i REP --page raw --csv`, keeps duration, DRAM bytes, pipe utilisation, launch geometry and the
per-issue stall reasons of the first kernel in the report.
"""
import argparse
import csv
import io
import json
import subprocess

KEEP = [
    "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
    "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active",
    "smsp__inst_executed.sum", "smsp__thread_inst_executed_per_inst_executed.ratio",
    "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active",
    "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread", "launch__grid_size",
    "launch__block_size", "launch__shared_mem_per_block_dynamic", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
    "sm__throughput.avg.pct_of_peak_sustained_elapsed", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
    "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "smsp__inst_executed_op_shared_ld.sum", "smsp__inst_executed_op_shared_st.sum",
    "lts__t_sector_hit_rate.pct", "pcie__read_bytes.sum", "pcie__write_bytes.sum",
]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("rep")
    ap.add_argument("out")
    ap.add_argument("--traffic")
    a = ap.parse_args()
    txt = subprocess.run(["ncu", "-i", a.rep, "--page", "raw", "--csv"], capture_output=True, text=True, check=True).stdout
    rows = list(csv.reader(io.StringIO(txt)))
    names, units, vals = rows[0], rows[1], rows[2]
    out = {}
    for n, u, v in zip(names, units, vals):
        if "." in n and "__" in n and not n.split(".")[0].count("__"):
            continue                                    # section-prefixed duplicates (FBSP.TriageCompute.x)
        if n in KEEP or (n.startswith("smsp__average_warps_issue_stalled_") and n.endswith("_per_issue_active.ratio")):
            try:
                x = float(v.replace(",", ""))
            except ValueError:
                x = v
            out[n] = {"value": x, "unit": u}
        elif n == "Kernel Name":
            out["kernel"] = v
    json.dump(out, open(a.out, "w"), indent=1)
    if a.traffic:
        def b(k):
            m = out[k]
            s = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}[m["unit"]]
            return m["value"] * s
        d = out["gpu__time_duration.sum"]
        ms = d["value"] * {"ns": 1e-6, "us": 1e-3, "ms": 1.0, "s": 1e3, "nsecond": 1e-6, "usecond": 1e-3, "msecond": 1.0, "second": 1e3}[d["unit"]]
        json.dump({"kernel": out.get("kernel"), "dram_bytes_read": b("dram__bytes_read.sum"), "dram_bytes_write": b("dram__bytes_write.sum"),
                   "dram_bytes_per_launch": b("dram__bytes_read.sum") + b("dram__bytes_write.sum"), "ncu_duration_ms": ms,
                   "source": a.rep}, open(a.traffic, "w"))
    for k in ("gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "smsp__issue_active.avg.pct_of_peak_sustained_active",
              "smsp__inst_executed.sum"):
        print(k, out.get(k))


if __name__ == "__main__":
    main()
