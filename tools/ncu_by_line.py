#!/usr/bin/env python
"""Attribute ncu's per-SASS-instruction samples to CUDA source lines / stage functions.

  ncu -i X.ncu-rep --page source --csv > source.csv
  python tools/ncu_by_line.py source.csv gmr_b200/libgmr_b200.so clip_kernelIfE [--top 40]

Joins the SASS page (address, executed instructions, stall samples) with
`nvdisasm --print-line-info-inline` of the same cubin (built with -lineinfo).  Each instruction is
attributed to its innermost source line and to the stage function whose line range holds it.
"""
import argparse
import collections
import csv
import os
import re
import subprocess
import tempfile

# stage functions of gmr_stages.cuh / gmr_clip.cuh by name -> found from the source text
FUNC_RE = re.compile(r"^\s*(?:template\s*<[^>]*>\s*)?(?:__device__|__global__)[^;{]*?\b(\w+)\s*\(")


def function_ranges(path):
    """[(first_line, name)] sorted; a function extends to the next one."""
    out = []
    try:
        lines = open(path).read().split("\n")
    except OSError:
        return out
    for i, l in enumerate(lines, 1):
        m = FUNC_RE.match(l)
        if m:
            out.append((i, m.group(1)))
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("source_csv")
    ap.add_argument("so")
    ap.add_argument("kernel", help="substring of the mangled kernel name")
    ap.add_argument("--top", type=int, default=30)
    ap.add_argument("--ranges", default="", help="label=file:first-last,... : samples and instructions of source-line ranges "
                    "(an inlined helper counts for the outermost line of that file in its inline chain)")
    args = ap.parse_args()

    tmp = tempfile.mkdtemp()
    subprocess.check_call(["cuobjdump", "-xelf", "all", os.path.abspath(args.so)], cwd=tmp, stdout=subprocess.DEVNULL)
    cubin = [f for f in os.listdir(tmp) if f.endswith(".cubin")][0]
    dis = subprocess.run(["nvdisasm", "--print-line-info-inline", os.path.join(tmp, cubin)], capture_output=True, text=True).stdout
    addr_line = {}
    in_k, pending = False, []
    for l in dis.split("\n"):
        if l.startswith("\t.section\t.text."):
            in_k = args.kernel in l
            pending = []
            continue
        if not in_k:
            continue
        m = re.match(r'\s*//## File "([^"]+)", line (\d+)', l)
        if m:
            pending.append((os.path.basename(m.group(1)), int(m.group(2))))
            continue
        m = re.match(r"\s*/\*([0-9a-f]{4,})\*/\s+(\S+)", l)
        if m:
            if pending:
                cur = pending[0]          # innermost first
                chain = list(pending)
                pending = []
            addr_line[int(m.group(1), 16)] = (cur, chain, m.group(2))
    ranges = {}
    here = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "gmr_b200", "csrc")
    for f in os.listdir(here):
        ranges[f] = function_ranges(os.path.join(here, f))

    def func_of(fl):
        f, ln = fl
        best = "?"
        for first, name in ranges.get(f, []):
            if first <= ln:
                best = name
        return best if f in ranges else f

    rows = list(csv.reader(open(args.source_csv)))
    hdr_i = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
    hdr = rows[hdr_i]
    col = {n: i for i, n in enumerate(hdr)}
    base = None
    inst_line = collections.Counter(); static_line = collections.Counter()
    by_line = collections.Counter(); by_func = collections.Counter(); inst_func = collections.Counter()
    noinst_func = collections.Counter(); thr_func = collections.Counter(); by_outer = collections.Counter()
    static_func = collections.Counter()
    STALLS = ["stall_wait", "stall_no_inst", "stall_short_sb", "stall_branch_resolving", "stall_not_selected", "stall_selected",
              "stall_long_sb", "stall_math", "stall_dispatch", "stall_mio"]
    stall_outer = collections.defaultdict(collections.Counter); inst_outer = collections.Counter()
    exc_line = collections.Counter(); wav_line = collections.Counter()
    rng = []
    for spec in filter(None, args.ranges.split(",")):
        label, rest = spec.split("=")
        f, ab = rest.split(":")
        a, b = ab.split("-")
        rng.append((label, f, int(a), int(b)))
    rng_smp = collections.Counter(); rng_inst = collections.Counter(); rng_thr = collections.Counter(); rng_fp64 = collections.Counter()
    total = 0
    for r in rows[hdr_i + 1:]:
        if len(r) < len(hdr):
            continue
        a = int(r[col["Address"]], 16) if r[col["Address"]].startswith("0x") else int(r[col["Address"]])
        if base is None:
            base = a
        off = a - base
        smp = int(r[col["# Samples"]] or 0)
        ins = int(r[col["Instructions Executed"]] or 0)
        thr = int(r[col["Thread Instructions Executed"]] or 0)
        noi = int(r[col["stall_no_inst"]] or 0) if "stall_no_inst" in col else 0
        cur, chain, opcode = addr_line.get(off, (("?", 0), [("?", 0)], ""))
        fn = func_of(cur)
        for label, f, a, b in rng:
            hit = [fl for fl in chain if fl[0] == f]
            if hit and a <= hit[-1][1] <= b:
                rng_smp[label] += smp; rng_inst[label] += ins; rng_thr[label] += thr
                if opcode.startswith(("DFMA", "DMUL", "DADD", "DSETP", "MUFU.RCP64H", "MUFU.RSQ64H")):
                    rng_fp64[label] += ins
                break
        # outermost stage function (the call made from gmr_clip.cuh)
        outer = fn
        for fl in chain:
            if fl[0] == "gmr_stages.cuh":
                outer = func_of(fl)
        inst_line[cur] += ins; static_line[cur] += 1
        by_line[cur] += smp; by_func[fn] += smp; inst_func[fn] += ins; noinst_func[fn] += noi; thr_func[fn] += thr
        by_outer[outer] += smp; static_func[outer] += 1; inst_outer[outer] += ins
        for st in STALLS:
            if st in col:
                stall_outer[outer][st] += int(r[col[st]] or 0)
        if "L1 Wavefronts Shared Excessive" in col:
            exc_line[cur] += int(r[col["L1 Wavefronts Shared Excessive"]] or 0)
            wav_line[cur] += int(r[col["L1 Wavefronts Shared"]] or 0)
        total += smp
    print(f"total samples {total}")
    print("\n-- by outermost stage function (samples %, static SASS instructions)")
    for fn, s in by_outer.most_common(20):
        print(f"{100 * s / total:6.2f}%  {static_func[fn]:6d}  {fn}")
    print("\n-- by innermost function: samples %, no_inst share, warp-instr executed, avg active threads")
    for fn, s in by_func.most_common(25):
        ie = inst_func[fn]
        print(f"{100 * s / total:6.2f}%  noinst {100 * noinst_func[fn] / max(s, 1):5.1f}%  inst {ie:14d}  thr/inst {thr_func[fn] / max(ie, 1):5.1f}  {fn}")
    print("\n-- by outermost stage function: warp-instr executed, share of the kernel's samples per stall reason (%)")
    print(" " * 40 + "".join(f"{st[6:][:9]:>10s}" for st in STALLS))
    for fn, s in by_outer.most_common(9):
        print(f"{fn:22s} {inst_outer[fn]:14d}   " + "".join(f"{100 * stall_outer[fn][st] / total:10.2f}" for st in STALLS))
    print("\n-- shared-memory wavefronts beyond the ideal (bank conflicts), top lines: excess, total")
    for (f, ln), x in exc_line.most_common(12):
        print(f"{x:14d} {wav_line[(f, ln)]:14d}  {f}:{ln}")
    if rng:
        print("\n-- source ranges: samples %, warp-instr executed, FP64 share, avg active threads")
        for label, f, a, b in rng:
            ie = rng_inst[label]
            print(f"{100 * rng_smp[label] / total:6.2f}%  inst {ie:14d}  fp64 {100 * rng_fp64[label] / max(ie, 1):5.1f}%  thr/inst {rng_thr[label] / max(ie, 1):5.1f}  {label} ({f}:{a}-{b})")
    print("\n-- top lines")
    for (f, ln), s in by_line.most_common(args.top):
        print(f"{100 * s / total:6.2f}%  inst {inst_line[(f, ln)]:13d}  static {static_line[(f, ln)]:4d}  {f}:{ln}")


if __name__ == "__main__":
    main()
