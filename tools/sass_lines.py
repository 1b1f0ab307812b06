#!/usr/bin/env python
"""Static SASS instruction counts of one kernel by source line range (a line counts for the outermost line of
`--file` in its inline chain).   python tools/sass_lines.py LIB KERNEL_SUBSTR label=a-b ... [--list a-b]"""
import collections
import os
import re
import subprocess
import sys
import tempfile

lib, kern = sys.argv[1], sys.argv[2]
specs = [a for a in sys.argv[3:] if "=" in a]
lst = [a for a in sys.argv[3:] if a.startswith("--list")]
fname = "gmr_stages.cuh"
tmp = tempfile.mkdtemp()
subprocess.check_call(["cuobjdump", "-xelf", "all", os.path.abspath(lib)], cwd=tmp, stdout=subprocess.DEVNULL)
cubin = [f for f in os.listdir(tmp) if f.endswith(".cubin")][0]
txt = subprocess.run(["nvdisasm", "--print-line-info-inline", os.path.join(tmp, cubin)], capture_output=True, text=True).stdout.split("\n")
in_k, chain, pending = False, [], []
cnt, ops = collections.Counter(), collections.defaultdict(collections.Counter)
total = 0
for l in txt:
    if l.startswith("\t.section\t.text."):
        in_k = kern in l
        pending = []
        continue
    if not in_k:
        continue
    m = re.match(r'\s*//## File "([^"]+)", line (\d+)', l)
    if m:
        pending.append((m.group(1).split("/")[-1], int(m.group(2))))
        continue
    m = re.match(r"\s*/\*([0-9a-f]{4,})\*/\s+(?:@!?U?P\d+\s+)?(\S+)", l)
    if m:
        if pending:
            chain, pending = list(pending), []
        total += 1
        hit = [fl for fl in chain if fl[0] == fname]
        if hit:
            cnt[hit[-1][1]] += 1
            ops[hit[-1][1]][m.group(2).split(".")[0]] += 1
print("total", total)
for s in specs:
    label, ab = s.split("=")
    a, b = map(int, ab.split("-"))
    o = collections.Counter()
    for ln in range(a, b + 1):
        o.update(ops[ln])
    print(f"{label:12s} {sum(o.values()):5d}  " + " ".join(f"{k}:{v}" for k, v in o.most_common(14)))
for s in lst:
    a, b = map(int, s.split("=")[1].split("-")) if "=" in s else (0, 0)
