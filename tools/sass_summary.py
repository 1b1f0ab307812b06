#!/usr/bin/env python
"""Static per-kernel SASS summary of the built library (runs without a GPU):

    python tools/sass_summary.py [gmr_b200/libgmr_b200.so] > profiles/r02_sass_summary.txt

Per kernel: arch, registers / static shared memory / stack / local loads and stores (cuobjdump --dump-resource-usage), code
size in instructions and bytes, and the histogram of the instruction classes that decide this code's speed: FP64
arithmetic (DFMA / DMUL / DADD / DSETP), MUFU (the f64 reciprocal / rsqrt seeds), shared-memory and global loads /
stores with their widths, constant-bank loads (LDC - what the compile-time layouts remove), shuffles, barriers and
warp syncs, branches.  The clip_kernel instantiations (one per compile-time layout and stream dtype) are listed one by
one and then the rest of the library's kernels.
"""
import collections
import os
import re
import subprocess
import sys

lib = sys.argv[1] if len(sys.argv) > 1 else os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "gmr_b200", "libgmr_b200.so")
CLASSES = ["DFMA", "DMUL", "DADD", "DSETP", "MUFU", "LDS", "STS", "LDG", "STG", "LDC", "LDCU", "ULDC", "SHFL", "BAR", "WARPSYNC",
           "BRA", "BSSY", "BSYNC", "ATOMG", "NANOSLEEP", "IMAD", "FSEL", "SEL", "MOV", "LDL", "STL"]


def demangle(names):
    out = subprocess.run(["cu++filt"] + names, capture_output=True, text=True).stdout.split("\n")
    return dict(zip(names, out))


def resources():
    txt = subprocess.run(["cuobjdump", "--dump-resource-usage", lib], capture_output=True, text=True).stdout
    res, arch = {}, None
    cur = None
    for l in txt.split("\n"):
        m = re.match(r"\s*arch = (\S+)", l)
        if m:
            arch = m.group(1)
        m = re.match(r"\s*Function (\S+):", l)
        if m:
            cur = m.group(1)
            continue
        if cur and "REG:" in l:
            res[cur] = dict(kv.split(":") for kv in l.split() if ":" in kv)
            cur = None
    return arch, res


def sass():
    txt = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True).stdout
    out, cur = collections.OrderedDict(), None
    for l in txt.split("\n"):
        m = re.match(r"\s*Function : (\S+)", l)
        if m:
            cur = m.group(1)
            out[cur] = {"n": 0, "ops": collections.Counter(), "w": collections.Counter(), "last": 0}
            continue
        m = re.match(r"\s*/\*([0-9a-f]{4,6})\*/\s+(?:@!?U?P\w+\s+)?([A-Z][A-Z0-9_]*)((?:\.[A-Z0-9_]+)*)", l)
        if m and cur:
            k = out[cur]
            k["n"] += 1
            k["last"] = int(m.group(1), 16)
            op, mods = m.group(2), m.group(3)
            k["ops"][op] += 1
            if op in ("LDS", "STS", "LDG", "STG"):
                w = "128" if ".128" in mods else "64" if ".64" in mods else "32"
                k["w"][f"{op}.{w}"] += 1
    return out


def main():
    arch, res = resources()
    ks = sass()
    names = demangle(list(ks))
    print(f"# {os.path.basename(lib)}: {os.path.getsize(lib)} bytes, arch = {arch}, {len(ks)} kernels   (tools/sass_summary.py)")
    print("# per kernel: instructions, code bytes, registers, static smem, stack, then instruction classes (static counts)")
    order = sorted(ks, key=lambda n: (0 if "clip_kernel" in n else 1, names[n]))
    for n in order:
        k, r = ks[n], res.get(n, {})
        full = names[n]
        short = (full[:full.rfind("(")] if full.endswith(")") else full).replace("gmr::", "").replace("(bool)", "")
        print(f"\n{short}")
        print(f"  instr {k['n']}  bytes {k['last'] + 16}  REG {r.get('REG', '?')}  SHARED {r.get('SHARED', '?')}  STACK {r.get('STACK', '?')}  "
              f"LOCAL {r.get('LOCAL', '?')}  local ld/st {k['ops']['LDL'] + k['ops']['STL']}")
        print("  " + "  ".join(f"{c} {k['ops'][c]}" for c in CLASSES if k["ops"][c]))
        if k["w"]:
            print("  widths: " + "  ".join(f"{a} {b}" for a, b in sorted(k["w"].items())))
        rest = sum(v for o, v in k["ops"].items() if o not in CLASSES)
        print(f"  other {rest}: " + " ".join(f"{o}:{v}" for o, v in k["ops"].most_common(40) if o not in CLASSES)[:200])


if __name__ == "__main__":
    main()
