#!/usr/bin/env python
"""configs[2] / configs[3] blocks of bench.py alone (development probe).   python tools/extras_quick.py [hands] [hetero]"""
import json
import os
import sys

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import bench_extras as bx  # noqa: E402

which = sys.argv[1:] or ["hands", "hetero"]
w = min(16, os.cpu_count() or 1)
if "hands" in which:
    print(json.dumps(bx.hands_block(0, w)), flush=True)
if "hetero" in which:
    print(json.dumps(bx.hetero_block(0, w)), flush=True)
