#!/usr/bin/env python
"""Device timing of the BVH -> frame stream pass (gmr_bvh_to_stream) on a LAFAN1-shaped batch, with the numpy
restatement of the reference loader timed beside it on a bounded sample.  Development probe; prints one JSON line.

    python tools/bvh_bench.py [--clips 4096 --frames 300]
"""
import argparse
import ctypes as C
import json
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
from gmr_b200 import _lib, bvh  # noqa: E402
from gmr_b200.abi import F32  # noqa: E402
from oracle import bvh_oracle as bo  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--clips", type=int, default=4096)
ap.add_argument("--frames", type=int, default=300)
ap.add_argument("--reps", type=int, default=5)
args = ap.parse_args()

z = np.load(os.path.join(os.path.dirname(__file__), "..", "tests", "golden", "bvh_lafan.npz"))
c0 = bvh.parse_bvh(str(z["bvh_text"]), is_text=True)
J, Cn, T = len(c0.names), args.clips, args.frames
N = Cn * T
rows, sp, sr = bvh.stream_layout(c0.names)
head, feet = bvh.height_rows(rows)
h = len(rows)
dev = torch.device("cuda", 0)
g = torch.Generator(device=dev).manual_seed(1)
t = (torch.arange(T, device=dev, dtype=torch.float64) / 30.0).view(1, T, 1, 1)
amp = torch.rand((Cn, 1, J, 3), generator=g, device=dev, dtype=torch.float64) * 150 + 20
ph = torch.rand((Cn, 1, J, 3), generator=g, device=dev, dtype=torch.float64) * 6.28
eul = ((amp * torch.sin(2 * np.pi * 0.7 * t + ph) + 180.0) % 360.0 - 180.0).reshape(N, J, 3).contiguous()
lpos = torch.from_numpy(c0.offsets).to(dev).view(1, J, 3).repeat(N, 1, 1).contiguous()
lpos[:, 0, 1] += 92.0
start = torch.arange(0, N + 1, T, dtype=torch.int32, device=dev)
i32 = lambda a: torch.from_numpy(np.ascontiguousarray(a, np.int32)).to(dev)      # noqa: E731
d_par, d_sp, d_sr = i32(c0.parents), i32(sp), i32(sr)
out = torch.empty((N, h, 7), dtype=torch.float32, device=dev)
hg = torch.empty((Cn,), dtype=torch.float64, device=dev)
p = lambda x: C.c_void_p(x.data_ptr())                                            # noqa: E731
feet4 = (list(feet) + [-1] * 4)[:4]


def run():
    s = torch.cuda.current_stream().cuda_stream
    _lib.check(_lib.lib().gmr_bvh_to_stream(0, C.c_void_p(s), p(lpos), p(eul), p(start), Cn, N, J, p(d_par),
                                            (C.c_int32 * 3)(*c0.order), p(d_sp), p(d_sr), h, F32, p(out), head,
                                            (C.c_int32 * 4)(*feet4), p(hg)), "gmr_bvh_to_stream")


for _ in range(3):
    run()
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(args.reps):
    run()
e1.record()
torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / args.reps
# CPU: numpy restatement of the reference loader (vectorised over frames - faster than the reference's per-frame
# Python loop) on the first clips, one core
nc = min(Cn, 64)
eh, lh = eul[:nc * T].cpu().numpy().reshape(nc, T, J, 3), lpos[:nc * T].cpu().numpy().reshape(nc, T, J, 3)
t0 = time.time()
for c in range(nc):
    ref, rh = bo.bvh_to_stream(lh[c], eh[c], c0.parents, c0.order, sp, sr, head, feet)
cpu_s = time.time() - t0
got = out[(nc - 1) * T:nc * T].cpu().numpy()
err = float(np.abs(got - ref.astype(np.float32)).max())
alg = J * 6 * 8 + h * 7 * 4
print(json.dumps({"metric": "bvh_frames_to_stream_per_sec", "value": N / ms * 1e3, "unit": "frames/s", "ms": ms,
                  "clips": Cn, "frames_per_clip": T, "bones": J, "rows": h, "algorithmic_bytes_per_frame": alg,
                  "achieved_GBps": N * alg / ms / 1e6, "cpu_numpy_frames_per_sec_1core": nc * T / cpu_s,
                  "max_abs_diff_vs_oracle_f32": err}))
