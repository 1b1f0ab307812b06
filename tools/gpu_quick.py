#!/usr/bin/env python
"""Quick device timing of the fused clip kernel (not the bench; a development probe)."""
import argparse
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
from gmr_b200 import model_pack, synth  # noqa: E402
from gmr_b200.engine import BatchedRetargeter, DeviceModel  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--src", default="bvh")
ap.add_argument("--robot", default="unitree_g1")
ap.add_argument("--B", type=int, default=4096)
ap.add_argument("--T", type=int, default=300)
ap.add_argument("--reps", type=int, default=3)
ap.add_argument("--seed", type=int, default=0)
args = ap.parse_args()

m, t = model_pack.build_tables(args.src, args.robot)
t0 = time.time()
human, heights, names = synth.synth_human_clips(m, t, args.src, B=args.B, T=args.T)
print(f"synth {time.time() - t0:.1f}s  human {human.shape} {human.nbytes / 1e6:.0f} MB", flush=True)
dm = DeviceModel.load(args.src, args.robot, names)
br = BatchedRetargeter(dm, args.B)
br.set_heights(heights)
print("launch", br.launch_info(), flush=True)
hd = torch.from_numpy(human).cuda()
out = torch.zeros((args.B, args.T, br.nq_max), dtype=torch.float64, device="cuda")
for rep in range(args.reps):
    br.reset()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    qpos, diag = br.run(hd, diag=True, out=out)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)
    it = diag["iters"].double().mean((0, 1)).cpu().numpy()
    print(f"rep {rep}: {ms:.1f} ms  {args.B * args.T / ms * 1e3:.3e} frames/s  iters/frame {it}  "
          f"nfact/frame {diag['nfact'].double().mean().item():.2f}  maxiter-hit {(diag['iters'] == 11).double().mean().item():.4f}", flush=True)
print("finite", torch.isfinite(qpos).all().item())
# per clip: solves (= solve_ik calls) and box-QP sweeps; the slowest chains of the launch
solves = diag["iters"].sum((1, 2)).cpu().numpy()
sweeps = diag["nfact"].sum(1).cpu().numpy()
order = np.argsort(-sweeps)
print(f"clips: solves mean {solves.mean():.0f} max {solves.max()}  sweeps mean {sweeps.mean():.0f} max {sweeps.max()}  "
      f"sweeps/solve overall {sweeps.sum() / solves.sum():.3f}")
print("slowest by sweeps (clip, solves, sweeps, sweeps/solve):",
      [(int(c), int(solves[c]), int(sweeps[c]), round(float(sweeps[c] / solves[c]), 3)) for c in order[:12]])
nf = diag["nfact"].cpu().numpy()
itf = diag["iters"].sum(-1).cpu().numpy()
print("sweeps per frame histogram of (sweeps - solves):", np.bincount(np.clip(nf - itf, 0, 12).ravel())[:13].tolist())
