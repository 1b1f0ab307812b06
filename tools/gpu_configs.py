#!/usr/bin/env python
"""Device timing of the BASELINE.json configs that are not the bench line (development probe, one GPU):

  configs[2]  unitree_g1_with_hands <- SMPL-X-shaped streams, 8192 clips x 300
  configs[3]  heterogeneous launch: every robot constructible for the source in ONE launch, instance = (clip, robot)
  configs[4]  dataset scale: ~10 M frames (33 334 clips x 300) on one GPU, uniform and ragged (lengths 60..300,
              longest first)

Prints one JSON line per case and writes them to gpurun_out/configs/configs.json.
"""
import argparse
import json
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
from gmr_b200 import model_pack, synth  # noqa: E402
from gmr_b200.engine import BatchedRetargeter, DeviceModel  # noqa: E402

SMPLX_ROBOTS = ["unitree_g1", "unitree_g1_with_hands", "booster_t1", "stanford_toddy", "fourier_n1", "engineai_pm01",
                "kuavo_s45", "hightorque_hi", "galaxea_r1pro", "booster_k1"]


def timed(br, human_dev, out, reps, T_of=None):
    best = None
    for _ in range(reps):
        br.reset()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        br.run(human_dev, T_of=T_of, out=out)
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1)
        best = ms if best is None else min(best, ms)
    return best


def case_single(src, robot, B, T, reps, workers, ragged=False, tag=""):
    m, t = model_pack.build_tables(src, robot)
    t0 = time.time()
    human, heights, names = synth.synth_human_clips(m, t, src, B=B, T=T, seed=0xB205, workers=workers)
    synth_s = time.time() - t0
    dm = DeviceModel.load(src, robot, names)
    br = BatchedRetargeter(dm, B)
    br.set_heights(heights)
    hd = torch.from_numpy(human).cuda()
    out = torch.zeros((B, T, br.nq_max), dtype=torch.float64, device="cuda")
    res = {"case": tag or f"{robot} <- {src}", "clips": B, "frames_per_clip": T, "launch": br.launch_info(),
           "synth_seconds": round(synth_s, 1)}
    ms = timed(br, hd, out, reps)
    res.update(ms=ms, frames=B * T, frames_per_s=B * T / ms * 1e3, finite=bool(torch.isfinite(out).all().item()))
    rows = [res]
    if ragged:
        T_of = np.random.default_rng(5).integers(60, T + 1, size=B).astype(np.int32)
        br.set_order(np.argsort(-T_of).astype(np.int32))
        ms = timed(br, hd, out, reps, T_of=T_of)
        r2 = dict(res)
        r2.update(case=res["case"] + ", ragged lengths 60..%d, longest first" % T, ms=ms, frames=int(T_of.sum()),
                  frames_per_s=float(T_of.sum()) / ms * 1e3)
        rows.append(r2)
    return rows


def case_hetero(per, T, reps, workers):
    models, humans, heights_all, model_of = [], [], [], []
    for i, r in enumerate(SMPLX_ROBOTS):
        m, t = model_pack.build_tables("smplx", r)
        human, heights, names = synth.synth_human_clips(m, t, "smplx", B=per, T=T, seed=0xB204 + i, workers=workers)
        models.append(DeviceModel.load("smplx", r, names))
        humans.append(human)
        heights_all += list(heights)
        model_of += [i] * per
    B = per * len(SMPLX_ROBOTS)
    # interleave the robots so that every CTA sees a mix
    perm = np.arange(B).reshape(len(SMPLX_ROBOTS), per).T.reshape(-1)
    human = np.concatenate(humans, 0)[perm]
    br = BatchedRetargeter(models, B, model_of_instance=list(np.asarray(model_of)[perm]))
    br.set_heights(list(np.asarray(heights_all)[perm]))
    hd = torch.from_numpy(human).cuda()
    out = torch.zeros((B, T, br.nq_max), dtype=torch.float64, device="cuda")
    ms = timed(br, hd, out, reps)
    return [{"case": "heterogeneous launch: %d robots x %d clips, smplx, instance = (clip, robot)" % (len(SMPLX_ROBOTS), per),
             "robots": SMPLX_ROBOTS, "clips": B, "frames_per_clip": T, "launch": br.launch_info(), "ms": ms,
             "frames": B * T, "frames_per_s": B * T / ms * 1e3, "finite": bool(torch.isfinite(out).all().item())}]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--reps", type=int, default=2)
    ap.add_argument("--workers", type=int, default=16)
    ap.add_argument("--skip-scale", action="store_true")
    args = ap.parse_args()
    rows = []
    rows += case_single("smplx", "unitree_g1_with_hands", 8192, 300, args.reps, args.workers, tag="configs[2]")
    rows += case_hetero(410, 300, args.reps, args.workers)
    if not args.skip_scale:
        rows += case_single("bvh", "unitree_g1", 33334, 300, args.reps, args.workers, ragged=True,
                            tag="configs[4] on one GPU: unitree_g1 <- bvh")
    os.makedirs("gpurun_out/configs", exist_ok=True)
    with open("gpurun_out/configs/configs.json", "w") as f:
        for r in rows:
            print(json.dumps(r), flush=True)
            f.write(json.dumps(r) + "\n")


if __name__ == "__main__":
    main()
