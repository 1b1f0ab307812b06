#!/usr/bin/env python
"""The three workloads of the G1 <- LAFAN1-shaped path on one GPU: frames/s, IK iterations per frame, fraction of
frames that end on max_iter, final task-space residuals; and CUDA-vs-oracle parity on a sample of each.

  robot-authored   gmr_b200.synth.synth_human_clips        (the bench workload: targets near-reachable)
  human-fk         gmr_b200.synth_human.synth_human_fk_clips (sinusoids on a human skeleton -> human FK)
  canonical        gmr_b200.synth_human.canonical_clips     (the reference's real clip out/test_canonical.bvh, tiled)

  python tools/workloads.py [--clips 4096] [--out gpurun_out/workloads.json]
"""
import argparse
import json
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
from gmr_b200 import model_pack, synth, synth_human  # noqa: E402
from gmr_b200.engine import BatchedRetargeter, DeviceModel  # noqa: E402
from oracle import oracle as oc  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--clips", type=int, default=4096)
    ap.add_argument("--frames", type=int, default=300)
    ap.add_argument("--out", default="gpurun_out/workloads.json")
    a = ap.parse_args()
    m, t = model_pack.build_tables("bvh", "unitree_g1")
    cores = len(os.sched_getaffinity(0))
    rows = []
    for name in ("robot-authored", "human-fk", "canonical"):
        if name == "robot-authored":
            human, heights, names = synth.synth_human_clips(m, t, "bvh", B=a.clips, T=a.frames, seed=0xB200 + 2, workers=cores)
        elif name == "human-fk":
            human, heights, names = synth_human.synth_human_fk_clips(m, t, "bvh", B=a.clips, T=a.frames)
        else:
            human, heights, names = synth_human.canonical_clips(a.clips)
        B, T = human.shape[:2]
        dm = DeviceModel.load("bvh", "unitree_g1", names)
        br = BatchedRetargeter(dm, B)
        br.set_heights(heights)
        hd = torch.from_numpy(human).cuda()
        out = torch.zeros((B, T, br.nq_max), dtype=torch.float64, device="cuda")
        _, diag = br.run(hd, diag=True, out=out)
        ms = []
        for _ in range(3):
            br.reset()
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(); br.run(hd, out=out); e1.record()
            torch.cuda.synchronize()
            ms.append(e0.elapsed_time(e1))
        it = diag["iters"].cpu().numpy()
        st = diag["status"].cpu().numpy()
        err = diag["err"].cpu().numpy()
        # parity on a sample: the first 32 clips against the C oracle
        n = min(32, B)
        ref = oc.run_clips(dm.desc, human[:n], scale=synth.clip_scales(t, heights[:n]), nthreads=cores, diag=True)
        q = out[:n, :, :m.nq].cpu().numpy()
        rows.append({"workload": name, "clips": B, "frames_per_clip": T, "ms_per_launch": float(np.mean(ms)),
                     "frames_per_s": B * T / float(np.mean(ms)) * 1e3, "iters_per_frame": float(it.sum(-1).mean()),
                     "frames_ending_on_max_iter": float(((st & 3) != 0).mean()), "qp_cap_frames": int(((st & 4) != 0).sum()),
                     "mean_error1": float(err[..., 0].mean()), "mean_error2": float(err[..., 1].mean()),
                     "parity_sample_clips": n, "max_abs_dqpos_vs_oracle": float(np.abs(q - ref["qpos"]).max()),
                     "iteration_counts_equal": bool((it[:n] == ref["iters"]).all()), "finite": bool(torch.isfinite(out).all().item())})
        print(json.dumps(rows[-1]), flush=True)
        br.close(); dm.close()
        del hd, out
    os.makedirs(os.path.dirname(a.out) or ".", exist_ok=True)
    json.dump(rows, open(a.out, "w"), indent=1)


if __name__ == "__main__":
    main()
