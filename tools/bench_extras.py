"""Extra blocks of bench.py's JSON line (not the headline): the other BASELINE.json configs on the same kernel.

  configs[0]  one clip x 1000 frames through the B = 1 shim GeneralMotionRetargeting.retarget(dict)
  configs[2]  unitree_g1_with_hands <- SMPL-X-shaped streams, 8192 clips x 300
  configs[3]  heterogeneous launch: every robot constructible for smplx in ONE launch, instance = (clip, robot)
  configs[4]  strong scaling: 33 334 clips x 300 (~10 M frames) cut across the ranks by gmr_b200.shard.shard_clips

Every number is timed on the device with CUDA events around whole launches, after one untimed launch.
"""
from __future__ import annotations

import json
import os
import time

import numpy as np
import torch

from gmr_b200 import model_pack, synth
from gmr_b200.engine import BatchedRetargeter, DeviceModel
from gmr_b200.shard import shard_clips

SMPLX_ROBOTS = ["unitree_g1", "unitree_g1_with_hands", "booster_t1", "stanford_toddy", "fourier_n1", "engineai_pm01",
                "kuavo_s45", "hightorque_hi", "galaxea_r1pro", "booster_k1"]
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _timed(br, human_dev, out, reps):
    br.reset()
    br.run(human_dev, out=out)                       # untimed
    ms = []
    for _ in range(reps):
        br.reset()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        br.run(human_dev, out=out)
        e1.record()
        torch.cuda.synchronize()
        ms.append(e0.elapsed_time(e1))
    return float(np.mean(ms))


def hands_block(device, workers, B=8192, T=300, reps=2):
    """configs[2]."""
    src, robot = "smplx", "unitree_g1_with_hands"
    m, t = model_pack.build_tables(src, robot)
    human, heights, names = synth.synth_human_clips(m, t, src, B=B, T=T, seed=0xB200 + 3, workers=workers)
    dm = DeviceModel.load(src, robot, names, device=device)
    br = BatchedRetargeter(dm, B)
    br.set_heights(heights)
    hd = torch.from_numpy(human).to(f"cuda:{device}")
    out = torch.zeros((B, T, br.nq_max), dtype=torch.float64, device=hd.device)
    ms = _timed(br, hd, out, reps)
    res = {"config": "BASELINE.json configs[2]: unitree_g1_with_hands (nv 49) <- synthetic SMPL-X-shaped streams (55 joints)",
           "clips": B, "frames_per_clip": T, "ms_per_launch": ms, "frames_per_s": B * T / ms * 1e3, "launch": br.launch_info(),
           "finite": bool(torch.isfinite(out).all().item())}
    br.close(); dm.close()
    return res


def hetero_block(device, workers, per=820, T=300, reps=2):
    """configs[3]: the same number of clips for each of the ten smplx robots, one launch."""
    models, humans, model_of, heights_all = [], [], [], []
    for i, r in enumerate(SMPLX_ROBOTS):
        m, t = model_pack.build_tables("smplx", r)
        human, heights, names = synth.synth_human_clips(m, t, "smplx", B=per, T=T, seed=0xB200 + 40 + i, workers=workers)
        models.append(DeviceModel.load("smplx", r, names, device=device))
        humans.append(human)
        heights_all += list(heights)
        model_of += [i] * per
    B = per * len(SMPLX_ROBOTS)
    br = BatchedRetargeter(models, B, model_of_instance=model_of)
    br.set_heights(heights_all)
    hd = torch.from_numpy(np.concatenate(humans, 0)).to(f"cuda:{device}")
    out = torch.zeros((B, T, br.nq_max), dtype=torch.float64, device=hd.device)
    ms = _timed(br, hd, out, reps)
    res = {"config": "BASELINE.json configs[3]: heterogeneous launch, instance = (clip, robot), ten smplx robots",
           "robots": SMPLX_ROBOTS, "clips": B, "frames_per_clip": T, "ms_per_launch": ms, "frames_per_s": B * T / ms * 1e3,
           "launch": br.launch_info(), "finite": bool(torch.isfinite(out).all().item())}
    br.close()
    for dm in models:
        dm.close()
    return res


def shim_block(T=1000):
    """configs[0] on the GPU side: ONE object, one frame dict per call - the teleoperation loop the reference's
    README quotes at 60-70 FPS (README.md:615-621)."""
    from gmr_b200.motion_retarget import GeneralMotionRetargeting
    m, t = model_pack.build_tables("bvh", "unitree_g1")
    human, heights, names = synth.synth_human_clips(m, t, "bvh", B=1, T=T, seed=0xB200 + 1, dtype=np.float64)
    frames = [{n: (human[0, i, k, :3], human[0, i, k, 3:]) for k, n in enumerate(names)} for i in range(T)]
    gmr = GeneralMotionRetargeting("bvh", "unitree_g1", actual_human_height=float(heights[0]))
    for f in frames[:50]:
        gmr.retarget(dict(f))
    gmr2 = GeneralMotionRetargeting("bvh", "unitree_g1", actual_human_height=float(heights[0]))
    t0 = time.perf_counter()
    for f in frames:
        gmr2.retarget(f)
    dt = time.perf_counter() - t0
    return {"config": "BASELINE.json configs[0]: unitree_g1 <- one synthetic LAFAN1-shaped clip x 1000 frames, one "
                      "GeneralMotionRetargeting.retarget(dict) call per frame (host dict in, host ndarray out, wall clock)",
            "frames": T, "frames_per_s": T / dt, "us_per_frame": dt / T * 1e6, "published_reference_fps": "60-70 (README.md:615-621)"}


def strong_block(br_factory, human_dev, rank, world, maxr, barrier, total_clips=33334, reps=2):
    """configs[4]: `total_clips` x T frames, cut across the ranks by shard_clips (longest-processing-time first; all
    clips have T frames here, so the cut is even).  The motions are this rank's bench clips tiled (clip g plays motion
    g mod B with its own seeded body height): distinct IK problems, no extra synthesis time.  Every rank runs its own
    shard, no collective in the data path; time = max over ranks."""
    B, T = human_dev.shape[0], human_dev.shape[1]
    mine = shard_clips(np.full(total_clips, T, np.int64), world)[rank]
    n = len(mine)
    idx = torch.from_numpy(np.asarray(mine) % B).to(human_dev.device)
    hs = human_dev[idx].contiguous()
    heights = np.random.default_rng(0xB204).uniform(1.55, 1.90, size=total_clips)[mine]
    br = br_factory(n)
    br.set_heights(heights)
    out = torch.zeros((n, T, br.nq_max), dtype=torch.float64, device=human_dev.device)
    br.reset(); br.run(hs, out=out)
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        br.reset()
        br.run(hs, out=out)
    e1.record()
    barrier()
    ms = maxr(e0.elapsed_time(e1)) / reps
    finite = bool(torch.isfinite(out).all().item())
    br.close()
    res = {"config": "BASELINE.json configs[4]: dataset-scale retarget sharded by clip (shard_clips), strong scaling",
           "total_clips": total_clips, "frames_per_clip": T, "total_frames": total_clips * T, "n_gpus": world,
           "clips_on_rank0": n, "ms_per_pass": ms, "frames_per_s": total_clips * T / ms * 1e3, "finite_rank0": finite}
    p = os.path.join(ROOT, "profiles", "strong_n1.json")
    if world == 1:
        res["efficiency"] = 1.0
    elif os.path.exists(p):
        with open(p) as f:
            n1 = json.load(f)
        if n1.get("total_clips") == total_clips and n1.get("frames_per_clip") == T:
            res["n1_frames_per_s"] = n1["frames_per_s"]
            res["n1_source"] = "profiles/strong_n1.json (N = 1 run of this block)"
            res["efficiency"] = res["frames_per_s"] / (world * n1["frames_per_s"])
    res["limiter"] = ("a launch is bounded below by its slowest clip's chain of solves; with total_clips / N clips per GPU "
                      "below ~8000 the GPU leaves its throughput regime (profiles/README.md, batch sweep)")
    return res
