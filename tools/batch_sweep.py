#!/usr/bin/env python
"""Launch time of the fused kernel against the number of clips (development probe): python tools/batch_sweep.py"""
import os, sys
import numpy as np, torch
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
from gmr_b200 import model_pack, synth
from gmr_b200.engine import BatchedRetargeter, DeviceModel
m, t = model_pack.build_tables("bvh", "unitree_g1")
Bmax = 16384
human, heights, names = synth.synth_human_clips(m, t, "bvh", B=Bmax, T=300, seed=0xB200 + 2, workers=len(os.sched_getaffinity(0)))
dm = DeviceModel.load("bvh", "unitree_g1", names)
hd = torch.from_numpy(human).cuda()
for B in (1, 148, 592, 1184, 2368, 3552, 4096, 4736, 8192, 16384):
    br = BatchedRetargeter(dm, B)
    br.set_heights(heights[:B])
    out = torch.zeros((B, 300, br.nq_max), dtype=torch.float64, device="cuda")
    ms = []
    for rep in range(3):
        br.reset(); torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); _, diag = br.run(hd[:B], out=out, diag=True); e1.record(); torch.cuda.synchronize()
        ms.append(e0.elapsed_time(e1))
    it = diag["iters"].sum(-1).sum(-1).cpu().numpy()
    print(f"B={B:6d}  {min(ms):8.2f} ms  {B*300/min(ms)*1e3/1e6:6.2f} M frames/s  solves: mean/clip {it.mean():.0f} max {it.max()}  "
          f"us per solve of the slowest clip if it alone bounded the launch {min(ms)*1e3/it.max():.2f}", flush=True)
    br.close()
