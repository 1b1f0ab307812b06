#!/usr/bin/env python
"""Development probe: CUDA-vs-oracle deviation of the fused path for a few robots (max |dqpos|, max error excess)."""
import os, sys
import numpy as np
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
from gmr_b200 import model_pack, synth
from gmr_b200.engine import BatchedRetargeter, DeviceModel
from oracle import oracle as oc
for src, robot in (("bvh", "fourier_n1"), ("bvh", "unitree_g1"), ("smplx", "unitree_g1_with_hands"), ("smplx", "galaxea_r1pro")):
    m, t = model_pack.build_tables(src, robot)
    human, heights, names = synth.synth_human_clips(m, t, src, B=48, T=40, seed=3)
    dm = DeviceModel.load(src, robot, names)
    br = BatchedRetargeter(dm, 48)
    br.set_heights(heights)
    qpos, diag = br.run(human, diag=True)
    ref = oc.run_clips(dm.desc, human, scale=synth.clip_scales(t, heights), nthreads=8)
    d = np.abs(qpos.cpu().numpy()[..., :m.nq] - ref["qpos"])
    e = diag["err"].cpu().numpy() - ref["err"]
    w = np.unravel_index(d.argmax(), d.shape)
    print(f"{robot:24s} max|dq| {d.max():.3e} at clip {w[0]} frame {w[1]} q[{w[2]}]  median-frame max {np.median(d.max(-1)):.2e}  "
          f"err excess max {e.max():.3e} min {e.min():.3e}  iters equal {(diag['iters'].cpu().numpy() == ref['iters']).all()}", flush=True)
