#!/usr/bin/env python
"""Third workload: the one real motion the reference ships, out/test_canonical.bvh (250 frames, 101-bone
Character-Creator rig), turned into a LAFAN1-shaped frame stream by the REFERENCE's own code (this container only):

  load_lafan1_file (utils/lafan1.py:8-71)  ->  _fill_synonyms, _synthesize_foot_mod, _synthesize_spine2,
  _apply_orientation_fix(preset) (scripts/bvh_to_robot.py:34-185, called in that order at :310-313)

The helper functions live in a script that imports the OpenGL viewer at module level, so their source (everything
between `def _quat_mul` and `if __name__`) is exec()'d as is.  The rig's file is Z-up while the loader assumes Y-up;
`--orient_fix auto` (an option of the reference script) stands it upright.  Output: tests/golden/canonical_clip.npz
with frames [250, 24, 7] in gmr_b200.synth.LAFAN1_BODIES order (bodies the rig does not have are copies of Hips; the
bvh -> g1 config does not read them), the loader's height guess and the preset used.  bench.py / tools read the npz
(the GPU box has no /root/reference) and tile it with per-clip heights and headings.

    python tools/make_canonical_clip.py [--gmr-root /root/reference] [--orient_fix auto]
"""
import argparse
import os
import sys
import types

import numpy as np

REPO = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..")
sys.path.insert(0, REPO)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gmr-root", default="/root/reference")
    ap.add_argument("--orient_fix", default="auto")
    ap.add_argument("--out", default=os.path.join(REPO, "tests", "golden", "canonical_clip.npz"))
    args = ap.parse_args()
    pkg = types.ModuleType("general_motion_retargeting")
    pkg.__path__ = [os.path.join(args.gmr_root, "general_motion_retargeting")]
    sys.modules.setdefault("general_motion_retargeting", pkg)
    from general_motion_retargeting.utils.lafan1 import load_lafan1_file
    text = open(os.path.join(args.gmr_root, "scripts", "bvh_to_robot.py")).read()
    beg, end = text.index("def _quat_mul"), text.index('if __name__ == "__main__":')
    ns = {"np": np, "sqrt": np.sqrt, "print": lambda *a, **k: None}
    exec(compile(text[beg:end], "scripts/bvh_to_robot.py", "exec"), ns)
    frames, height = load_lafan1_file(os.path.join(args.gmr_root, "out", "test_canonical.bvh"))
    frames = ns["_fill_synonyms"](frames)
    frames = ns["_synthesize_foot_mod"](frames)
    frames = ns["_synthesize_spine2"](frames)
    frames = ns["_apply_orientation_fix"](frames, args.orient_fix)
    from gmr_b200 import model_pack
    from gmr_b200.synth import LAFAN1_BODIES
    m, t = model_pack.build_tables("bvh", "unitree_g1")
    missing = [b for b in t.scaled_names if b not in frames[0]]
    if missing:
        raise SystemExit(f"bodies the bvh -> g1 config reads are missing after the synonym fill: {missing}")
    arr = np.zeros((len(frames), len(LAFAN1_BODIES), 7))
    have = []
    for k, n in enumerate(LAFAN1_BODIES):
        src = n if n in frames[0] else "Hips"
        have.append(n in frames[0])
        for i, f in enumerate(frames):
            arr[i, k, :3], arr[i, k, 3:] = f[src][0], f[src][1]
    np.savez_compressed(args.out, frames=arr, names=np.array(LAFAN1_BODIES), present=np.array(have), height=np.array(float(height)),
                        orient_fix=np.array(args.orient_fix), source=np.array("out/test_canonical.bvh"),
                        scaled_names=np.array(list(t.scaled_names)))
    z = arr[:, LAFAN1_BODIES.index("Hips"), 2]
    print("wrote", args.out, arr.shape, "height", float(height), "hips z", float(z.min()), float(z.max()),
          "absent:", [n for n, h in zip(LAFAN1_BODIES, have) if not h])


if __name__ == "__main__":
    main()
