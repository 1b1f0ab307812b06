#!/usr/bin/env python
"""Generate tests/golden/tail_*.npz by EXECUTING the reference dataset drivers' own post-processing lines
(this container only; needs /root/reference and CPU torch).

The tail of a driver is script code, not a function: scripts/smplx_to_robot_dataset.py:93-146 and
scripts/bvh_to_robot_dataset.py:106-150 take ``qpos_list`` (the stacked results of retarget()), build the
reference's ``KinematicsModel``, swap the root quaternion to xyzw IN PLACE, run the float32 FK with an identity
root (``local_body_pos``), optionally shift the root height by the lowest body point of the clip and zero the root
XY of frame 0, and fill ``motion_data``.  This generator cuts exactly those lines out of the reference scripts
(from ``device = "cuda:0"`` to the closing brace of ``motion_data = {...}``), replaces the device string by
"cpu" and nothing else, and exec()s them on qpos produced by the repo's CPU oracle for seeded synthetic clips.
What is pinned is therefore the reference's own tail arithmetic (numpy in-place ops + its torch FK), fed with
known qpos; the IK that produces qpos is outside this file (tests/golden/ik_*.npz, tools/make_golden_ik.py).

    python tools/make_golden_tail.py [--gmr-root /root/reference]
"""
import argparse
import importlib
import os
import sys
import textwrap
import types

import numpy as np

REPO = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..")
sys.path.insert(0, REPO)

CASES = [  # (tag, driver script, source, robot, clip lengths, seed)
    ("bvh_unitree_g1", "scripts/bvh_to_robot_dataset.py", "bvh", "unitree_g1", [17, 5, 30, 1, 22], 9),
    ("smplx_unitree_g1", "scripts/smplx_to_robot_dataset.py", "smplx", "unitree_g1", [12, 3, 25], 10),
    ("smplx_booster_t1", "scripts/smplx_to_robot_dataset.py", "smplx", "booster_t1", [9, 14], 12),
]


def tail_source(gmr_root, script):
    """The reference driver's tail, verbatim: from `device = "cuda:0"` to the end of the motion_data literal."""
    lines = open(os.path.join(gmr_root, script)).read().split("\n")
    beg = next(i for i, l in enumerate(lines) if l.strip() == 'device = "cuda:0"')
    md = next(i for i, l in enumerate(lines) if i > beg and l.strip().startswith("motion_data = {"))
    end = next(i for i, l in enumerate(lines) if i > md and l.strip() == "}")
    src = textwrap.dedent("\n".join(lines[beg:end + 1]))
    assert src.count("cuda:0") == 1
    return src.replace('"cuda:0"', '"cpu"'), (beg + 1, end + 1)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gmr-root", default="/root/reference")
    ap.add_argument("--out", default=os.path.join(REPO, "tests", "golden"))
    args = ap.parse_args()
    import torch
    from gmr_b200 import model_pack, params, synth
    from gmr_b200.abi import ModelDesc
    from oracle import oracle as oc
    params.set_gmr_root(args.gmr_root)
    pkg = types.ModuleType("general_motion_retargeting")
    pkg.__path__ = [os.path.join(args.gmr_root, "general_motion_retargeting")]
    sys.modules["general_motion_retargeting"] = pkg
    km_mod = importlib.import_module("general_motion_retargeting.kinematics_model")
    oc.build()
    for tag, script, src, robot, lengths, seed in CASES:
        code, span = tail_source(args.gmr_root, script)
        m, t = model_pack.build_tables(src, robot)
        human, heights, names = synth.synth_human_clips(m, t, src, B=len(lengths), T=max(lengths), seed=seed)
        desc = ModelDesc(m, t, names)
        qpos = oc.run_clips(desc, human, T_of=np.array(lengths, np.int32), scale=synth.clip_scales(t, heights), nthreads=4)["qpos"]
        rec = {"lengths": np.array(lengths), "script": np.array(script), "lines": np.array(span), "robot": np.array(robot),
               "source": np.array(src)}
        for i, n in enumerate(lengths):
            q = qpos[i, :n, :m.nq].copy()
            rec[f"qpos_{i}"] = q.copy()
            ns = {"np": np, "torch": torch, "KinematicsModel": km_mod.KinematicsModel, "qpos_list": q,
                  "retargeter": types.SimpleNamespace(xml_file=str(params.ROBOT_XML_DICT[robot])),
                  "retarget": types.SimpleNamespace(xml_file=str(params.ROBOT_XML_DICT[robot])),
                  "aligned_fps": 30.0, "src_fps": 30, "smplx_file_path": "<golden>", "log_memory": lambda *a, **k: None}
            # the reference's lines, on CPU, as the body of a function (the smplx tail holds a `return`)
            fn = "def __tail__():\n" + textwrap.indent(code, "    ") + "\n    return motion_data\n"
            exec(compile(fn, script, "exec"), ns)
            md = ns["__tail__"]()
            for k in ("root_pos", "root_rot", "dof_pos", "local_body_pos"):
                rec[f"{k}_{i}"] = np.array(md[k])
            rec[f"fps_{i}"] = np.array(md["fps"])
            rec["link_body_list"] = np.array(md["link_body_list"])
        path = os.path.join(args.out, f"tail_{tag}.npz")
        np.savez_compressed(path, **rec)
        print("wrote", path, f"(reference lines {span[0]}-{span[1]} of {script})")


if __name__ == "__main__":
    main()
