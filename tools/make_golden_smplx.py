#!/usr/bin/env python
"""Generate tests/golden/smplx_prep_*.npz by running the REFERENCE's own SMPL-X stream preparation
(general_motion_retargeting/utils/smpl.py:109-198 ``get_smplx_data_offline_fast``, unmodified) in this container.

The SMPL-X body model itself (licensed files, the ``smplx`` package) is absent, so the function is fed the three
arrays it reads from the model output - ``global_orient``, ``full_pose``, ``joints`` (float32 torch tensors, as
``smplx`` returns them) - with synthetic values; ``smplx`` / ``smplx.joint_names`` are stubbed with the public
joint list.  What is pinned: fps alignment (SLERP of rotation vectors, linear interpolation of joints) and the
global-orientation chain down the kinematic tree.

    python tools/make_golden_smplx.py [--gmr-root /root/reference]
"""
import argparse
import os
import sys
import types

import numpy as np
import torch

REPO = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..")
sys.path.insert(0, REPO)
from gmr_b200.smplx_prep import SMPLX_JOINT_NAMES, SMPLX_PARENTS  # noqa: E402


def synth(T, seed, n_joint_out=127):
    rng = np.random.default_rng(seed)
    J = len(SMPLX_PARENTS)
    t = np.arange(T)[:, None, None] / 120.0
    amp = rng.uniform(0.05, 1.2, (1, J, 3))
    pose = amp * np.sin(2 * np.pi * rng.uniform(0.3, 3.0, (1, J, 3)) * t + rng.uniform(0, 6.28, (1, J, 3)))
    pose[:, 5] *= 1e-4                       # a joint that stays inside the small-angle series
    pose[:, 7] = pose[0:1, 7]                # a joint that does not move (dot > 0.9995: linear branch)
    pose[:, 9, 0] += 3.0 + 2.5 * t[:, 0, 0] * 40      # sweeps through pi and beyond: canonical sign in as_rotvec
    k = np.arange(T)
    pose[:, 13] = 0.0                        # axis-angle wrap: nearly the same rotation, opposite quaternion hemisphere
    pose[:, 13, 1] = np.where(k % 2 == 0, 1.0, -1.0) * (np.pi - 0.03 - 0.001 * k)
    pose[:, 11] *= 6.0                       # fast: large steps between source frames
    pose = pose.astype(np.float32)
    joints = (rng.normal(0, 0.4, (1, n_joint_out, 3)) + 0.2 * np.sin(3 * t + rng.uniform(0, 6, (1, n_joint_out, 3)))).astype(np.float32)
    return pose, joints


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gmr-root", default="/root/reference")
    args = ap.parse_args()
    smplx = types.ModuleType("smplx")
    jn = types.ModuleType("smplx.joint_names")
    jn.JOINT_NAMES = list(SMPLX_JOINT_NAMES) + [f"extra_{i}" for i in range(72)]
    smplx.joint_names = jn
    sys.modules["smplx"], sys.modules["smplx.joint_names"] = smplx, jn
    pkg = types.ModuleType("general_motion_retargeting")
    pkg.__path__ = [os.path.join(args.gmr_root, "general_motion_retargeting")]
    sys.modules["general_motion_retargeting"] = pkg
    from general_motion_retargeting.utils.smpl import get_smplx_data_offline_fast

    gold = os.path.join(REPO, "tests", "golden")
    for name, T, src_fps, tgt_fps, seed in (("skip4", 41, 120.0, 30, 1), ("native", 12, 30.0, 30, 2), ("skip1", 7, 50.0, 30, 3)):
        pose, joints = synth(T, seed)
        J = pose.shape[1]
        out = types.SimpleNamespace(global_orient=torch.from_numpy(pose[:, 0:1].copy()),        # [T,1,3]
                                    full_pose=torch.from_numpy(pose.reshape(T, J * 3).copy()),
                                    joints=torch.from_numpy(joints.copy()))
        model = types.SimpleNamespace(parents=np.asarray(SMPLX_PARENTS))
        data = {"mocap_frame_rate": np.array(src_fps), "pose_body": np.zeros((T, 63), np.float32)}
        frames, fps = get_smplx_data_offline_fast(data, model, out, tgt_fps=tgt_fps)
        rows = list(frames[0].keys())
        assert rows == list(SMPLX_JOINT_NAMES)
        pos = np.stack([np.stack([np.asarray(fr[n][0], np.float64) for n in rows]) for fr in frames])
        quat = np.stack([np.stack([np.asarray(fr[n][1], np.float64) for n in rows]) for fr in frames])
        np.savez_compressed(os.path.join(gold, f"smplx_prep_{name}.npz"), pose=pose, joints=joints, src_fps=np.float64(src_fps),
                            tgt_fps=np.int32(tgt_fps), pos=pos, quat=quat, aligned_fps=np.float64(fps),
                            pos_dtype=np.array(str(np.asarray(frames[0][rows[0]][0]).dtype)))
        print(name, "T", T, "->", len(frames), "aligned_fps", fps, "pos dtype", np.asarray(frames[0][rows[0]][0]).dtype)


if __name__ == "__main__":
    main()
