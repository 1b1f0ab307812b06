"""SMPL-X model output -> human frame stream: the step right before the retargeter in the reference's SMPL-X
drivers.

Reference: ``get_smplx_data_offline_fast`` (general_motion_retargeting/utils/smpl.py:109-198) - per-frame,
per-joint Python loops over scipy ``Rotation`` objects: fps alignment (SLERP of each joint's rotation vector
between the two neighbouring source frames, linear interpolation of the joint positions, :124-174) and the
global-orientation chain ``R_i = R_parent(i) * R(rotvec_i)`` down the kinematic tree (:178-196).
The body-model forward itself (``smplx.create`` / ``body_model(...)``, :11-43) needs the licensed SMPL-X files
and stays outside: this module takes what it returns (``global_orient``, ``full_pose``, ``joints``).

Here one CUDA pass (``gmr_smplx_to_stream``: ``csrc/gmr_smplx.cuh``, thread = output frame) does both for all
frames of all clips and writes the ``[frames, 55, 7]`` stream (pos xyz, quat wxyz) the retargeter reads.
"""
from __future__ import annotations

import ctypes as C

import numpy as np

_HAND = ["index", "middle", "pinky", "ring", "thumb"]
# smplx.joint_names.JOINT_NAMES[:55] (public SMPL-X joint order) and the model's kinematic tree
SMPLX_JOINT_NAMES = [
    "pelvis", "left_hip", "right_hip", "spine1", "left_knee", "right_knee", "spine2", "left_ankle",
    "right_ankle", "spine3", "left_foot", "right_foot", "neck", "left_collar", "right_collar", "head",
    "left_shoulder", "right_shoulder", "left_elbow", "right_elbow", "left_wrist", "right_wrist",
    "jaw", "left_eye_smplhf", "right_eye_smplhf",
] + [f"left_{f}{i}" for f in _HAND for i in (1, 2, 3)] + [f"right_{f}{i}" for f in _HAND for i in (1, 2, 3)]
SMPLX_PARENTS = [-1, 0, 0, 0, 1, 2, 3, 4, 5, 6, 7, 8, 9, 9, 9, 12, 13, 14, 16, 17, 18, 19, 15, 15, 15,
                 20, 25, 26, 20, 28, 29, 20, 31, 32, 20, 34, 35, 20, 37, 38,
                 21, 40, 41, 21, 43, 44, 21, 46, 47, 21, 49, 50, 21, 52, 53]


def alignment(num_frames: int, src_fps: float, tgt_fps: float):
    """smpl.py:118-131,172-174: -> (target_time float64 [T_out] in source frames or None when the clip is used as is,
    aligned_fps)."""
    frame_skip = int(src_fps / tgt_fps)
    if tgt_fps < src_fps:
        new_num = num_frames // frame_skip
        target = np.linspace(0, num_frames - 1, new_num)
        return target, (new_num / num_frames * src_fps if num_frames else tgt_fps)
    return None, tgt_fps


def smplx_to_stream(clips, tgt_fps: float = 30, device: int = 0, dtype=np.float32, parents=None, stream=None):
    """clips: list of dicts {full_pose [T,J*3] or [T,J,3] float32 rotation vectors (joint 0 = global_orient),
    joints [T,>=J,3] float32, fps}.  -> (streams: list of [T_out_i, J, 7] torch CUDA tensors, aligned fps per clip)."""
    import torch

    from . import _lib
    from .abi import F32, F64
    parents = np.asarray(SMPLX_PARENTS if parents is None else parents, np.int32)
    J = len(parents)
    if not clips:
        return [], []
    tdev = torch.device("cuda", device)
    poses, joints, src_start, times, aligned, out_start, fps_out = [], [], [0], [], [], [0], []
    Jall = None
    for c in clips:
        p = np.ascontiguousarray(c["full_pose"], np.float32)
        T = p.shape[0]
        p = p.reshape(T, -1, 3)
        if p.shape[1] < J:
            raise ValueError(f"full_pose carries {p.shape[1]} joints, the tree has {J}")
        jt = np.ascontiguousarray(c["joints"], np.float32)
        if jt.shape[0] != T or jt.shape[1] < J:
            raise ValueError("joints must be [T, >= J, 3]")
        Jall = jt.shape[1] if Jall is None else Jall
        if jt.shape[1] != Jall:
            raise ValueError("clips of one batch must carry the same number of joints")
        tt, f = alignment(T, float(c["fps"]), tgt_fps)
        if tt is None:
            tt, al = np.arange(T, dtype=np.float64), 0
        else:
            al = 1
        poses.append(p[:, :J]); joints.append(jt)
        times.append(tt); aligned.append(np.full(len(tt), al, np.int32))
        src_start.append(src_start[-1] + T); out_start.append(out_start[-1] + len(tt)); fps_out.append(f)
    N, N_out = src_start[-1], out_start[-1]
    d_pose = torch.from_numpy(np.concatenate(poses, 0)).to(tdev)
    d_joints = torch.from_numpy(np.concatenate(joints, 0)).to(tdev)
    clip_of = np.repeat(np.arange(len(clips)), np.diff(out_start)).astype(np.int32)
    i32 = lambda a: torch.from_numpy(np.ascontiguousarray(a, np.int32)).to(tdev)      # noqa: E731
    d_src, d_clip, d_al, d_par = i32(src_start), i32(clip_of), i32(np.concatenate(aligned) if N_out else np.zeros(0)), i32(parents)
    d_time = torch.from_numpy(np.concatenate(times) if N_out else np.zeros(0)).to(tdev)
    tdt = torch.float32 if np.dtype(dtype) == np.float32 else torch.float64
    out = torch.empty((N_out, J, 7), dtype=tdt, device=tdev)
    ptr = lambda t: C.c_void_p(t.data_ptr())                                           # noqa: E731
    s = stream if stream is not None else torch.cuda.current_stream(tdev).cuda_stream
    _lib.check(_lib.lib().gmr_smplx_to_stream(
        device, C.c_void_p(s), ptr(d_pose), ptr(d_joints), int(Jall), ptr(d_src), ptr(d_clip), ptr(d_time), ptr(d_al),
        int(N), int(N_out), int(J), ptr(d_par), F32 if tdt == torch.float32 else F64, ptr(out)), "gmr_smplx_to_stream")
    return [out[out_start[i]:out_start[i + 1]] for i in range(len(clips))], fps_out
