"""Synthetic LAFAN1-shaped / SMPL-X-shaped human joint streams (host-side data generator).

There is no network and no mocap data on the build or GPU box, so benchmark and parity
inputs are generated.  A stream has the layout the reference's loaders produce
(utils/lafan1.py:17-40: 22 LAFAN1 joints + LeftFootMod/RightFootMod, metres, Z-up;
utils/smpl.py:178-196: 55 SMPL-X joints), i.e. ``[B, T, h, 7]`` = position xyz +
orientation wxyz per body, 30 fps.

To give the IK a retargeting-like workload (feasible whole-body poses, joints that run
into their limits now and then, a human whose proportions differ from the robot's), the
motion is authored on the *robot* and mapped back through the inverse of the config's
scale/offset tables:

1. a smooth random robot trajectory q(t): root random walk (xy speed <= 1.5 m/s, yaw
   rate <= 1 rad/s, small roll/pitch/height sway) + per-joint sums of 3 sinusoids
   (0.2-2 Hz) around a mid-range pose, deliberately overshooting joint ranges a little
   (clipped) so that limits become active;
2. robot FK -> pose of the robot frame matched to each scaled human body (table 1);
3. a per-clip "human": limb-length factors 1 +- 8 % about the root, body height
   U(1.55, 1.90) m, smooth positional noise (1.5 cm) and orientation noise (3 deg);
4. inverse of offset (motion_retarget.py:234-250) and scale (:209-232) -> raw human body
   poses; bodies the config does not use are filled with blends of used ones (ignored
   by the hot path).

Everything is float64 numpy, seeded, deterministic; the stream is returned as float32
(what the engine ingests) so oracle and GPU read identical values.
"""
from __future__ import annotations

import numpy as np

from .ik_config import RetargetTables
from .mjcf import JNT_FREE, JNT_HINGE, JNT_SLIDE, RobotModel

LAFAN1_JOINTS = [
    "Hips", "LeftUpLeg", "LeftLeg", "LeftFoot", "LeftToe", "RightUpLeg", "RightLeg", "RightFoot",
    "RightToe", "Spine", "Spine1", "Spine2", "Neck", "Head", "LeftShoulder", "LeftArm",
    "LeftForeArm", "LeftHand", "RightShoulder", "RightArm", "RightForeArm", "RightHand",
]
LAFAN1_BODIES = LAFAN1_JOINTS + ["LeftFootMod", "RightFootMod"]          # h = 24

_HAND = ["index", "middle", "pinky", "ring", "thumb"]
SMPLX_BODIES = [
    "pelvis", "left_hip", "right_hip", "spine1", "left_knee", "right_knee", "spine2", "left_ankle",
    "right_ankle", "spine3", "left_foot", "right_foot", "neck", "left_collar", "right_collar", "head",
    "left_shoulder", "right_shoulder", "left_elbow", "right_elbow", "left_wrist", "right_wrist",
    "jaw", "left_eye_smplhf", "right_eye_smplhf",
] + [f"left_{f}{i}" for f in _HAND for i in (1, 2, 3)] + [f"right_{f}{i}" for f in _HAND for i in (1, 2, 3)]   # h = 55

HUMAN_BODIES = {"bvh": LAFAN1_BODIES, "smplx": SMPLX_BODIES}


def human_body_names(src_human: str, tables: RetargetTables) -> list:
    """Frame layout for a source: the loader's body list when known (bvh, smplx), else
    (fbx / OptiTrack rigs with their own naming) just the bodies the config uses."""
    names = HUMAN_BODIES.get(src_human)
    if names is None or any(b not in names for b in tables.scaled_names):
        return list(tables.scaled_names)
    return list(names)


# ---------------------------------------------------------------- quaternion helpers (batched, wxyz)
def _qmul(a, b):
    aw, ax, ay, az = np.moveaxis(a, -1, 0)
    bw, bx, by, bz = np.moveaxis(b, -1, 0)
    return np.stack([aw * bw - ax * bx - ay * by - az * bz,
                     aw * bx + ax * bw + ay * bz - az * by,
                     aw * by - ax * bz + ay * bw + az * bx,
                     aw * bz + ax * by - ay * bx + az * bw], -1)


def _qconj(q):
    return q * np.array([1.0, -1.0, -1.0, -1.0])


def _qrot(q, v):
    w = q[..., :1]
    u = q[..., 1:]
    t = 2.0 * np.cross(u, v)
    return v + w * t + np.cross(u, t)


def _axis_angle(axis, ang):
    h = 0.5 * ang[..., None]
    return np.concatenate([np.cos(h), np.sin(h) * axis], -1)


def batch_fk(model: RobotModel, q: np.ndarray):
    """Vectorised mj_kinematics: q [N,nq] -> xpos [N,nbody,3], xquat [N,nbody,4]."""
    N = q.shape[0]
    xpos = np.zeros((N, model.nbody, 3))
    xquat = np.zeros((N, model.nbody, 4))
    xquat[:, 0, 0] = 1.0
    for b in range(1, model.nbody):
        p = model.body_parentid[b]
        jadr, jnum = model.body_jntadr[b], model.body_jntnum[b]
        if jnum == 1 and model.jnt_type[jadr] == JNT_FREE:
            qa = model.jnt_qposadr[jadr]
            pos = q[:, qa:qa + 3].copy()
            quat = q[:, qa + 3:qa + 7] / np.linalg.norm(q[:, qa + 3:qa + 7], axis=-1, keepdims=True)
        else:
            pos = xpos[:, p] + _qrot(xquat[:, p], model.body_pos[b])
            quat = _qmul(xquat[:, p], np.broadcast_to(model.body_quat[b], (N, 4)))
            for j in range(jadr, jadr + jnum):
                qa = model.jnt_qposadr[j]
                val = q[:, qa] - model.qpos0[qa]
                if model.jnt_type[j] == JNT_SLIDE:
                    pos = pos + _qrot(quat, model.jnt_axis[j]) * val[:, None]
                elif model.jnt_type[j] == JNT_HINGE:
                    quat = _qmul(quat, _axis_angle(model.jnt_axis[j], val))
        xpos[:, b] = pos
        xquat[:, b] = quat / np.linalg.norm(quat, axis=-1, keepdims=True)
    return xpos, xquat


def _smooth_noise(rng, shape_bt, n_comp, fps, f_lo, f_hi):
    """Sum of n_comp sinusoids per channel, unit-ish amplitude; shape_bt = (B, T, C)."""
    B, T, Cn = shape_bt
    t = np.arange(T) / fps
    f = rng.uniform(f_lo, f_hi, size=(B, 1, Cn, n_comp))
    ph = rng.uniform(0, 2 * np.pi, size=(B, 1, Cn, n_comp))
    a = rng.uniform(0.3, 1.0, size=(B, 1, Cn, n_comp))
    return (a * np.sin(2 * np.pi * f * t[None, :, None, None] + ph)).sum(-1) / np.sqrt(n_comp)


def synth_robot_motion(model: RobotModel, B: int, T: int, rng, fps: float = 30.0,
                       overshoot: float = 0.15) -> np.ndarray:
    """Smooth random robot trajectories q [B,T,nq] (see module docstring, step 1)."""
    q = np.tile(model.qpos0, (B, T, 1))
    t = np.arange(T) / fps
    for j in range(model.njnt):
        qa = model.jnt_qposadr[j]
        typ = model.jnt_type[j]
        if typ == JNT_FREE:
            speed = rng.uniform(0.0, 1.5, size=(B, 1))
            yaw_rate = _smooth_noise(rng, (B, T, 1), 3, fps, 0.05, 0.4)[..., 0] * 0.6
            yaw = rng.uniform(-np.pi, np.pi, size=(B, 1)) + np.cumsum(yaw_rate, 1) / fps
            gait = 0.6 + 0.4 * _smooth_noise(rng, (B, T, 1), 2, fps, 0.05, 0.3)[..., 0]
            vx, vy = speed * gait * np.cos(yaw), speed * gait * np.sin(yaw)
            q[:, :, qa + 0] = model.qpos0[qa + 0] + np.cumsum(vx, 1) / fps
            q[:, :, qa + 1] = model.qpos0[qa + 1] + np.cumsum(vy, 1) / fps
            sway = _smooth_noise(rng, (B, T, 3), 3, fps, 0.2, 2.0)
            q[:, :, qa + 2] = model.qpos0[qa + 2] - 0.04 + 0.03 * sway[..., 0]
            roll, pitch = 0.08 * sway[..., 1], 0.05 + 0.10 * sway[..., 2]
            qz = _axis_angle(np.array([0.0, 0.0, 1.0]), yaw)
            qy = _axis_angle(np.array([0.0, 1.0, 0.0]), pitch)
            qx = _axis_angle(np.array([1.0, 0.0, 0.0]), roll)
            q[:, :, qa + 3:qa + 7] = _qmul(_qmul(qz, qy), qx)
        else:
            if model.jnt_limited[j]:
                lo, hi = model.jnt_range[j]
            elif typ == JNT_SLIDE:
                lo, hi = -2.0, 2.0
            else:
                lo, hi = -np.pi, np.pi
            mid = 0.5 * (lo + hi)
            half = 0.5 * (hi - lo)
            centre = np.clip(0.0, lo + 0.25 * half, hi - 0.25 * half) if lo < 0 < hi else mid
            amp = np.minimum(0.6, half * (1.0 + overshoot)) * rng.uniform(0.2, 1.0, size=(B, 1))
            s = _smooth_noise(rng, (B, T, 1), 3, fps, 0.2, 2.0)[..., 0]
            val = centre + amp * s
            if model.jnt_limited[j]:
                val = np.clip(val, lo, hi)
            q[:, :, qa] = val
    del t
    return q


def synth_human_clips(model: RobotModel, tables: RetargetTables, src_human: str, B: int, T: int,
                      seed: int = 0xB200, fps: float = 30.0, dtype=np.float32, chunk: int = 256,
                      workers: int = 1, out=None):
    """-> (human [B,T,h,7] dtype, heights [B] float64, human_names list).
    Chunks of `chunk` clips are seeded independently ([seed, first clip]), so the result does not
    depend on `workers` (processes used to generate chunks in parallel).  `out` may be a
    preallocated [B,T,h,7] array (e.g. the numpy view of a pinned torch tensor)."""
    names = human_body_names(src_human, tables)
    h = len(names)
    s = tables.n_scaled
    idx = [names.index(b) for b in tables.scaled_names]
    # robot frame matched to each scaled body (table 1; fall back to table 2)
    body_of = np.zeros(s, np.int64)
    for i in range(s):
        found = False
        for k in range(2):
            hit = np.nonzero(tables.task_src[k] == i)[0]
            if len(hit) and tables.task_body[k][hit[0]] >= 0:
                body_of[i] = tables.task_body[k][hit[0]]
                found = True
                break
        if not found:
            raise KeyError(tables.scaled_names[i])
    if out is None:
        out = np.zeros((B, T, h, 7), dtype)
    assert out.shape == (B, T, h, 7)
    heights = np.zeros(B)
    jobs = [(model, tables, body_of, idx, h, seed, c0, min(B, c0 + chunk), T, fps, np.dtype(dtype)) for c0 in range(0, B, chunk)]
    if workers > 1 and len(jobs) > 1:
        import multiprocessing as mp
        with mp.get_context("fork").Pool(min(workers, len(jobs))) as pool:
            for c0, c1, arr, hts in pool.imap_unordered(_synth_chunk, jobs):
                out[c0:c1] = arr
                heights[c0:c1] = hts
    else:
        for job in jobs:
            c0, c1, arr, hts = _synth_chunk(job)
            out[c0:c1] = arr
            heights[c0:c1] = hts
    return out, heights, list(names)


def _synth_chunk(job):
    """One independently seeded chunk of clips -> (c0, c1, human [nb,T,h,7], heights [nb])."""
    model, tables, body_of, idx, h, seed, c0, c1, T, fps, dtype = job
    s = tables.n_scaled
    if True:
        nb = c1 - c0
        rng = np.random.default_rng([seed, c0])
        q = synth_robot_motion(model, nb, T, rng, fps)
        xpos, xquat = batch_fk(model, q.reshape(nb * T, model.nq))
        tp = xpos[:, body_of].reshape(nb, T, s, 3)
        tq = xquat[:, body_of].reshape(nb, T, s, 4)
        # per-clip human: limb-length factors about the root + smooth noise
        limb = 1.0 + 0.08 * rng.uniform(-1, 1, size=(nb, 1, s, 1))
        limb[:, :, 0] = 1.0
        root = tp[:, :, :1]
        tp = root + (tp - root) * limb
        tp = tp + 0.015 * _smooth_noise(rng, (nb, T, s * 3), 2, fps, 0.3, 3.0).reshape(nb, T, s, 3)
        rv = np.deg2rad(3.0) * _smooth_noise(rng, (nb, T, s * 3), 2, fps, 0.3, 3.0).reshape(nb, T, s, 3)
        ang = np.linalg.norm(rv, axis=-1)
        axis = rv / np.maximum(ang, 1e-12)[..., None]
        tq = _qmul(tq, _axis_angle(axis, ang))
        # invert offset: uq = qh * rot_off ; target_pos = scaled_pos + R(uq) pos_off
        qh = _qmul(tq, _qconj(tables.rot_off))
        scaled_pos = tp - _qrot(tq, tables.pos_off)
        # invert scale
        hts = rng.uniform(1.55, 1.90, size=nb)
        sc = tables.scale[None, :] * (hts / tables.human_height_assumption)[:, None]     # [nb,s]
        p_root = scaled_pos[:, :, :1] / sc[:, None, :1, None]
        p = (scaled_pos - scaled_pos[:, :, :1]) / sc[:, None, :, None] + p_root
        p[:, :, 0] = p_root[:, :, 0]
        # unused bodies: deterministic blends of used ones (never read by the hot path)
        full_p = np.zeros((nb, T, h, 3))
        full_q = np.zeros((nb, T, h, 4))
        for k in range(h):
            a, b2 = k % s, (3 * k + 1) % s
            full_p[:, :, k] = 0.5 * (p[:, :, a] + p[:, :, b2])
            full_q[:, :, k] = qh[:, :, a]
        full_p[:, :, idx] = p
        full_q[:, :, idx] = qh
        # sign-continuous quaternions along time
        dots = np.sum(full_q[:, 1:] * full_q[:, :-1], axis=-1)
        flip = np.cumprod(np.where(dots < 0, -1.0, 1.0), axis=1)
        full_q[:, 1:] *= flip[..., None]
        arr = np.empty((nb, T, h, 7), dtype)
        arr[..., :3] = full_p
        arr[..., 3:] = full_q
    return c0, c1, arr, hts


def clip_scales(tables: RetargetTables, heights: np.ndarray) -> np.ndarray:
    """Per-clip human_scale_table with actual_human_height folded in (motion_retarget.py:36-43)."""
    return tables.scale[None, :] * (np.asarray(heights, np.float64) / tables.human_height_assumption)[:, None]
