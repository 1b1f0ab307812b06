"""CPU restatement (numpy, float64) of the reference's SMPL-X stream preparation.

TEST INFRASTRUCTURE ONLY: imported by tests/ and tools/, never by the product path (gmr_b200/smplx_prep.py calls
the CUDA kernel of csrc/gmr_smplx.cuh through the C ABI and fails loudly without it).

Parity status: PINNED - tests/golden/smplx_prep_*.npz hold the outputs of the reference's own
``get_smplx_data_offline_fast`` (general_motion_retargeting/utils/smpl.py:109-198, run in the build container by
tools/make_golden_smplx.py with scipy 1.18) and tests/test_smplx_oracle.py checks this restatement against them.

Follows smpl.py:78-107 (slerp), :118-174 (fps alignment), :178-196 (orientation chain) and, for the scipy
``Rotation`` calls those lines make, scipy's published formulas (spatial/transform/_rotation_xp.py of 1.18:
from_rotvec :159-179, as_rotvec :336-354, compose_quat :1114-1127, _normalize_quaternion, _quat_canonical :969-978,
Rotation.__mul__ normalises the product) and scipy.interpolate.interp1d(kind='linear') on float32 samples.
Quaternions are xyzw inside (scipy's storage), wxyz on output (as_quat(scalar_first=True), smpl.py:194).
"""
import numpy as np


def from_rotvec(v):
    v = np.asarray(v, np.float64)
    angle = np.sqrt(np.sum(v * v, axis=-1, keepdims=True))
    small = angle <= 1e-3
    a2 = angle * angle
    scale = np.where(small, 0.5 - a2 / 48 + a2 * a2 / 3840, np.sin(angle / 2) / (angle + small))
    return np.concatenate([v * scale, np.cos(angle / 2)], axis=-1)


def normalize(q):
    return q / np.sqrt(np.sum(q * q, axis=-1, keepdims=True))


def compose(p, q):
    cx = p[..., 1] * q[..., 2] - p[..., 2] * q[..., 1]
    cy = p[..., 2] * q[..., 0] - p[..., 0] * q[..., 2]
    cz = p[..., 0] * q[..., 1] - p[..., 1] * q[..., 0]
    x = p[..., 3] * q[..., 0] + q[..., 3] * p[..., 0] + cx
    y = p[..., 3] * q[..., 1] + q[..., 3] * p[..., 1] + cy
    z = p[..., 3] * q[..., 2] + q[..., 3] * p[..., 2] + cz
    w = p[..., 3] * q[..., 3] - p[..., 0] * q[..., 0] - p[..., 1] * q[..., 1] - p[..., 2] * q[..., 2]
    return normalize(np.stack([x, y, z, w], axis=-1))


def as_rotvec(q):
    neg = (q[..., 3] < 0) | ((q[..., 3] == 0) & ((q[..., 0] < 0) | ((q[..., 0] == 0) & ((q[..., 1] < 0) | ((q[..., 1] == 0) & (q[..., 2] < 0))))))
    q = np.where(neg[..., None], -q, q)
    axn = np.sqrt(np.sum(q[..., :3] ** 2, axis=-1, keepdims=True))
    angle = 2 * np.arctan2(axn, q[..., 3:4])
    small = angle <= 1e-3
    a2 = angle * angle
    scale = np.where(small, 2 + a2 / 12 + 7 * a2 * a2 / 2880, angle / (np.sin(angle / 2.0) + small))
    return scale * q[..., :3]


def slerp(q1, q2, t):
    """smpl.py:78-107 on xyzw quaternion arrays [...,4]; t broadcastable [...,1]."""
    q1, q2 = normalize(q1), normalize(q2)
    dot = np.sum(q1 * q2, axis=-1, keepdims=True)
    flip = dot < 0.0
    q2 = np.where(flip, -q2, q2)
    dot = np.where(flip, -dot, dot)
    lin = q1 + t * (q2 - q1)
    dc = np.clip(dot, -1.0, 1.0)
    th0 = np.arccos(dc)
    th = th0 * t
    s0 = np.cos(th) - dot * np.sin(th) / np.where(dot > 0.9995, 1.0, np.sin(th0))
    s1 = np.sin(th) / np.where(dot > 0.9995, 1.0, np.sin(th0))
    return normalize(np.where(dot > 0.9995, lin, s0 * q1 + s1 * q2))          # R.from_quat normalises


def interp_linear_f32(y, x_new):
    """scipy.interpolate.interp1d(arange(T), y, kind='linear')(x_new) for float32 y [T,...] (the samples are
    promoted to float64 before the slope is taken: checked against the golden vectors)."""
    y = np.asarray(y, np.float64)
    T = y.shape[0]
    idx = np.clip(np.searchsorted(np.arange(T), x_new), 1, T - 1)
    lo, hi = idx - 1, idx
    shp = (-1,) + (1,) * (y.ndim - 1)
    slope = (y[hi] - y[lo]) / (hi - lo).reshape(shp)
    return slope * (x_new - lo).reshape(shp) + y[lo]


def smplx_to_stream(full_pose, joints, parents, target_time=None):
    """full_pose [T,J,3] float32 rotvecs (joint 0 = global_orient), joints [T,>=J,3] float32.
    target_time: None = frames as they are (smpl.py:172-174) else float64 [T_out] source-frame times (:124-171).
    -> (pos [T_out,J,3] float64, quat wxyz [T_out,J,4])."""
    J = len(parents)
    pose = np.asarray(full_pose)[:, :J]
    T = pose.shape[0]
    if target_time is None:
        rv = pose.astype(np.float64)
        pos = np.asarray(joints)[:, :J].astype(np.float64)
    else:
        tt = np.asarray(target_time, np.float64)
        i1 = np.floor(tt).astype(np.int64)
        i2 = np.minimum(i1 + 1, T - 1)
        alpha = (tt - i1)[:, None, None]
        rv = as_rotvec(slerp(from_rotvec(pose[i1]), from_rotvec(pose[i2]), alpha))
        pos = interp_linear_f32(np.asarray(joints, np.float32), tt)[:, :J]
    lq = from_rotvec(rv)
    gq = np.zeros_like(lq)
    gq[:, 0] = lq[:, 0]
    for i in range(1, J):
        gq[:, i] = compose(gq[:, parents[i]], lq[:, i])
    return pos, np.concatenate([gq[..., 3:4], gq[..., :3]], axis=-1)
