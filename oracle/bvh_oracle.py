"""CPU restatement (numpy, float64) of the reference's BVH loader after the text parse.

TEST INFRASTRUCTURE ONLY: imported by tests/ and tools/, never by the product path (gmr_b200/bvh.py calls the
CUDA kernels of csrc/gmr_bvh.cuh through the C ABI and fails loudly without them).

Parity status: PINNED - tests/golden/bvh_*.npz hold the outputs of the reference's own ``load_lafan1_file``
(general_motion_retargeting/utils/lafan1.py:8-71, run in the build container by tools/make_golden_bvh.py) on
committed synthetic BVH files, and tests/test_bvh_oracle.py checks this restatement against them.

Follows, in order:
  utils/lafan_vendor/extract.py:162-163   radians -> euler_to_quat -> remove_quat_discontinuities
  utils/lafan_vendor/utils.py:40-75       angle_axis_to_quat, euler_to_quat  (q = q0 * (q1 * q2))
  utils/lafan_vendor/utils.py:251-269     remove_quat_discontinuities (sequential in time)
  utils/lafan_vendor/utils.py:83-103      quat_fk
  utils/lafan_vendor/utils.py:132-163     quat_mul, quat_mul_vec
  utils/lafan1.py:20-39                   Y-up cm -> Z-up m, FootMod rows
  utils/lafan1.py:42-69                   height heuristic
"""
import numpy as np

# scipy 1.18 Rotation.from_matrix([[1,0,0],[0,0,-1],[0,1,0]]).as_quat(scalar_first=True)  (lafan1.py:20-21)
ROT_W = float.fromhex("0x1.6a09e667f3bccp-1")
ROT_QUAT = np.array([ROT_W, ROT_W, 0.0, 0.0])


def quat_mul(x, y):
    """utils.py:132-148 (operand order of every product kept)."""
    x0, x1, x2, x3 = x[..., 0], x[..., 1], x[..., 2], x[..., 3]
    y0, y1, y2, y3 = y[..., 0], y[..., 1], y[..., 2], y[..., 3]
    return np.stack([y0 * x0 - y1 * x1 - y2 * x2 - y3 * x3,
                     y0 * x1 + y1 * x0 - y2 * x3 + y3 * x2,
                     y0 * x2 + y1 * x3 + y2 * x0 - y3 * x1,
                     y0 * x3 - y1 * x2 + y2 * x1 + y3 * x0], axis=-1)


def quat_mul_vec(q, v):
    """utils.py:151-163."""
    t = 2.0 * np.cross(q[..., 1:], v)
    return v + q[..., 0:1] * t + np.cross(q[..., 1:], t)


def euler_to_quat(e_rad, order):
    """utils.py:56-75; order = axis index (0/1/2) of each of the three channels."""
    qs = []
    for k in range(3):
        half = e_rad[..., k] / 2.0
        q = np.zeros(e_rad.shape[:-1] + (4,))
        q[..., 0] = np.cos(half)
        q[..., 1 + order[k]] = np.sin(half)
        qs.append(q)
    return quat_mul(qs[0], quat_mul(qs[1], qs[2]))


def remove_quat_discontinuities(q):
    """utils.py:251-269: frame i is negated when that brings it closer to the (already processed) frame i-1."""
    q = q.copy()
    for i in range(1, q.shape[0]):
        d = np.sum(q[i - 1] * q[i], axis=-1)
        flip = d < -d
        q[i][flip] = -q[i][flip]
    return q


def quat_fk(lrot, lpos, parents):
    """utils.py:83-103."""
    gr = np.zeros_like(lrot)
    gp = np.zeros_like(lpos)
    gr[..., 0, :], gp[..., 0, :] = lrot[..., 0, :], lpos[..., 0, :]
    for i in range(1, len(parents)):
        p = parents[i]
        gp[..., i, :] = quat_mul_vec(gr[..., p, :], lpos[..., i, :]) + gp[..., p, :]
        gr[..., i, :] = quat_mul(gr[..., p, :], lrot[..., i, :])
    return gr, gp


def height_from_last_frame(last, head, feet):
    """lafan1.py:42-69 on the rows of the last frame ([h,7]); head = row of "Head" or -1, feet = candidate rows."""
    z = last[:, 2]
    if z.shape[0] == 0:
        return 1.75
    if head >= 0:
        hgt = z[head] - (min(z[f] for f in feet) if len(feet) else z.min())
    else:
        hgt = z.max() - z.min()
    if not np.isfinite(hgt) or hgt < 0.9 or hgt > 2.3:
        hgt = 1.75
    return float(hgt)


def bvh_to_stream(positions, eulers_deg, parents, order, src_pos, src_rot, head=-1, feet=()):
    """positions, eulers_deg [T,J,3] -> ([T,h,7] float64 stream, height)."""
    T = positions.shape[0]
    lrot = remove_quat_discontinuities(euler_to_quat(np.radians(eulers_deg), order))
    gr, gp = quat_fk(lrot, positions, parents)
    rot = quat_mul(np.broadcast_to(ROT_QUAT, gr.shape), gr)                       # lafan1.py:31
    pos = np.stack([gp[..., 0], -gp[..., 2], gp[..., 1]], axis=-1) / 100.0         # lafan1.py:32 (p @ R.T / 100)
    out = np.concatenate([pos[:, src_pos], rot[:, src_rot]], axis=-1)
    hgt = height_from_last_frame(out[-1], head, list(feet)) if T else 1.75
    return out, hgt
