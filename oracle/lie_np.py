"""TEST INFRASTRUCTURE (oracle) - not part of the product path.

numpy float64 restatement of the Lie-group maths the reference gets from the
third-party package ``mink`` (kevinzakka/mink, unpinned in the reference's setup.py:16;
0.0.x series of mid-2025).  mink's source is not under /root/reference and cannot be
installed here, so its published formulas are restated and
anchored on the reference's call sites:

* ``task.compute_error``    <- motion_retarget.py:191,198  : e = log(T_b^-1 T_t), [v; w]
* ``task.compute_jacobian`` <- inside mink.solve_ik, motion_retarget.py:147 :
  J = -jlog(T_t^-1 T_b) J_body = -Jl^-1(e) J_body
* ``mink.SE3.from_rotation_and_translation(mink.SO3(q), p)`` <- motion_retarget.py:130,136

PARITY UNPINNED: no golden vector of mink's exists in the reference tree; the formulas
are validated by independent means in tests/ (finite differences, scipy logm, scipy
Rotation).

Conventions: quaternions wxyz; SE3 tangent = [linear(3), angular(3)].
"""
import numpy as np

EPS = 1e-10     # mink.lie.utils.get_epsilon(float64)


def skew(w):
    return np.array([[0.0, -w[2], w[1]], [w[2], 0.0, -w[0]], [-w[1], w[0], 0.0]])


def quat_mul(a, b):
    aw, ax, ay, az = a
    bw, bx, by, bz = b
    return np.array([
        aw * bw - ax * bx - ay * by - az * bz,
        aw * bx + ax * bw + ay * bz - az * by,
        aw * by - ax * bz + ay * bw + az * bx,
        aw * bz + ax * by - ay * bx + az * bw,
    ])


def quat_conj(q):
    return np.array([q[0], -q[1], -q[2], -q[3]])


def quat_to_mat(q):
    w, x, y, z = q
    return np.array([
        [1 - 2 * (y * y + z * z), 2 * (x * y - w * z), 2 * (x * z + w * y)],
        [2 * (x * y + w * z), 1 - 2 * (x * x + z * z), 2 * (y * z - w * x)],
        [2 * (x * z - w * y), 2 * (y * z + w * x), 1 - 2 * (x * x + y * y)],
    ])


def quat_rotate(q, v):
    return quat_to_mat(q) @ v


def normalize4(q):
    """mju_normalize4: (1,0,0,0) for a zero quaternion, else q/|q|."""
    n = np.linalg.norm(q)
    if n < 1e-15:
        return np.array([1.0, 0.0, 0.0, 0.0])
    return q / n


def so3_log(q):
    """Rotation vector of a unit quaternion (mink SO3.log)."""
    w = q[0]
    xyz = q[1:]
    norm_sq = xyz @ xyz
    if norm_sq < EPS:
        factor = 2.0 / w - 2.0 / 3.0 * norm_sq / w ** 3
    else:
        norm = np.sqrt(norm_sq)
        if abs(w) < EPS:
            factor = (1.0 if w > 0 else -1.0) * np.pi / norm
        else:
            factor = 2.0 * np.arctan2(-norm if w < 0 else norm, abs(w)) / norm
    return factor * xyz


def so3_exp(w):
    th = np.linalg.norm(w)
    if th < 1e-12:
        q = np.array([1.0, 0.5 * w[0], 0.5 * w[1], 0.5 * w[2]])
        return q / np.linalg.norm(q)
    s = np.sin(0.5 * th) / th
    return np.array([np.cos(0.5 * th), s * w[0], s * w[1], s * w[2]])


def se3_log(q, p):
    """log of T = (R(q), p): [V^-1 p ; w]  (mink SE3.log)."""
    w = so3_log(q)
    th2 = w @ w
    W = skew(w)
    W2 = W @ W
    if th2 < EPS:
        Vinv = np.eye(3) - 0.5 * W + W2 / 12.0
    else:
        th = np.sqrt(th2)
        half = 0.5 * th
        Vinv = np.eye(3) - 0.5 * W + ((1.0 - 0.5 * th * np.cos(half) / np.sin(half)) / th2) * W2
    return np.concatenate([Vinv @ p, w])


def so3_ljacinv(w):
    th = np.sqrt(w @ w)
    if th < EPS:
        t2 = th * th
        A = (1.0 / 12.0) * (1.0 + t2 / 60.0 * (1.0 + t2 / 42.0 * (1.0 + t2 / 40.0)))
    else:
        A = (1.0 / th ** 2) * (1.0 - (th * np.sin(th) / (2.0 * (1.0 - np.cos(th)))))
    W = skew(w)
    return np.eye(3) - 0.5 * W + A * (W @ W)


def _se3_Q(c):
    th2 = c[3:] @ c[3:]
    A = 0.5
    if th2 < EPS:
        B = 1.0 / 6.0 + 1.0 / 120.0 * th2
        C = -1.0 / 24.0 + 1.0 / 720.0 * th2
        D = -1.0 / 60.0
    else:
        th = np.sqrt(th2)
        s, co = np.sin(th), np.cos(th)
        B = (th - s) / (th2 * th)
        C = (1.0 - th2 / 2.0 - co) / (th2 * th2)
        D = (2.0 * th - 3.0 * s + th * co) / (2.0 * th2 * th2 * th)
    V = skew(c[:3])
    W = skew(c[3:])
    VW = V @ W
    WV = VW.T
    WVW = WV @ W
    VWW = VW @ W
    return A * V + B * (WV + VW + WVW) - C * (VWW - VWW.T - 3.0 * WVW) + D * (WVW @ W + W @ WVW)


def se3_ljacinv(c):
    """Inverse left Jacobian of SE3 at tangent c=[v;w] (mink SE3.ljacinv).

    Identity when |w|^2 < eps, exactly as mink does (it drops the -1/2 [v]x block there).
    """
    w = c[3:]
    if w @ w < EPS:
        return np.eye(6)
    Q = _se3_Q(c)
    Ji = so3_ljacinv(w)
    out = np.zeros((6, 6))
    out[:3, :3] = Ji
    out[:3, 3:] = -Ji @ Q @ Ji
    out[3:, 3:] = Ji
    return out


def se3_exp(c):
    """exp of tangent [v;w] -> (q, p).  Used only by tests (finite differences)."""
    v, w = c[:3], c[3:]
    th2 = w @ w
    W = skew(w)
    if th2 < 1e-16:
        V = np.eye(3) + 0.5 * W
    else:
        th = np.sqrt(th2)
        V = np.eye(3) + (1 - np.cos(th)) / th2 * W + (th - np.sin(th)) / (th2 * th) * (W @ W)
    return so3_exp(w), V @ v


def se3_mul(qa, pa, qb, pb):
    return quat_mul(qa, qb), quat_rotate(qa, pb) + pa


def se3_inv(q, p):
    qi = quat_conj(q)
    return qi, -quat_rotate(qi, p)
