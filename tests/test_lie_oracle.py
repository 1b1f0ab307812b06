"""Oracle self-checks: the restated mink Lie maths against independent means."""
import numpy as np
import pytest
from scipy.linalg import logm
from scipy.spatial.transform import Rotation as R

from oracle import lie_np as lie


def _rand_T(rng, ang_scale=1.0, pos_scale=0.5):
    w = rng.normal(size=3) * ang_scale
    return lie.so3_exp(w), rng.normal(size=3) * pos_scale


def test_so3_log_matches_scipy():
    rng = np.random.default_rng(0)
    for _ in range(200):
        q = rng.normal(size=4)
        q /= np.linalg.norm(q)
        ref = R.from_quat(q, scalar_first=True).as_rotvec()
        np.testing.assert_allclose(lie.so3_log(q), ref, atol=1e-12)


def test_se3_log_matches_matrix_log():
    rng = np.random.default_rng(1)
    for _ in range(50):
        q, p = _rand_T(rng)
        T = np.eye(4)
        T[:3, :3] = lie.quat_to_mat(q)
        T[:3, 3] = p
        L = np.real(logm(T))
        ref = np.array([L[0, 3], L[1, 3], L[2, 3], L[2, 1], L[0, 2], L[1, 0]])
        np.testing.assert_allclose(lie.se3_log(q, p), ref, atol=1e-9)


def test_exp_log_roundtrip():
    rng = np.random.default_rng(2)
    for _ in range(100):
        c = rng.normal(size=6) * 0.8
        q, p = lie.se3_exp(c)
        np.testing.assert_allclose(lie.se3_log(q, p), c, atol=1e-12)


def test_task_jacobian_factor_is_exact_derivative():
    """-Jl^-1(e) must equal d/d(delta) log(exp(-delta) T) at 0, i.e. the derivative of the task
    error w.r.t. a body-frame twist of the robot frame (verified invariant)."""
    rng = np.random.default_rng(3)
    for _ in range(20):
        q, p = _rand_T(rng, 0.7, 0.3)
        e = lie.se3_log(q, p)
        J = -lie.se3_ljacinv(e)
        h = 1e-6
        Jfd = np.zeros((6, 6))
        for k in range(6):
            d = np.zeros(6)
            d[k] = h
            qp, pp = lie.se3_mul(*lie.se3_exp(-d), q, p)
            qm, pm = lie.se3_mul(*lie.se3_exp(d), q, p)
            Jfd[:, k] = (lie.se3_log(qp, pp) - lie.se3_log(qm, pm)) / (2 * h)
        np.testing.assert_allclose(J, Jfd, atol=5e-9)


def test_ljacinv_small_angle_is_identity_like_mink():
    c = np.array([0.3, -0.2, 0.1, 1e-6, 0.0, 0.0])
    np.testing.assert_array_equal(lie.se3_ljacinv(c), np.eye(6))
