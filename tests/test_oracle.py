"""CPU: the oracle against (a) golden vectors produced by the reference's own code,
(b) independent numerical checks, (c) its second implementation."""
import glob
import os

import numpy as np
import pytest
from scipy.optimize import lsq_linear

from gmr_b200 import model_pack, synth
from gmr_b200.abi import ModelDesc
from oracle import gmr_oracle_np as onp
from oracle import oracle as oc

PAIRS = [("bvh", "unitree_g1"), ("smplx", "unitree_g1_with_hands"), ("smplx", "galaxea_r1pro"),
         ("bvh", "engineai_pm01"), ("smplx", "hightorque_hi")]


def _desc(src, robot):
    m, t = model_pack.build_tables(src, robot)
    return m, t, ModelDesc(m, t, synth.human_body_names(src, t))


def _random_q(m, rng):
    q = m.qpos0.copy()
    for j in range(m.njnt):
        qa = m.jnt_qposadr[j]
        if m.jnt_type[j] == 0:
            q[qa:qa + 3] += rng.normal(size=3) * 0.3
            w = rng.normal(size=4)
            q[qa + 3:qa + 7] = w / np.linalg.norm(w)
        elif m.jnt_limited[j]:
            q[qa] = rng.uniform(*m.jnt_range[j])
        else:
            q[qa] = rng.uniform(-1, 1)
    return q


# ---------------------------------------------------------------- golden: reference target prep
def test_target_prep_matches_reference_golden(golden_dir):
    files = sorted(glob.glob(os.path.join(golden_dir, "target_prep_*.npz")))
    assert len(files) >= 18
    for f in files:
        src, robot = os.path.basename(f)[len("target_prep_"):-4].split("_", 1)
        g = np.load(f)
        m, t = model_pack.build_tables(src, robot)
        names = [str(n) for n in g["names"]]
        d = ModelDesc(m, t, names)
        human, heights = g["human"], g["heights"]
        for b in range(human.shape[0]):
            scale = synth.clip_scales(t, heights[b:b + 1])[0]
            assert list(g[f"task_frames1_b{b}"]) == t.task_frame_names[0]
            assert list(g[f"task_frames2_b{b}"]) == t.task_frame_names[1]
            np.testing.assert_array_equal(g[f"task_costs1_b{b}"][:, 0], t.task_wpos[0])
            np.testing.assert_array_equal(g[f"task_costs2_b{b}"][:, 1], t.task_wrot[1])
            for ground in (0, 1):
                ref_scaled = g[f"scaled_b{b}_g{ground}"]
                for ti in range(human.shape[1]):
                    out = oc.prepare(d, human[b, ti], scale, bool(ground))
                    np.testing.assert_allclose(out, ref_scaled[ti], atol=2e-14, rtol=0)
                    for k in range(2):
                        if not t.use_table[k]:
                            continue
                        tgt = g[f"target{k + 1}_b{b}_g{ground}"][ti]
                        np.testing.assert_allclose(out[t.task_src[k]], tgt, atol=2e-14, rtol=0)
                # numpy restatement too (one frame)
                o = onp.OracleRetargeter(m, t, actual_human_height=float(heights[b]))
                idx = [names.index(n) for n in t.scaled_names]
                np.testing.assert_allclose(o.prepare(human[b, 0][idx], bool(ground)), ref_scaled[0], atol=2e-14)


# ---------------------------------------------------------------- golden: reference KinematicsModel FK
def test_fk_matches_reference_kinematics_model(golden_dir):
    files = sorted(glob.glob(os.path.join(golden_dir, "fk_*.npz")))
    assert len(files) >= 8
    for f in files:
        robot = os.path.basename(f)[3:-4]
        g = np.load(f)
        m = model_pack.load_robot(robot)
        assert [str(n) for n in g["body_names"]] == m.body_names[1:]
        np.testing.assert_array_equal(g["parent"][1:] + 1, m.body_parentid[2:])
        _, t = model_pack.build_tables("smplx", robot)
        d = ModelDesc(m, t)
        for n in range(g["dof_pos"].shape[0]):
            q = m.qpos0.copy()
            q[:3] = g["root_pos"][n]
            q[3:7] = g["root_rot_xyzw"][n][[3, 0, 1, 2]]
            q[7:] = g["dof_pos"][n]
            xpos, xquat = oc.fk(d, q)
            np.testing.assert_allclose(xpos[1:], g["body_pos"][n], atol=5e-6)
            ref_q = g["body_rot_xyzw"][n][:, [3, 0, 1, 2]]
            sign = np.sign(np.sum(ref_q * xquat[1:], axis=-1, keepdims=True))
            np.testing.assert_allclose(xquat[1:], sign * ref_q, atol=5e-6)
            k = onp.mj_kinematics(m, q)
            np.testing.assert_allclose(k.xpos, xpos, atol=1e-13)
            np.testing.assert_allclose(k.xquat, xquat, atol=1e-13)


# ---------------------------------------------------------------- independent checks
@pytest.mark.parametrize("src,robot", PAIRS)
def test_task_jacobian_is_derivative_of_error(src, robot):
    m, t, d = _desc(src, robot)
    rng = np.random.default_rng(5)
    q = _random_q(m, rng)
    k = onp.mj_kinematics(m, q)
    for body in sorted(set(int(b) for b in t.task_body[0]))[:6]:
        tq = onp.lie.so3_exp(rng.normal(size=3) * 0.4)
        tq = onp.lie.quat_mul(k.xquat[body], tq)
        tp = k.xpos[body] + rng.normal(size=3) * 0.1
        J = onp.frame_task_jacobian(m, k, body, tq, tp)
        Jfd = np.zeros_like(J)
        h = 1e-6
        for i in range(m.nv):
            dq = np.zeros(m.nv)
            dq[i] = h
            kp = onp.mj_kinematics(m, onp.mj_integrate_pos(m, q, dq, 1.0))
            km = onp.mj_kinematics(m, onp.mj_integrate_pos(m, q, -dq, 1.0))
            Jfd[:, i] = (onp.frame_task_error(kp, body, tq, tp) - onp.frame_task_error(km, body, tq, tp)) / (2 * h)
        np.testing.assert_allclose(J, Jfd, atol=2e-8)
        Jc, ec = oc.task_jacobian(d, q, body, np.concatenate([tp, tq]))
        np.testing.assert_allclose(Jc, J, atol=1e-12)
        np.testing.assert_allclose(ec, onp.frame_task_error(k, body, tq, tp), atol=1e-13)


def test_box_qp_matches_bvls_and_kkt():
    rng = np.random.default_rng(9)
    for trial in range(40):
        n = int(rng.integers(5, 50))
        A = rng.normal(size=(n + 10, n))
        H = A.T @ A + 0.5 * np.eye(n)
        c = rng.normal(size=n) * 5
        lo = -np.abs(rng.normal(size=n)) * 0.3
        hi = np.abs(rng.normal(size=n)) * 0.3
        free = rng.random(n) < 0.2
        lo[free], hi[free] = -np.inf, np.inf
        x, nf = oc.box_qp(H, c, lo, hi)
        xn, _ = onp.solve_box_qp(H, c, lo, hi)
        np.testing.assert_allclose(x, xn, atol=1e-11)
        # KKT
        g = H @ x + c
        assert np.all(x >= lo - 1e-9) and np.all(x <= hi + 1e-9)
        at_lo, at_hi = np.isclose(x, lo, atol=1e-9), np.isclose(x, hi, atol=1e-9)
        inner = ~(at_lo | at_hi)
        assert np.all(np.abs(g[inner]) < 1e-7)
        assert np.all(g[at_lo] > -1e-7) and np.all(g[at_hi] < 1e-7)
        # independent solver: bounded least squares on the Cholesky factor
        L = np.linalg.cholesky(H)
        res = lsq_linear(L.T, -np.linalg.solve(L, c), bounds=(lo, hi), method="bvls", tol=1e-14)
        np.testing.assert_allclose(x, res.x, atol=1e-7)


@pytest.mark.parametrize("src,robot", PAIRS)
def test_c_oracle_matches_numpy_oracle_on_clip(src, robot):
    m, t, d = _desc(src, robot)
    human, heights, names = synth.synth_human_clips(m, t, src, B=2, T=12, seed=3)
    r = oc.run_clips(d, human, scale=synth.clip_scales(t, heights))
    idx = [names.index(n) for n in t.scaled_names]
    for b in range(2):
        o = onp.OracleRetargeter(m, t, actual_human_height=float(heights[b]))
        qs = np.stack([o.retarget(human[b, i][idx].astype(np.float64)) for i in range(human.shape[1])])
        np.testing.assert_allclose(qs, r["qpos"][b], atol=1e-10)
        for k in range(2):
            if t.use_table[k]:
                assert o.stats["iters"][k] == list(r["iters"][b, :, k])


def test_build_qp_c_matches_numpy():
    m, t, d = _desc("bvh", "unitree_g1")
    rng = np.random.default_rng(2)
    q = _random_q(m, rng)
    human, heights, names = synth.synth_human_clips(m, t, "bvh", B=1, T=1, seed=5, dtype=np.float64)
    scaled = oc.prepare(d, human[0, 0], synth.clip_scales(t, heights)[0])
    o = onp.OracleRetargeter(m, t)
    k = onp.mj_kinematics(m, q)
    for stage in range(2):
        H, c, lo, hi, e = oc.build_qp(d, q, stage, scaled)
        Hn, cn, lon, hin = onp.build_ik(m, k, q, o._tasks(stage, scaled), 0.5)
        np.testing.assert_allclose(H, Hn, rtol=1e-12, atol=1e-9)
        np.testing.assert_allclose(c, cn, rtol=1e-12, atol=1e-9)
        np.testing.assert_array_equal(lo, lon)
        np.testing.assert_array_equal(hi, hin)


def test_threads_give_identical_results():
    m, t, d = _desc("bvh", "unitree_g1")
    human, heights, _ = synth.synth_human_clips(m, t, "bvh", B=6, T=10, seed=1)
    T_of = np.array([10, 3, 7, 1, 10, 5], np.int32)
    a = oc.run_clips(d, human, T_of=T_of, scale=synth.clip_scales(t, heights), nthreads=1)
    b = oc.run_clips(d, human, T_of=T_of, scale=synth.clip_scales(t, heights), nthreads=4)
    np.testing.assert_array_equal(a["qpos"], b["qpos"])
    assert np.all(a["qpos"][1, 3:] == 0)
