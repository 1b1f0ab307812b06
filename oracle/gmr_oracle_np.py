"""TEST INFRASTRUCTURE (oracle) - not part of the product path.

numpy float64 restatement of the reference hot path
``GeneralMotionRetargeting(src, robot).retarget(human_data)``
(reference: general_motion_retargeting/motion_retarget.py:13-270), with the
arithmetic the reference delegates to mink / MuJoCo / daqp restated from their
published algorithms:

    mujoco (unpinned, setup.py:17)   mj_kinematics, mj_jacBody, mj_integratePos
    mink   (unpinned, setup.py:16)   FrameTask, Configuration, solve_ik, ConfigurationLimit
    daqp   (not listed; default solver, motion_retarget.py:18)   exact dense QP

PARITY UNPINNED: none of those packages is installable here and the reference has
no test that pins a number on this path.  What *is* pinned to reference code:
target preparation (tests/golden/target_prep_*.npz, produced by running the
reference's own scale/offset methods) and forward kinematics
(tests/golden/fk_*.npz, produced by the reference's KinematicsModel).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
legs may import this module.
"""
from __future__ import annotations

import numpy as np

from . import lie_np as lie

JNT_FREE, JNT_SLIDE, JNT_HINGE = 0, 2, 3


# ----------------------------------------------------------------------------
# MuJoCo restatement
# ----------------------------------------------------------------------------
class Kin:
    """Result of mj_kinematics for one q."""
    __slots__ = ("xpos", "xquat", "xmat", "dof_axis", "dof_anchor")


def mj_kinematics(m, q) -> Kin:
    """SURVEY 8c item 2.  World = body 0 at origin."""
    nb = m.nbody
    k = Kin()
    k.xpos = np.zeros((nb, 3))
    k.xquat = np.zeros((nb, 4))
    k.xquat[0, 0] = 1.0
    k.xmat = np.zeros((nb, 3, 3))
    k.xmat[0] = np.eye(3)
    k.dof_axis = np.zeros((m.nv, 3))      # world axis of each dof (free: trans = world xyz, rot = body xyz)
    k.dof_anchor = np.zeros((m.nv, 3))    # world anchor of rotational dofs
    for b in range(1, nb):
        p = m.body_parentid[b]
        jadr, jnum = m.body_jntadr[b], m.body_jntnum[b]
        if jnum == 1 and m.jnt_type[jadr] == JNT_FREE:
            qa = m.jnt_qposadr[jadr]
            xpos = q[qa:qa + 3].copy()
            xquat = lie.normalize4(q[qa + 3:qa + 7])
            da = m.jnt_dofadr[jadr]
            R = lie.quat_to_mat(xquat)
            k.dof_axis[da:da + 3] = np.eye(3)
            k.dof_axis[da + 3:da + 6] = R.T          # rows = body axes in world
            k.dof_anchor[da:da + 6] = xpos
        else:
            xpos = k.xpos[p] + k.xmat[p] @ m.body_pos[b]
            xquat = lie.quat_mul(k.xquat[p], m.body_quat[b])
            for j in range(jadr, jadr + jnum):
                qa, da = m.jnt_qposadr[j], m.jnt_dofadr[j]
                R = lie.quat_to_mat(xquat)
                axis_w = R @ m.jnt_axis[j]
                anchor = xpos + R @ m.jnt_pos[j]
                k.dof_axis[da] = axis_w
                k.dof_anchor[da] = anchor
                val = q[qa] - m.qpos0[qa]
                if m.jnt_type[j] == JNT_SLIDE:
                    xpos = xpos + axis_w * val
                elif m.jnt_type[j] == JNT_HINGE:
                    half = 0.5 * val
                    qloc = np.concatenate([[np.cos(half)], np.sin(half) * m.jnt_axis[j]])
                    xquat = lie.quat_mul(xquat, qloc)
                    # off-centre rotation correction (zero for every shipped robot: jnt_pos = 0)
                    xpos = anchor - lie.quat_to_mat(xquat) @ m.jnt_pos[j]
                else:
                    raise ValueError("unsupported joint type")
            xquat = lie.normalize4(xquat)
        k.xpos[b] = xpos
        k.xquat[b] = xquat
        k.xmat[b] = lie.quat_to_mat(xquat)
    return k


def mj_jac_body(m, k: Kin, body: int):
    """mj_jacBody: world-aligned Jacobian at the body origin; returns (jacp[3,nv], jacr[3,nv])."""
    jacp = np.zeros((3, m.nv))
    jacr = np.zeros((3, m.nv))
    point = k.xpos[body]
    b = body
    while b != 0:
        jadr, jnum = m.body_jntadr[b], m.body_jntnum[b]
        for j in range(jadr, jadr + jnum):
            da = m.jnt_dofadr[j]
            t = m.jnt_type[j]
            if t == JNT_FREE:
                jacp[:, da:da + 3] = np.eye(3)
                for i in range(3):
                    a = k.dof_axis[da + 3 + i]
                    jacr[:, da + 3 + i] = a
                    jacp[:, da + 3 + i] = np.cross(a, point - k.dof_anchor[da + 3 + i])
            elif t == JNT_SLIDE:
                jacp[:, da] = k.dof_axis[da]
            else:
                a = k.dof_axis[da]
                jacr[:, da] = a
                jacp[:, da] = np.cross(a, point - k.dof_anchor[da])
        b = m.body_parentid[b]
    return jacp, jacr


def mj_integrate_pos(m, q, v, dt):
    """mj_integratePos (in place on a copy)."""
    q = q.copy()
    for j in range(m.njnt):
        qa, da = m.jnt_qposadr[j], m.jnt_dofadr[j]
        if m.jnt_type[j] == JNT_FREE:
            q[qa:qa + 3] += dt * v[da:da + 3]
            w = v[da + 3:da + 6]
            n = np.linalg.norm(w)
            quat = lie.normalize4(q[qa + 3:qa + 7])
            if n < 1e-15:
                axis, n = np.array([1.0, 0.0, 0.0]), 0.0
            else:
                axis = w / n
            ang = dt * n
            if ang != 0.0:
                qrot = np.concatenate([[np.cos(0.5 * ang)], np.sin(0.5 * ang) * axis])
                quat = lie.quat_mul(quat, qrot)
            q[qa + 3:qa + 7] = quat
        else:
            q[qa] += dt * v[da]
    return q


# ----------------------------------------------------------------------------
# mink restatement
# ----------------------------------------------------------------------------
def frame_task_error(k: Kin, body: int, tq, tp):
    """FrameTask.compute_error: log(T_b^-1 T_t)."""
    qi, pi = lie.se3_inv(k.xquat[body], k.xpos[body])
    qbt, pbt = lie.se3_mul(qi, pi, tq, tp)
    return lie.se3_log(qbt, pbt)


def frame_task_jacobian(m, k: Kin, body: int, tq, tp):
    """FrameTask.compute_jacobian: -jlog(T_t^-1 T_b) @ body Jacobian  (6 x nv)."""
    jacp, jacr = mj_jac_body(m, k, body)
    Rt = k.xmat[body].T
    jb = np.vstack([Rt @ jacp, Rt @ jacr])
    qti, pti = lie.se3_inv(tq, tp)
    qtb, ptb = lie.se3_mul(qti, pti, k.xquat[body], k.xpos[body])
    log_tb = lie.se3_log(qtb, ptb)
    jlog = lie.se3_ljacinv(-log_tb)          # rjacinv(x) = ljacinv(-x)
    return -jlog @ jb


def configuration_limit(m, q, gain=0.95, vmax=None, dt=None):
    """mink ConfigurationLimit (default limit set of solve_ik): box on limited hinge/slide dofs.
    vmax [nv] (optional, not used by the reference - motion_retarget.py:147-149 passes no limits): mink
    VelocityLimit, |dq| <= vmax * dt, intersected with the position box.

    Returns (lo[nv], hi[nv]) with +-inf where unconstrained:  lo <= dq <= hi.
    """
    lo = np.full(m.nv, -np.inf)
    hi = np.full(m.nv, np.inf)
    for j in range(m.njnt):
        if m.jnt_type[j] == JNT_FREE or not m.jnt_limited[j]:
            continue
        qa, da = m.jnt_qposadr[j], m.jnt_dofadr[j]
        hi[da] = gain * (m.jnt_range[j, 1] - q[qa])
        lo[da] = -gain * (q[qa] - m.jnt_range[j, 0])
    if vmax is not None:
        hi = np.minimum(hi, np.asarray(vmax) * dt)
        lo = np.maximum(lo, -np.asarray(vmax) * dt)
    return lo, hi


def build_ik(m, k: Kin, q, tasks, damping, vmax=None):
    """mink build_ik: H, c, lo, hi.  tasks = list of (body, wpos, wrot, tq, tp)."""
    H = np.eye(m.nv) * damping
    c = np.zeros(m.nv)
    for body, wpos, wrot, tq, tp in tasks:
        J = frame_task_jacobian(m, k, body, tq, tp)
        e = frame_task_error(k, body, tq, tp)
        W = np.diag([wpos] * 3 + [wrot] * 3)
        minus_gain_error = -1.0 * e
        WJ = W @ J
        We = W @ minus_gain_error
        mu = 1.0 * (We @ We)                  # lm_damping = 1 (motion_retarget.py:88,106)
        H += WJ.T @ WJ + mu * np.eye(m.nv)
        c += -We @ WJ
    lo, hi = configuration_limit(m, q, vmax=vmax, dt=m.timestep)
    return H, c, lo, hi


# ----------------------------------------------------------------------------
# exact box-constrained QP (stands in for daqp; unique minimiser of a strictly convex QP)
# ----------------------------------------------------------------------------
def solve_box_qp(H, c, lo, hi, ptol=1e-10, dtol=1e-9, max_iter=200):
    """min 1/2 x'Hx + c'x  s.t. lo <= x <= hi, H SPD.  Block principal pivoting
    (Judice & Pires) with Murty single-pivot fallback.  Returns (x, n_factorisations)."""
    n = len(c)
    state = np.zeros(n, np.int8)              # 0 free, -1 at lower, +1 at upper
    x = np.zeros(n)
    best = n + 1
    p = 3
    for it in range(max_iter):
        F = state == 0
        x = np.where(state < 0, lo, np.where(state > 0, hi, 0.0))
        if F.any():
            rhs = -c[F] - H[np.ix_(F, ~F)] @ x[~F]
            x[F] = np.linalg.solve(H[np.ix_(F, F)], rhs)
        g = H @ x + c
        bad = np.zeros(n, bool)
        bad |= F & ((x < lo - ptol) | (x > hi + ptol))
        bad |= (state < 0) & (g < -dtol)
        bad |= (state > 0) & (g > dtol)
        nbad = int(bad.sum())
        if nbad == 0:
            return x, it + 1
        if nbad < best:
            best, p = nbad, 3
            flip = bad
        elif p > 0:
            p -= 1
            flip = bad
        else:
            flip = np.zeros(n, bool)
            flip[np.nonzero(bad)[0].max()] = True
        for i in np.nonzero(flip)[0]:
            if state[i] == 0:
                state[i] = -1 if x[i] < lo[i] else 1
            else:
                state[i] = 0
    raise RuntimeError("box QP did not terminate")


def solve_ik(m, k: Kin, q, tasks, dt, damping, vmax=None):
    H, c, lo, hi = build_ik(m, k, q, tasks, damping, vmax=vmax)
    dq, nfac = solve_box_qp(H, c, lo, hi)
    return dq / dt, nfac


# ----------------------------------------------------------------------------
# the reference class, restated (motion_retarget.py:10-270)
# ----------------------------------------------------------------------------
class OracleRetargeter:
    """Same constructor / ``retarget`` contract as the reference class, driven by the
    flattened tables (gmr_b200.mjcf.RobotModel + gmr_b200.ik_config.RetargetTables)."""

    def __init__(self, model, tables, actual_human_height=None, solver="daqp", damping=5e-1,
                 verbose=False, dof_vmax=None):
        self.m = model
        self.t = tables
        ratio = 1.0 if actual_human_height is None else actual_human_height / tables.human_height_assumption
        self.scale = tables.scale * ratio                       # :36-43
        self.damping = damping
        self.dof_vmax = dof_vmax                                 # optional velocity box (None = the reference)
        self.max_iter = 10                                       # :56
        self.q = model.qpos0.copy()                              # mink.Configuration(model), :75
        self.kin = mj_kinematics(model, self.q)
        self.targets = [None, None]
        self.scaled_human_data = None
        self.stats = {"solves": 0, "factorisations": 0, "iters": [[], []], "err": [[], []]}

    # -- Kernel-1 part ------------------------------------------------------
    def prepare(self, frame, offset_to_ground=False):
        """frame: [h,7] array in the order of ``human_names`` given by ``index`` below,
        or an [s,7] array already gathered in ``tables.scaled_names`` order."""
        t = self.t
        s = t.n_scaled
        assert frame.shape == (s, 7)
        root_pos = frame[0, :3]
        out = np.zeros((s, 7))
        scaled_root = self.scale[0] * root_pos                   # :214
        for i in range(s):
            if i == 0:
                pos = scaled_root
            else:
                pos = (frame[i, :3] - root_pos) * self.scale[i] + scaled_root   # :222-230
            # offset (:234-250): scipy normalises the input quaternion
            qn = frame[i, 3:7] / np.linalg.norm(frame[i, 3:7])
            uq = lie.quat_mul(qn, t.rot_off[i])
            uq = uq / np.linalg.norm(uq)
            out[i, :3] = pos + lie.quat_rotate(uq, t.pos_off[i])
            out[i, 3:] = uq
        if offset_to_ground:                                      # :252-270
            feet = t.is_foot.astype(bool)
            lowest = np.inf
            for i in range(s):
                if feet[i] and out[i, 2] < lowest:
                    lowest = out[i, 2]
            out[:, 2] = out[:, 2] - lowest + 0.1
        return out

    def _tasks(self, stage, scaled):
        t = self.t
        return [(int(t.task_body[stage][i]), t.task_wpos[stage][i], t.task_wrot[stage][i],
                 scaled[t.task_src[stage][i], 3:7], scaled[t.task_src[stage][i], :3])
                for i in range(len(t.task_body[stage]))]

    def error(self, tasks):
        return float(np.linalg.norm(np.concatenate(
            [frame_task_error(self.kin, b, tq, tp) for b, _, _, tq, tp in tasks])))

    def _step(self, tasks):
        dt = self.m.timestep
        v, nfac = solve_ik(self.m, self.kin, self.q, tasks, dt, self.damping, vmax=self.dof_vmax)
        self.q = mj_integrate_pos(self.m, self.q, v, dt)
        self.kin = mj_kinematics(self.m, self.q)
        self.stats["solves"] += 1
        self.stats["factorisations"] += nfac

    def retarget(self, scaled_frame, offset_to_ground=False):
        """scaled_frame: [s,7] raw human (pos, wxyz) gathered in ``tables.scaled_names`` order."""
        scaled = self.prepare(np.asarray(scaled_frame, np.float64), offset_to_ground)
        self.scaled_human_data = scaled
        for stage in range(2):
            if not self.t.use_table[stage]:
                continue
            tasks = self._tasks(stage, scaled)
            curr = self.error(tasks)
            self._step(tasks)
            nxt = self.error(tasks)
            it = 0
            while curr - nxt > 0.001 and it < self.max_iter:      # :153,172
                curr = nxt
                self._step(tasks)
                nxt = self.error(tasks)
                it += 1
            self.stats["iters"][stage].append(it + 1)
            self.stats["err"][stage].append(nxt)
        return self.q.copy()
