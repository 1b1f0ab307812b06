"""CPU: the static schedule gmr_model_create builds for the fused kernel's box-QP solver (DevModel.aba_*).

A host-only model (device = GMR_HOST_ONLY, no CUDA call) exposes the tables through gmr_model_table; this test
walks them exactly as csrc/gmr_stages.cuh::qp_solve_aba does - same block array, same byte offsets, same order of
passes and 8-lane groups, the chains' accumulators and velocities in per-group registers - with random task stiffnesses on the real kinematic trees, and compares the
result (solution and multipliers of locked dofs) with a dense solve of the same QP.  All ten (source, robot) pairs,
both stages, plus an edited table with two task frames on one dof.
"""
import copy
import ctypes as C

import numpy as np
import pytest

PAIRS = [("bvh", "unitree_g1"), ("smplx", "unitree_g1_with_hands"), ("smplx", "galaxea_r1pro"),
         ("bvh", "engineai_pm01"), ("smplx", "hightorque_hi"), ("smplx", "kuavo_s45"), ("smplx", "booster_k1"),
         ("bvh", "booster_t1"), ("smplx", "stanford_toddy"), ("bvh", "fourier_n1")]
INFO, UP, CHILD, DN, MERGE, TASK_DOF, DOF_PARENT, TASKMASK = 0, 2, 3, 5, 6, 7, 8, 9


class HostModel:
    def __init__(self, desc):
        from gmr_b200 import _lib
        self.L = _lib.lib()
        self.h = C.c_void_p()
        _lib.check(self.L.gmr_model_create(desc.ref, -1, C.byref(self.h)), "gmr_model_create (host only)")

    def table(self, which, stage):
        n = self.L.gmr_model_table(self.h, which, stage, None, 0)
        assert n >= 0, self.L.gmr_last_error()
        a = np.zeros(max(n, 1), np.int32)
        self.L.gmr_model_table(self.h, which, stage, a.ctypes.data_as(C.c_void_p), n)
        return a[:n]

    def close(self):
        self.L.gmr_model_destroy(self.h)


def _sym_pack(K):
    return np.array([K[i, j] for i in range(6) for j in range(i + 1)])


def _sym_unpack(v):
    K = np.zeros((6, 6))
    k = 0
    for i in range(6):
        for j in range(i + 1):
            K[i, j] = K[j, i] = v[k]
            k += 1
    return K


def _walk(hm, stage, S, K, g, d, locked, bound):
    """qp_solve_aba's sweeps for a fixed set of locked dofs -> (x, multipliers, dofs swept).
    Mirrors the kernel: four 8-lane groups, each with a REGISTER copy of its chain's accumulator (rows, p, U of the
    last dof) in the upward sweep and of the link velocity in the downward sweep; shared memory (XA, Ud, v) is only
    touched where the kernel touches it, and starts as NaN so that anything read before it is written shows up."""
    nv, m, nup, ndn, nmerge, rootX, rootdof, kkt, doubles, KS, XS, SS = hm.table(INFO, stage)
    up = hm.table(UP, stage).reshape(-1, 8)
    ch = hm.table(CHILD, stage).reshape(-1, 2)
    dn = hm.table(DN, stage).reshape(-1, 4)
    merge = hm.table(MERGE, stage).reshape(-1, 2)
    task_dof = hm.table(TASK_DOF, stage)
    tmask = hm.table(TASKMASK, stage).astype(np.uint32)
    assert len(up) == 4 * nup and len(dn) == 4 * ndn and len(merge) == nmerge and (KS, XS, SS) == (27, 42, 6)
    XA = np.full(doubles, np.nan)
    for t in range(m):
        XA[KS * t:KS * t + 21] = _sym_pack(K[t])
        XA[KS * t + 21:KS * t + 27] = g[t]
    for dst, src in merge:
        XA[KS * dst:KS * dst + KS] += XA[KS * src:KS * src + KS]
    Sarr = np.zeros(SS * nv)
    for k in range(nv):
        Sarr[SS * k:SS * k + 6] = S[k]
    Ud = np.full(8 * nv, np.nan)
    v = np.full(6 * nv, np.nan)
    x = np.where(locked, bound, 0.0)
    parent = hm.table(DOF_PARENT, stage)
    seen = set()
    reg = [dict(rows=np.full((6, 6), np.nan), p=np.full(6, np.nan), dof=-1) for _ in range(4)]
    stored, raw = set(), set()
    child_of = {}
    for ps in range(nup):
        writes = []                                       # shared-memory stores become visible after the pass (__syncwarp)
        for grp in range(4):
            c, flags, So, Uo, own, cb, ce, Xo = up[ps * 4 + grp]
            if c < 0:
                continue
            assert c not in seen and So == c * SS * 8 and Uo == c * 64 and Xo % 16 == 0
            seen.add(int(c))
            r = reg[grp]
            if flags & 1:                                 # the registers hold X^a of the chain's previous dof
                assert r["dof"] >= 0 and parent[r["dof"]] == c, "carried accumulator is not a child of this dof"
                rows, p = r["rows"].copy(), r["p"].copy()
            else:
                rows, p = np.zeros((6, 6)), np.zeros(6)
            if own >= 0:
                rows = rows + _sym_unpack(XA[own // 8:own // 8 + 21])
                p = p + XA[own // 8 + 21:own // 8 + 27]
            for Xc, udc in ch[cb:ce]:
                assert Xc in stored and parent[udc // 64] == c
                rows = rows + XA[Xc // 8:Xc // 8 + 36].reshape(6, 6)
                p = p + XA[Xc // 8 + 36:Xc // 8 + 42]
            if flags & 4:
                writes.append((XA, Xo // 8, np.concatenate([rows.ravel(), p])))
                raw.add(int(Xo))
            Sc = Sarr[So // 8:So // 8 + 6]
            U = rows @ Sc
            if locked[c]:
                al, be = 0.0, -x[c]
            else:
                D = Sc @ U + d
                al, be = 1.0 / D, (Sc @ p) / D
            writes.append((Ud, Uo // 8, np.concatenate([U, [al, be]])))
            rows, p = rows - np.outer(U, U) * al, p - U * be
            if flags & 2:
                writes.append((XA, Xo // 8, np.concatenate([rows.ravel(), p])))
                stored.add(int(Xo))
            reg[grp] = dict(rows=rows, p=p, dof=int(c))
        for tgt, off, val in writes:
            tgt[off:off + val.size] = val
    if rootX >= 0:
        assert rootX in raw
        I = XA[rootX // 8:rootX // 8 + 36].reshape(6, 6)
        p = XA[rootX // 8 + 36:rootX // 8 + 42]
        Sb = np.stack([S[rootdof + b] for b in range(6)], 1)
        xb = np.linalg.solve(Sb.T @ I @ Sb + d * np.eye(6), -Sb.T @ p)
        x[rootdof:rootdof + 6] = xb
        v[6 * (rootdof + 5):6 * (rootdof + 5) + 6] = Sb @ xb
    vreg = [dict(v=np.full(6, np.nan), dof=-1) for _ in range(4)]
    down_seen = set()
    for ps in range(ndn):
        writes = []
        for grp in range(4):
            c, flags, vpo, _ = dn[ps * 4 + grp]
            if c < 0:
                continue
            assert c not in down_seen
            down_seen.add(int(c))
            if flags & 1:
                vp = vreg[grp]["v"]
            else:
                vp = v[vpo // 8:vpo // 8 + 6].copy() if vpo >= 0 else np.zeros(6)
            xc = -(Ud[8 * c:8 * c + 6] @ vp) * Ud[8 * c + 6] - Ud[8 * c + 7]
            vc = vp + Sarr[SS * c:SS * c + 6] * xc
            vreg[grp] = dict(v=vc, dof=int(c))
            if flags & 2:
                writes.append((6 * c, vc))
            x[c] = xc
        for off, val in writes:
            v[off:off + 6] = val
    assert np.isfinite(x).all()
    skipped = set(range(rootdof, rootdof + 6)) if rootX >= 0 else set()
    assert down_seen == seen - skipped
    # multipliers the KKT check forms for locked dofs
    f = np.zeros((m, 6))
    for t in range(m):
        a = task_dof[t]
        Kt, gt = _sym_unpack(XA[KS * t:KS * t + 21]), XA[KS * t + 21:KS * t + 27]
        f[t] = gt + (Kt @ v[6 * a:6 * a + 6] if a >= 0 else 0.0)
    assert np.isfinite(f).all()
    lam = np.zeros(nv)
    for i in np.nonzero(locked)[0]:
        F = sum((f[t] for t in range(m) if (int(tmask[i]) & int(kkt)) >> t & 1), np.zeros(6))
        lam[i] = S[i] @ F + d * x[i]
    return x, lam, seen


def _dense(parent, task_dof, S, K, g, d):
    nv = len(parent)
    H, c = d * np.eye(nv), np.zeros(nv)
    for t, a in enumerate(task_dof):
        anc = []
        while a >= 0:
            anc.append(a)
            a = parent[a]
        for k in anc:
            c[k] += S[k] @ g[t]
            for l in anc:
                H[k, l] += S[k] @ K[t] @ S[l]
    return H, c


def _check(desc, rng, stages):
    hm = HostModel(desc)
    try:
        for stage in stages:
            info = hm.table(INFO, stage)
            nv, m, rootX, rootdof = info[0], info[1], info[5], info[6]
            parent = hm.table(DOF_PARENT, stage)
            task_dof = hm.table(TASK_DOF, stage)
            S = rng.normal(size=(nv, 6))
            K = []
            for _ in range(m):
                A = rng.normal(size=(6, 6))
                K.append(A.T @ A)
            g = rng.normal(size=(m, 6))
            d = 0.5 + rng.random()
            H, c = _dense(parent, task_dof, S, K, g, d)
            for trial in range(3):
                locked = np.zeros(nv, bool)
                cand = np.array([i for i in range(nv) if not (rootX >= 0 and rootdof <= i < rootdof + 6)])
                if trial:
                    locked[rng.choice(cand, size=min(5, len(cand)), replace=False)] = True
                bound = rng.normal(size=nv) * 0.1
                x, lam, seen = _walk(hm, stage, S, K, g, d, locked, bound)
                free = ~locked
                ref = np.where(locked, bound, 0.0)
                ref[free] = np.linalg.solve(H[np.ix_(free, free)], -c[free] - H[np.ix_(free, locked)] @ bound[locked])
                np.testing.assert_allclose(x, ref, atol=1e-10)
                np.testing.assert_allclose(lam[locked], (H @ ref + c)[locked], atol=1e-8)
                # the schedule holds exactly the dofs with a task below them (minus the five upper base dofs)
                below = {i for i in range(nv) if H[i, i] > d + 1e-12}
                skipped = set(range(rootdof, rootdof + 5)) if rootX >= 0 else set()
                assert seen == below - skipped
    finally:
        hm.close()


@pytest.mark.parametrize("src,robot", PAIRS)
def test_schedule_solves_the_qp(src, robot):
    from gmr_b200 import model_pack, synth
    from gmr_b200.abi import ModelDesc
    m, t = model_pack.build_tables(src, robot)
    desc = ModelDesc(m, t, synth.human_body_names(src, t))
    _check(desc, np.random.default_rng(PAIRS.index((src, robot))), [k for k in range(2) if t.use_table[k]])


def test_schedule_with_two_tasks_on_one_dof():
    from gmr_b200 import model_pack, synth
    from gmr_b200.abi import ModelDesc
    m, t = model_pack.build_tables("bvh", "unitree_g1")
    t2 = copy.deepcopy(t)
    k = t2.task_frame_names[0].index("left_knee_link")
    t2.task_body[0] = np.array(t2.task_body[0]).copy()
    t2.task_body[0][k] = m.body_names.index("left_toe_link")        # welded to left_ankle_roll_link, which has a task
    desc = ModelDesc(m, t2, synth.human_body_names("bvh", t2))
    hm = HostModel(desc)
    assert hm.table(INFO, 0)[4] == 1 and hm.table(INFO, 1)[4] == 0     # one merge in stage 1 only
    hm.close()
    _check(desc, np.random.default_rng(3), [0])


def test_host_only_model_cannot_back_a_batch():
    from gmr_b200 import _lib, model_pack, synth
    from gmr_b200.abi import ModelDesc
    m, t = model_pack.build_tables("bvh", "unitree_g1")
    hm = HostModel(ModelDesc(m, t, synth.human_body_names("bvh", t)))
    arr = (C.c_void_p * 1)(hm.h)
    out = C.c_void_p()
    assert _lib.lib().gmr_batch_create(arr, 1, None, 4, None, C.byref(out)) != 0
    hm.close()


@pytest.mark.parametrize("src,robot", PAIRS)
def test_fk_sweeps_every_body_the_ik_reads(src, robot):
    """DevModel.fk_body: the world, every body whose joints have a task frame below them, every task frame, closed
    under `parent` (finger bodies of a robot whose tasks end at the wrists stay out)."""
    from gmr_b200 import model_pack, synth
    from gmr_b200.abi import ModelDesc
    m, t = model_pack.build_tables(src, robot)
    hm = HostModel(ModelDesc(m, t, synth.human_body_names(src, t)))
    fk = list(hm.table(10, 0))
    hm.close()
    assert fk == sorted(set(fk)) and 0 not in fk          # the world's pose is a constant, written apart
    # owners of scheduled joints: a joint with some task frame in its subtree
    below = {int(b) for k in range(2) if t.use_table[k] for b in t.task_body[k]}
    for b in sorted(below, reverse=True):
        while b > 0:
            below.add(int(b))
            b = int(m.body_parentid[b])
    need = {b for b in below if m.body_jntnum[b] > 0}
    for k in range(2):
        if t.use_table[k]:
            need |= {int(b) for b in t.task_body[k]}
    for b in sorted(need, reverse=True):
        while b > 0:
            need.add(int(b))
            b = int(m.body_parentid[b])
    assert set(fk) == need - {0}
