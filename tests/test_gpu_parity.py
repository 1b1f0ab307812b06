"""GPU parity: every stage and the fused clip kernel, through the C ABI, against the oracle."""
import glob
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

PAIRS = [("bvh", "unitree_g1"), ("smplx", "unitree_g1_with_hands"), ("smplx", "galaxea_r1pro"),
         ("bvh", "engineai_pm01"), ("smplx", "hightorque_hi"), ("smplx", "kuavo_s45"), ("smplx", "booster_k1"),
         ("bvh", "booster_t1"), ("smplx", "stanford_toddy"), ("bvh", "fourier_n1")]

QPOS_TOL = 1e-3      # rad, north_star
ROOT_TOL = 1e-4      # m, north_star


def _setup(src, robot, B, T, seed=3):
    import torch  # noqa: F401
    from gmr_b200 import model_pack, synth
    from gmr_b200.engine import BatchedRetargeter, DeviceModel
    m, t = model_pack.build_tables(src, robot)
    human, heights, names = synth.synth_human_clips(m, t, src, B=B, T=T, seed=seed)
    dm = DeviceModel.load(src, robot, names)
    br = BatchedRetargeter(dm, B)
    br.set_heights(heights)
    return m, t, dm, br, human, heights


def _random_q(m, rng, B):
    q = np.tile(m.qpos0, (B, 1))
    for j in range(m.njnt):
        qa = m.jnt_qposadr[j]
        if m.jnt_type[j] == 0:
            q[:, qa:qa + 3] += rng.normal(size=(B, 3)) * 0.3
            w = rng.normal(size=(B, 4))
            q[:, qa + 3:qa + 7] = w / np.linalg.norm(w, axis=-1, keepdims=True)
        elif m.jnt_limited[j]:
            q[:, qa] = rng.uniform(m.jnt_range[j, 0], m.jnt_range[j, 1], size=B)
        else:
            q[:, qa] = rng.uniform(-1, 1, size=B)
    return q


@pytest.mark.parametrize("src,robot", PAIRS)
def test_kernel1_target_prep(src, robot):
    from gmr_b200 import synth
    from oracle import oracle as oc
    m, t, dm, br, human, heights = _setup(src, robot, 8, 3)
    sc = synth.clip_scales(t, heights)
    for ground in (False, True):
        out = br.target_prep(human[:, 1], ground).cpu().numpy()
        for b in range(8):
            ref = oc.prepare(dm.desc, human[b, 1].astype(np.float64), sc[b], ground)
            np.testing.assert_allclose(out[b, :t.n_scaled], ref, atol=1e-13)


def test_kernel1_matches_reference_golden(golden_dir):
    """Against vectors produced by the reference's own update_targets (tools/make_golden.py)."""
    from gmr_b200 import model_pack, synth
    from gmr_b200.engine import BatchedRetargeter, DeviceModel
    for f in sorted(glob.glob(os.path.join(golden_dir, "target_prep_*.npz"))):
        src, robot = os.path.basename(f)[len("target_prep_"):-4].split("_", 1)
        g = np.load(f)
        names = [str(n) for n in g["names"]]
        dm = DeviceModel.load(src, robot, names)
        t = dm.desc.tables
        human = g["human"]                      # [2, T, h, 7] float64
        B, T = human.shape[:2]
        br = BatchedRetargeter(dm, B)
        br.set_heights(g["heights"])
        for ground in (0, 1):
            for ti in range(T):
                out = br.target_prep(human[:, ti], bool(ground)).cpu().numpy()
                for b in range(B):
                    np.testing.assert_allclose(out[b, :t.n_scaled], g[f"scaled_b{b}_g{ground}"][ti], atol=1e-13)


@pytest.mark.parametrize("src,robot", PAIRS)
def test_kernel2_fk(src, robot):
    from oracle import oracle as oc
    m, t, dm, br, human, heights = _setup(src, robot, 16, 1)
    q = _random_q(m, np.random.default_rng(1), 16)
    xpos, xquat = br.fk(q)
    xpos, xquat = xpos.cpu().numpy(), xquat.cpu().numpy()
    for b in range(16):
        rp, rq = oc.fk(dm.desc, q[b])
        np.testing.assert_allclose(xpos[b, :m.nbody], rp, atol=1e-13)
        np.testing.assert_allclose(xquat[b, :m.nbody], rq, atol=1e-13)


@pytest.mark.parametrize("src,robot", PAIRS)
def test_kernel23_build_qp(src, robot):
    from gmr_b200 import synth
    from oracle import oracle as oc
    B = 16
    m, t, dm, br, human, heights = _setup(src, robot, B, 2)
    q = _random_q(m, np.random.default_rng(2), B)
    # moderate errors: start from qpos0-ish joint values too
    q[B // 2:, 7 if m.jnt_type[0] == 0 else 0:] *= 0.3
    br.reset(q)
    scaled = br.target_prep(human[:, 0]).cpu().numpy()
    for stage in range(2):
        if not t.use_table[stage]:
            continue
        H, c, lo, hi, e = [x.cpu().numpy() for x in br.build_qp(stage, scaled)]
        for b in range(B):
            Hr, cr, lor, hir, er = oc.build_qp(dm.desc, q[b], stage, scaled[b, :t.n_scaled])
            scale = np.abs(Hr).max()
            np.testing.assert_allclose(H[b, :m.nv, :m.nv], Hr, atol=1e-11 * scale)
            np.testing.assert_allclose(c[b, :m.nv], cr, atol=1e-11 * max(1.0, np.abs(cr).max()))
            np.testing.assert_allclose(e[b, :len(er)], er, atol=1e-12)
            np.testing.assert_array_equal(np.isinf(lo[b, :m.nv]), np.isinf(lor))
            fin = np.isfinite(lor)
            np.testing.assert_allclose(lo[b, :m.nv][fin], lor[fin], atol=1e-14)
            np.testing.assert_allclose(hi[b, :m.nv][fin], hir[fin], atol=1e-14)


@pytest.mark.parametrize("src,robot", PAIRS)
def test_kernel3_tree_qp_solve_ik(src, robot):
    """The tree-structured (L'L, no fill-in) box QP of the fused path against the oracle's dense
    active-set solve of the oracle's own H, c, box - from states pushed against the joint limits."""
    from oracle import oracle as oc
    B = 32
    m, t, dm, br, human, heights = _setup(src, robot, B, 2)
    rng = np.random.default_rng(4)
    q = _random_q(m, rng, B)
    for j in range(m.njnt):          # park a third of the limited joints on / next to a limit
        if m.jnt_type[j] != 0 and m.jnt_limited[j]:
            qa = m.jnt_qposadr[j]
            pick = rng.random(B) < 0.35
            side = rng.random(B) < 0.5
            eps = rng.uniform(0, 0.02, size=B)
            q[pick, qa] = np.where(side, m.jnt_range[j, 0] + eps, m.jnt_range[j, 1] - eps)[pick]
    br.reset(q)
    scaled = br.target_prep(human[:, 0]).cpu().numpy()
    n_active = 0
    for stage in range(2):
        if not t.use_table[stage]:
            continue
        dq, nf = br.solve_ik(stage, scaled)
        dq, nf = dq.cpu().numpy(), nf.cpu().numpy()
        assert (nf > 0).all()
        for b in range(B):
            Hr, cr, lor, hir, _ = oc.build_qp(dm.desc, q[b], stage, scaled[b, :t.n_scaled])
            xr, _ = oc.box_qp(Hr, cr, lor, hir)
            np.testing.assert_allclose(dq[b, :m.nv], xr, atol=1e-9)
            n_active += int(((xr <= lor) | (xr >= hir)).sum())
    assert n_active > 0, "test did not exercise the active set"


@pytest.mark.parametrize("n", [5, 27, 35, 49, 64])
def test_kernel3_qp_solve(n):
    from gmr_b200.engine import qp_solve
    from oracle import oracle as oc
    rng = np.random.default_rng(n)
    N = 64
    Hs, cs, los, his = [], [], [], []
    for i in range(N):
        A = rng.normal(size=(n + 10, n))
        Hs.append(A.T @ A + 0.5 * np.eye(n))
        cs.append(rng.normal(size=n) * (5 if i % 2 else 0.5))
        lo = -np.abs(rng.normal(size=n)) * 0.3
        hi = np.abs(rng.normal(size=n)) * 0.3
        free = rng.random(n) < 0.2
        lo[free], hi[free] = -np.inf, np.inf
        los.append(lo)
        his.append(hi)
    x, nf = qp_solve(np.array(Hs), np.array(cs), np.array(los), np.array(his))
    x, nf = x.cpu().numpy(), nf.cpu().numpy()
    assert (nf > 0).all()
    for i in range(N):
        xr, nfr = oc.box_qp(Hs[i], cs[i], los[i], his[i])
        np.testing.assert_allclose(x[i], xr, atol=1e-10)
        assert nf[i] == nfr


def _check_clip(qpos, ref, m, tag=""):
    d = np.abs(qpos - ref)
    free = m.jnt_type[0] == 0
    if free:
        assert d[..., :3].max() < ROOT_TOL, f"{tag} root pos err {d[..., :3].max()}"
        # quaternion sign-insensitive
        qa, qb = qpos[..., 3:7], ref[..., 3:7]
        dq = np.minimum(np.abs(qa - qb).max(-1), np.abs(qa + qb).max(-1))
        assert dq.max() < QPOS_TOL, f"{tag} root quat err {dq.max()}"
        assert d[..., 7:].max() < QPOS_TOL, f"{tag} joint err {d[..., 7:].max()}"
    else:
        assert d.max() < QPOS_TOL, f"{tag} qpos err {d.max()}"
    return d.max()


@pytest.mark.parametrize("src,robot", PAIRS)
def test_fused_clip_matches_oracle(src, robot):
    from gmr_b200 import synth
    from oracle import oracle as oc
    B, T = 48, 40
    m, t, dm, br, human, heights = _setup(src, robot, B, T)
    qpos, diag = br.run(human, diag=True)
    qpos = qpos.cpu().numpy()
    ref = oc.run_clips(dm.desc, human, scale=synth.clip_scales(t, heights), nthreads=8)
    worst = _check_clip(qpos[..., :m.nq], ref["qpos"], m, f"{src}->{robot}")
    # float64 against float64: typical frames agree to 1e-10..1e-13; a few frames next to a change of the active set
    # amplify rounding to ~4e-8 (tools/ab_parity.py) - three to four orders inside north_star's 1e-3 rad / 1e-4 m
    assert worst < 1e-6, f"float64 path should agree far below tolerance, got {worst}"
    np.testing.assert_array_equal(diag["iters"].cpu().numpy(), ref["iters"])
    np.testing.assert_allclose(diag["err"].cpu().numpy(), ref["err"], atol=1e-6)
    assert (diag["nfact"].cpu().numpy() > 0).all()
    # task-space residual no worse than the reference (north_star)
    assert (diag["err"].cpu().numpy() <= ref["err"] + 1e-6).all()


def test_step_equals_run_and_state_persists():
    B, T = 16, 12
    m, t, dm, br, human, heights = _setup("bvh", "unitree_g1", B, T)
    full = br.run(human).cpu().numpy()
    br.reset()
    a = br.run(human[:, :5]).cpu().numpy()
    b = br.run(human[:, 5:]).cpu().numpy()
    np.testing.assert_array_equal(np.concatenate([a, b], 1), full)
    br.reset()
    steps = np.stack([br.step(human[:, i]).cpu().numpy() for i in range(T)], 1)
    np.testing.assert_array_equal(steps, full)
    np.testing.assert_array_equal(br.state().cpu().numpy(), full[:, -1])


def test_host_pipeline_is_bit_identical_to_one_launch():
    """gmr_batch_run_host: zero-copy (0) and chunked H2D / retarget / D2H over three streams (>= 1) == one
    launch on resident data."""
    import torch
    B, T = 40, 37
    m, t, dm, br, human, heights = _setup("bvh", "unitree_g1", B, T)
    full = br.run(human).cpu().numpy()
    T_of = np.random.default_rng(1).integers(1, T + 1, size=B).astype(np.int32)
    for chunks in (None, 0, 1, 3, 37):
        hh = torch.from_numpy(human).pin_memory()
        out = br.retarget_clips_host(hh, n_chunks=chunks)
        torch.cuda.synchronize()
        np.testing.assert_array_equal(out.numpy(), full)
    with pytest.raises(Exception, match="pinned"):
        br.retarget_clips_host(torch.from_numpy(human), n_chunks=0)               # zero-copy refuses pageable memory
    br.reset()
    ragged = br.run(human, T_of=T_of).cpu().numpy()
    out = br.retarget_clips_host(torch.from_numpy(human.astype(np.float32)).pin_memory(), T_of=T_of, n_chunks=0)
    torch.cuda.synchronize()
    br.reset()
    r32 = br.run(human.astype(np.float32), T_of=T_of).cpu().numpy()
    for b in range(B):
        np.testing.assert_array_equal(out.numpy()[b, :T_of[b]], r32[b, :T_of[b]])
    out = br.retarget_clips_host(torch.from_numpy(human), T_of=T_of, n_chunks=4)     # pageable host memory works too
    torch.cuda.synchronize()
    for b in range(B):
        np.testing.assert_array_equal(out.numpy()[b, :T_of[b]], ragged[b, :T_of[b]])


def test_ragged_lengths_and_order():
    from gmr_b200 import synth
    from oracle import oracle as oc
    B, T = 24, 20
    m, t, dm, br, human, heights = _setup("smplx", "unitree_g1", B, T)
    T_of = np.random.default_rng(0).integers(1, T + 1, size=B).astype(np.int32)
    br.set_order(np.argsort(-T_of).astype(np.int32))
    qpos = br.run(human, T_of=T_of).cpu().numpy()
    ref = oc.run_clips(dm.desc, human, T_of=T_of, scale=synth.clip_scales(t, heights), nthreads=8)
    for b in range(B):
        _check_clip(qpos[b, :T_of[b], :m.nq], ref["qpos"][b, :T_of[b]], m)
        assert (qpos[b, T_of[b]:] == 0).all()


def test_offset_to_ground_and_f64_stream():
    from gmr_b200 import synth
    from oracle import oracle as oc
    B, T = 8, 10
    m, t, dm, br, human, heights = _setup("smplx", "booster_t1", B, T)
    h64 = human.astype(np.float64) + 1e-9
    qpos = br.run(h64, offset_to_ground=True).cpu().numpy()
    ref = oc.run_clips(dm.desc, h64, scale=synth.clip_scales(t, heights), offset_to_ground=True, nthreads=8)
    assert _check_clip(qpos[..., :m.nq], ref["qpos"], m) < 1e-6


def test_heterogeneous_launch():
    """config 4: one launch, instance = (clip, robot)."""
    from gmr_b200 import model_pack, synth
    from gmr_b200.engine import BatchedRetargeter, DeviceModel
    from oracle import oracle as oc
    robots = ["unitree_g1", "unitree_g1_with_hands", "booster_t1", "stanford_toddy", "fourier_n1", "engineai_pm01",
              "kuavo_s45", "hightorque_hi", "galaxea_r1pro", "booster_k1"]          # every robot constructible for smplx
    per, T = 5, 12
    models, humans, refs, model_of, heights_all = [], [], [], [], []
    for i, r in enumerate(robots):
        m, t = model_pack.build_tables("smplx", r)
        human, heights, names = synth.synth_human_clips(m, t, "smplx", B=per, T=T, seed=10 + i)
        dm = DeviceModel.load("smplx", r, names)
        models.append(dm)
        humans.append(human)
        heights_all += list(heights)
        model_of += [i] * per
        refs.append(oc.run_clips(dm.desc, human, scale=synth.clip_scales(t, heights), nthreads=8)["qpos"])
    br = BatchedRetargeter(models, per * len(robots), model_of_instance=model_of)
    br.set_heights(heights_all)
    qpos = br.run(np.concatenate(humans, 0)).cpu().numpy()
    for i, r in enumerate(robots):
        m = models[i].desc.model
        _check_clip(qpos[i * per:(i + 1) * per, :, :m.nq], refs[i], m, r)


def test_fk_tail_matches_reference_golden(golden_dir):
    import torch
    from gmr_b200 import model_pack
    from gmr_b200.engine import DeviceModel
    for f in sorted(glob.glob(os.path.join(golden_dir, "fk_*.npz"))):
        robot = os.path.basename(f)[3:-4]
        g = np.load(f)
        dm = DeviceModel.load("smplx", robot)
        dev = torch.device("cuda", 0)
        bp, br = dm.fk_bodies_f32(torch.from_numpy(g["root_pos"]).to(dev), torch.from_numpy(g["root_rot_xyzw"]).to(dev),
                                  torch.from_numpy(g["dof_pos"]).to(dev))
        np.testing.assert_allclose(bp.cpu().numpy(), g["body_pos"], atol=2e-6)
        np.testing.assert_allclose(br.cpu().numpy(), g["body_rot_xyzw"], atol=2e-6)
        n = g["dof_pos"].shape[0]
        ident = torch.zeros((n, 4), device=dev)
        ident[:, 3] = 1
        lbp, _ = dm.fk_bodies_f32(torch.zeros((n, 3), device=dev), ident, torch.from_numpy(g["dof_pos"]).to(dev), want_rot=False)
        np.testing.assert_allclose(lbp.cpu().numpy(), g["local_body_pos"], atol=2e-6)


def test_full_size_bench_workload():
    """BASELINE.json configs[1] at full size (4096 clips x 300 frames, unitree_g1 <- bvh): a random sample of
    clips against the oracle, and the size-independent properties of the whole result - two launches agree bit
    for bit whatever order the work queue served the frame ranges in, limited joints stay inside their range
    (ConfigurationLimit with gain 0.95 never leaves it), the floating-base quaternion is unit, every frame
    closed both stages."""
    from gmr_b200 import synth
    from oracle import oracle as oc
    B, T = 4096, 300
    m, t, dm, br, human, heights = _setup("bvh", "unitree_g1", B, T, seed=0xB201)
    qpos, diag = br.run(human, diag=True)
    qpos = qpos.cpu().numpy()
    iters = diag["iters"].cpu().numpy()
    pick = np.sort(np.random.default_rng(7).choice(B, size=16, replace=False))
    ref = oc.run_clips(dm.desc, human[pick], scale=synth.clip_scales(t, heights)[pick], nthreads=8)
    worst = _check_clip(qpos[pick][..., :m.nq], ref["qpos"], m, "4096x300 sample")
    assert worst < 1e-6
    np.testing.assert_array_equal(iters[pick], ref["iters"])
    assert np.isfinite(qpos).all()
    assert (iters >= 1).all() and (iters <= 11).all()
    assert (diag["nfact"].cpu().numpy() >= iters.sum(-1)).all()          # at least one solver sweep per solve_ik
    for j in range(m.njnt):
        qa = m.jnt_qposadr[j]
        if m.jnt_type[j] == 0:
            np.testing.assert_allclose(np.linalg.norm(qpos[..., qa + 3:qa + 7], axis=-1), 1.0, atol=1e-12)
        elif m.jnt_limited[j]:
            assert qpos[..., qa].min() >= m.jnt_range[j, 0] - 1e-12 and qpos[..., qa].max() <= m.jnt_range[j, 1] + 1e-12
    br.reset()
    again = br.run(human).cpu().numpy()
    np.testing.assert_array_equal(again, qpos)


@pytest.mark.parametrize("T", [1, 31, 32, 33, 65])
def test_frame_range_boundaries(T):
    """Clip lengths around the work queue's 32-frame unit: the warm start crosses units through the batch state."""
    from gmr_b200 import synth
    from oracle import oracle as oc
    B = 12
    m, t, dm, br, human, heights = _setup("bvh", "unitree_g1", B, T, seed=21)
    qpos = br.run(human).cpu().numpy()
    ref = oc.run_clips(dm.desc, human, scale=synth.clip_scales(t, heights), nthreads=8)
    assert _check_clip(qpos[..., :m.nq], ref["qpos"], m, f"T={T}") < 1e-6
    np.testing.assert_array_equal(br.state().cpu().numpy()[:, :m.nq], qpos[:, -1, :m.nq])


def test_two_tasks_on_one_dof():
    """Two task frames that hang on the same dof (a jointed body and a body welded to it): the solver sums their
    stiffness blocks before its sweeps.  None of the shipped configs does this, so the table is edited here:
    stage-1 task `left_knee_link` moves to `left_toe_link`, which is welded to `left_ankle_roll_link` (task 3)."""
    import copy
    from gmr_b200 import model_pack, synth
    from gmr_b200.abi import ModelDesc
    from gmr_b200.engine import BatchedRetargeter, DeviceModel
    from oracle import oracle as oc
    m, t = model_pack.build_tables("bvh", "unitree_g1")
    B, T = 16, 24
    human, heights, names = synth.synth_human_clips(m, t, "bvh", B=B, T=T, seed=9)
    t2 = copy.deepcopy(t)
    toe, ankle = m.body_names.index("left_toe_link"), m.body_names.index("left_ankle_roll_link")
    assert m.body_jntnum[toe] == 0 and m.body_parentid[toe] == ankle
    k = t2.task_frame_names[0].index("left_knee_link")
    assert ankle in list(t2.task_body[0])
    t2.task_body[0] = np.array(t2.task_body[0]).copy()
    t2.task_body[0][k] = toe
    t2.task_frame_names[0][k] = "left_toe_link"
    dm = DeviceModel(ModelDesc(m, t2, names))
    br = BatchedRetargeter(dm, B)
    br.set_heights(heights)
    qpos, diag = br.run(human, diag=True)
    ref = oc.run_clips(dm.desc, human, scale=synth.clip_scales(t2, heights), nthreads=8)
    assert _check_clip(qpos.cpu().numpy()[..., :m.nq], ref["qpos"], m, "merged tasks") < 1e-6
    np.testing.assert_array_equal(diag["iters"].cpu().numpy(), ref["iters"])


def test_csr_stream_is_bit_identical_to_the_dense_layout():
    """gmr_batch_run_csr / gmr_batch_run_host_csr (clips back to back, no padding) == gmr_batch_run on [B,T_max,h,7]."""
    import torch
    B, T = 21, 45
    m, t, dm, br, human, heights = _setup("bvh", "unitree_g1", B, T, seed=8)
    T_of = np.random.default_rng(2).integers(0, T + 1, size=B).astype(np.int32)
    T_of[3], T_of[7] = T, 1                                          # longest, shortest, and (maybe) empty clips
    T_of[11] = 0
    br.set_order(np.argsort(-T_of, kind="stable").astype(np.int32))
    dense, ddiag = br.run(human, T_of=T_of, diag=True)
    dense = dense.cpu().numpy()
    cs = np.concatenate([[0], np.cumsum(T_of)]).astype(np.int64)
    flat = np.concatenate([human[b, :T_of[b]] for b in range(B)])
    br.reset()
    got, gdiag = br.run_csr(flat, cs, diag=True)
    got = got.cpu().numpy()
    for b in range(B):
        np.testing.assert_array_equal(got[cs[b]:cs[b + 1]], dense[b, :T_of[b]])
        np.testing.assert_array_equal(gdiag["iters"][cs[b]:cs[b + 1]].cpu().numpy(), ddiag["iters"][b, :T_of[b]].cpu().numpy())
        np.testing.assert_array_equal(gdiag["status"][cs[b]:cs[b + 1]].cpu().numpy(), ddiag["status"][b, :T_of[b]].cpu().numpy())
    for pinned in (True, False):                                     # zero-copy / staged host forms
        hh = torch.from_numpy(flat)
        q, st = br.retarget_clips_host_csr(hh.pin_memory() if pinned else hh,
                                           cs, qpos_host=None if pinned else torch.zeros((len(flat), br.nq_max), dtype=torch.float64),
                                           status_host=None if pinned else torch.zeros((len(flat),), dtype=torch.int32))
        torch.cuda.synchronize()
        np.testing.assert_array_equal(q.numpy(), got)
        np.testing.assert_array_equal(st.numpy(), gdiag["status"].cpu().numpy())


def test_status_flags():
    """gmr_diag.status_dev: max_iter bits == the loop counters, bad input / output bits on a NaN frame only."""
    from gmr_b200 import abi
    B, T = 32, 30
    m, t, dm, br, human, heights = _setup("bvh", "unitree_g1", B, T, seed=13)
    h = human.copy()
    h[5, 12, dm.desc.human_names.index(t.scaled_names[2]), 1] = np.inf      # a body the config reads
    qpos, diag = br.run(h, diag=True)
    st, it = diag["status"].cpu().numpy(), diag["iters"].cpu().numpy()
    ok = np.ones((B, T), bool)
    ok[5, 12:] = False                                               # the poisoned tail of clip 5
    for k, bit in enumerate((abi.STATUS_MAX_ITER1, abi.STATUS_MAX_ITER2)):
        hit = ((st & bit) != 0) & ok
        assert (it[..., k][hit] == 11).all()                         # the counter ran out: 1 + max_iter solves
        # 11 solves without the bit = the 11th solve converged; fewer solves never carry the bit
        assert not (((st & bit) != 0) & ok & (it[..., k] < 11)).any()
    assert ((st & abi.STATUS_MAX_ITER1) != 0)[:, 0].any()            # cold starts run out of iterations
    bad_in = (st & abi.STATUS_BAD_INPUT) != 0
    assert bad_in[5, 12] and bad_in.sum() == 1
    bad_out = (st & abi.STATUS_BAD_OUTPUT) != 0
    assert bad_out[5, 12:].all() and not bad_out[5, :12].any() and bad_out.sum() == T - 12
    assert not (st & abi.STATUS_QP_CAP).any()
    assert np.isfinite(qpos.cpu().numpy()[np.arange(B) != 5]).all()


def test_velocity_limit_box():
    """Optional |dq| <= vmax dt box (gmr_model_desc.dof_vmax; off by default = the reference): CUDA == oracle with the
    same box, every step inside it, and NULL == +inf == the unlimited run bit for bit."""
    import torch
    from gmr_b200 import model_pack, synth
    from gmr_b200.abi import ModelDesc
    from gmr_b200.engine import BatchedRetargeter, DeviceModel
    from oracle.gmr_oracle_np import OracleRetargeter
    B, T = 4, 10
    m, t = model_pack.build_tables("bvh", "unitree_g1")
    human, heights, names = synth.synth_human_clips(m, t, "bvh", B=B, T=T, seed=17, dtype=np.float64)
    idx = [names.index(b) for b in t.scaled_names]
    vmax = 2.0                                                       # rad/s: 0.004 rad per solve at dt = 0.002
    runs = {}
    for key, vm in (("off", None), ("inf", np.full(m.nv, np.inf)), ("on", vmax)):
        dm = DeviceModel(ModelDesc(m, t, names, dof_vmax=vm))
        br = BatchedRetargeter(dm, B)
        br.set_heights(heights)
        runs[key] = br.run(human).cpu().numpy()[..., :m.nq]
        if key == "on":
            scaled = br.target_prep(human[:, 0])
            br.reset()
            dq, _ = br.solve_ik(0, scaled)
            assert np.abs(dq.cpu().numpy()[:, 6:m.nv]).max() <= vmax * m.timestep + 1e-12
            vfull = dm.desc.dof_vmax
        br.close(); dm.close()
    np.testing.assert_array_equal(runs["off"], runs["inf"])
    assert np.abs(runs["on"] - runs["off"]).max() > 1e-3             # the box binds on a cold start
    for b in range(B):
        ref = OracleRetargeter(m, t, actual_human_height=float(heights[b]), dof_vmax=vfull)
        for i in range(T):
            np.testing.assert_allclose(runs["on"][b, i], ref.retarget(human[b, i][idx]), atol=1e-7)
    with pytest.raises(Exception, match="free joint"):
        DeviceModel(ModelDesc(m, t, names, dof_vmax=np.full(m.nv, 1.0)))


def test_two_batches_of_different_footprint_alternate():
    """The dynamic shared-memory opt-in of clip_kernel is per function and process-wide: creating a smaller batch
    (other robot, narrower CTA) must not break the next launch of a larger one."""
    mA, tA, dmA, brA, humanA, _ = _setup("smplx", "unitree_g1_with_hands", 64, 6, seed=1)
    first = brA.run(humanA).cpu().numpy()
    mB, tB, dmB, brB, humanB, _ = _setup("smplx", "galaxea_r1pro", 3, 6, seed=2)      # 3 warps per CTA, small blocks
    small = brB.run(humanB).cpu().numpy()
    brA.reset()
    np.testing.assert_array_equal(brA.run(humanA).cpu().numpy(), first)
    brB.reset()
    np.testing.assert_array_equal(brB.run(humanB).cpu().numpy(), small)


@pytest.mark.parametrize("src,robot", [("bvh", "unitree_g1"), ("smplx", "unitree_g1_with_hands"), ("smplx", "galaxea_r1pro")])
def test_compile_time_layout_kernels_equal_the_generic_one(src, robot, monkeypatch):
    """A single-model batch of a known robot runs the kernel compiled for its layout (csrc/gmr_layouts.cuh); the
    generic kernel (layout passed at run time) must give the same bits - same code, same arithmetic order."""
    B, T = 40, 33
    m, t, dm, br, human, heights = _setup(src, robot, B, T, seed=21)
    assert br.launch_info()["layout_id"] >= 0, "known robot did not get its compile-time layout"
    fixed, dfix = br.run(human, diag=True)
    monkeypatch.setenv("GMR_DEBUG_RUNTIME_LAYOUT", "1")
    from gmr_b200.engine import BatchedRetargeter
    br2 = BatchedRetargeter(dm, B)
    br2.set_heights(heights)
    assert br2.launch_info()["layout_id"] == -1
    generic, dgen = br2.run(human, diag=True)
    np.testing.assert_array_equal(fixed.cpu().numpy(), generic.cpu().numpy())
    np.testing.assert_array_equal(dfix["iters"].cpu().numpy(), dgen["iters"].cpu().numpy())
    np.testing.assert_array_equal(dfix["err"].cpu().numpy(), dgen["err"].cpu().numpy())


def test_set_order_rejects_non_permutations():
    m, t, dm, br, human, heights = _setup("bvh", "unitree_g1", 6, 2)
    with pytest.raises(ValueError):
        br.set_order([0, 1, 1, 3, 4, 5])
    br.set_order([5, 4, 3, 2, 1, 0])


def test_short_math_matches_libm():
    """rcp_pos / rsqrt_pos / atan_ratio of csrc/gmr_device.cuh (hardware seed + two Newton steps; degree-39 odd
    polynomial) against numpy over the ranges the kernels feed them: a few ulp."""
    import ctypes as C
    import torch
    from gmr_b200 import _lib
    rng = np.random.default_rng(0)
    x = np.concatenate([10.0 ** rng.uniform(-12, 12, 20000), rng.uniform(0.5, 2.0, 20000), [1e-10, 1.0, 0.5, 4.0]])
    y = np.concatenate([x[:20000] * 10.0 ** rng.uniform(-6, 6, 20000), rng.uniform(0.5, 2.0, 20000), [1e-10, 1.0, 3.0, 4.0]])
    xd, yd = torch.from_numpy(x).cuda(), torch.from_numpy(y).cuda()
    out = torch.zeros((len(x), 3), dtype=torch.float64, device="cuda")
    _lib.check(_lib.lib().gmr_probe_math(0, len(x), C.c_void_p(xd.data_ptr()), C.c_void_p(yd.data_ptr()), C.c_void_p(out.data_ptr())),
               "gmr_probe_math")
    torch.cuda.synchronize()
    o = out.cpu().numpy()
    assert np.abs(o[:, 0] * x - 1.0).max() < 5e-16
    assert np.abs(o[:, 1] * np.sqrt(x) - 1.0).max() < 5e-16
    ref = np.arctan(np.minimum(x, y) / np.maximum(x, y))
    assert np.abs(o[:, 2] - ref).max() < 4.5e-16


@pytest.mark.parametrize("kind", ["human-fk-bvh", "human-fk-smplx", "canonical"])
def test_other_workloads_match_oracle(kind):
    """The human-authored generator (human skeleton -> human FK; unreachable targets, large residuals)
    and the reference's real clip (out/test_canonical.bvh through the reference's own loader, tiled): CUDA == oracle,
    identical iteration counts."""
    from gmr_b200 import model_pack, synth, synth_human
    from gmr_b200.engine import BatchedRetargeter, DeviceModel
    from oracle import oracle as oc
    src = "smplx" if kind.endswith("smplx") else "bvh"
    m, t = model_pack.build_tables(src, "unitree_g1")
    if kind == "canonical":
        human, heights, names = synth_human.canonical_clips(6)
        human = human[:, :60]
    else:
        human, heights, names = synth_human.synth_human_fk_clips(m, t, src, B=12, T=40, seed=3)
    dm = DeviceModel.load(src, "unitree_g1", names)
    br = BatchedRetargeter(dm, human.shape[0])
    br.set_heights(heights)
    qpos, diag = br.run(human, diag=True)
    ref = oc.run_clips(dm.desc, human, scale=synth.clip_scales(t, heights), nthreads=8, diag=True)
    assert _check_clip(qpos.cpu().numpy()[..., :m.nq], ref["qpos"], m, kind) < 1e-6
    np.testing.assert_array_equal(diag["iters"].cpu().numpy(), ref["iters"])
    assert diag["err"].cpu().numpy()[:, 5:].mean() > 0.5          # residual-heavy, unlike the robot-authored clips
