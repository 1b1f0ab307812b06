#!/usr/bin/env python
"""Pin the IK core: dump golden vectors of the REAL reference (mink + mujoco + daqp) into tests/golden/ik_*.npz.

Needs a machine where the reference's third-party stack is installed:

    pip install mink mujoco daqp "qpsolvers[daqp]" scipy rich
    python tools/make_golden_ik.py --gmr-root <Zudva/GMR checkout> [--frames 50]

It cannot run in the build container or on the GPU box (no network, packages absent: DESIGN.md section 3); until
its output is committed, tests/test_ik_golden.py skips with "PARITY UNPINNED" and the IK rows of SURVEY.md 8a
(a7-a10) stay "partial".  Nothing is stubbed: ``general_motion_retargeting.motion_retarget`` is imported as is
(only the package __init__, which pulls in the OpenGL viewer, is bypassed) and driven exactly like the dataset
drivers drive it (scripts/bvh_to_robot_dataset.py:84-102): one object per clip, frames in order, ``retarget(frame)``.

Per case (50 seeded synthetic frames of the repo's generator, float64, so that every box can regenerate the
inputs bit for bit) the file holds
  human [T,h,7], names, height                          the frames fed
  qpos [T,nq]                                           retarget()'s return value per frame (motion_retarget.py:185)
  err [T,2], iters [T,2]                                error1()/error2() after the frame, solve_ik calls per stage
  solve_frame/stage [S], solve_q [S,nq], solve_err [S]  configuration and stage error BEFORE each mink.solve_ik call
  solve_dq [S,nv]                                       its result v * dt (motion_retarget.py:147-150)
  qp_index [Q], qp_H, qp_c, qp_lo, qp_hi, qp_G, qp_h    mink.build_ik's QP of the first Q solves (H, c, box)
  task_err_first [m1,6]                                 FrameTask.compute_error of every stage-1 task before solve 0
  versions                                              mink / mujoco / daqp / qpsolvers / numpy / scipy versions
"""
import argparse
import importlib
import json
import os
import sys
import types

import numpy as np

REPO = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..")
sys.path.insert(0, REPO)

CASES = [("bvh", "unitree_g1", 31), ("smplx", "unitree_g1_with_hands", 32), ("smplx", "galaxea_r1pro", 33),
         ("bvh", "engineai_pm01", 34)]
N_QP = 8


def import_reference(gmr_root):
    """general_motion_retargeting.motion_retarget with the real mink/mujoco, without the package __init__."""
    pkg = types.ModuleType("general_motion_retargeting")
    pkg.__path__ = [os.path.join(gmr_root, "general_motion_retargeting")]
    sys.modules["general_motion_retargeting"] = pkg
    return importlib.import_module("general_motion_retargeting.motion_retarget")


def versions():
    out = {}
    for name in ("mink", "mujoco", "daqp", "qpsolvers", "numpy", "scipy"):
        try:
            mod = importlib.import_module(name)
            out[name] = getattr(mod, "__version__", None)
            if out[name] is None:
                from importlib import metadata
                out[name] = metadata.version(name)
        except Exception as e:          # recorded, not hidden
            out[name] = f"unavailable: {e}"
    return out


def box_from_inequalities(G, h, nv):
    """mink ConfigurationLimit rows [P; -P] dq <= [hi; -lo] -> per-dof box (rows with a single entry)."""
    lo, hi = np.full(nv, -np.inf), np.full(nv, np.inf)
    if G is None:
        return lo, hi
    G = np.asarray(G.todense() if hasattr(G, "todense") else G, np.float64)
    for row, rhs in zip(G, np.asarray(h, np.float64)):
        nz = np.nonzero(row)[0]
        if len(nz) != 1:
            raise SystemExit("inequality row is not a box row; the reference is not using ConfigurationLimit only")
        j, g = int(nz[0]), row[nz[0]]
        if g > 0:
            hi[j] = min(hi[j], rhs / g)
        else:
            lo[j] = max(lo[j], rhs / g)
    return lo, hi


def run_case(ref, mink, src, robot, seed, T):
    from gmr_b200 import model_pack, synth
    m, t = model_pack.build_tables(src, robot)
    human, heights, names = synth.synth_human_clips(m, t, src, B=1, T=T, seed=seed, dtype=np.float64)
    human, height = human[0], float(heights[0])
    gmr = ref.GeneralMotionRetargeting(src_human=src, tgt_robot=robot, actual_human_height=height)
    nq, nv = gmr.model.nq, gmr.model.nv
    assert (nq, nv) == (m.nq, m.nv), "flattened model and MuJoCo disagree on nq/nv"
    rec = {k: [] for k in ("solve_frame", "solve_stage", "solve_q", "solve_err", "solve_dq", "qp_index", "qp_H", "qp_c",
                           "qp_lo", "qp_hi")}
    qp_raw = []
    state = {"frame": 0, "task_err_first": None}
    real_solve = mink.solve_ik

    def recording_solve_ik(configuration, tasks, dt, solver, damping, *a, **kw):
        stage = 0 if tasks is gmr.tasks1 else 1
        s = len(rec["solve_frame"])
        if s == 0:
            state["task_err_first"] = np.array([tk.compute_error(configuration) for tk in tasks])
        rec["solve_frame"].append(state["frame"])
        rec["solve_stage"].append(stage)
        rec["solve_q"].append(configuration.q.copy())
        rec["solve_err"].append(gmr.error1() if stage == 0 else gmr.error2())
        if s < N_QP:
            try:
                prob = mink.build_ik(configuration, tasks, dt, damping)
            except TypeError:
                prob = mink.build_ik(configuration, tasks, dt, damping, None)
            P = np.asarray(prob.P.todense() if hasattr(prob.P, "todense") else prob.P, np.float64)
            lo, hi = box_from_inequalities(prob.G, prob.h, nv)
            rec["qp_index"].append(s)
            rec["qp_H"].append(P); rec["qp_c"].append(np.asarray(prob.q, np.float64)); rec["qp_lo"].append(lo); rec["qp_hi"].append(hi)
            qp_raw.append((None if prob.G is None else np.asarray(prob.G.todense() if hasattr(prob.G, "todense") else prob.G),
                           None if prob.h is None else np.asarray(prob.h)))
        v = real_solve(configuration, tasks, dt, solver, damping, *a, **kw)
        rec["solve_dq"].append(np.asarray(v, np.float64) * dt)
        return v

    mink.solve_ik = recording_solve_ik
    try:
        qpos, err, iters = [], [], []
        for i in range(T):
            state["frame"] = i
            before = len(rec["solve_frame"])
            frame = {n: (human[i, k, :3].tolist(), human[i, k, 3:].tolist()) for k, n in enumerate(names)}
            qpos.append(gmr.retarget(frame))
            stages = np.array(rec["solve_stage"][before:])
            iters.append([int((stages == 0).sum()), int((stages == 1).sum())])
            err.append([gmr.error1() if gmr.use_ik_match_table1 else 0.0, gmr.error2() if gmr.use_ik_match_table2 else 0.0])
    finally:
        mink.solve_ik = real_solve
    out = {"source": np.array(src), "robot": np.array(robot), "names": np.array(names), "height": np.array(height),
           "human": human, "qpos": np.array(qpos), "err": np.array(err), "iters": np.array(iters, np.int32),
           "task_err_first": state["task_err_first"], "versions": np.array(json.dumps(versions())),
           "solver": np.array(gmr.solver), "damping": np.array(gmr.damping), "timestep": np.array(gmr.model.opt.timestep)}
    for k, v in rec.items():
        out[k] = np.array(v)
    nrow = max((g.shape[0] for g, _ in qp_raw if g is not None), default=0)
    out["qp_G"] = np.array([np.zeros((nrow, nv)) if g is None else np.vstack([g, np.zeros((nrow - g.shape[0], nv))]) for g, _ in qp_raw])
    out["qp_h"] = np.array([np.full(nrow, np.inf) if h is None else np.concatenate([h, np.full(nrow - len(h), np.inf)]) for _, h in qp_raw])
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gmr-root", default=os.environ.get("GMR_ROOT", "/root/reference"))
    ap.add_argument("--out", default=os.path.join(REPO, "tests", "golden"))
    ap.add_argument("--frames", type=int, default=50)
    args = ap.parse_args()
    try:
        import mink
        import mujoco  # noqa: F401
        import daqp  # noqa: F401
        import qpsolvers  # noqa: F401
    except ImportError as e:
        raise SystemExit(f"make_golden_ik.py needs the reference's real stack (pip install mink mujoco daqp qpsolvers): {e}\n"
                         "IK-core parity stays UNPINNED until this script has run somewhere and its npz files are committed.")
    from gmr_b200 import params
    params.set_gmr_root(args.gmr_root)
    ref = import_reference(args.gmr_root)
    os.makedirs(args.out, exist_ok=True)
    for src, robot, seed in CASES:
        out = run_case(ref, mink, src, robot, seed, args.frames)
        path = os.path.join(args.out, f"ik_{src}_{robot}.npz")
        np.savez_compressed(path, **out)
        print("wrote", path, "solves", len(out["solve_frame"]), "versions", str(out["versions"]))


if __name__ == "__main__":
    main()
