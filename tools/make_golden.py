#!/usr/bin/env python
"""Generate tests/golden/*.npz by running the REFERENCE's own Python code (this container only).

What can be run of the reference without its missing third-party packages:

* target preparation - ``GeneralMotionRetargeting.__init__``, ``setup_retarget_configuration``
  and ``update_targets`` (motion_retarget.py:13-136, 203-270) run unmodified; ``mink`` and
  ``mujoco`` are replaced by recording stubs, so what is pinned is the scale/offset/ground
  arithmetic (numpy + scipy Rotation) and which target each FrameTask receives;
* the FK tail - ``KinematicsModel.forward_kinematics`` (kinematics_model.py:213-246) on CPU torch.

The IK solve itself (mink/mujoco/daqp) cannot be run here: parity for it stays unpinned.

    python tools/make_golden.py [--gmr-root /root/reference]
"""
import argparse
import importlib
import os
import sys
import types

import numpy as np

REPO = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..")
sys.path.insert(0, REPO)


def install_reference(gmr_root):
    """Make ``general_motion_retargeting.<module>`` importable without running its __init__."""
    pkg = types.ModuleType("general_motion_retargeting")
    pkg.__path__ = [os.path.join(gmr_root, "general_motion_retargeting")]
    sys.modules["general_motion_retargeting"] = pkg

    mink = types.ModuleType("mink")

    class SO3:
        def __init__(self, wxyz):
            self.wxyz = np.array(wxyz, dtype=np.float64)

    class SE3:
        def __init__(self, rot, pos):
            self.rot, self.pos = rot, np.array(pos, dtype=np.float64)

        @staticmethod
        def from_rotation_and_translation(rot, pos):
            return SE3(rot, pos)

    class FrameTask:
        def __init__(self, frame_name, frame_type, position_cost, orientation_cost, lm_damping=0.0, gain=1.0):
            self.frame_name, self.frame_type = frame_name, frame_type
            self.position_cost, self.orientation_cost, self.lm_damping = position_cost, orientation_cost, lm_damping
            self.target = None

        def set_target(self, T):
            self.target = T

    class Configuration:
        def __init__(self, model):
            self.model = model

    mink.SO3, mink.SE3, mink.FrameTask, mink.Configuration = SO3, SE3, FrameTask, Configuration
    sys.modules["mink"] = mink

    mj = types.ModuleType("mujoco")

    class MjModel:
        @staticmethod
        def from_xml_path(path):
            return MjModel()

    mj.MjModel = MjModel
    sys.modules["mujoco"] = mj
    return importlib.import_module("general_motion_retargeting.motion_retarget")


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gmr-root", default="/root/reference")
    ap.add_argument("--out", default=os.path.join(REPO, "tests", "golden"))
    args = ap.parse_args()

    from gmr_b200 import params, model_pack, synth
    params.set_gmr_root(args.gmr_root)
    ref = install_reference(args.gmr_root)
    os.makedirs(args.out, exist_ok=True)

    # ---- target preparation ------------------------------------------------
    for src, robot in model_pack.constructible_pairs():
        model, tables = model_pack.build_tables(src, robot)
        human, heights, names = synth.synth_human_clips(model, tables, src, B=2, T=6, seed=7, dtype=np.float64)
        rec = {"human": human, "heights": heights, "names": np.array(names)}
        for b in range(2):
            gmr = ref.GeneralMotionRetargeting(src, robot, actual_human_height=float(heights[b]))
            for ground in (False, True):
                scaled_all, tgt1_all, tgt2_all = [], [], []
                for t in range(human.shape[1]):
                    frame = {n: (human[b, t, i, :3].tolist(), human[b, t, i, 3:].tolist()) for i, n in enumerate(names)}
                    gmr.update_targets(frame, offset_to_ground=ground)
                    sh = gmr.scaled_human_data
                    scaled_all.append(np.array([np.concatenate([sh[n][0], sh[n][1]]) for n in tables.scaled_names]))
                    for k, (tasks, store) in enumerate(((gmr.tasks1, tgt1_all), (gmr.tasks2, tgt2_all))):
                        use = (gmr.use_ik_match_table1, gmr.use_ik_match_table2)[k]
                        if use:
                            store.append(np.array([np.concatenate([tk.target.pos, tk.target.rot.wxyz]) for tk in tasks]))
                tag = f"b{b}_g{int(ground)}"
                rec[f"scaled_{tag}"] = np.array(scaled_all)
                rec[f"target1_{tag}"] = np.array(tgt1_all)
                rec[f"target2_{tag}"] = np.array(tgt2_all)
            rec[f"task_frames1_b{b}"] = np.array([tk.frame_name for tk in gmr.tasks1])
            rec[f"task_frames2_b{b}"] = np.array([tk.frame_name for tk in gmr.tasks2])
            rec[f"task_costs1_b{b}"] = np.array([[tk.position_cost, tk.orientation_cost] for tk in gmr.tasks1], float)
            rec[f"task_costs2_b{b}"] = np.array([[tk.position_cost, tk.orientation_cost] for tk in gmr.tasks2], float)
        path = os.path.join(args.out, f"target_prep_{src}_{robot}.npz")
        np.savez_compressed(path, **rec)
        print("wrote", path)

    # ---- FK tail -----------------------------------------------------------
    import torch
    km_mod = importlib.import_module("general_motion_retargeting.kinematics_model")
    rng = np.random.default_rng(11)
    for robot, xml in params.ROBOT_XML_DICT.items():
        try:
            km = km_mod.KinematicsModel(str(xml), device="cpu")
        except Exception as e:     # pm01 / bhl wrappers have no <worldbody>/<body> (reference limitation)
            print(f"KinematicsModel cannot parse {robot}: {type(e).__name__}: {e}")
            continue
        N = 5
        lo, hi = km.get_dof_limits()
        u = torch.from_numpy(rng.uniform(0, 1, size=(N, km.num_dof))).float()
        dof = lo + (hi - lo) * u if len(lo) == km.num_dof else (u - 0.5)
        root_pos = torch.from_numpy(rng.normal(size=(N, 3))).float()
        rq = rng.normal(size=(N, 4))
        rq /= np.linalg.norm(rq, axis=-1, keepdims=True)
        root_rot = torch.from_numpy(rq).float()
        bp, br = km.forward_kinematics(root_pos, root_rot, dof)
        ident_pos = torch.zeros((N, 3))
        ident_rot = torch.zeros((N, 4)); ident_rot[:, -1] = 1.0
        lbp, _ = km.forward_kinematics(ident_pos, ident_rot, dof)
        path = os.path.join(args.out, f"fk_{robot}.npz")
        np.savez_compressed(path, body_names=np.array(km.body_names), root_pos=root_pos.numpy(),
                            root_rot_xyzw=root_rot.numpy(), dof_pos=dof.numpy(), body_pos=bp.numpy(),
                            body_rot_xyzw=br.numpy(), local_body_pos=lbp.numpy(),
                            parent=km.parent_indices.numpy())
        print("wrote", path, tuple(bp.shape))


if __name__ == "__main__":
    main()
