"""Batch form of the reference's dataset drivers (scripts/bvh_to_robot_dataset.py:84-156,
scripts/smplx_to_robot_dataset.py:79-146): retarget every clip of a folder in one launch, then the
drivers' tail - quaternion wxyz -> xyzw, float32 FK of the robot bodies with identity root
(``local_body_pos``), optional height adjust and root-XY zeroing - and one motion dict per clip with the
reference's keys {fps, root_pos, root_rot (xyzw), dof_pos, local_body_pos, link_body_list}.

The reference builds a new retargeter per file (cold start per clip) and loops over frames in Python;
here all clips go through ``gmr_batch_run_host`` at once and the FK tail is one ``gmr_fk_bodies_f32`` launch over
all frames of all clips.
"""
from __future__ import annotations

import os
import pickle

import numpy as np
import torch

from .engine import BatchedRetargeter, DeviceModel
from .shard import work_order


def link_body_list(model) -> list:
    """KinematicsModel.body_names (kinematics_model.py:101-163): bodies below the worldbody, document order."""
    return list(model.body_names[1:])


def retarget_clips(src_human: str, tgt_robot: str, clips, heights=None, human_names=None, fps=30.0,
                   height_adjust: bool = False, root_origin_offset: bool = False, offset_to_ground: bool = False,
                   device: int = 0):
    """clips: list of [T_i, h, 7] arrays (pos xyz + quat wxyz per body, frames in order).
    heights: actual_human_height per clip (or None).  fps: scalar or per clip.
    height_adjust / root_origin_offset: the HEIGHT_ADJUST / ROOT_ORIGIN_OFFSET switches of the reference
    drivers (smplx driver: both True, :118-131; bvh driver: both False, :127-142).
    -> list of motion dicts (numpy), one per clip, in input order."""
    B = len(clips)
    if B == 0:
        return []
    lengths = np.array([c.shape[0] for c in clips], np.int32)
    T = int(lengths.max())
    h = clips[0].shape[1]
    dm = DeviceModel.load(src_human, tgt_robot, human_names, device=device)
    model = dm.desc.model
    if h != dm.desc.c.n_human:
        raise ValueError(f"clips carry {h} bodies per frame, the model expects {dm.desc.c.n_human}")
    if model.jnt_type[0] != 0:
        # the reference drivers slice qpos[:3] / [3:7] / [7:] whatever the robot; for a robot without a floating
        # base (galaxea_r1pro) that slicing is meaningless - refuse rather than write garbage
        raise ValueError(f"{tgt_robot} has no floating base: the dataset format (root_pos/root_rot/dof_pos) does not apply")
    br = BatchedRetargeter(dm, B, device=device)
    br.set_heights(heights if heights is not None else [None] * B)
    br.set_order(work_order(lengths))
    human = torch.zeros((B, T, h, 7), dtype=torch.float32, pin_memory=True)
    hn = human.numpy()
    for i, c in enumerate(clips):
        hn[i, :lengths[i]] = c
        hn[i, lengths[i]:, :, 3] = 1.0                       # padding frames: identity quaternions (never read)
    qpos_host = br.retarget_clips_host(human, T_of=lengths, offset_to_ground=offset_to_ground)
    torch.cuda.synchronize(device)
    nq = model.nq
    qpos = qpos_host.numpy()[:, :, :nq]

    # ---- driver tail on the device, all valid frames of all clips at once
    tdev = torch.device("cuda", device)
    valid = np.arange(T)[None, :] < lengths[:, None]                  # [B,T]
    flat = torch.from_numpy(qpos[valid]).to(tdev)                      # [N,nq] float64
    root_pos = flat[:, :3].clone()
    root_rot_xyzw = flat[:, [4, 5, 6, 3]].contiguous()                 # wxyz -> xyzw (:112 / :103)
    dof_pos = flat[:, 7:].contiguous()
    N = flat.shape[0]
    ident = torch.zeros((N, 4), dtype=torch.float32, device=tdev)
    ident[:, 3] = 1.0
    local_body_pos, _ = dm.fk_bodies_f32(torch.zeros((N, 3), dtype=torch.float32, device=tdev), ident, dof_pos.float(),
                                         want_rot=False)
    clip_of = torch.from_numpy(np.repeat(np.arange(B), lengths)).to(tdev)
    if height_adjust:
        # lowest body point of the whole clip goes to z = 0 (smplx_to_robot_dataset.py:118-126)
        body_pos, _ = dm.fk_bodies_f32(root_pos.float(), root_rot_xyzw.float(), dof_pos.float(), want_rot=False)
        low = torch.full((B,), float("inf"), dtype=torch.float32, device=tdev)
        low = low.scatter_reduce(0, clip_of, body_pos[..., 2].amin(dim=1), reduce="amin")
        root_pos[:, 2] -= low.double()[clip_of]
    if root_origin_offset:
        first = torch.from_numpy(np.concatenate([[0], np.cumsum(lengths)[:-1]])).to(tdev)
        root_pos[:, :2] -= root_pos[first][clip_of][:, :2]               # :128-131
    root_pos, root_rot_xyzw, dof_pos = root_pos.cpu().numpy(), root_rot_xyzw.cpu().numpy(), dof_pos.cpu().numpy()
    local_body_pos = local_body_pos.cpu().numpy()
    names = link_body_list(model)
    fps = np.broadcast_to(np.asarray(fps, np.float64), (B,))
    out, o = [], 0
    for i in range(B):
        n = int(lengths[i])
        out.append({"fps": float(fps[i]) if float(fps[i]) != int(fps[i]) else int(fps[i]),
                    "root_pos": root_pos[o:o + n].copy(), "root_rot": root_rot_xyzw[o:o + n].copy(),
                    "dof_pos": dof_pos[o:o + n].copy(), "local_body_pos": local_body_pos[o:o + n].copy(),
                    "link_body_list": list(names)})
        o += n
    br.close()
    dm.close()
    return out


def save_motion(path: str, motion: dict) -> None:
    """The drivers' pickle (:144-156 / :134-146)."""
    d = os.path.dirname(path)
    if d:
        os.makedirs(d, exist_ok=True)
    with open(path, "wb") as f:
        pickle.dump(motion, f)


def pending_files(src_folder: str, tgt_folder: str, exts, override: bool = False):
    """File walk + skip-if-exists of the drivers (bvh_to_robot_dataset.py:59-72): [(src_path, tgt_path)]."""
    todo = []
    for root, _, files in os.walk(src_folder):
        for name in sorted(files):
            if not name.lower().endswith(tuple(exts)):
                continue
            src = os.path.join(root, name)
            rel = os.path.relpath(src, src_folder)
            tgt = os.path.join(tgt_folder, os.path.splitext(rel)[0] + ".pkl")
            if os.path.exists(tgt) and not override:
                continue
            todo.append((src, tgt))
    return todo


# key robot body <- human body pairs of the reference's error log (scripts/fbx_to_robot.py:1148-1153)
KEY_BODIES = (("pelvis", "Hips"), ("left_wrist_yaw_link", "LeftHand"), ("right_wrist_yaw_link", "RightHand"))
ERRORS_CSV_HEADER = "frame,error1,error2,pelvis_pos_err,left_hand_pos_err,right_hand_pos_err"


def clip_diagnostics(dm: DeviceModel, human_clip, qpos, height=None, offset_to_ground: bool = False, key_bodies=KEY_BODIES):
    """Per-frame diagnostics of one retargeted clip, all frames at once (frames become the instances of a batch):
    what scripts/fbx_to_robot.py:1183-1212 computes after each retarget() call -
    error1() / error2() (motion_retarget.py:188-200: unweighted task-error norms of stage 1 / 2 at the FINAL
    configuration of the frame) and |robot body - scaled human body| for the key bodies (NaN where either is missing).
    human_clip [T,h,7], qpos [T,nq] -> dict of float64 arrays [T]: error1, error2, and one per key body."""
    human_clip = np.asarray(human_clip)
    T = human_clip.shape[0]
    model, tables = dm.desc.model, dm.desc.tables
    br = BatchedRetargeter(dm, T, device=dm.device)
    br.set_heights([height] * T)
    q = np.zeros((T, br.nq_max))
    q[:, :model.nq] = np.asarray(qpos)[:, :model.nq]
    br.reset(q)
    scaled = br.target_prep(human_clip, offset_to_ground=offset_to_ground)          # scaled_human_data of every frame
    xpos, _ = br.fk()
    out = {}
    for stage in (0, 1):
        n = dm.desc.c.n_task[stage]
        if tables.use_table[stage] and n:
            e = br.build_qp(stage, scaled)[4][:, :n]
            out[f"error{stage + 1}"] = torch.linalg.vector_norm(e.reshape(T, -1), dim=1).cpu().numpy()
        else:
            out[f"error{stage + 1}"] = np.zeros(T)
    scaled, xpos = scaled.cpu().numpy(), xpos.cpu().numpy()
    for robot_body, human_body in key_bodies:
        if robot_body in model.body_names and human_body in tables.scaled_names:
            d = xpos[:, model.body_names.index(robot_body)] - scaled[:, list(tables.scaled_names).index(human_body), :3]
            out[robot_body] = np.linalg.norm(d, axis=1)
        else:
            out[robot_body] = np.full(T, np.nan)
    br.close()
    return out


def write_errors_csv(path: str, diag: dict, key_bodies=KEY_BODIES) -> None:
    """The --errors_csv file of scripts/fbx_to_robot.py (:982-990 header on creation, appended rows :1211-1212)."""
    new = not os.path.exists(path)
    with open(path, "a") as f:
        if new:
            f.write(ERRORS_CSV_HEADER + "\n")
        cols = [diag["error1"], diag["error2"]] + [diag[rb] for rb, _ in key_bodies]
        for i in range(len(cols[0])):
            f.write(f"{i}," + ",".join(str(float(c[i])) for c in cols) + "\n")
