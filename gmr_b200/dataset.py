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


def driver_tail(dm: DeviceModel, qpos, lengths, height_adjust: bool = False, root_origin_offset: bool = False):
    """The reference drivers' post-processing of the stacked retarget() results, for all clips at once
    (scripts/bvh_to_robot_dataset.py:110-142, scripts/smplx_to_robot_dataset.py:97-131).
    qpos [N,nq] float64 (clips back to back, clip i = ``lengths[i]`` rows) ->
    dict(root_pos [N,3] f64, root_rot [N,4] f64 xyzw, dof_pos [N,nq-7] f64, local_body_pos [N,nb,3] f32).
    Pinned by tests/golden/tail_*.npz (the reference's own lines run on CPU, tools/make_golden_tail.py)."""
    lengths = np.asarray(lengths, np.int64)
    B = len(lengths)
    tdev = torch.device("cuda", dm.device)
    flat = torch.as_tensor(np.ascontiguousarray(qpos, np.float64)).to(tdev)
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
    return {"root_pos": root_pos.cpu().numpy(), "root_rot": root_rot_xyzw.cpu().numpy(), "dof_pos": dof_pos.cpu().numpy(),
            "local_body_pos": local_body_pos.cpu().numpy()}


def retarget_clips(src_human: str, tgt_robot: str, clips, heights=None, human_names=None, fps=30.0,
                   height_adjust: bool = False, root_origin_offset: bool = False, offset_to_ground: bool = False,
                   device: int = 0, stream_dtype=np.float64, return_status: bool = False, dof_vmax=None):
    """clips: list of [T_i, h, 7] arrays (pos xyz + quat wxyz per body, frames in order).
    heights: actual_human_height per clip (or None).  fps: scalar or per clip.
    height_adjust / root_origin_offset: the HEIGHT_ADJUST / ROOT_ORIGIN_OFFSET switches of the reference
    drivers (smplx driver: both True, :118-131; bvh driver: both False, :127-142).
    stream_dtype: float64 feeds the solver the frames the reference's retarget() sees (its loaders produce float64);
    float32 halves the bytes that cross the bus at the price of ~1e-7 m in the targets.
    -> list of motion dicts (numpy), one per clip, in input order; a clip the engine flagged as failed (non-finite
    input or output, QP search cap: abi.STATUS_FAILED) yields None - the drivers skip and report it like the
    reference's per-file try/except (scripts/bvh_to_robot_dataset.py:75-80).  With return_status also the per-clip
    OR of the frames' GMR_STATUS_* bits.
    The clips go to the device as one CSR stream (back to back, nothing padded to the longest clip)."""
    from .abi import STATUS_FAILED
    B = len(clips)
    if B == 0:
        return ([], np.zeros(0, np.int32)) if return_status else []
    lengths = np.array([c.shape[0] for c in clips], np.int64)
    h = clips[0].shape[1]
    from . import model_pack
    from .abi import ModelDesc
    from .ik_config import build_retarget_tables
    from .synth import human_body_names
    model, cfg = model_pack.load_pair(src_human, tgt_robot)
    tables = build_retarget_tables(model, cfg, src_human, tgt_robot)
    if human_names is None:
        human_names = human_body_names(src_human, tables)
    dm = DeviceModel(ModelDesc(model, tables, human_names, dof_vmax=dof_vmax), device)
    if h != dm.desc.c.n_human:
        raise ValueError(f"clips carry {h} bodies per frame, the model expects {dm.desc.c.n_human}")
    if model.jnt_type[0] != 0:
        # the reference drivers slice qpos[:3] / [3:7] / [7:] whatever the robot; for a robot without a floating
        # base (galaxea_r1pro) that slicing is meaningless - refuse rather than write garbage
        raise ValueError(f"{tgt_robot} has no floating base: the dataset format (root_pos/root_rot/dof_pos) does not apply")
    br = BatchedRetargeter(dm, B, device=device)
    br.set_heights(heights if heights is not None else [None] * B)
    br.set_order(work_order(lengths))
    clip_start = np.concatenate([[0], np.cumsum(lengths)]).astype(np.int64)
    N = int(clip_start[-1])
    human = torch.empty((N, h, 7), dtype=torch.float64 if np.dtype(stream_dtype) == np.float64 else torch.float32, pin_memory=True)
    hn = human.numpy()
    for i, c in enumerate(clips):
        hn[clip_start[i]:clip_start[i + 1]] = c
    qpos_host, status_host = br.retarget_clips_host_csr(human, clip_start, offset_to_ground=offset_to_ground)
    torch.cuda.synchronize(device)
    nq = model.nq
    qpos = qpos_host.numpy()[:, :nq]
    status = np.bitwise_or.reduceat(status_host.numpy(), clip_start[:-1][lengths > 0]) if (lengths > 0).any() else np.zeros(0, np.int32)
    clip_status = np.zeros(B, np.int32)
    clip_status[lengths > 0] = status
    tail = driver_tail(dm, qpos, lengths, height_adjust, root_origin_offset)
    names = link_body_list(model)
    fps = np.broadcast_to(np.asarray(fps, np.float64), (B,))
    out = []
    for i in range(B):
        a, b = int(clip_start[i]), int(clip_start[i + 1])
        if clip_status[i] & STATUS_FAILED:
            out.append(None)
            continue
        out.append({"fps": float(fps[i]) if float(fps[i]) != int(fps[i]) else int(fps[i]),
                    "root_pos": tail["root_pos"][a:b].copy(), "root_rot": tail["root_rot"][a:b].copy(),
                    "dof_pos": tail["dof_pos"][a:b].copy(), "local_body_pos": tail["local_body_pos"][a:b].copy(),
                    "link_body_list": list(names)})
    br.close()
    dm.close()
    return (out, clip_status) if return_status else out


def describe_status(bits: int) -> str:
    from . import abi
    names = [(abi.STATUS_MAX_ITER1, "stage 1 ended on max_iter"), (abi.STATUS_MAX_ITER2, "stage 2 ended on max_iter"),
             (abi.STATUS_QP_CAP, "QP active-set search hit its cap"), (abi.STATUS_BAD_INPUT, "non-finite input frame"),
             (abi.STATUS_BAD_OUTPUT, "non-finite qpos")]
    return ", ".join(n for b, n in names if bits & b) or "ok"


def save_motion_pt(path: str, motion: dict) -> None:
    """The .pt form of a motion file (scripts/convert_motion_pkl_to_pt.py:39-56,89-91): the same dict with every
    ndarray as a torch tensor, root_rot kept in xyzw, everything else verbatim, written with torch.save."""
    d = os.path.dirname(path)
    if d:
        os.makedirs(d, exist_ok=True)
    torch.save({k: (torch.from_numpy(v) if isinstance(v, np.ndarray) else v) for k, v in motion.items()}, path)


def load_motion(path: str) -> dict:
    """.pkl or .pt motion file -> dict of numpy arrays (the converter's load_any + to_pickle_dict, :26-56)."""
    if path.endswith(".pt"):
        obj = torch.load(path, map_location="cpu", weights_only=False)
        if not isinstance(obj, dict):
            raise TypeError("Expected dict in .pt file")
        return {k: (v.detach().cpu().numpy() if isinstance(v, torch.Tensor) else v) for k, v in obj.items()}
    with open(path, "rb") as f:
        return pickle.load(f)


def rank_info():
    """(rank, world_size, local device) of this process when launched under torchrun, else (0, 1, None)."""
    return (int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1")),
            int(os.environ["LOCAL_RANK"]) if "LOCAL_RANK" in os.environ else None)


def shard_files(todo, rank: int, world: int):
    """This rank's share of the (src, tgt) file list: the reference hands files to a process pool
    (scripts/smplx_to_robot_dataset.py:241-242); here a rank is a GPU and the list is cut by
    gmr_b200.shard.shard_clips on the files' sizes (bytes stand in for frames before anything is parsed), every
    rank computing the same partition.  No communication: each rank loads, retargets and writes its own files."""
    if world <= 1:
        return list(todo)
    from .shard import shard_clips
    sizes = [os.path.getsize(s) for s, _ in todo]
    mine = shard_clips(sizes, world)[rank]
    return [todo[i] for i in sorted(mine.tolist())]


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
