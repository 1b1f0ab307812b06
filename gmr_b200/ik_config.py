"""IK config JSON -> flat retargeting tables (host side, one-time per (source, robot)).

Restates what ``GeneralMotionRetargeting.__init__`` / ``setup_retarget_configuration``
do with the JSON (reference: general_motion_retargeting/motion_retarget.py:29-114) and
what ``update_targets`` reads from it every frame (:117-136), as index tables:

* ``scaled`` list  = bodies of ``human_scale_table`` (root first, then table order);
  per entry: scale, table-1 position offset minus ground (``:91``), table-1 rotation
  offset (``:92``), foot flag (name contains "Foot"/"foot", ``:261``).
  Table-2 offsets are built by the reference but never applied (``:121``) - kept only
  for the Python attribute mirror.
* per stage ``k`` in {1,2}: active tasks (pos_w != 0 or rot_w != 0, ``:82,100``):
  robot body id, index into ``scaled``, position cost, orientation cost.

Errors mirror the reference: unknown bodies raise ``KeyError``.
"""
from __future__ import annotations

import copy
import json
from dataclasses import dataclass

import numpy as np

from .mjcf import RobotModel


@dataclass
class RetargetTables:
    src_human: str
    tgt_robot: str
    human_root_name: str
    robot_root_name: str
    ground_height: float
    human_height_assumption: float
    use_table: tuple                 # (bool, bool)
    scaled_names: list               # [s] human body names, root first
    scale: np.ndarray                # [s] base scale (ratio NOT folded in)
    pos_off: np.ndarray              # [s,3] table-1 pos offset - ground*z
    rot_off: np.ndarray              # [s,4] table-1 rot offset wxyz, unit (scipy normalises)
    is_foot: np.ndarray              # [s] int32
    task_body: list                  # 2 x [m_k] int32 robot body ids
    task_src: list                   # 2 x [m_k] int32 index into scaled list
    task_wpos: list                  # 2 x [m_k] float64
    task_wrot: list                  # 2 x [m_k] float64
    task_frame_names: list           # 2 x [m_k] robot body names
    ik_config: dict                  # the parsed JSON (unscaled), for attribute mirrors

    @property
    def n_scaled(self) -> int:
        return len(self.scaled_names)


def load_ik_config(path) -> dict:
    with open(path) as f:
        return json.load(f)


def build_retarget_tables(model: RobotModel, ik_config: dict, src_human: str = "",
                          tgt_robot: str = "") -> RetargetTables:
    cfg = copy.deepcopy(ik_config)
    root = cfg["human_root_name"]
    scale_table = cfg["human_scale_table"]
    ground = float(cfg["ground_height"])
    if root not in scale_table:
        raise KeyError(root)        # reference: human_scale_table[human_root_name], :214

    tables = (cfg["ik_match_table1"], cfg["ik_match_table2"])

    # table-1 offsets, keyed by HUMAN body name (reference :90-94)
    off1 = {}
    for frame_name, entry in tables[0].items():
        human_body, pw, rw, p_off, r_off = entry
        if pw != 0 or rw != 0:
            off1[human_body] = (np.asarray(p_off, np.float64), np.asarray(r_off, np.float64))

    # scaled list: the reference's output dict order is root first, then the order in
    # which scale-table bodies appear in the *input frame*.  Order has no numerical
    # effect; we use scale-table order.
    scaled_names = [root] + [b for b in scale_table if b != root]
    s = len(scaled_names)
    scale = np.array([scale_table[b] for b in scaled_names], np.float64)
    pos_off = np.zeros((s, 3))
    rot_off = np.zeros((s, 4))
    is_foot = np.zeros(s, np.int32)
    for i, b in enumerate(scaled_names):
        if b not in off1:
            raise KeyError(b)       # reference: rot_offsets[body_name], :241
        p, r = off1[b]
        pos_off[i] = p - ground * np.array([0.0, 0.0, 1.0])
        n = np.linalg.norm(r)
        if n == 0:
            raise ValueError(f"zero rotation offset for {b}")
        rot_off[i] = r / n          # scipy Rotation.from_quat normalises
        is_foot[i] = int("Foot" in b or "foot" in b)

    task_body, task_src, task_wpos, task_wrot, task_names = [], [], [], [], []
    for k in range(2):
        tb, ts, wp, wr, names, seen = [], [], [], [], [], set()
        for frame_name, entry in tables[k].items():
            human_body, pw, rw, _p, _r = entry
            if pw == 0 and rw == 0:
                continue
            if human_body in seen:
                # reference keeps one task per human body in human_body_to_task (:90);
                # a second frame on the same body would never get a target (TargetNotSet)
                raise ValueError(f"table{k + 1}: human body {human_body!r} mapped twice")
            seen.add(human_body)
            use = bool(cfg[f"use_ik_match_table{k + 1}"])
            if use:
                bid = model.body_id(frame_name)       # KeyError if the robot lacks it
                if human_body not in scaled_names:
                    raise KeyError(human_body)        # reference: human_data[body_name], :129
                src = scaled_names.index(human_body)
            else:
                bid = model.body_names.index(frame_name) if frame_name in model.body_names else -1
                src = scaled_names.index(human_body) if human_body in scaled_names else -1
            tb.append(bid)
            ts.append(src)
            wp.append(float(pw))
            wr.append(float(rw))
            names.append(frame_name)
        task_body.append(np.array(tb, np.int32))
        task_src.append(np.array(ts, np.int32))
        task_wpos.append(np.array(wp, np.float64))
        task_wrot.append(np.array(wr, np.float64))
        task_names.append(names)

    return RetargetTables(
        src_human=src_human, tgt_robot=tgt_robot,
        human_root_name=root, robot_root_name=cfg["robot_root_name"],
        ground_height=ground, human_height_assumption=float(cfg["human_height_assumption"]),
        use_table=(bool(cfg["use_ik_match_table1"]), bool(cfg["use_ik_match_table2"])),
        scaled_names=scaled_names, scale=scale, pos_off=pos_off, rot_off=rot_off, is_foot=is_foot,
        task_body=task_body, task_src=task_src, task_wpos=task_wpos, task_wrot=task_wrot,
        task_frame_names=task_names, ik_config=cfg,
    )
