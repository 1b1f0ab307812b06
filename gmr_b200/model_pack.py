"""Resolve (source, robot) -> (RobotModel, ik_config dict), from a GMR checkout or the bundled pack.

The pack (``gmr_b200/data/model_pack.json``) holds *derived* tables: the flattened
kinematic tree of each registered robot (output of ``gmr_b200.mjcf.load_mjcf``) and the
parsed IK config of each constructible (source, robot) pair.  It exists because the
GPU box has no GMR checkout; ``tools/build_model_pack.py`` regenerates it and
``tests/test_model_pack.py`` checks it against a fresh parse when a checkout is present.
"""
from __future__ import annotations

import functools
import json
import os

from . import params
from .ik_config import RetargetTables, build_retarget_tables, load_ik_config
from .mjcf import RobotModel, load_mjcf


@functools.lru_cache(maxsize=1)
def _pack() -> dict:
    if not params.PACK_PATH.exists():
        raise FileNotFoundError(
            f"model pack {params.PACK_PATH} missing and GMR_ROOT not set; "
            "run tools/build_model_pack.py or export GMR_ROOT=<GMR checkout>")
    with open(params.PACK_PATH) as f:
        return json.load(f)


@functools.lru_cache(maxsize=64)
def _robot_from_root(root: str, tgt_robot: str) -> RobotModel:
    return load_mjcf(os.fspath(params.ROBOT_XML_DICT[tgt_robot]), name=tgt_robot)


def load_robot(tgt_robot: str) -> RobotModel:
    if tgt_robot not in params.ROBOT_XML_DICT:
        raise KeyError(tgt_robot)           # reference: ROBOT_XML_DICT[tgt_robot], motion_retarget.py:24
    root = params.get_gmr_root()
    if root is not None:
        return _robot_from_root(os.fspath(root), tgt_robot)
    robots = _pack()["robots"]
    if tgt_robot not in robots:
        raise KeyError(tgt_robot)
    return RobotModel.from_dict(robots[tgt_robot])


def load_config(src_human: str, tgt_robot: str) -> dict:
    cfg_path = params.IK_CONFIG_DICT[src_human][tgt_robot]   # KeyError like the reference (:30)
    root = params.get_gmr_root()
    if root is not None:
        return load_ik_config(cfg_path)                      # FileNotFoundError like the reference
    configs = _pack()["ik_configs"]
    key = cfg_path.name
    if key not in configs:
        raise FileNotFoundError(str(cfg_path))
    return json.loads(json.dumps(configs[key]))              # private copy


def load_pair(src_human: str, tgt_robot: str) -> tuple:
    model = load_robot(tgt_robot)
    cfg = load_config(src_human, tgt_robot)
    return model, cfg


def build_tables(src_human: str, tgt_robot: str) -> tuple:
    model, cfg = load_pair(src_human, tgt_robot)
    return model, build_retarget_tables(model, cfg, src_human, tgt_robot)


def constructible_pairs() -> list:
    root = params.get_gmr_root()
    if root is not None:
        return params.constructible_pairs(root)
    return [tuple(p) for p in _pack()["pairs"]]
