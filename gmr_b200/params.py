"""Registry: robot key -> MJCF file, (source, robot) -> IK config, base body, camera distance.

Mirrors the public names of the reference registry
(general_motion_retargeting/params.py:7-77: ``ROBOT_XML_DICT``, ``IK_CONFIG_DICT``,
``ROBOT_BASE_DICT``, ``VIEWER_CAM_DISTANCE_DICT``, ``IK_CONFIG_ROOT``, ``ASSET_ROOT``)
so a script written against the reference keeps working.

Where the files come from
-------------------------
The engine never needs the XML/JSON text at run time, only the flattened tables.
Two sources are supported:

* a GMR checkout: set ``GMR_ROOT`` (or call ``set_gmr_root``) to a directory that
  holds ``assets/`` and ``general_motion_retargeting/ik_configs/``; files are then
  parsed by ``gmr_b200.mjcf`` / ``gmr_b200.ik_config`` exactly as the reference
  would read them;
* the bundled model pack ``gmr_b200/data/model_pack.json`` (derived tables for
  every constructible (source, robot) pair, written by ``tools/build_model_pack.py``)
  which is what runs on a machine without a GMR checkout.
"""
from __future__ import annotations

import os
import pathlib

HERE = pathlib.Path(__file__).parent
PACK_PATH = HERE / "data" / "model_pack.json"

# robot key -> (MJCF path under assets/, base body, viewer camera distance, config stem)
_ROBOTS = {
    "unitree_g1":             ("unitree_g1/g1_mocap_29dof.xml",             "pelvis",      2.0, "g1"),
    "unitree_g1_with_hands":  ("unitree_g1/g1_mocap_29dof_with_hands.xml",  "pelvis",      2.0, "g1"),
    "booster_t1":             ("booster_t1/t1_mocap.xml",                   "Waist",       2.0, "t1"),
    "stanford_toddy":         ("stanford_toddy/toddy_mocap.xml",            "waist_link",  1.0, "toddy"),
    "fourier_n1":             ("fourier_n1/n1_mocap.xml",                   "base_link",   2.0, "n1"),
    "engineai_pm01":          ("engineai_pm01/pm_v2.xml",                   "LINK_BASE",   2.0, "pm01"),
    "kuavo_s45":              ("kuavo_s45/biped_s45_collision.xml",         "base_link",   3.0, "kuavo"),
    "hightorque_hi":          ("hightorque_hi/hi_25dof.xml",                "base_link",   2.0, "hi"),
    "galaxea_r1pro":          ("galaxea_r1pro/r1_pro.xml",                  "torso_link4", 3.0, "r1pro"),
    "berkeley_humanoid_lite": ("berkeley_humanoid_lite/bhl_scene.xml",      "pelvis",      2.0, "bhl"),
    "booster_k1":             ("booster_k1/K1_serial.xml",                  "Trunk",       2.0, "k1"),
}

# human source -> robots that have a "<source>_to_<stem>.json" entry in the reference registry
_SOURCES = {
    "smplx": ["unitree_g1", "unitree_g1_with_hands", "booster_t1", "stanford_toddy", "fourier_n1",
              "engineai_pm01", "kuavo_s45", "hightorque_hi", "galaxea_r1pro",
              "berkeley_humanoid_lite", "booster_k1"],
    "bvh": ["unitree_g1", "unitree_g1_with_hands", "booster_t1", "fourier_n1", "stanford_toddy",
            "engineai_pm01"],
    "fbx": ["unitree_g1", "unitree_g1_with_hands"],
}


def _default_root() -> pathlib.Path | None:
    env = os.environ.get("GMR_ROOT")
    if env:
        return pathlib.Path(env)
    return None


_gmr_root: pathlib.Path | None = _default_root()


def set_gmr_root(path) -> None:
    """Point the registry at a GMR checkout (directory holding ``assets/``)."""
    global _gmr_root
    _gmr_root = pathlib.Path(path) if path is not None else None
    _rebuild()


def get_gmr_root() -> pathlib.Path | None:
    return _gmr_root


ROBOT_XML_DICT: dict = {}
IK_CONFIG_DICT: dict = {}
ROBOT_BASE_DICT = {k: v[1] for k, v in _ROBOTS.items()}
VIEWER_CAM_DISTANCE_DICT = {k: v[2] for k, v in _ROBOTS.items()}
ASSET_ROOT: pathlib.Path | None = None
IK_CONFIG_ROOT: pathlib.Path | None = None


def _rebuild() -> None:
    global ASSET_ROOT, IK_CONFIG_ROOT
    root = _gmr_root if _gmr_root is not None else pathlib.Path("<model_pack>")
    ASSET_ROOT = root / "assets"
    IK_CONFIG_ROOT = root / "general_motion_retargeting" / "ik_configs"
    ROBOT_XML_DICT.clear()
    ROBOT_XML_DICT.update({k: ASSET_ROOT / v[0] for k, v in _ROBOTS.items()})
    IK_CONFIG_DICT.clear()
    for src, robots in _SOURCES.items():
        IK_CONFIG_DICT[src] = {r: IK_CONFIG_ROOT / f"{src}_to_{_ROBOTS[r][3]}.json" for r in robots}


_rebuild()


def constructible_pairs(root=None) -> list:
    """(source, robot) pairs whose IK config file exists under ``root`` (a GMR checkout)."""
    root = pathlib.Path(root) if root is not None else _gmr_root
    out = []
    for src, robots in _SOURCES.items():
        for r in robots:
            cfg = root / "general_motion_retargeting" / "ik_configs" / f"{src}_to_{_ROBOTS[r][3]}.json"
            xml = root / "assets" / _ROBOTS[r][0]
            if cfg.exists() and xml.exists():
                out.append((src, r))
    return out
