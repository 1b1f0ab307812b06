"""MJCF -> flat kinematic-tree tables (host side, one-time per robot).

Replaces the reference's ``mj.MjModel.from_xml_path(xml)`` call
(reference: general_motion_retargeting/motion_retarget.py:27) for the part of the
model the retargeting hot path actually reads: the kinematic tree, joint axes,
joint ranges and ``qpos0``.  Geometry, inertia, actuators, sensors, assets and
keyframes are ignored (the reference ignores them too: the IK loop only calls
``mj_kinematics`` / ``mj_jacBody`` / ``mj_integratePos``).

MuJoCo compile semantics that are reproduced:

* ``<include file=...>`` is textual, resolved relative to the including file;
* ``<default>`` classes nest and inherit; an element takes its explicit
  ``class``, else the nearest ancestor body's ``childclass``, else ``main``;
  ``<freejoint>`` never reads defaults;
* bodies are numbered depth-first in document order, world is body 0;
  joints are numbered in the same traversal, document order within a body;
* joint ``axis`` and body ``quat`` are normalised;
* ``limited`` = explicit true/false, else (``autolimits``, default true)
  ``range[0] < range[1]``;
* ``qpos0``: free joint = body ``pos`` + ``quat``; hinge/slide = ``ref``.

Only what the shipped robots use is supported (body ``pos``/``quat``; free,
slide and hinge joints; ``angle="radian"``); anything else raises
``MjcfError`` instead of silently producing a different tree.
"""
from __future__ import annotations

import os
import xml.etree.ElementTree as ET
from dataclasses import dataclass, field

import numpy as np

# MuJoCo mjtJoint enum values
JNT_FREE, JNT_BALL, JNT_SLIDE, JNT_HINGE = 0, 1, 2, 3


class MjcfError(ValueError):
    pass


@dataclass
class RobotModel:
    """Flat tables, MuJoCo naming.  All float arrays are float64."""

    name: str
    nbody: int                      # including world (body 0)
    nq: int
    nv: int
    njnt: int
    timestep: float
    body_names: list
    body_parentid: np.ndarray       # [nbody] int32 (world -> 0)
    body_pos: np.ndarray            # [nbody,3]
    body_quat: np.ndarray           # [nbody,4] wxyz, unit
    body_jntadr: np.ndarray         # [nbody] int32 (-1 if none)
    body_jntnum: np.ndarray         # [nbody] int32
    body_depth: np.ndarray          # [nbody] int32, world = 0
    jnt_names: list
    jnt_type: np.ndarray            # [njnt] int32 (mjtJoint)
    jnt_bodyid: np.ndarray          # [njnt] int32
    jnt_qposadr: np.ndarray         # [njnt] int32
    jnt_dofadr: np.ndarray          # [njnt] int32
    jnt_axis: np.ndarray            # [njnt,3] unit, body frame
    jnt_pos: np.ndarray             # [njnt,3] anchor, body frame
    jnt_limited: np.ndarray         # [njnt] int32 0/1
    jnt_range: np.ndarray           # [njnt,2]
    qpos0: np.ndarray               # [nq]
    dof_bodyid: np.ndarray          # [nv] int32
    dof_jntid: np.ndarray           # [nv] int32
    source_files: list = field(default_factory=list)

    def body_id(self, name: str) -> int:
        try:
            return self.body_names.index(name)
        except ValueError:
            raise KeyError(f"body {name!r} not in model {self.name!r}") from None

    # ---- (de)serialisation for the bundled model pack -------------------
    _ARRAYS = ("body_parentid", "body_pos", "body_quat", "body_jntadr", "body_jntnum",
               "body_depth", "jnt_type", "jnt_bodyid", "jnt_qposadr", "jnt_dofadr",
               "jnt_axis", "jnt_pos", "jnt_limited", "jnt_range", "qpos0",
               "dof_bodyid", "dof_jntid")

    def to_dict(self) -> dict:
        d = {"name": self.name, "nbody": self.nbody, "nq": self.nq, "nv": self.nv,
             "njnt": self.njnt, "timestep": self.timestep,
             "body_names": list(self.body_names), "jnt_names": list(self.jnt_names)}
        for k in self._ARRAYS:
            d[k] = getattr(self, k).tolist()
        return d

    @classmethod
    def from_dict(cls, d: dict) -> "RobotModel":
        kw = {k: d[k] for k in ("name", "nbody", "nq", "nv", "njnt", "timestep",
                                "body_names", "jnt_names")}
        for k in cls._ARRAYS:
            a = np.asarray(d[k])
            if a.dtype.kind in "iu" or k in ("body_parentid", "body_jntadr", "body_jntnum",
                                             "body_depth", "jnt_type", "jnt_bodyid",
                                             "jnt_qposadr", "jnt_dofadr", "jnt_limited",
                                             "dof_bodyid", "dof_jntid"):
                a = a.astype(np.int32)
            else:
                a = a.astype(np.float64)
            kw[k] = a
        kw["jnt_axis"] = kw["jnt_axis"].reshape(-1, 3)
        kw["jnt_pos"] = kw["jnt_pos"].reshape(-1, 3)
        kw["jnt_range"] = kw["jnt_range"].reshape(-1, 2)
        return cls(**kw)


# --------------------------------------------------------------------------
# XML loading with textual <include>
# --------------------------------------------------------------------------
def _load_with_includes(path: str, files: list, depth: int = 0) -> ET.Element:
    if depth > 16:
        raise MjcfError(f"<include> nesting too deep at {path}")
    root = ET.parse(path).getroot()
    if root.tag != "mujoco":
        raise MjcfError(f"{path}: root element is <{root.tag}>, expected <mujoco>")
    files.append(path)
    base = os.path.dirname(path)

    def expand(elem: ET.Element) -> None:
        new_children = []
        for child in list(elem):
            if child.tag == "include":
                inc_file = child.get("file")
                if inc_file is None:
                    raise MjcfError(f"{path}: <include> without file")
                inc_root = _load_with_includes(os.path.join(base, inc_file), files, depth + 1)
                new_children.extend(list(inc_root))
            else:
                expand(child)
                new_children.append(child)
        for child in list(elem):
            elem.remove(child)
        for child in new_children:
            elem.append(child)

    expand(root)
    return root


def _floats(text: str, n: int, what: str) -> np.ndarray:
    a = np.array([float(t) for t in text.split()], dtype=np.float64)
    if a.shape != (n,):
        raise MjcfError(f"{what}: expected {n} numbers, got {text!r}")
    return a


def _parse_bool(text: str, what: str) -> bool:
    if text == "true":
        return True
    if text == "false":
        return False
    raise MjcfError(f"{what}: expected true/false, got {text!r}")


_JOINT_TYPES = {"free": JNT_FREE, "ball": JNT_BALL, "slide": JNT_SLIDE, "hinge": JNT_HINGE}
_UNSUPPORTED_ORIENT = ("euler", "axisangle", "xyaxes", "zaxis")


def _collect_joint_defaults(root: ET.Element) -> dict:
    """class name -> merged <joint> attribute dict (inheritance resolved)."""
    classes: dict = {"main": {}}

    def walk(elem: ET.Element, cls_name: str, inherited: dict) -> None:
        attrs = dict(inherited)
        for j in elem.findall("joint"):
            attrs.update(j.attrib)
        # several top-level <default> blocks all contribute to "main"
        merged = dict(classes.get(cls_name, {}))
        merged.update(attrs)
        classes[cls_name] = merged
        for sub in elem.findall("default"):
            sub_name = sub.get("class")
            if sub_name is None:
                raise MjcfError("nested <default> without class")
            walk(sub, sub_name, classes[cls_name])

    for top in root.findall("default"):
        walk(top, top.get("class") or "main", classes["main"])
    return classes


def load_mjcf(xml_path: str, name: str | None = None) -> RobotModel:
    xml_path = os.fspath(xml_path)
    files: list = []
    root = _load_with_includes(xml_path, files)

    autolimits = True
    for comp in root.findall("compiler"):
        angle = comp.get("angle")
        if angle is not None and angle != "radian":
            raise MjcfError(f"{xml_path}: compiler angle={angle!r} unsupported (radian only)")
        if comp.get("coordinate", "local") != "local":
            raise MjcfError("compiler coordinate=global unsupported")
        if comp.get("autolimits") is not None:
            autolimits = _parse_bool(comp.get("autolimits"), "compiler autolimits")
        if comp.get("fusestatic", "false") != "false":
            raise MjcfError("compiler fusestatic=true unsupported")
    if not any(c.get("angle") == "radian" for c in root.findall("compiler")):
        raise MjcfError(f"{xml_path}: compiler angle must be 'radian' (MuJoCo default is degree)")

    timestep = 0.002  # MuJoCo default opt.timestep
    for opt in root.findall("option"):
        if opt.get("timestep") is not None:
            timestep = float(opt.get("timestep"))

    defaults = _collect_joint_defaults(root)

    body_names = ["world"]
    body_parent = [0]
    body_pos = [np.zeros(3)]
    body_quat = [np.array([1.0, 0, 0, 0])]
    body_jntadr = [-1]
    body_jntnum = [0]
    body_depth = [0]
    jnt = dict(names=[], type=[], body=[], qadr=[], dadr=[], axis=[], pos=[], limited=[],
               range=[], ref=[])
    qpos0: list = []
    nq = nv = 0

    def add_body(elem: ET.Element, parent: int, childclass: str | None, depth: int) -> None:
        nonlocal nq, nv
        for bad in _UNSUPPORTED_ORIENT:
            if elem.get(bad) is not None:
                raise MjcfError(f"body {elem.get('name')!r}: orientation attribute {bad!r} unsupported")
        bid = len(body_names)
        bname = elem.get("name") or f"body{bid}"
        pos = _floats(elem.get("pos", "0 0 0"), 3, f"body {bname} pos")
        quat = _floats(elem.get("quat", "1 0 0 0"), 4, f"body {bname} quat")
        qn = np.linalg.norm(quat)
        if qn < 1e-15:
            raise MjcfError(f"body {bname}: zero quaternion")
        quat = quat / qn
        body_names.append(bname)
        body_parent.append(parent)
        body_pos.append(pos)
        body_quat.append(quat)
        body_depth.append(depth)
        cc = elem.get("childclass", childclass)

        first_j = len(jnt["names"])
        count = 0
        for child in elem:
            if child.tag == "freejoint":
                attrs = {"type": "free"}
                jname = child.get("name")
            elif child.tag == "joint":
                cls = child.get("class") or cc or "main"
                if cls not in defaults:
                    raise MjcfError(f"joint {child.get('name')!r}: unknown default class {cls!r}")
                attrs = dict(defaults[cls])
                attrs.update(child.attrib)
                jname = child.get("name")
            else:
                continue
            jid = len(jnt["names"])
            jtype = _JOINT_TYPES.get(attrs.get("type", "hinge"))
            if jtype is None:
                raise MjcfError(f"joint {jname!r}: unknown type {attrs.get('type')!r}")
            if jtype == JNT_BALL:
                raise MjcfError(f"joint {jname!r}: ball joints unsupported")
            axis = _floats(attrs.get("axis", "0 0 1"), 3, f"joint {jname} axis")
            jpos = _floats(attrs.get("pos", "0 0 0"), 3, f"joint {jname} pos")
            rng = _floats(attrs.get("range", "0 0"), 2, f"joint {jname} range")
            ref = float(attrs.get("ref", "0"))
            lim_attr = attrs.get("limited", "auto")
            if jtype == JNT_FREE:
                if count != 0 or parent != 0:
                    raise MjcfError(f"free joint {jname!r} must be the only joint of a top-level body")
                limited = False
                axis = np.array([0.0, 0.0, 1.0])
                jpos = np.zeros(3)
            else:
                an = np.linalg.norm(axis)
                if an < 1e-15:
                    raise MjcfError(f"joint {jname!r}: zero axis")
                axis = axis / an
                if lim_attr == "auto":
                    if not autolimits and (rng[0] != 0 or rng[1] != 0):
                        raise MjcfError(f"joint {jname!r}: range without limited and autolimits=false")
                    limited = bool(autolimits and rng[0] < rng[1])
                else:
                    limited = _parse_bool(lim_attr, f"joint {jname} limited")
            jnt["names"].append(jname or f"joint{jid}")
            jnt["type"].append(jtype)
            jnt["body"].append(bid)
            jnt["qadr"].append(nq)
            jnt["dadr"].append(nv)
            jnt["axis"].append(axis)
            jnt["pos"].append(jpos)
            jnt["limited"].append(int(limited))
            jnt["range"].append(rng)
            if jtype == JNT_FREE:
                qpos0.extend(pos.tolist() + quat.tolist())
                nq += 7
                nv += 6
            else:
                qpos0.append(ref)
                nq += 1
                nv += 1
            count += 1
        if count > 1 and any(t == JNT_FREE for t in jnt["type"][first_j:]):
            raise MjcfError(f"body {bname}: free joint combined with other joints")
        body_jntadr.append(first_j if count else -1)
        body_jntnum.append(count)
        for child in elem.findall("body"):
            add_body(child, bid, cc, depth + 1)

    for wb in root.findall("worldbody"):
        for top in wb.findall("body"):
            add_body(top, 0, wb.get("childclass"), 1)

    njnt = len(jnt["names"])
    dof_body = np.zeros(nv, np.int32)
    dof_jnt = np.zeros(nv, np.int32)
    for j in range(njnt):
        w = 6 if jnt["type"][j] == JNT_FREE else 1
        dof_body[jnt["dadr"][j]:jnt["dadr"][j] + w] = jnt["body"][j]
        dof_jnt[jnt["dadr"][j]:jnt["dadr"][j] + w] = j

    model_name = name or root.get("model") or os.path.basename(xml_path)
    return RobotModel(
        name=model_name, nbody=len(body_names), nq=nq, nv=nv, njnt=njnt, timestep=timestep,
        body_names=body_names,
        body_parentid=np.array(body_parent, np.int32),
        body_pos=np.array(body_pos, np.float64).reshape(-1, 3),
        body_quat=np.array(body_quat, np.float64).reshape(-1, 4),
        body_jntadr=np.array(body_jntadr, np.int32),
        body_jntnum=np.array(body_jntnum, np.int32),
        body_depth=np.array(body_depth, np.int32),
        jnt_names=jnt["names"],
        jnt_type=np.array(jnt["type"], np.int32),
        jnt_bodyid=np.array(jnt["body"], np.int32),
        jnt_qposadr=np.array(jnt["qadr"], np.int32),
        jnt_dofadr=np.array(jnt["dadr"], np.int32),
        jnt_axis=np.array(jnt["axis"], np.float64).reshape(-1, 3),
        jnt_pos=np.array(jnt["pos"], np.float64).reshape(-1, 3),
        jnt_limited=np.array(jnt["limited"], np.int32),
        jnt_range=np.array(jnt["range"], np.float64).reshape(-1, 2),
        qpos0=np.array(qpos0, np.float64),
        dof_bodyid=dof_body, dof_jntid=dof_jnt,
        source_files=files,
    )
