"""ctypes mirror of include/gmr_b200.h (struct layouts + descriptor builder).

Host-side glue only: turns (RobotModel, RetargetTables, human body order) into a
``gmr_model_desc`` whose pointers reference numpy arrays kept alive by the returned
``ModelDesc`` object.
"""
from __future__ import annotations

import ctypes as C

import numpy as np

from .ik_config import RetargetTables
from .mjcf import RobotModel

ABI_VERSION = 2
MAX_NV, MAX_NBODY, MAX_TASKS, MAX_SCALED = 64, 64, 32, 32
F32, F64 = 0, 1

_pi = C.POINTER(C.c_int32)
_pd = C.POINTER(C.c_double)


class gmr_model_desc(C.Structure):
    _fields_ = [
        ("abi_version", C.c_int32),
        ("nbody", C.c_int32), ("njnt", C.c_int32), ("nq", C.c_int32), ("nv", C.c_int32),
        ("body_parentid", _pi), ("body_pos", _pd), ("body_quat", _pd),
        ("body_jntadr", _pi), ("body_jntnum", _pi),
        ("jnt_type", _pi), ("jnt_qposadr", _pi), ("jnt_dofadr", _pi),
        ("jnt_axis", _pd), ("jnt_pos", _pd), ("jnt_limited", _pi), ("jnt_range", _pd),
        ("qpos0", _pd), ("timestep", C.c_double),
        ("n_human", C.c_int32), ("n_scaled", C.c_int32),
        ("scaled_human_index", _pi), ("scale", _pd), ("pos_off", _pd), ("rot_off", _pd),
        ("is_foot", _pi),
        ("use_table", C.c_int32 * 2), ("n_task", C.c_int32 * 2),
        ("task_body", _pi * 2), ("task_src", _pi * 2), ("task_wpos", _pd * 2), ("task_wrot", _pd * 2),
        ("damping", C.c_double), ("max_iter", C.c_int32), ("converge_eps", C.c_double),
        ("limit_gain", C.c_double), ("lm_damping", C.c_double),
        ("dof_vmax", _pd),
    ]


class gmr_diag(C.Structure):
    _fields_ = [("err_dev", C.c_void_p), ("iters_dev", C.c_void_p), ("nfact_dev", C.c_void_p), ("status_dev", C.c_void_p)]


# gmr_diag.status_dev bits (include/gmr_b200.h)
STATUS_MAX_ITER1, STATUS_MAX_ITER2, STATUS_QP_CAP, STATUS_BAD_INPUT, STATUS_BAD_OUTPUT = 1, 2, 4, 8, 16
STATUS_FAILED = STATUS_QP_CAP | STATUS_BAD_INPUT | STATUS_BAD_OUTPUT      # what makes a driver skip a clip


class ModelDesc:
    """Owns the numpy arrays a ``gmr_model_desc`` points to."""

    def __init__(self, model: RobotModel, tables: RetargetTables, human_names=None,
                 damping: float = 0.5, max_iter: int = 10, dof_vmax=None):
        """dof_vmax: optional velocity limit per dof in rad/s or m/s ([nv], +inf = none; a scalar applies to every
        hinge/slide dof).  None = off, the reference's behaviour (motion_retarget.py:147-149 passes no limits)."""
        self.model, self.tables = model, tables
        if human_names is None:
            human_names = list(tables.scaled_names)
        self.human_names = list(human_names)
        if model.nv > MAX_NV or model.nbody > MAX_NBODY or tables.n_scaled > MAX_SCALED:
            raise ValueError("model exceeds kernel limits")
        for k in range(2):
            if tables.use_table[k] and len(tables.task_body[k]) > MAX_TASKS:
                raise ValueError("too many tasks")
        idx = []
        for b in tables.scaled_names:
            if b not in self.human_names:
                raise KeyError(b)          # reference: human_data[human_root_name] / [body_name]
            idx.append(self.human_names.index(b))
        self._keep = []
        d = gmr_model_desc()
        d.abi_version = ABI_VERSION
        d.nbody, d.njnt, d.nq, d.nv = model.nbody, model.njnt, model.nq, model.nv

        def ai(a):
            a = np.ascontiguousarray(a, np.int32)
            self._keep.append(a)
            return a.ctypes.data_as(_pi)

        def ad(a):
            a = np.ascontiguousarray(a, np.float64)
            self._keep.append(a)
            return a.ctypes.data_as(_pd)

        d.body_parentid = ai(model.body_parentid)
        d.body_pos, d.body_quat = ad(model.body_pos), ad(model.body_quat)
        d.body_jntadr, d.body_jntnum = ai(model.body_jntadr), ai(model.body_jntnum)
        d.jnt_type, d.jnt_qposadr, d.jnt_dofadr = ai(model.jnt_type), ai(model.jnt_qposadr), ai(model.jnt_dofadr)
        d.jnt_axis, d.jnt_pos = ad(model.jnt_axis), ad(model.jnt_pos)
        d.jnt_limited, d.jnt_range = ai(model.jnt_limited), ad(model.jnt_range)
        d.qpos0, d.timestep = ad(model.qpos0), float(model.timestep)
        d.n_human, d.n_scaled = len(self.human_names), tables.n_scaled
        d.scaled_human_index = ai(idx)
        d.scale, d.pos_off, d.rot_off = ad(tables.scale), ad(tables.pos_off), ad(tables.rot_off)
        d.is_foot = ai(tables.is_foot)
        for k in range(2):
            use = bool(tables.use_table[k])
            d.use_table[k] = int(use)
            d.n_task[k] = len(tables.task_body[k]) if use else 0
            d.task_body[k] = ai(tables.task_body[k] if use else [0])
            d.task_src[k] = ai(tables.task_src[k] if use else [0])
            d.task_wpos[k] = ad(tables.task_wpos[k] if use else [0.0])
            d.task_wrot[k] = ad(tables.task_wrot[k] if use else [0.0])
        d.damping, d.max_iter = float(damping), int(max_iter)
        d.converge_eps, d.limit_gain, d.lm_damping = 0.001, 0.95, 1.0
        self.dof_vmax = None
        if dof_vmax is not None:
            vm = np.full(model.nv, np.inf)
            if np.ndim(dof_vmax) == 0:
                for j in range(model.njnt):
                    if model.jnt_type[j] != 0:
                        vm[model.jnt_dofadr[j]] = float(dof_vmax)
            else:
                vm[:] = np.asarray(dof_vmax, np.float64)
            self.dof_vmax = vm
            d.dof_vmax = ad(vm)
        self.c = d

    @property
    def ref(self):
        return C.byref(self.c)
