"""TEST INFRASTRUCTURE (oracle) - ctypes front end of oracle/liboracle.so (gmr_oracle.c).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs
may import this module.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

from gmr_b200.abi import F32, F64, ModelDesc, gmr_model_desc

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "liboracle.so")
_lib = None


def build(force: bool = False) -> str:
    src = os.path.join(HERE, "gmr_oracle.c")
    hdr = os.path.join(HERE, "..", "include", "gmr_b200.h")
    stale = (not os.path.exists(LIB_PATH)
             or (os.path.exists(src) and os.path.getmtime(src) > os.path.getmtime(LIB_PATH))
             or (os.path.exists(hdr) and os.path.getmtime(hdr) > os.path.getmtime(LIB_PATH)))
    if force or stale:
        subprocess.check_call(["make", "-C", HERE, "CC=gcc", "-B", "liboracle.so"],
                              stdout=subprocess.DEVNULL)
    return LIB_PATH


def lib():
    global _lib
    if _lib is None:
        build()
        _lib = C.CDLL(LIB_PATH)
        _lib.oracle_run_clips.restype = C.c_int
    return _lib


def _p(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def run_clips(desc: ModelDesc, human: np.ndarray, T_of=None, scale=None, offset_to_ground=False,
              q_state=None, nthreads: int = 1, diag: bool = True):
    """human [B,T,h,7] float32/float64 -> dict(qpos [B,T,nq], err [B,T,2], iters [B,T,2], nfact [B,T])."""
    human = np.ascontiguousarray(human)
    assert human.ndim == 4 and human.shape[2] == desc.c.n_human and human.shape[3] == 7
    dtype = {np.dtype(np.float32): F32, np.dtype(np.float64): F64}[human.dtype]
    B, T = human.shape[:2]
    nq = desc.c.nq
    qpos = np.zeros((B, T, nq))
    err = np.zeros((B, T, 2)) if diag else None
    iters = np.zeros((B, T, 2), np.int32) if diag else None
    nfact = np.zeros((B, T), np.int32) if diag else None
    if T_of is not None:
        T_of = np.ascontiguousarray(T_of, np.int32)
    if scale is not None:
        scale = np.ascontiguousarray(scale, np.float64)
        assert scale.shape == (B, desc.c.n_scaled)
    if q_state is not None:
        assert q_state.dtype == np.float64 and q_state.shape == (B, nq) and q_state.flags.c_contiguous
    rc = lib().oracle_run_clips(desc.ref, B, T, _p(T_of), _p(human), dtype, _p(scale),
                                int(bool(offset_to_ground)), _p(q_state), _p(qpos), _p(err), _p(iters),
                                _p(nfact), int(nthreads))
    if rc != 0:
        raise RuntimeError(f"oracle_run_clips failed rc={rc}")
    return {"qpos": qpos, "err": err, "iters": iters, "nfact": nfact}


def prepare(desc: ModelDesc, frame: np.ndarray, scale=None, offset_to_ground=False) -> np.ndarray:
    frame = np.ascontiguousarray(frame, np.float64)
    assert frame.shape == (desc.c.n_human, 7)
    sc = np.ascontiguousarray(desc.tables.scale if scale is None else scale, np.float64)
    out = np.zeros((desc.c.n_scaled, 7))
    lib().oracle_prepare(desc.ref, _p(sc), _p(frame), int(bool(offset_to_ground)), _p(out))
    return out


def fk(desc: ModelDesc, q: np.ndarray):
    q = np.ascontiguousarray(q, np.float64)
    xpos = np.zeros((desc.c.nbody, 3))
    xquat = np.zeros((desc.c.nbody, 4))
    lib().oracle_fk_export(desc.ref, _p(q), _p(xpos), _p(xquat))
    return xpos, xquat


def build_qp(desc: ModelDesc, q: np.ndarray, stage: int, scaled: np.ndarray):
    q = np.ascontiguousarray(q, np.float64)
    scaled = np.ascontiguousarray(scaled, np.float64)
    nv = desc.c.nv
    H = np.zeros((nv, nv)); c = np.zeros(nv); lo = np.zeros(nv); hi = np.zeros(nv)
    e = np.zeros((desc.c.n_task[stage], 6))
    lib().oracle_build_qp_export(desc.ref, _p(q), int(stage), _p(scaled), _p(H), _p(c), _p(lo), _p(hi), _p(e))
    return H, c, lo, hi, e


def task_jacobian(desc: ModelDesc, q: np.ndarray, body: int, target7: np.ndarray):
    q = np.ascontiguousarray(q, np.float64)
    target7 = np.ascontiguousarray(target7, np.float64)
    J = np.zeros((6, desc.c.nv)); e = np.zeros(6)
    lib().oracle_task_jacobian_export(desc.ref, _p(q), int(body), _p(target7), _p(J), _p(e))
    return J, e


def box_qp(H, c, lo, hi):
    H = np.ascontiguousarray(H, np.float64); c = np.ascontiguousarray(c, np.float64)
    lo = np.ascontiguousarray(lo, np.float64); hi = np.ascontiguousarray(hi, np.float64)
    n = len(c)
    x = np.zeros(n); nf = C.c_int(0)
    rc = lib().oracle_box_qp(n, _p(H), _p(c), _p(lo), _p(hi), _p(x), C.byref(nf))
    if rc != 0:
        raise RuntimeError(f"oracle_box_qp rc={rc}")
    return x, nf.value


def integrate(desc: ModelDesc, q: np.ndarray, dq: np.ndarray) -> np.ndarray:
    q = np.array(q, np.float64); dq = np.ascontiguousarray(dq, np.float64)
    lib().oracle_integrate_export(desc.ref, _p(q), _p(dq))
    return q
