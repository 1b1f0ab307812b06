"""Loader of the CUDA engine (gmr_b200/libgmr_b200.so, built in-tree by gmr_b200.build).

There is no CPU fallback: if the library is missing or does not load, importing the engine
raises.  Only the C ABI of include/gmr_b200.h is used (ctypes; plain pointers and sizes).
"""
from __future__ import annotations

import ctypes as C
import os

from .abi import ABI_VERSION, gmr_diag, gmr_model_desc

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("GMR_B200_LIB") or os.path.join(HERE, "libgmr_b200.so")

EXPORTS = [
    "gmr_last_error", "gmr_abi_version", "gmr_model_create", "gmr_model_destroy", "gmr_batch_create",
    "gmr_batch_destroy", "gmr_batch_reset", "gmr_batch_set_scale", "gmr_batch_get_state",
    "gmr_batch_set_order", "gmr_batch_launch_info", "gmr_batch_dims", "gmr_batch_step", "gmr_batch_run",
    "gmr_target_prep", "gmr_fk", "gmr_build_qp", "gmr_qp_solve", "gmr_fk_bodies_f32", "gmr_probe_fma_peak", "gmr_solve_ik", "gmr_batch_run_host",
    "gmr_bvh_to_stream", "gmr_smplx_to_stream", "gmr_model_table", "gmr_batch_run_csr", "gmr_batch_run_host_csr",
    "gmr_batch_step_host", "gmr_probe_math", "gmr_model_layout", "gmr_batch_layout_id", "gmr_bvh_to_stream_static",
]

_lib = None


class EngineError(RuntimeError):
    pass


def lib():
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise EngineError(f"{LIB_PATH} not found: build it with `python -m gmr_b200.build` "
                          "(there is no CPU fallback)")
    L = C.CDLL(LIB_PATH)
    for name in EXPORTS:
        if not hasattr(L, name):
            raise EngineError(f"{LIB_PATH} does not export {name}")
    L.gmr_last_error.restype = C.c_char_p
    vp, i32 = C.c_void_p, C.c_int32
    L.gmr_model_create.argtypes = [C.POINTER(gmr_model_desc), C.c_int, C.POINTER(vp)]
    L.gmr_model_destroy.argtypes = [vp]
    L.gmr_model_table.argtypes = [vp, C.c_int, C.c_int, vp, C.c_int]
    L.gmr_model_layout.argtypes = [vp, vp, C.c_int]
    L.gmr_batch_layout_id.argtypes = [vp]
    L.gmr_batch_create.argtypes = [C.POINTER(vp), C.c_int, vp, C.c_int, vp, C.POINTER(vp)]
    L.gmr_batch_destroy.argtypes = [vp]
    L.gmr_batch_reset.argtypes = [vp, vp]
    L.gmr_batch_set_scale.argtypes = [vp, vp]
    L.gmr_batch_get_state.argtypes = [vp, vp]
    L.gmr_batch_set_order.argtypes = [vp, vp]
    L.gmr_batch_launch_info.argtypes = [vp] + [C.POINTER(i32)] * 4
    L.gmr_batch_dims.argtypes = [vp] + [C.POINTER(i32)] * 5
    L.gmr_batch_step.argtypes = [vp, vp, C.c_int, C.c_int, vp, vp, C.POINTER(gmr_diag)]
    L.gmr_batch_run.argtypes = [vp, vp, C.c_int, vp, C.c_int, C.c_int, vp, C.POINTER(gmr_diag)]
    L.gmr_target_prep.argtypes = [vp, vp, C.c_int, C.c_int, vp]
    L.gmr_fk.argtypes = [vp, vp, vp, vp]
    L.gmr_build_qp.argtypes = [vp, C.c_int, vp, vp, vp, vp, vp, vp]
    L.gmr_batch_run_host.argtypes = [vp, vp, C.c_int, vp, C.c_int, C.c_int, C.c_int, C.c_int, vp]
    L.gmr_batch_run_csr.argtypes = [vp, vp, C.c_int, vp, C.c_int, C.c_int, vp, C.POINTER(gmr_diag)]
    L.gmr_batch_run_host_csr.argtypes = [vp, vp, C.c_int, vp, C.c_int, C.c_int, vp, vp]
    L.gmr_batch_step_host.argtypes = [vp, vp, C.c_int, C.c_int, vp, vp, C.c_int]
    L.gmr_solve_ik.argtypes = [vp, C.c_int, vp, vp, vp]
    L.gmr_qp_solve.argtypes = [C.c_int, vp, C.c_int, C.c_int, vp, vp, vp, vp, vp, vp]
    L.gmr_fk_bodies_f32.argtypes = [vp, vp, C.c_int, vp, vp, vp, vp, vp]
    L.gmr_probe_fma_peak.argtypes = [C.c_int, C.c_int, C.POINTER(C.c_double)]
    L.gmr_probe_math.argtypes = [C.c_int, C.c_int, vp, vp, vp]
    L.gmr_smplx_to_stream.argtypes = [C.c_int, vp, vp, vp, C.c_int, vp, vp, vp, vp, C.c_int, C.c_int, C.c_int, vp, C.c_int, vp]
    L.gmr_bvh_to_stream.argtypes = [C.c_int, vp, vp, vp, vp, C.c_int, C.c_int, C.c_int, vp, C.POINTER(i32), vp, vp, C.c_int,
                                    C.c_int, vp, C.c_int, C.POINTER(i32), vp]
    L.gmr_bvh_to_stream_static.argtypes = L.gmr_bvh_to_stream.argtypes
    if L.gmr_abi_version() != ABI_VERSION:
        raise EngineError("ABI version mismatch between gmr_b200/abi.py and libgmr_b200.so")
    _lib = L
    return L


def check(rc: int, what: str) -> None:
    if rc != 0:
        msg = lib().gmr_last_error()
        raise EngineError(f"{what} failed ({rc}): {msg.decode() if msg else ''}")
