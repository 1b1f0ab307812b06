"""Host-side front end of the CUDA engine: device memory and streams through torch, every
compute call through the C ABI (gmr_b200/_lib.py -> libgmr_b200.so).

``BatchedRetargeter`` is the batch form of the reference class
(general_motion_retargeting/motion_retarget.py:10-185): B independent instances, each
equivalent to one ``GeneralMotionRetargeting`` object fed its clip frame by frame.
"""
from __future__ import annotations

import ctypes as C

import numpy as np
import torch

from . import _lib, model_pack
from .abi import F32, F64, ModelDesc, gmr_diag
from .ik_config import build_retarget_tables
from .synth import human_body_names


def _ptr(t):
    return None if t is None else C.c_void_p(t.data_ptr())


class DeviceModel:
    """gmr_model handle + the host tables it was built from."""

    def __init__(self, desc: ModelDesc, device: int = 0):
        self.desc = desc
        self.device = int(device)
        self._h = C.c_void_p()
        _lib.check(_lib.lib().gmr_model_create(desc.ref, self.device, C.byref(self._h)), "gmr_model_create")

    @classmethod
    def load(cls, src_human: str, tgt_robot: str, human_names=None, device: int = 0, damping: float = 0.5,
             ik_config: dict | None = None):
        model, cfg = model_pack.load_pair(src_human, tgt_robot)
        if ik_config is not None:
            cfg = ik_config
        tables = build_retarget_tables(model, cfg, src_human, tgt_robot)
        if human_names is None:
            human_names = human_body_names(src_human, tables)
        return cls(ModelDesc(model, tables, human_names, damping=damping), device)

    @property
    def handle(self):
        return self._h

    def close(self):
        if self._h:
            _lib.lib().gmr_model_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def fk_bodies_f32(self, root_pos, root_rot_xyzw, dof_pos, want_rot: bool = True):
        """FK tail (KinematicsModel.forward_kinematics semantics): float32 torch tensors on the device."""
        N = root_pos.shape[0]
        nb = self.desc.model.nbody - 1
        dev = root_pos.device
        root_pos = root_pos.contiguous().float()
        root_rot_xyzw = root_rot_xyzw.contiguous().float()
        dof_pos = dof_pos.contiguous().float()
        body_pos = torch.empty((N, nb, 3), dtype=torch.float32, device=dev)
        body_rot = torch.empty((N, nb, 4), dtype=torch.float32, device=dev) if want_rot else None
        st = torch.cuda.current_stream(dev).cuda_stream
        _lib.check(_lib.lib().gmr_fk_bodies_f32(self._h, C.c_void_p(st), N, _ptr(root_pos), _ptr(root_rot_xyzw),
                                                _ptr(dof_pos), _ptr(body_pos), _ptr(body_rot)), "gmr_fk_bodies_f32")
        return body_pos, body_rot


class BatchedRetargeter:
    def __init__(self, models, B: int, model_of_instance=None, device: int | None = None):
        """models: DeviceModel or list of DeviceModel (heterogeneous launch)."""
        if isinstance(models, DeviceModel):
            models = [models]
        self.models = list(models)
        self.device = self.models[0].device if device is None else int(device)
        self.tdev = torch.device("cuda", self.device)
        self.B = int(B)
        self.stream = torch.cuda.current_stream(self.tdev)
        arr = (C.c_void_p * len(self.models))(*[m.handle for m in self.models])
        mo = None
        if model_of_instance is not None:
            self.model_of = np.ascontiguousarray(model_of_instance, np.int32)
            mo = self.model_of.ctypes.data_as(C.c_void_p)
        else:
            self.model_of = np.zeros(self.B, np.int32)
        self._h = C.c_void_p()
        _lib.check(_lib.lib().gmr_batch_create(arr, len(self.models), mo, self.B, C.c_void_p(self.stream.cuda_stream),
                                               C.byref(self._h)), "gmr_batch_create")
        d = [C.c_int32() for _ in range(5)]
        _lib.lib().gmr_batch_dims(self._h, *[C.byref(x) for x in d])
        _, self.nq_max, self.s_max, self.h_max, self.nbody_max = [x.value for x in d]
        self.nv_max = max(m.desc.c.nv for m in self.models)
        self.m_max = max(1, max(max(m.desc.c.n_task[0], m.desc.c.n_task[1]) for m in self.models))

    # ---------------------------------------------------------------- state
    def close(self):
        if self._h:
            _lib.lib().gmr_batch_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def launch_info(self) -> dict:
        d = [C.c_int32() for _ in range(4)]
        _lib.lib().gmr_batch_launch_info(self._h, *[C.byref(x) for x in d])
        info = dict(zip(("grid", "warps_per_cta", "smem_bytes_per_cta", "smem_bytes_per_instance"), [x.value for x in d]))
        info["layout_id"] = int(_lib.lib().gmr_batch_layout_id(self._h))
        return info

    def reset(self, q0=None):
        q0 = None if q0 is None else self._dev(q0, torch.float64, (self.B, self.nq_max))
        _lib.check(_lib.lib().gmr_batch_reset(self._h, _ptr(q0)), "gmr_batch_reset")

    def set_scale(self, scale=None):
        """scale [B, s_max]: per-instance human_scale_table (ratio folded in); None = config table."""
        scale = None if scale is None else self._dev(scale, torch.float64, (self.B, self.s_max))
        _lib.check(_lib.lib().gmr_batch_set_scale(self._h, _ptr(scale)), "gmr_batch_set_scale")

    def set_heights(self, heights):
        """actual_human_height per instance (motion_retarget.py:36-43)."""
        sc = np.zeros((self.B, self.s_max))
        for i in range(self.B):
            t = self.models[self.model_of[i]].desc.tables
            ratio = 1.0 if heights is None or heights[i] is None else float(heights[i]) / t.human_height_assumption
            sc[i, :t.n_scaled] = t.scale * ratio
        self.set_scale(sc)

    def set_order(self, order=None):
        """Processing order of the work queue: a permutation of range(B) (longest clips first), or None."""
        if order is not None:
            o = torch.as_tensor(order).to("cpu", torch.int64).reshape(-1)
            if o.numel() != self.B or not torch.equal(torch.sort(o).values, torch.arange(self.B)):
                raise ValueError("order must be a permutation of range(B): a duplicate would make two warps race on one "
                                 "clip and leave another one untouched")
        order = None if order is None else self._dev(order, torch.int32, (self.B,))
        _lib.check(_lib.lib().gmr_batch_set_order(self._h, _ptr(order)), "gmr_batch_set_order")

    def state(self) -> torch.Tensor:
        q = torch.empty((self.B, self.nq_max), dtype=torch.float64, device=self.tdev)
        _lib.check(_lib.lib().gmr_batch_get_state(self._h, _ptr(q)), "gmr_batch_get_state")
        return q

    def _dev(self, x, dtype, shape=None):
        t = torch.as_tensor(x)
        t = t.to(device=self.tdev, dtype=dtype, non_blocking=True).contiguous()
        if shape is not None and tuple(t.shape) != tuple(shape):
            raise ValueError(f"expected shape {tuple(shape)}, got {tuple(t.shape)}")
        return t

    def _human(self, human, ndim):
        t = torch.as_tensor(human)
        if t.dtype not in (torch.float32, torch.float64):
            t = t.double()
        t = t.to(self.tdev, non_blocking=True).contiguous()
        if t.dim() != ndim or t.shape[0] != self.B or t.shape[-2] != self.h_max or t.shape[-1] != 7:
            raise ValueError(f"human stream must be [B={self.B},{'T,' if ndim == 4 else ''}h={self.h_max},7], got {tuple(t.shape)}")
        return t, (F32 if t.dtype == torch.float32 else F64)

    def _diag(self, T, want):
        if not want:
            return None, {}
        out = {"err": torch.zeros((self.B, T, 2), dtype=torch.float64, device=self.tdev),
               "iters": torch.zeros((self.B, T, 2), dtype=torch.int32, device=self.tdev),
               "nfact": torch.zeros((self.B, T), dtype=torch.int32, device=self.tdev),
               "status": torch.zeros((self.B, T), dtype=torch.int32, device=self.tdev)}
        d = gmr_diag(out["err"].data_ptr(), out["iters"].data_ptr(), out["nfact"].data_ptr(), out["status"].data_ptr())
        return d, out

    # ---------------------------------------------------------------- hot path
    def run(self, human, T_of=None, offset_to_ground=False, diag=False, out=None):
        """human [B,T,h,7] (float32/float64; host or device) -> qpos [B,T,nq_max] float64 on the device.
        Instance state persists, so consecutive calls continue the clips (streaming in T-chunks)."""
        h, dt = self._human(human, 4)
        T = h.shape[1]
        t_of = None if T_of is None else self._dev(T_of, torch.int32, (self.B,))
        qpos = out if out is not None else torch.zeros((self.B, T, self.nq_max), dtype=torch.float64, device=self.tdev)
        d, dout = self._diag(T, diag)
        _lib.check(_lib.lib().gmr_batch_run(self._h, _ptr(h), dt, _ptr(t_of), T, int(bool(offset_to_ground)), _ptr(qpos),
                                            C.byref(d) if d is not None else None), "gmr_batch_run")
        return (qpos, dout) if diag else qpos

    def retarget_clips_host(self, human_host, qpos_host=None, T_of=None, offset_to_ground=False, cold_start=True,
                            n_chunks=None):
        """Whole clips from HOST buffers to a HOST result: the call the dataset drivers make
        (scripts/bvh_to_robot_dataset.py:84-113: new retargeter per clip -> per-frame retarget ->
        stacked qpos).  human_host [B,T,h,7] float32/float64 torch CPU tensor (pinned for full overlap);
        returns qpos [B,T,nq_max] float64 on the host (pinned; pass qpos_host to reuse it).
        One C-ABI call (gmr_batch_run_host).  n_chunks = 0: zero-copy - one launch whose warps read each frame
        from, and write each qpos row to, the pinned host buffers directly (needs both pinned).  n_chunks >= 1:
        the frames are cut into n_chunks ranges and H2D, retargeting and D2H of consecutive ranges overlap on
        three streams through device staging (any host memory).  None: zero-copy when both buffers are pinned,
        else 3 ranges.  Either way the result is bit-identical to run().
        Asynchronous on the batch's stream - synchronise it before reading the result."""
        human_host = torch.as_tensor(human_host)
        if human_host.is_cuda:
            raise ValueError("retarget_clips_host takes host buffers; use run() for device tensors")
        if human_host.dtype not in (torch.float32, torch.float64):
            human_host = human_host.double()
        human_host = human_host.contiguous()
        if human_host.dim() != 4 or human_host.shape[0] != self.B or human_host.shape[2] != self.h_max or human_host.shape[3] != 7:
            raise ValueError(f"human stream must be [B={self.B},T,h={self.h_max},7], got {tuple(human_host.shape)}")
        T = human_host.shape[1]
        if qpos_host is None:
            qpos_host = torch.zeros((self.B, T, self.nq_max), dtype=torch.float64, pin_memory=True)
        if qpos_host.dtype != torch.float64 or tuple(qpos_host.shape) != (self.B, T, self.nq_max) or not qpos_host.is_contiguous():
            raise ValueError("qpos_host must be a contiguous float64 [B,T,nq_max] host tensor")
        t_of = None
        if T_of is not None:
            t_of = np.ascontiguousarray(T_of, np.int32)
            if t_of.shape != (self.B,):
                raise ValueError("T_of must have one entry per clip")
        if n_chunks is None:
            n_chunks = 0 if (human_host.is_pinned() and qpos_host.is_pinned()) else 3
        _lib.check(_lib.lib().gmr_batch_run_host(
            self._h, C.c_void_p(human_host.data_ptr()), F32 if human_host.dtype == torch.float32 else F64,
            None if t_of is None else t_of.ctypes.data_as(C.c_void_p), T, int(bool(offset_to_ground)), int(bool(cold_start)),
            int(n_chunks), C.c_void_p(qpos_host.data_ptr())), "gmr_batch_run_host")
        # the launch is asynchronous and torch's pinned-memory allocator does not know the kernel uses these blocks:
        # every pair stays referenced until an event recorded behind its launch has completed
        ev = torch.cuda.Event()
        ev.record(self.stream)
        self._inflight = [x for x in getattr(self, "_inflight", []) if not x[0].query()]
        self._inflight.append((ev, human_host, qpos_host))
        return qpos_host

    def run_csr(self, human, clip_start, offset_to_ground=False, diag=False, out=None):
        """CSR form of run(): human [N,h,7] holds the clips back to back, clip i = rows clip_start[i]:clip_start[i+1]
        (clip_start: B+1 int64, starting at 0).  -> qpos [N,nq_max] on the device (+ diag arrays with N rows).
        Nothing is padded to the longest clip."""
        t = torch.as_tensor(human)
        if t.dtype not in (torch.float32, torch.float64):
            t = t.double()
        t = t.to(self.tdev, non_blocking=True).contiguous()
        cs = np.ascontiguousarray(clip_start, np.int64)
        if cs.shape != (self.B + 1,) or cs[0] != 0 or (np.diff(cs) < 0).any():
            raise ValueError("clip_start must be B+1 non-decreasing int64 offsets starting at 0")
        N = int(cs[-1])
        if t.dim() != 3 or t.shape[0] != N or t.shape[1] != self.h_max or t.shape[2] != 7:
            raise ValueError(f"human stream must be [N={N},h={self.h_max},7], got {tuple(t.shape)}")
        qpos = out if out is not None else torch.zeros((N, self.nq_max), dtype=torch.float64, device=self.tdev)
        dout, d = {}, None
        if diag:
            dout = {"err": torch.zeros((N, 2), dtype=torch.float64, device=self.tdev),
                    "iters": torch.zeros((N, 2), dtype=torch.int32, device=self.tdev),
                    "nfact": torch.zeros((N,), dtype=torch.int32, device=self.tdev),
                    "status": torch.zeros((N,), dtype=torch.int32, device=self.tdev)}
            d = gmr_diag(dout["err"].data_ptr(), dout["iters"].data_ptr(), dout["nfact"].data_ptr(), dout["status"].data_ptr())
        cs_dev = torch.from_numpy(cs).to(self.tdev)
        T_max = int(np.diff(cs).max()) if self.B else 0
        if N:
            _lib.check(_lib.lib().gmr_batch_run_csr(self._h, _ptr(t), F32 if t.dtype == torch.float32 else F64, _ptr(cs_dev),
                                                    max(T_max, 1), int(bool(offset_to_ground)), _ptr(qpos),
                                                    C.byref(d) if d is not None else None), "gmr_batch_run_csr")
        return (qpos, dout) if diag else qpos

    def retarget_clips_host_csr(self, human_host, clip_start, qpos_host=None, status_host=None, offset_to_ground=False,
                                cold_start=True):
        """CSR form of retarget_clips_host(): human_host [N,h,7] (torch CPU tensor; pinned = zero-copy), clip_start
        B+1 int64 -> (qpos_host [N,nq_max] float64, status_host [N] int32 GMR_STATUS_* bits), both pinned.
        One C-ABI call (gmr_batch_run_host_csr); asynchronous on the batch's stream."""
        human_host = torch.as_tensor(human_host)
        if human_host.is_cuda:
            raise ValueError("retarget_clips_host_csr takes host buffers; use run_csr() for device tensors")
        if human_host.dtype not in (torch.float32, torch.float64):
            human_host = human_host.double()
        human_host = human_host.contiguous()
        cs = np.ascontiguousarray(clip_start, np.int64)
        if cs.shape != (self.B + 1,) or cs[0] != 0 or (np.diff(cs) < 0).any():
            raise ValueError("clip_start must be B+1 non-decreasing int64 offsets starting at 0")
        N = int(cs[-1])
        if human_host.dim() != 3 or human_host.shape[0] != N or human_host.shape[1] != self.h_max or human_host.shape[2] != 7:
            raise ValueError(f"human stream must be [N={N},h={self.h_max},7], got {tuple(human_host.shape)}")
        if qpos_host is None:
            qpos_host = torch.zeros((N, self.nq_max), dtype=torch.float64, pin_memory=True)
        if status_host is None:
            status_host = torch.zeros((N,), dtype=torch.int32, pin_memory=True)
        if qpos_host.dtype != torch.float64 or tuple(qpos_host.shape) != (N, self.nq_max) or not qpos_host.is_contiguous():
            raise ValueError("qpos_host must be a contiguous float64 [N,nq_max] host tensor")
        if status_host.dtype != torch.int32 or tuple(status_host.shape) != (N,):
            raise ValueError("status_host must be an int32 [N] host tensor")
        _lib.check(_lib.lib().gmr_batch_run_host_csr(
            self._h, C.c_void_p(human_host.data_ptr()), F32 if human_host.dtype == torch.float32 else F64,
            cs.ctypes.data_as(C.c_void_p), int(bool(offset_to_ground)), int(bool(cold_start)),
            C.c_void_p(qpos_host.data_ptr()), C.c_void_p(status_host.data_ptr())), "gmr_batch_run_host_csr")
        ev = torch.cuda.Event()
        ev.record(self.stream)
        self._inflight = [x for x in getattr(self, "_inflight", []) if not x[0].query()]
        self._inflight.append((ev, human_host, qpos_host, status_host, cs))
        return qpos_host, status_host

    def step_host(self, human_host, qpos_host, scaled_host=None, offset_to_ground=False, sync=True):
        """One retarget() for every instance straight from / to pinned host tensors (gmr_batch_step_host): no staging
        copies, one launch; with sync the result is readable on return."""
        _lib.check(_lib.lib().gmr_batch_step_host(
            self._h, C.c_void_p(human_host.data_ptr()), F32 if human_host.dtype == torch.float32 else F64,
            int(bool(offset_to_ground)), C.c_void_p(qpos_host.data_ptr()),
            None if scaled_host is None else C.c_void_p(scaled_host.data_ptr()), int(bool(sync))), "gmr_batch_step_host")

    def step(self, human, offset_to_ground=False, diag=False, want_scaled=False):
        """One retarget() call for every instance: human [B,h,7] -> qpos [B,nq_max]."""
        h, dt = self._human(human, 3)
        qpos = torch.zeros((self.B, self.nq_max), dtype=torch.float64, device=self.tdev)
        scaled = torch.zeros((self.B, self.s_max, 7), dtype=torch.float64, device=self.tdev) if want_scaled else None
        d, dout = self._diag(1, diag)
        _lib.check(_lib.lib().gmr_batch_step(self._h, _ptr(h), dt, int(bool(offset_to_ground)), _ptr(qpos), _ptr(scaled),
                                             C.byref(d) if d is not None else None), "gmr_batch_step")
        res = [qpos]
        if want_scaled:
            res.append(scaled)
        if diag:
            res.append(dout)
        return res[0] if len(res) == 1 else tuple(res)

    # ---------------------------------------------------------------- stage entry points (parity tests)
    def target_prep(self, human, offset_to_ground=False):
        h, dt = self._human(human, 3)
        out = torch.zeros((self.B, self.s_max, 7), dtype=torch.float64, device=self.tdev)
        _lib.check(_lib.lib().gmr_target_prep(self._h, _ptr(h), dt, int(bool(offset_to_ground)), _ptr(out)), "gmr_target_prep")
        return out

    def fk(self, q=None):
        q = None if q is None else self._dev(q, torch.float64, (self.B, self.nq_max))
        xpos = torch.zeros((self.B, self.nbody_max, 3), dtype=torch.float64, device=self.tdev)
        xquat = torch.zeros((self.B, self.nbody_max, 4), dtype=torch.float64, device=self.tdev)
        _lib.check(_lib.lib().gmr_fk(self._h, _ptr(q), _ptr(xpos), _ptr(xquat)), "gmr_fk")
        return xpos, xquat

    def build_qp(self, stage, scaled):
        scaled = self._dev(scaled, torch.float64, (self.B, self.s_max, 7))
        n = self.nv_max
        e = torch.zeros((self.B, self.m_max, 6), dtype=torch.float64, device=self.tdev)
        H = torch.zeros((self.B, n, n), dtype=torch.float64, device=self.tdev)
        c, lo, hi = (torch.zeros((self.B, n), dtype=torch.float64, device=self.tdev) for _ in range(3))
        _lib.check(_lib.lib().gmr_build_qp(self._h, int(stage), _ptr(scaled), _ptr(e), _ptr(H), _ptr(c), _ptr(lo), _ptr(hi)),
                   "gmr_build_qp")
        return H, c, lo, hi, e


def _solve_ik(self, stage, scaled):
    """One solve_ik on the batch state -> (dq [B,nv_max], nfact [B]) (tree-structured QP of the fused path)."""
    scaled = self._dev(scaled, torch.float64, (self.B, self.s_max, 7))
    dq = torch.zeros((self.B, self.nv_max), dtype=torch.float64, device=self.tdev)
    nf = torch.zeros((self.B,), dtype=torch.int32, device=self.tdev)
    _lib.check(_lib.lib().gmr_solve_ik(self._h, int(stage), _ptr(scaled), _ptr(dq), _ptr(nf)), "gmr_solve_ik")
    return dq, nf


BatchedRetargeter.solve_ik = _solve_ik


def qp_solve(H, c, lo, hi, device: int = 0):
    """Batched exact box QP on the device: H [N,n,n], c/lo/hi [N,n] -> x [N,n], nfact [N]."""
    dev = torch.device("cuda", device)
    H = torch.as_tensor(H).to(dev, torch.float64).contiguous()
    c = torch.as_tensor(c).to(dev, torch.float64).contiguous()
    lo = torch.as_tensor(lo).to(dev, torch.float64).contiguous()
    hi = torch.as_tensor(hi).to(dev, torch.float64).contiguous()
    N, n = c.shape
    x = torch.zeros((N, n), dtype=torch.float64, device=dev)
    nf = torch.zeros((N,), dtype=torch.int32, device=dev)
    st = torch.cuda.current_stream(dev).cuda_stream
    _lib.check(_lib.lib().gmr_qp_solve(device, C.c_void_p(st), N, n, _ptr(H), _ptr(c), _ptr(lo), _ptr(hi), _ptr(x), _ptr(nf)),
               "gmr_qp_solve")
    return x, nf


def probe_fma_peak(device: int = 0, fp64: bool = True) -> float:
    """Measured CUDA-core FMA peak in TFLOP/s (compute roofline denominator)."""
    out = C.c_double(0.0)
    _lib.check(_lib.lib().gmr_probe_fma_peak(int(device), int(bool(fp64)), C.byref(out)), "gmr_probe_fma_peak")
    return out.value
