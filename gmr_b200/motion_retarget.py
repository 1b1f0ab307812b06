"""Drop-in for the reference class ``GeneralMotionRetargeting``
(general_motion_retargeting/motion_retarget.py:10-270): same constructor, same ``retarget`` /
``update_targets`` / ``error1`` / ``error2`` / ``scale_human_data`` / ``offset_human_data``
methods, same attributes the reference's scripts touch - but every number is
produced by the CUDA engine through the C ABI (a batch of one instance).  There is no CPU
fallback: without a CUDA device or without libgmr_b200.so construction raises.

Behaviour kept on purpose: table-1 offsets feed both stages; the scale table
and ``pos_offsets1`` are re-read every frame, so mutating them after construction takes effect
(scripts/fbx_to_robot.py:1003-1041), while mutating ``ik_match_table*[..][1]`` weights does not;
the caller's frame dict is overwritten with ndarrays (:203-206); the state persists across calls
(warm start) and there is no reset - construct a new object per clip.
"""
from __future__ import annotations

import copy

import numpy as np
import torch

from . import model_pack, params
from .abi import ModelDesc
from .engine import BatchedRetargeter, DeviceModel
from .ik_config import build_retarget_tables


class _Task:
    """The slice of mink.FrameTask the reference's scripts read (frame_name, costs, compute_error)."""

    def __init__(self, owner, stage, index, frame_name, pos_cost, rot_cost):
        self._owner, self._stage, self._index = owner, stage, index
        self.frame_name, self.frame_type = frame_name, "body"
        self.position_cost, self.orientation_cost, self.lm_damping = pos_cost, rot_cost, 1

    def compute_error(self, configuration=None):
        """FrameTask.compute_error: e = log(T_body^-1 T_target) as [v(3), omega(3)]."""
        return self._owner._task_errors(self._stage)[self._index]


class _Body:
    def __init__(self, i, name):
        self.id, self.name = i, name


class _Opt:
    def __init__(self, timestep):
        self.timestep = timestep


class _Model:
    """The slice of mj.MjModel the scripts read: nq, nv, nbody, body(name).id, opt.timestep."""

    def __init__(self, rm):
        self._rm = rm
        self.nq, self.nv, self.nbody, self.njnt = rm.nq, rm.nv, rm.nbody, rm.njnt
        self.opt = _Opt(rm.timestep)
        self.qpos0 = rm.qpos0.copy()

    def body(self, key):
        if isinstance(key, str):
            return _Body(self._rm.body_id(key), key)
        return _Body(int(key), self._rm.body_names[int(key)])


class _Data:
    def __init__(self, owner):
        self._owner = owner

    @property
    def qpos(self):
        return self._owner._qpos()

    @property
    def xpos(self):
        return self._owner._fk()[0]

    @property
    def xquat(self):
        return self._owner._fk()[1]


class _Configuration:
    def __init__(self, owner, model):
        self.model, self.data = model, _Data(owner)

    @property
    def q(self):
        return self.data.qpos


class GeneralMotionRetargeting:
    """General Motion Retargeting (GMR) on the B200 engine."""

    def __init__(self, src_human: str, tgt_robot: str, actual_human_height: float = None, solver: str = "daqp",
                 damping: float = 5e-1, verbose: bool = False, device: int = 0) -> None:
        if not torch.cuda.is_available():
            raise RuntimeError("gmr_b200 needs a CUDA device (no CPU fallback)")
        self.xml_file = str(params.ROBOT_XML_DICT[tgt_robot])           # KeyError like the reference (:24)
        self._robot, ik_config = model_pack.load_pair(src_human, tgt_robot)
        self._src, self._tgt, self._device = src_human, tgt_robot, int(device)
        if verbose:
            print("Use robot model: ", self.xml_file)
            print("Use IK config: ", params.IK_CONFIG_DICT[src_human][tgt_robot])
        self._base_config = copy.deepcopy(ik_config)
        ratio = actual_human_height / ik_config["human_height_assumption"] if actual_human_height is not None else 1.0
        for key in ik_config["human_scale_table"].keys():
            ik_config["human_scale_table"][key] = ik_config["human_scale_table"][key] * ratio
        self.ik_match_table1 = ik_config["ik_match_table1"]
        self.ik_match_table2 = ik_config["ik_match_table2"]
        self.human_root_name = ik_config["human_root_name"]
        self.robot_root_name = ik_config["robot_root_name"]
        self.use_ik_match_table1 = ik_config["use_ik_match_table1"]
        self.use_ik_match_table2 = ik_config["use_ik_match_table2"]
        self.human_scale_table = ik_config["human_scale_table"]
        self.ground = ik_config["ground_height"] * np.array([0, 0, 1])
        self.max_iter = 10
        self.solver = solver        # any exact QP solver returns the same unique minimiser; kept for the signature
        self.damping = damping
        self.model = _Model(self._robot)
        self.setup_retarget_configuration()

    # ------------------------------------------------------------------ construction
    def setup_retarget_configuration(self):
        from scipy.spatial.transform import Rotation as R
        tables = build_retarget_tables(self._robot, self._base_config, self._src, self._tgt)
        self._tables = tables
        self.pos_offsets1, self.rot_offsets1, self.pos_offsets2, self.rot_offsets2 = {}, {}, {}, {}
        self.human_body_to_task1, self.human_body_to_task2 = {}, {}
        self.tasks1, self.tasks2 = [], []
        self.task_errors1, self.task_errors2 = {}, {}
        for k, (table, tasks, by_body, pos_o, rot_o, errs) in enumerate((
                (self.ik_match_table1, self.tasks1, self.human_body_to_task1, self.pos_offsets1, self.rot_offsets1, self.task_errors1),
                (self.ik_match_table2, self.tasks2, self.human_body_to_task2, self.pos_offsets2, self.rot_offsets2, self.task_errors2))):
            for frame_name, entry in table.items():
                body_name, pos_weight, rot_weight, pos_offset, rot_offset = entry
                if pos_weight != 0 or rot_weight != 0:
                    task = _Task(self, k, len(tasks), frame_name, pos_weight, rot_weight)
                    by_body[body_name] = task
                    pos_o[body_name] = np.array(pos_offset) - self.ground
                    rot_o[body_name] = R.from_quat(rot_offset, scalar_first=True)
                    tasks.append(task)
                    errs[task] = []
        self._names = list(tables.scaled_names)       # frame layout of this object: the bodies the config reads
        self._build_engine()
        self.configuration = _Configuration(self, self.model)
        self.scaled_human_data = None
        self._last_scaled = None

    def _offsets_key(self):
        return tuple(tuple(np.asarray(self.pos_offsets1[b], np.float64).tolist()) for b in self._names)

    def _build_engine(self, keep_state=None):
        t = self._tables
        for i, b in enumerate(self._names):           # pos_offsets1 may have been edited by the caller
            t.pos_off[i] = np.asarray(self.pos_offsets1[b], np.float64)
        self._dm = DeviceModel(ModelDesc(self._robot, t, self._names, damping=self.damping, max_iter=self.max_iter),
                               self._device)
        self._br = BatchedRetargeter(self._dm, 1)
        if keep_state is not None:
            self._br.reset(keep_state)
        self._key = self._offsets_key()
        self._scale_key = None
        # page-locked buffers of the per-frame call, allocated once: the kernel reads the frame from, and writes
        # qpos and the scaled targets to, host memory directly (gmr_batch_step_host: one launch, one sync per frame)
        s, nq = len(self._names), self._br.nq_max
        self._frame = torch.zeros((1, s, 7), dtype=torch.float64, pin_memory=True)
        self._out = torch.zeros((nq + self._br.s_max * 7,), dtype=torch.float64, pin_memory=True)
        self._frame_np, self._out_np = self._frame.numpy(), self._out.numpy()
        self._qpos_view = self._out[:nq].view(1, nq)
        self._scaled_view = self._out[nq:].view(1, self._br.s_max, 7)

    def _sync_tables(self):
        if self._offsets_key() != self._key:
            self._build_engine(keep_state=self._br.state())
        key = tuple(self.human_scale_table[b] for b in self._names)     # re-read every frame like the reference (:120)
        if key != self._scale_key:
            self._br.set_scale(np.array([key], np.float64))
            self._scale_key = key

    # ------------------------------------------------------------------ hot path
    def _pack(self, human_data):
        frame = self._frame_np
        for i, b in enumerate(self._names):
            pos, quat = human_data[b]                 # KeyError like the reference (:213 / :129)
            frame[0, i, :3] = pos
            frame[0, i, 3:] = quat
        return self._frame

    def _unpack_scaled(self, scaled):
        out = {}
        for i, b in enumerate(self._names):
            out[b] = [scaled[i, :3].copy(), scaled[i, 3:].copy()]
        return out

    def update_targets(self, human_data, offset_to_ground=False):
        human_data = self.to_numpy(human_data)
        self._sync_tables()
        scaled = self._br.target_prep(self._pack(human_data), offset_to_ground)
        self._last_scaled = scaled
        self.scaled_human_data = self._unpack_scaled(scaled[0].cpu().numpy())

    def retarget(self, human_data, offset_to_ground=False):
        human_data = self.to_numpy(human_data)
        self._sync_tables()
        self._br.step_host(self._pack(human_data), self._qpos_view, self._scaled_view, offset_to_ground, sync=True)
        nq = self.model.nq
        self._last_scaled = self._scaled_view.clone()
        self.scaled_human_data = self._unpack_scaled(self._out_np[self._br.nq_max:].reshape(-1, 7))
        return self._out_np[:nq].copy()

    # ------------------------------------------------------------------ diagnostics
    def _qpos(self):
        return self._br.state()[0, :self.model.nq].cpu().numpy()

    def _fk(self):
        xpos, xquat = self._br.fk()
        return xpos[0, :self.model.nbody].cpu().numpy(), xquat[0, :self.model.nbody].cpu().numpy()

    def _task_errors(self, stage):
        if self._last_scaled is None:
            raise RuntimeError("target not set: call update_targets()/retarget() first")   # mink.TargetNotSet
        e = self._br.build_qp(stage, self._last_scaled)[4]
        n = len(self.tasks1 if stage == 0 else self.tasks2)
        return e[0, :n].cpu().numpy()

    def error1(self):
        return float(np.linalg.norm(self._task_errors(0).ravel()))

    def error2(self):
        return float(np.linalg.norm(self._task_errors(1).ravel()))

    # ------------------------------------------------------------------ helpers with reference semantics
    def to_numpy(self, human_data):
        for body_name in human_data.keys():
            human_data[body_name] = [np.asarray(human_data[body_name][0]), np.asarray(human_data[body_name][1])]
        return human_data

    def scale_human_data(self, human_data, human_root_name, human_scale_table):
        """motion_retarget.py:209-232, evaluated by Kernel 1 with zero offsets for the bodies given."""
        names = [human_root_name] + [b for b in human_data if b in human_scale_table and b != human_root_name]
        out = self._prep_on_device(human_data, names, [human_scale_table[b] for b in names], None, None)
        return {b: (out[b][0], human_data[b][1]) for b in names}      # rotations pass through untouched (:228-230)

    def offset_human_data(self, human_data, pos_offsets, rot_offsets):
        """motion_retarget.py:234-250 (scale 1 about the first body = identity scaling)."""
        names = list(human_data.keys())
        return self._prep_on_device(human_data, names, [1.0] * len(names), pos_offsets, rot_offsets)

    def _prep_on_device(self, human_data, names, scale, pos_offsets, rot_offsets):
        from .ik_config import RetargetTables
        s = len(names)
        pos_off, rot_off = np.zeros((s, 3)), np.tile([1.0, 0, 0, 0], (s, 1))
        if pos_offsets is not None:
            for i, b in enumerate(names):
                pos_off[i] = pos_offsets[b]
                rot_off[i] = rot_offsets[b].as_quat(scalar_first=True)
        t = self._tables
        key = (tuple(names), pos_off.tobytes(), rot_off.tobytes())
        cache = self.__dict__.setdefault("_prep_cache", {})
        if key not in cache:
            if len(cache) >= 4:                            # a handful of layouts at most (fbx_to_robot.py:1003-1041)
                _, (odm, obr) = cache.popitem()
                obr.close(); odm.close()
            tt = RetargetTables(t.src_human, t.tgt_robot, names[0], t.robot_root_name, 0.0, t.human_height_assumption,
                                (False, False), names, np.ones(s), pos_off, rot_off,
                                np.zeros(s, np.int32), [np.zeros(0, np.int32)] * 2, [np.zeros(0, np.int32)] * 2,
                                [np.zeros(0)] * 2, [np.zeros(0)] * 2, [[], []], {})
            dm = DeviceModel(ModelDesc(self._robot, tt, names), self._device)
            cache[key] = (dm, BatchedRetargeter(dm, 1))
        dm, br = cache[key]
        br.set_scale(np.asarray(scale, np.float64)[None])
        frame = np.zeros((1, s, 7))
        for i, b in enumerate(names):
            frame[0, i, :3], frame[0, i, 3:] = human_data[b][0], human_data[b][1]
        out = br.target_prep(frame, False)[0].cpu().numpy()
        return {b: [out[i, :3].copy(), out[i, 3:].copy()] for i, b in enumerate(names)}
