"""BVH -> human frame stream: the step right before the retargeter in the reference's BVH drivers
(SURVEY.md 8f rank 1).

Reference: ``load_lafan1_file`` (general_motion_retargeting/utils/lafan1.py:8-71) =
``read_bvh`` (utils/lafan_vendor/extract.py:43-166: text -> local positions + Euler angles -> quaternions,
sign continuity) -> ``quat_fk`` (utils/lafan_vendor/utils.py:83-103) -> per frame / per bone Python loop that
turns Y-up centimetres into Z-up metres (lafan1.py:20-33), synthesises ``Left/RightFootMod`` (:35-39) and
guesses the human height from the last frame (:42-69).

Here the text parse stays on the CPU (``parse_bvh``, one pass over the tokens); everything after it runs on the
GPU for all frames of all clips of a folder at once (``gmr_bvh_to_stream``: ``csrc/gmr_bvh.cuh``) and writes the
``[frames, h, 7]`` stream (pos xyz, quat wxyz) the retargeter reads - no per-frame Python objects.
"""
from __future__ import annotations

import ctypes as C
import re
from dataclasses import dataclass

import numpy as np

_AXIS = {"Xrotation": 0, "Yrotation": 1, "Zrotation": 2}
_WORD = re.compile(r"\w+")
FOOT_MODS = (("LeftFootMod", "LeftFoot", "LeftToe"), ("RightFootMod", "RightFoot", "RightToe"))   # lafan1.py:35-39


@dataclass
class BvhClip:
    """What extract.py:read_bvh leaves before the Euler -> quaternion conversion."""
    names: list            # J bone names, file order (End Sites dropped)
    parents: np.ndarray    # [J] int32, -1 for the root
    offsets: np.ndarray    # [J,3] float64
    positions: np.ndarray  # [T,J,3] float64 local positions (offsets where the file has no position channels)
    eulers: np.ndarray     # [T,J,3] float64 degrees, in channel order
    order: tuple           # axis (0/1/2) of each of the three rotation channels
    frame_time: float


def parse_bvh(path_or_text, is_text: bool = False) -> BvhClip:
    """Same reading of a BVH file as extract.py:43-166 (start/end unused by the loader):
    * bones in order of appearance, parent = enclosing bone, ``End Site`` blocks contribute nothing (:87-95,124-126);
    * a bone's name is the leading ``\\w+`` of what follows ROOT/JOINT (:74,115);
    * the rotation order comes from the first CHANNELS line whose rotation triple is all rotations (:103-112);
    * the channel count that shapes the motion rows is the LAST one declared (:101, used :149-160):
      3 -> only the root carries a position (LAFAN1), 6 -> every bone does;
    * positions default to the offsets (:134)."""
    text = path_or_text if is_text else open(path_or_text, "r").read()
    lines = text.split("\n")
    names, parents, offsets = [], [], []
    active, end_site, order, channels = -1, False, None, None
    n_frames, frame_time = None, None
    li = 0
    while li < len(lines):
        line = lines[li]
        li += 1
        s = line.strip()
        if not s or s.startswith("HIERARCHY"):
            continue
        if s.startswith("MOTION"):
            continue
        head = s.split(None, 1)
        key = head[0]
        if key in ("ROOT", "JOINT"):
            m = _WORD.match(head[1]) if len(head) > 1 else None
            if m is None:
                continue
            names.append(m.group(0))
            offsets.append([0.0, 0.0, 0.0])
            parents.append(active)
            active = len(parents) - 1
        elif key == "{":
            continue
        elif key == "}":
            if end_site:
                end_site = False
            else:
                active = parents[active]
        elif key == "OFFSET":
            if not end_site:
                offsets[active] = [float(v) for v in head[1].split()[:3]]
        elif key == "CHANNELS":
            tok = head[1].split()
            channels = int(tok[0])
            if order is None:
                first = 0 if channels == 3 else 3
                parts = tok[1 + first:1 + first + 3]
                if len(parts) == 3 and all(p in _AXIS for p in parts):
                    order = tuple(_AXIS[p] for p in parts)
        elif key == "End" and s.startswith("End Site"):
            end_site = True
        elif key == "Frames:":
            n_frames = int(head[1])
        elif key == "Frame" and s.startswith("Frame Time:"):
            frame_time = float(s.split(":", 1)[1])
            break
    if n_frames is None or order is None or channels is None:
        raise ValueError("not a BVH file: missing Frames / CHANNELS")
    J = len(names)
    offsets = np.asarray(offsets, np.float64).reshape(J, 3)
    rows = [r for r in (l.strip() for l in lines[li:]) if r]
    data = np.array([np.array(r.split(), dtype=np.float64) for r in rows[:n_frames]]) if n_frames else np.zeros((0, 0))
    if data.shape[0] != n_frames:
        raise ValueError(f"BVH declares {n_frames} frames, found {data.shape[0]} motion rows")
    positions = np.repeat(offsets[None], n_frames, axis=0)
    if channels == 3:
        positions[:, 0] = data[:, 0:3]
        eulers = data[:, 3:].reshape(n_frames, J, 3)
    elif channels == 6:
        d = data.reshape(n_frames, J, 6)
        positions = d[:, :, 0:3].copy()
        eulers = d[:, :, 3:6].copy()
    else:
        raise ValueError(f"unsupported channel layout ({channels} channels per bone)")
    return BvhClip(names, np.asarray(parents, np.int32), offsets, positions, np.ascontiguousarray(eulers), order,
                   1.0 / 30.0 if frame_time is None else frame_time)


def stream_layout(names):
    """Rows of the frame stream = the keys of the reference's per-frame dict, in its insertion order
    (lafan1.py:29-39): every bone, then LeftFootMod / RightFootMod when their sources exist.
    -> (row names, src_pos[h], src_rot[h]) with src_* = bone index whose position / orientation the row takes."""
    idx = {}
    for i, n in enumerate(names):
        idx[n] = i                      # a repeated name keeps its first slot and takes the last bone's values
    rows = list(idx.keys())
    src_pos = [idx[n] for n in rows]
    src_rot = list(src_pos)
    for mod, foot, toe in FOOT_MODS:
        if foot in idx and toe in idx:
            rows.append(mod)
            src_pos.append(idx[foot])
            src_rot.append(idx[toe])
    return rows, np.asarray(src_pos, np.int32), np.asarray(src_rot, np.int32)


def height_rows(rows):
    """Row indices the height heuristic looks at (lafan1.py:42-61): (head, [foot candidates])."""
    pos = {n: i for i, n in enumerate(rows)}
    head = pos.get("Head", -1)
    feet = [pos[k] for k in ("LeftFootMod", "RightFootMod", "LeftFoot", "RightFoot") if k in pos]
    return head, feet


def load_bvh_clips(clips, device: int = 0, dtype=np.float32, stream=None):
    """clips: list of BvhClip sharing one skeleton.  -> (streams: list of [T_i,h,7] torch CUDA tensors (views of
    one buffer), row names, heights: float64 numpy [C]).  One launch sequence for all frames of all clips."""
    import torch

    from . import _lib
    from .abi import F32, F64
    if not clips:
        return [], [], np.zeros(0)
    c0 = clips[0]
    J = len(c0.names)
    for c in clips[1:]:
        if list(c.names) != list(c0.names) or not np.array_equal(c.parents, c0.parents) or tuple(c.order) != tuple(c0.order):
            raise ValueError("clips of one batch must share the skeleton and the rotation order")
    rows, src_pos, src_rot = stream_layout(c0.names)
    head, feet = height_rows(rows)
    lengths = np.array([c.positions.shape[0] for c in clips], np.int64)
    start = np.concatenate([[0], np.cumsum(lengths)]).astype(np.int32)
    N, h, Cn = int(start[-1]), len(rows), len(clips)
    tdev = torch.device("cuda", device)
    lpos = torch.from_numpy(np.concatenate([c.positions for c in clips], axis=0).reshape(N, J, 3)).to(tdev)
    eul = torch.from_numpy(np.concatenate([c.eulers for c in clips], axis=0).reshape(N, J, 3)).to(tdev)
    tdt = torch.float32 if np.dtype(dtype) == np.float32 else torch.float64
    out = torch.empty((N, h, 7), dtype=tdt, device=tdev)
    heights = torch.empty((Cn,), dtype=torch.float64, device=tdev)
    i32 = lambda a: torch.from_numpy(np.ascontiguousarray(a, np.int32)).to(tdev)      # noqa: E731
    d_start, d_par, d_sp, d_sr = i32(start), i32(c0.parents), i32(src_pos), i32(src_rot)
    feet4 = (list(feet) + [-1] * 4)[:4]
    ptr = lambda t: C.c_void_p(t.data_ptr())                                           # noqa: E731
    s = stream if stream is not None else torch.cuda.current_stream(tdev).cuda_stream
    _lib.check(_lib.lib().gmr_bvh_to_stream(
        device, C.c_void_p(s), ptr(lpos), ptr(eul), ptr(d_start), Cn, N, J, ptr(d_par),
        (C.c_int32 * 3)(*[int(a) for a in c0.order]), ptr(d_sp), ptr(d_sr), h,
        F32 if tdt == torch.float32 else F64, ptr(out), int(head), (C.c_int32 * 4)(*feet4), ptr(heights)),
        "gmr_bvh_to_stream")
    hs = heights.cpu().numpy()
    return [out[int(start[i]):int(start[i + 1])] for i in range(Cn)], rows, hs


def load_lafan1_file(bvh_file, device: int = 0):
    """Drop-in for utils/lafan1.py:8-71: -> (frames: list of {bone: (pos[3], quat_wxyz[4])}, human_height)."""
    streams, rows, hs = load_bvh_clips([parse_bvh(bvh_file)], device=device, dtype=np.float64)
    arr = streams[0].cpu().numpy()
    frames = [{n: (arr[t, k, :3].copy(), arr[t, k, 3:].copy()) for k, n in enumerate(rows)} for t in range(arr.shape[0])]
    return frames, float(hs[0]) if len(frames) else 1.75
