#!/usr/bin/env python
"""Drop-in for the reference's scripts/bvh_to_robot_dataset.py (same flags, same output pickles), with the
per-file / per-frame Python loop replaced by ONE batched launch over every clip of the folder.

  python scripts/bvh_to_robot_dataset.py --src_folder <bvh or npz folder> --tgt_folder <out> --robot unitree_g1 [--override]

Inputs: *.bvh (text parsed on the CPU by gmr_b200.bvh.parse_bvh, then Euler -> quaternion, sign continuity, FK,
axis change, FootMod rows and the height guess of utils/lafan1.py:8-71 on the GPU for all files at once:
gmr_bvh_to_stream; --reference_loader uses the reference's own Python loader instead, needs GMR_ROOT) or *.npz frame streams
{frames [T,h,7] (pos xyz, quat wxyz), names [h], height, fps} as written by tools/bvh_to_npz.py.
Outputs: <tgt_folder>/<relative path>.pkl with {root_pos, root_rot (xyzw), dof_pos, local_body_pos, fps,
link_body_list} (reference :144-156); existing outputs are skipped unless --override (:70-72); --pt also writes the
.pt form (scripts/convert_motion_pkl_to_pt.py).  The frame stream is float64, like the frames the reference feeds
to retarget() (--stream_f32 halves the bus traffic at ~1e-7 m in the targets).

Several GPUs:  torchrun --nproc-per-node N scripts/bvh_to_robot_dataset.py ...   - the file list is cut across the
ranks (balanced by size, gmr_b200.dataset.shard_files), rank r works on GPU LOCAL_RANK and writes its own pickles;
there is no communication between ranks (the reference: one process per file, smplx_to_robot_dataset.py:241-242).
A clip the engine flags as failed (non-finite input / output, QP search cap) is reported and not written - the
batch form of the reference's per-file try/except (:75-80).
"""
import argparse
import os
import sys
import types

import numpy as np

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
from gmr_b200 import bvh, dataset, params  # noqa: E402


def reference_bvh_loader():
    root = params.get_gmr_root()
    if root is None:
        raise SystemExit("*.bvh input needs the reference's BVH parser: export GMR_ROOT=<GMR checkout> "
                         "(or convert to *.npz frame streams first)")
    # import the loader sub-package without running the reference's package __init__ (which imports mink/mujoco)
    pkg = types.ModuleType("general_motion_retargeting")
    pkg.__path__ = [os.path.join(str(root), "general_motion_retargeting")]
    sys.modules.setdefault("general_motion_retargeting", pkg)
    from general_motion_retargeting.utils.lafan1 import load_lafan1_file
    return load_lafan1_file


def frames_to_array(frames, names=None):
    """list of {body: (pos, quat_wxyz)} -> ([T,h,7], names)"""
    if names is None:
        names = list(frames[0].keys())
    arr = np.zeros((len(frames), len(names), 7), np.float64)
    for t, fr in enumerate(frames):
        for k, n in enumerate(names):
            arr[t, k, :3], arr[t, k, 3:] = fr[n][0], fr[n][1]
    return arr, names


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--src_folder", required=True, type=str, help="Folder containing BVH motion files to load.")
    ap.add_argument("--tgt_folder", default="../../motion_data/LAFAN1_g1_gmr", help="Folder to save the retargeted motion files.")
    ap.add_argument("--robot", default="unitree_g1")
    ap.add_argument("--override", default=False, action="store_true")
    ap.add_argument("--target_fps", default=30, type=int)
    ap.add_argument("--device", default=None, type=int, help="CUDA device (default: LOCAL_RANK under torchrun, else 0)")
    ap.add_argument("--pt", default=False, action="store_true", help="also write <name>.pt (convert_motion_pkl_to_pt.py format)")
    ap.add_argument("--stream_f32", default=False, action="store_true", help="float32 frame stream instead of float64")
    ap.add_argument("--reference_loader", default=False, action="store_true", help="parse *.bvh with the reference's Python loader")
    args = ap.parse_args()

    rank, world, local = dataset.rank_info()
    if args.device is None:
        args.device = local if local is not None else 0
    sdt = np.float32 if args.stream_f32 else np.float64
    # the partition is made on the whole file list, before the skip-if-exists filter: ranks that start at different
    # moments (or a re-run after an interruption) must agree on who owns which file
    mine = dataset.shard_files(dataset.pending_files(args.src_folder, args.tgt_folder, (".bvh", ".npz"), True), rank, world)
    todo = [(s, t) for s, t in mine if args.override or not os.path.exists(t)]
    if not todo:
        print(f"[rank {rank}/{world}] Nothing to do.")
        return
    load_bvh = None
    clips, heights, names, fps, kept = [], [], None, [], []
    parsed = []                                          # (BvhClip, tgt): converted on the GPU in one go below
    for src, tgt in todo:
        try:
            if src.lower().endswith(".bvh") and not args.reference_loader:
                parsed.append((bvh.parse_bvh(src), tgt))
                continue
            if src.lower().endswith(".npz"):
                z = np.load(src, allow_pickle=False)
                arr, nm, hgt, f = z["frames"].astype(sdt), [str(x) for x in z["names"]], float(z["height"]), float(z["fps"])
            else:
                load_bvh = load_bvh or reference_bvh_loader()
                frames, hgt = load_bvh(src)
                arr, nm = frames_to_array(frames, names)
                f = float(args.target_fps)
        except Exception as e:                           # reference: print and continue (:75-80)
            print(f"Error loading {src}: {e}")
            continue
        if names is None:
            names = nm
        elif nm != names:
            print(f"Skipping {src}: body list differs from the first file's")
            continue
        clips.append(arr); heights.append(hgt); fps.append(f); kept.append(tgt)
    # *.bvh files: one device pass per skeleton (files of one dataset share it)
    while parsed:
        c0 = parsed[0][0]
        same = [(c, t) for c, t in parsed if list(c.names) == list(c0.names) and np.array_equal(c.parents, c0.parents) and c.order == c0.order]
        parsed = [(c, t) for c, t in parsed if not any(c is c2 for c2, _ in same)]
        streams, rows, hs = bvh.load_bvh_clips([c for c, _ in same], device=args.device, dtype=sdt)
        if names is None:
            names = rows
        elif rows != names:
            print(f"Skipping {len(same)} file(s): body list differs from the first file's")
            continue
        for (c, tgt), s, hgt in zip(same, streams, hs):
            clips.append(s.cpu().numpy()); heights.append(float(hgt)); fps.append(float(args.target_fps)); kept.append(tgt)
    motions, status = dataset.retarget_clips("bvh", args.robot, clips, heights=heights, human_names=names, fps=fps,
                                             height_adjust=False, root_origin_offset=False, device=args.device,
                                             stream_dtype=sdt, return_status=True)
    saved = 0
    for tgt, m, st in zip(kept, motions, status):
        if m is None:
            print(f"Error retargeting {tgt}: {dataset.describe_status(int(st))} - skipped")
            continue
        dataset.save_motion(tgt, m)
        if args.pt:
            dataset.save_motion_pt(os.path.splitext(tgt)[0] + ".pt", m)
        saved += 1
    print(f"[rank {rank}/{world}] Done. {saved} file(s) saved to ", args.tgt_folder)


if __name__ == "__main__":
    main()
