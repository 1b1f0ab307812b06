#!/usr/bin/env python
"""Drop-in for the reference's scripts/smplx_to_robot_dataset.py (same flags, same output pickles), with the
mp.Pool of per-file / per-frame Python loops replaced by one device pass over every clip of the folder:
fps alignment + orientation chain (utils/smpl.py:109-198 -> gmr_smplx_to_stream), batched retargeting
(gmr_batch_run_host) and the driver tail (FK tail, HEIGHT_ADJUST, ROOT_ORIGIN_OFFSET, :93-146).

  python scripts/smplx_to_robot_dataset.py --src_folder <npz folder> --tgt_folder <out> --robot unitree_g1 [--override]

Inputs: *.npz with the SMPL-X body-model OUTPUT of each clip - ``full_pose`` [T,165] (or [T,55,3]) rotation
vectors, ``joints`` [T,>=55,3], ``mocap_frame_rate``, ``betas`` - i.e. what ``load_smplx_file`` (utils/smpl.py:11-43)
computes with the licensed SMPL-X model files, which this repository cannot ship or run.  ``*_stagei.npz`` files are
skipped as in the reference (:213-214); existing outputs are skipped unless --override.
"""
import argparse
import os
import sys

import numpy as np

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
from gmr_b200 import dataset, smplx_prep  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--robot", default="unitree_g1")
    ap.add_argument("--src_folder", type=str, required=True)
    ap.add_argument("--tgt_folder", type=str, required=True)
    ap.add_argument("--override", default=False, action="store_true")
    ap.add_argument("--num_cpus", default=4, type=int, help="accepted for compatibility; the work runs on the GPU")
    ap.add_argument("--device", default=0, type=int)
    args = ap.parse_args()

    todo = [(s, t) for s, t in dataset.pending_files(args.src_folder, args.tgt_folder, (".npz",), args.override)
            if not s.endswith("_stagei.npz")]
    if not todo:
        print("Nothing to do.")
        return
    clips, heights, kept = [], [], []
    for src, tgt in todo:
        try:
            z = np.load(src, allow_pickle=True)
            betas = np.asarray(z["betas"])
            clips.append({"full_pose": np.asarray(z["full_pose"], np.float32), "joints": np.asarray(z["joints"], np.float32),
                          "fps": float(np.asarray(z["mocap_frame_rate"]).item())})
            heights.append(1.66 + 0.1 * float(betas[0] if betas.ndim == 1 else betas[0, 0]))       # utils/smpl.py:38-41
            kept.append(tgt)
        except Exception as e:                                   # reference: print and carry on (:66-68)
            print(f"Error loading {src}: {e}")
    streams, fps = smplx_prep.smplx_to_stream(clips, tgt_fps=30, device=args.device, dtype=np.float32)
    keep = [i for i, s in enumerate(streams) if s.shape[0] > 0]
    motions = dataset.retarget_clips("smplx", args.robot, [streams[i].cpu().numpy() for i in keep],
                                     heights=[heights[i] for i in keep], human_names=list(smplx_prep.SMPLX_JOINT_NAMES),
                                     fps=[fps[i] for i in keep], height_adjust=True, root_origin_offset=True, device=args.device)
    for i, m in zip(keep, motions):
        dataset.save_motion(kept[i], m)
    print(f"Done. {len(motions)} file(s) saved to {args.tgt_folder}")


if __name__ == "__main__":
    main()
