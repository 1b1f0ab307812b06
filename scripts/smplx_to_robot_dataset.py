#!/usr/bin/env python
"""Drop-in for the reference's scripts/smplx_to_robot_dataset.py (same flags, same file selection, same output
pickles), with the mp.Pool of per-file / per-frame Python loops replaced by one device pass over every clip:
fps alignment + orientation chain (utils/smpl.py:109-198 -> gmr_smplx_to_stream), batched retargeting
(gmr_batch_run_host_csr) and the driver tail (FK tail, HEIGHT_ADJUST, ROOT_ORIGIN_OFFSET, :93-146).

  python scripts/smplx_to_robot_dataset.py --src_folder <folder> --tgt_folder <out> --robot unitree_g1 [--override]
  torchrun --nproc-per-node N scripts/smplx_to_robot_dataset.py ...      (file list cut across N GPUs, no communication)

Inputs: *.npz / *.pkl (np.load(allow_pickle=True), as in the reference) with the SMPL-X body-model OUTPUT of each
clip - ``full_pose`` [T,165] (or [T,55,3]) rotation vectors, ``joints`` [T,>=55,3], ``mocap_frame_rate``, ``betas`` -
i.e. what ``load_smplx_file`` (utils/smpl.py:11-43) computes with the licensed SMPL-X model files, which this
repository cannot ship or run.
File selection = the reference's (:193-234): natural sort inside each directory, ``*_stagei.npz`` skipped, existing
outputs skipped unless --override, the hard-motion lists (assets/hard_motions/{0,1}.txt of the GMR checkout when
GMR_ROOT is set, else the copy bundled in gmr_b200/data/hard_motions.txt) and names containing BMLrub / EKUT / crawl /
_lie / upstairs / downstairs excluded.
"""
import argparse
import os
import re
import sys

import numpy as np

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
from gmr_b200 import dataset, params, smplx_prep  # noqa: E402


EXCLUDE_FILE_CONTENT = ["BMLrub", "EKUT", "crawl", "_lie", "upstairs", "downstairs"]        # reference :222


def natural_key(name: str):
    """natsorted()'s default order: digit runs compare as numbers."""
    return [int(p) if p.isdigit() else p for p in re.split(r"(\d+)", name)]


def hard_motions() -> set:
    """Motion names of the reference's exclusion lists (:193-208: the `Motion:` lines of assets/hard_motions/{0,1}.txt,
    up to the first comma, extension dropped)."""
    root = params.get_gmr_root()
    names = set()
    if root is not None and (root / "assets" / "hard_motions").is_dir():
        for fn in ("0.txt", "1.txt"):
            with open(root / "assets" / "hard_motions" / fn) as f:
                for line in f:
                    if "Motion:" in line:
                        names.add(line.split(":")[1].strip().split(",")[0].strip().split(".")[0])
        return names
    bundled = os.path.join(os.path.dirname(os.path.abspath(dataset.__file__)), "data", "hard_motions.txt")
    if os.path.exists(bundled):
        with open(bundled) as f:
            names = {l.strip() for l in f if l.strip()}
    return names


def select_files(src_folder, tgt_folder, override, quiet=False):
    """The reference's walk (:210-234) -> [(src, tgt)]."""
    say = (lambda *a: None) if quiet else print
    out = []
    for dirpath, _, filenames in os.walk(src_folder):
        for filename in sorted(filenames, key=natural_key):
            if filename.endswith("_stagei.npz"):
                continue
            if filename.endswith((".pkl", ".npz")):
                src = os.path.join(dirpath, filename)
                tgt = src.replace(src_folder, tgt_folder).replace(".npz", ".pkl")
                if not os.path.exists(tgt) or override:
                    out.append((src, tgt))
    say("full args_list:", len(out))
    hard = hard_motions()
    keep = []
    for src, tgt in out:
        motion_name = src.split("/")[-1].split(".")[0]
        if motion_name in hard or any(c in motion_name for c in EXCLUDE_FILE_CONTENT):
            continue
        keep.append((src, tgt))
    say("new args_list:", len(keep))
    return keep


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--robot", default="unitree_g1")
    ap.add_argument("--src_folder", type=str, required=True)
    ap.add_argument("--tgt_folder", type=str, required=True)
    ap.add_argument("--override", default=False, action="store_true")
    ap.add_argument("--num_cpus", default=4, type=int, help="accepted for compatibility; the work runs on the GPU(s)")
    ap.add_argument("--device", default=None, type=int, help="CUDA device (default: LOCAL_RANK under torchrun, else 0)")
    ap.add_argument("--pt", default=False, action="store_true", help="also write <name>.pt (convert_motion_pkl_to_pt.py format)")
    args = ap.parse_args()

    rank, world, local = dataset.rank_info()
    if args.device is None:
        args.device = local if local is not None else 0
    if rank == 0:
        select_files(args.src_folder, args.tgt_folder, args.override)           # the reference's two counts
    # the partition is made on the whole selection, before the skip-if-exists filter, so that ranks starting at
    # different moments (or a re-run) agree on who owns which file
    mine = dataset.shard_files(select_files(args.src_folder, args.tgt_folder, True, quiet=True), rank, world)
    todo = [(s, t) for s, t in mine if args.override or not os.path.exists(t)]
    print(f"[rank {rank}/{world}] Total number of files to process: {len(todo)}")
    if not todo:
        return
    clips, heights, kept = [], [], []
    for src, tgt in todo:
        try:
            z = np.load(src, allow_pickle=True)
            if isinstance(z, np.ndarray) and z.dtype == object:
                z = z.item()
            betas = np.asarray(z["betas"])
            clips.append({"full_pose": np.asarray(z["full_pose"], np.float32), "joints": np.asarray(z["joints"], np.float32),
                          "fps": float(np.asarray(z["mocap_frame_rate"]).item())})
            heights.append(1.66 + 0.1 * float(betas[0] if betas.ndim == 1 else betas[0, 0]))       # utils/smpl.py:38-41
            kept.append(tgt)
        except Exception as e:                                   # reference: print and carry on (:66-68)
            print(f"Error loading {src}: {e}")
    streams, fps = smplx_prep.smplx_to_stream(clips, tgt_fps=30, device=args.device, dtype=np.float64)
    keep = [i for i, s in enumerate(streams) if s.shape[0] > 0]
    motions, status = dataset.retarget_clips("smplx", args.robot, [streams[i].cpu().numpy() for i in keep],
                                             heights=[heights[i] for i in keep], human_names=list(smplx_prep.SMPLX_JOINT_NAMES),
                                             fps=[fps[i] for i in keep], height_adjust=True, root_origin_offset=True,
                                             device=args.device, return_status=True)
    saved = 0
    for i, m, st in zip(keep, motions, status):
        if m is None:
            print(f"Error processing {kept[i]}: {dataset.describe_status(int(st))} - skipped")
            continue
        dataset.save_motion(kept[i], m)
        if args.pt:
            dataset.save_motion_pt(os.path.splitext(kept[i])[0] + ".pt", m)
        saved += 1
    print(f"[rank {rank}/{world}] Done. {saved} file(s) saved to {args.tgt_folder}")


if __name__ == "__main__":
    main()
