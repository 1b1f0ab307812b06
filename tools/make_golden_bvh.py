#!/usr/bin/env python
"""Generate tests/golden/bvh_*.npz by running the REFERENCE's own BVH loader (this container only).

Writes small synthetic BVH files (LAFAN1-shaped skeleton and two odd ones), runs
``general_motion_retargeting.utils.lafan1.load_lafan1_file`` (utils/lafan1.py:8-71, numpy + scipy only) and
``lafan_vendor.extract.read_bvh`` on them unmodified, and stores the BVH text together with what the reference
returned: local quaternions / positions after read_bvh, the per-frame dicts as a [T,h,7] array, the height.

    python tools/make_golden_bvh.py [--gmr-root /root/reference]
"""
import argparse
import os
import sys
import tempfile
import types

import numpy as np

REPO = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..")
sys.path.insert(0, REPO)

# (name, parent, offset cm) - LAFAN1 bone names and topology, offsets of a ~1.7 m person, Y up
LAFAN = [
    ("Hips", -1, (0, 0, 0)),
    ("LeftUpLeg", 0, (10.3, -1.8, 0.5)), ("LeftLeg", 1, (0, -43.5, 0)), ("LeftFoot", 2, (0, -41.9, 0)), ("LeftToe", 3, (0, -7.5, 14.0)),
    ("RightUpLeg", 0, (-10.3, -1.8, 0.5)), ("RightLeg", 5, (0, -43.5, 0)), ("RightFoot", 6, (0, -41.9, 0)), ("RightToe", 7, (0, -7.5, 14.0)),
    ("Spine", 0, (0, 6.9, -1.6)), ("Spine1", 9, (0, 12.6, 0)), ("Spine2", 10, (0, 12.3, 0)), ("Neck", 11, (0, 25.8, 0)), ("Head", 12, (0, 11.8, 1.3)),
    ("LeftShoulder", 11, (3.8, 19.6, 0)), ("LeftArm", 14, (11.3, 0, 0)), ("LeftForeArm", 15, (33.0, 0, 0)), ("LeftHand", 16, (25.2, 0, 0)),
    ("RightShoulder", 11, (-3.8, 19.6, 0)), ("RightArm", 18, (-11.3, 0, 0)), ("RightForeArm", 19, (-33.0, 0, 0)), ("RightHand", 20, (-25.2, 0, 0)),
]
# no Head, six channels everywhere, xyz order, a repeated bone name, far too tall -> clamped height
ODD = [
    ("Root", -1, (0, 0, 0)), ("Stem", 0, (0, 120.0, 0)), ("Tip", 1, (0, 150.0, 5.0)), ("Arm", 1, (40.0, 10.0, 0)),
    ("Arm", 3, (35.0, 0, 0)), ("Leg", 0, (12.0, -80.0, 0)),
]
# Head present, no foot bones -> head minus the lowest joint
NOFEET = [
    ("Hips", -1, (0, 0, 0)), ("Spine", 0, (0, 30.0, 0)), ("Head", 1, (0, 45.0, 2.0)), ("Tail", 0, (0, -85.0, 3.0)),
]


def write_bvh(path, skel, n_frames, seed, all_six=False, order="ZYX", root_height=92.0, swing=200.0):
    rng = np.random.default_rng(seed)
    J = len(skel)
    kids = {i: [k for k in range(J) if skel[k][1] == i] for i in range(J)}
    rot_names = " ".join(f"{a}rotation" for a in order)
    out = ["HIERARCHY"]

    def emit(i, ind):
        name, par, off = skel[i]
        t = "\t" * ind
        out.append(f"{t}{'ROOT' if par < 0 else 'JOINT'} {name}")
        out.append(t + "{")
        out.append(f"{t}\tOFFSET {off[0]:.6f} {off[1]:.6f} {off[2]:.6f}")
        if par < 0 or all_six:
            out.append(f"{t}\tCHANNELS 6 Xposition Yposition Zposition {rot_names}")
        else:
            out.append(f"{t}\tCHANNELS 3 {rot_names}")
        if kids[i]:
            for k in kids[i]:
                emit(k, ind + 1)
        else:
            out.append(f"{t}\tEnd Site")
            out.append(t + "\t{")
            out.append(f"{t}\t\tOFFSET 0.000000 5.000000 0.000000")
            out.append(t + "\t}")
        out.append(t + "}")

    # file order must be depth-first: re-index the skeleton so that emit() visits bones in list order
    emit(0, 0)
    out.append("MOTION")
    out.append(f"Frames: {n_frames}")
    out.append("Frame Time: 0.033333")
    # smooth motion with a few large swings (angles cross +-180 degrees: quaternion sign flips)
    ph = rng.uniform(0, 2 * np.pi, size=(J, 3))
    amp = rng.uniform(5.0, 60.0, size=(J, 3))
    amp[rng.integers(0, J, size=3), rng.integers(0, 3, size=3)] = swing
    for f in range(n_frames):
        tt = f / 30.0
        ang = amp * np.sin(2 * np.pi * 0.9 * tt + ph) + (swing * tt if J > 4 else 0.0) * (np.arange(J)[:, None] == 1)
        ang = (ang + 180.0) % 360.0 - 180.0              # files store wrapped angles: the half-angle jumps, q -> ~ -q
        root = np.array([30.0 * np.sin(tt), root_height + 2.0 * np.cos(3 * tt), 80.0 * tt])
        vals = []
        for i in range(J):
            if i == 0 or all_six:
                p = root if i == 0 else np.asarray(skel[i][2], float) + rng.normal(0, 0.3, 3)
                vals += [f"{v:.6f}" for v in p]
            vals += [f"{v:.6f}" for v in ang[i]]
        out.append(" ".join(vals))
    open(path, "w").write("\n".join(out) + "\n")


def dfs_order(skel):
    """Re-list a skeleton depth-first (the order a BVH file stores bones in)."""
    J = len(skel)
    kids = {i: [k for k in range(J) if skel[k][1] == i] for i in range(J)}
    order = []

    def go(i):
        order.append(i)
        for k in kids[i]:
            go(k)
    go(0)
    new = {o: n for n, o in enumerate(order)}
    return [(skel[o][0], -1 if skel[o][1] < 0 else new[skel[o][1]], skel[o][2]) for o in order]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gmr-root", default="/root/reference")
    args = ap.parse_args()
    pkg = types.ModuleType("general_motion_retargeting")
    pkg.__path__ = [os.path.join(args.gmr_root, "general_motion_retargeting")]
    sys.modules["general_motion_retargeting"] = pkg
    from general_motion_retargeting.utils.lafan1 import load_lafan1_file
    from general_motion_retargeting.utils.lafan_vendor.extract import read_bvh

    cases = [
        ("lafan", dict(skel=dfs_order(LAFAN), n_frames=24, seed=1)),
        ("odd", dict(skel=dfs_order(ODD), n_frames=9, seed=2, all_six=True, order="XYZ", root_height=10.0)),
        ("nofeet", dict(skel=dfs_order(NOFEET), n_frames=5, seed=3, order="YXZ", root_height=88.0)),
    ]
    gold = os.path.join(REPO, "tests", "golden")
    for name, kw in cases:
        with tempfile.TemporaryDirectory() as td:
            path = os.path.join(td, name + ".bvh")
            write_bvh(path, **kw)
            text = open(path).read()
            anim = read_bvh(path)
            frames, height = load_lafan1_file(path)
        rows = list(frames[0].keys())
        arr = np.zeros((len(frames), len(rows), 7))
        for t, fr in enumerate(frames):
            assert list(fr.keys()) == rows
            for k, n in enumerate(rows):
                arr[t, k, :3], arr[t, k, 3:] = fr[n][0], fr[n][1]
        np.savez_compressed(os.path.join(gold, f"bvh_{name}.npz"), bvh_text=np.array(text), rows=np.array(rows),
                            frames=arr, height=np.float64(height), local_quats=np.asarray(anim.quats),
                            local_pos=np.asarray(anim.pos), parents=np.asarray(anim.parents, np.int32),
                            bones=np.array(anim.bones))
        flips = int((np.sum(anim.quats[1:] * anim.quats[:-1], axis=-1) < 0).sum())
        print(f"{name}: T={arr.shape[0]} h={arr.shape[1]} height={height:.4f} rows[-2:]={rows[-2:]} residual sign flips={flips}")


if __name__ == "__main__":
    main()
