#!/usr/bin/env python
"""Regenerate gmr_b200/data/model_pack.json from a GMR checkout.

    python tools/build_model_pack.py [--gmr-root /root/reference]

The pack holds derived tables only (flattened kinematic trees + parsed IK configs);
see gmr_b200/model_pack.py.
"""
import argparse
import json
import os
import sys

sys.path.insert(0, os.path.join(os.path.dirname(__file__), ".."))

from gmr_b200 import params  # noqa: E402
from gmr_b200.mjcf import load_mjcf, MjcfError  # noqa: E402
from gmr_b200.ik_config import load_ik_config, build_retarget_tables  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gmr-root", default=os.environ.get("GMR_ROOT", "/root/reference"))
    args = ap.parse_args()
    params.set_gmr_root(args.gmr_root)

    robots, configs, pairs = {}, {}, []
    for robot, xml in params.ROBOT_XML_DICT.items():
        try:
            m = load_mjcf(os.fspath(xml), name=robot)
        except (MjcfError, FileNotFoundError) as e:
            print(f"skip robot {robot}: {e}")
            continue
        robots[robot] = m.to_dict()
        print(f"{robot:24s} nbody={m.nbody - 1:3d} nq={m.nq:3d} nv={m.nv:3d} "
              f"limited={int(m.jnt_limited.sum()):3d} depth={int(m.body_depth.max())}")
    for src, robot in params.constructible_pairs(args.gmr_root):
        path = params.IK_CONFIG_DICT[src][robot]
        cfg = load_ik_config(path)
        m = load_mjcf(os.fspath(params.ROBOT_XML_DICT[robot]), name=robot)
        t = build_retarget_tables(m, cfg, src, robot)     # validates body names
        configs[path.name] = cfg
        pairs.append([src, robot])
        print(f"{src:6s}->{robot:24s} s={t.n_scaled} m1={len(t.task_body[0])} m2={len(t.task_body[1])}")
    out = {"format": 1, "robots": robots, "ik_configs": configs, "pairs": pairs}
    os.makedirs(os.path.dirname(params.PACK_PATH), exist_ok=True)
    with open(params.PACK_PATH, "w") as f:
        json.dump(out, f, separators=(",", ":"))
    print("wrote", params.PACK_PATH, os.path.getsize(params.PACK_PATH), "bytes")
    # motion names of the smplx driver's exclusion lists (scripts/smplx_to_robot_dataset.py:193-208), derived list
    names = set()
    for fn in ("0.txt", "1.txt"):
        with open(os.path.join(args.gmr_root, "assets", "hard_motions", fn)) as f:
            for line in f:
                if "Motion:" in line:
                    names.add(line.split(":")[1].strip().split(",")[0].strip().split(".")[0])
    hm = os.path.join(os.path.dirname(params.PACK_PATH), "hard_motions.txt")
    with open(hm, "w") as f:
        f.write("\n".join(sorted(names)) + "\n")
    print("wrote", hm, len(names), "names")


if __name__ == "__main__":
    main()
