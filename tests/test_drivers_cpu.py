"""CPU: host logic of the dataset drivers - file selection of the smplx driver (reference
scripts/smplx_to_robot_dataset.py:193-234), the cut of a file list across ranks, the .pt writer
(scripts/convert_motion_pkl_to_pt.py:39-56) and the golden files of the driver tail."""
import glob
import importlib.util
import os
import pickle

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _script(name):
    spec = importlib.util.spec_from_file_location(name, os.path.join(ROOT, "scripts", name + ".py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def test_smplx_driver_selects_files_like_the_reference(tmp_path, capsys):
    drv = _script("smplx_to_robot_dataset")
    hard = sorted(drv.hard_motions())
    assert len(hard) > 200 and "ACCAD_Male2MartialArtsExtended_c3d_Extended_3_stageii" in hard
    src, tgt = tmp_path / "src", tmp_path / "tgt"
    (src / "a").mkdir(parents=True)
    names = ["walk_10_stageii.npz", "walk_2_stageii.npz", "walk_1_stageii.pkl", "walk_3_stagei.npz", "notes.txt",
             "BMLrub_jog_stageii.npz", "crawl_1_stageii.npz", "sit_lie_down_stageii.npz", "upstairs_4_stageii.npz",
             hard[0] + ".npz", "done_stageii.npz"]
    for n in names:
        (src / "a" / n).write_text("x")
    (tgt / "a").mkdir(parents=True)
    (tgt / "a" / "done_stageii.pkl").write_text("done")
    sel = drv.select_files(str(src), str(tgt), False)
    got = [os.path.basename(s) for s, _ in sel]
    assert got == ["walk_1_stageii.pkl", "walk_2_stageii.npz", "walk_10_stageii.npz"]        # natural order, filters applied
    assert all(t.startswith(str(tgt)) and t.endswith(".pkl") for _, t in sel)
    out = capsys.readouterr().out
    assert "full args_list: 8" in out and "new args_list: 3" in out
    assert len(drv.select_files(str(src), str(tgt), True)) == 4                              # --override brings done_ back
    assert drv.natural_key("a10b") > drv.natural_key("a9b")


def test_file_list_is_cut_across_ranks_without_overlap(tmp_path, monkeypatch):
    from gmr_b200 import dataset
    todo = []
    rng = np.random.default_rng(0)
    for i in range(23):
        p = tmp_path / f"f{i}.bvh"
        p.write_bytes(b"x" * int(rng.integers(100, 5000)))
        todo.append((str(p), str(tmp_path / f"f{i}.pkl")))
    for world in (1, 2, 4, 8):
        parts = [dataset.shard_files(todo, r, world) for r in range(world)]
        flat = [x for p in parts for x in p]
        assert sorted(flat) == sorted(todo) and len(set(flat)) == len(todo)
        if world > 1:
            loads = [sum(os.path.getsize(s) for s, _ in p) for p in parts]
            assert max(loads) - min(loads) <= 5000
    monkeypatch.setenv("RANK", "3"); monkeypatch.setenv("WORLD_SIZE", "8"); monkeypatch.setenv("LOCAL_RANK", "3")
    assert dataset.rank_info() == (3, 8, 3)
    monkeypatch.delenv("RANK"); monkeypatch.delenv("WORLD_SIZE"); monkeypatch.delenv("LOCAL_RANK")
    assert dataset.rank_info() == (0, 1, None)


def test_pt_writer_round_trip(tmp_path):
    import torch
    from gmr_b200 import dataset
    motion = {"fps": 30, "root_pos": np.random.rand(5, 3), "root_rot": np.random.rand(5, 4), "dof_pos": np.random.rand(5, 29),
              "local_body_pos": np.random.rand(5, 38, 3).astype(np.float32), "link_body_list": ["a", "b"]}
    p = str(tmp_path / "x" / "m.pt")
    dataset.save_motion_pt(p, motion)
    raw = torch.load(p, weights_only=False)
    assert isinstance(raw["root_pos"], torch.Tensor) and raw["local_body_pos"].dtype == torch.float32
    assert raw["fps"] == 30 and raw["link_body_list"] == ["a", "b"]
    back = dataset.load_motion(p)
    for k, v in motion.items():
        if isinstance(v, np.ndarray):
            np.testing.assert_array_equal(back[k], v)
    dataset.save_motion(str(tmp_path / "x" / "m.pkl"), motion)
    np.testing.assert_array_equal(dataset.load_motion(str(tmp_path / "x" / "m.pkl"))["dof_pos"], motion["dof_pos"])
    assert pickle.load(open(tmp_path / "x" / "m.pkl", "rb"))["fps"] == 30


@pytest.mark.parametrize("path", sorted(glob.glob(os.path.join(ROOT, "tests", "golden", "tail_*.npz"))))
def test_tail_golden_files_are_what_the_reference_lines_define(path):
    """Properties of the reference's tail that hold whatever the FK: the root quaternion is qpos[3:7] rolled to xyzw,
    dof_pos is qpos[7:], the smplx driver zeroes the root XY of frame 0 and lifts the clip by a constant."""
    g = np.load(path, allow_pickle=False)
    smplx = "smplx" in str(g["script"])
    for i, n in enumerate(g["lengths"]):
        q = g[f"qpos_{i}"]
        np.testing.assert_array_equal(g[f"root_rot_{i}"], q[:, [4, 5, 6, 3]])
        np.testing.assert_array_equal(g[f"dof_pos_{i}"], q[:, 7:])
        assert g[f"local_body_pos_{i}"].shape == (n, len(g["link_body_list"]), 3)
        rp = g[f"root_pos_{i}"]
        if smplx:
            assert rp[0, 0] == 0 and rp[0, 1] == 0
            dz = rp[:, 2] - q[:, 2]
            assert np.ptp(dz) < 1e-12
            np.testing.assert_allclose(rp[:, :2], q[:, :2] - q[0, :2], atol=1e-15)
        else:
            np.testing.assert_array_equal(rp, q[:, :3])
