"""Dataset-driver tail (gmr_b200/dataset.py) against a numpy restatement of the reference drivers' post-processing
(scripts/smplx_to_robot_dataset.py:93-146) and the structural checks of the reference's scripts/smoke_test.py."""
import os
import pickle

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _clips(src, robot, lengths, seed=9):
    from gmr_b200 import model_pack, synth
    m, t = model_pack.build_tables(src, robot)
    human, heights, names = synth.synth_human_clips(m, t, src, B=len(lengths), T=max(lengths), seed=seed)
    return m, t, [human[i, :n] for i, n in enumerate(lengths)], heights, names, human


def _golden_tail(tag):
    z = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", f"tail_{tag}.npz"), allow_pickle=False)
    return {k: z[k] for k in z.files}


@pytest.mark.parametrize("tag,adjust", [("bvh_unitree_g1", False), ("smplx_unitree_g1", True), ("smplx_booster_t1", True)])
def test_driver_tail_matches_the_reference_lines(tag, adjust):
    """dataset.driver_tail on known qpos against tests/golden/tail_*.npz = the reference drivers' OWN post-processing
    lines executed on CPU (tools/make_golden_tail.py: scripts/bvh_to_robot_dataset.py:107-151 with HEIGHT_ADJUST
    off, scripts/smplx_to_robot_dataset.py:93-141 with HEIGHT_ADJUST and ROOT_ORIGIN_OFFSET on)."""
    from gmr_b200 import dataset
    from gmr_b200.engine import DeviceModel
    g = _golden_tail(tag)
    lengths = [int(n) for n in g["lengths"]]
    dm = DeviceModel.load(str(g["source"]), str(g["robot"]))
    qpos = np.concatenate([g[f"qpos_{i}"] for i in range(len(lengths))])
    tail = dataset.driver_tail(dm, qpos, lengths, height_adjust=adjust, root_origin_offset=adjust)
    assert list(g["link_body_list"]) == dataset.link_body_list(dm.desc.model)
    o = 0
    for i, n in enumerate(lengths):
        np.testing.assert_array_equal(tail["root_rot"][o:o + n], g[f"root_rot_{i}"])            # a permutation: exact
        np.testing.assert_array_equal(tail["dof_pos"][o:o + n], g[f"dof_pos_{i}"])
        assert tail["local_body_pos"].dtype == np.float32 and g[f"local_body_pos_{i}"].dtype == np.float32
        np.testing.assert_allclose(tail["local_body_pos"][o:o + n], g[f"local_body_pos_{i}"], atol=2e-6)   # float32 FK chain
        np.testing.assert_allclose(tail["root_pos"][o:o + n], g[f"root_pos_{i}"], atol=2e-6 if adjust else 0)
        if adjust:
            assert tail["root_pos"][o, 0] == 0 and tail["root_pos"][o, 1] == 0
        o += n
    dm.close()


def test_motion_dicts_match_the_reference_tail(tmp_path):
    """Whole driver path (CSR host call + tail) on the clips the golden file was made from: qpos from the CUDA
    path, tail from the reference's lines on the oracle's qpos; structural checks of the reference's smoke_test.py."""
    from gmr_b200 import dataset
    g = _golden_tail("bvh_unitree_g1")
    lengths = [int(n) for n in g["lengths"]]
    m, t, clips, heights, names, human = _clips("bvh", "unitree_g1", lengths)
    motions, status = dataset.retarget_clips("bvh", "unitree_g1", clips, heights=heights, human_names=names, fps=30,
                                             return_status=True)
    assert len(motions) == len(lengths) and not (status & 28).any()
    for i, (mo, n) in enumerate(zip(motions, lengths)):
        # smoke_test.py:19-72 structural checks
        assert set(mo) == {"fps", "root_pos", "root_rot", "dof_pos", "local_body_pos", "link_body_list"}
        assert mo["root_pos"].shape == (n, 3) and mo["root_rot"].shape == (n, 4) and mo["dof_pos"].shape == (n, m.nq - 7)
        assert mo["local_body_pos"].shape == (n, m.nbody - 1, 3) and mo["local_body_pos"].dtype == np.float32
        assert len(mo["link_body_list"]) == m.nbody - 1 and mo["fps"] == 30
        norms = np.linalg.norm(mo["root_rot"], axis=-1)
        assert (norms > 0.5).all() and (norms < 1.5).all()
        np.testing.assert_allclose(mo["root_rot"], g[f"root_rot_{i}"], atol=1e-7)
        np.testing.assert_allclose(mo["dof_pos"], g[f"dof_pos_{i}"], atol=1e-7)
        np.testing.assert_allclose(mo["root_pos"], g[f"root_pos_{i}"], atol=1e-7)
        np.testing.assert_allclose(mo["local_body_pos"], g[f"local_body_pos_{i}"], atol=1e-5)
    p = tmp_path / "a" / "b.pkl"
    dataset.save_motion(str(p), motions[0])
    back = pickle.load(open(p, "rb"))
    np.testing.assert_array_equal(back["dof_pos"], motions[0]["dof_pos"])
    # .pt form (scripts/convert_motion_pkl_to_pt.py): tensors, xyzw kept, other fields verbatim
    import torch
    dataset.save_motion_pt(str(tmp_path / "a" / "b.pt"), motions[0])
    pt = torch.load(tmp_path / "a" / "b.pt", weights_only=False)
    assert isinstance(pt["root_rot"], torch.Tensor) and pt["link_body_list"] == motions[0]["link_body_list"] and pt["fps"] == 30
    np.testing.assert_array_equal(dataset.load_motion(str(tmp_path / "a" / "b.pt"))["root_rot"], motions[0]["root_rot"])


def test_failed_clip_is_flagged_and_skipped():
    """A non-finite frame poisons its own clip only; the driver gets None for it and the status bits say why."""
    from gmr_b200 import abi, dataset
    lengths = [9, 12, 7]
    m, t, clips, heights, names, human = _clips("bvh", "unitree_g1", lengths, seed=4)
    ok = dataset.retarget_clips("bvh", "unitree_g1", clips, heights=heights, human_names=names)
    bad = [c.copy() for c in clips]
    bad[1][5, names.index(t.scaled_names[3]), 0] = np.nan      # a body the config reads (the others are dropped, :219-220)
    motions, status = dataset.retarget_clips("bvh", "unitree_g1", bad, heights=heights, human_names=names, return_status=True)
    assert motions[1] is None and (status[1] & abi.STATUS_BAD_INPUT) and (status[1] & abi.STATUS_BAD_OUTPUT)
    assert "non-finite" in dataset.describe_status(int(status[1]))
    for i in (0, 2):
        assert not status[i] & abi.STATUS_FAILED
        np.testing.assert_array_equal(motions[i]["dof_pos"], ok[i]["dof_pos"])


def test_pending_files_skip_logic(tmp_path):
    from gmr_b200 import dataset
    src, tgt = tmp_path / "src", tmp_path / "tgt"
    (src / "sub").mkdir(parents=True)
    for n in ("b.bvh", "a.bvh", "sub/c.npz", "notes.txt"):
        (src / n).write_text("x")
    (tgt / "sub").mkdir(parents=True)
    (tgt / "a.pkl").write_text("done")
    todo = dataset.pending_files(str(src), str(tgt), (".bvh", ".npz"))
    assert [os.path.relpath(s, src) for s, _ in todo] == ["b.bvh", "sub/c.npz"]
    assert len(dataset.pending_files(str(src), str(tgt), (".bvh", ".npz"), override=True)) == 3


def test_robot_without_floating_base_is_refused():
    from gmr_b200 import dataset
    m, t, clips, heights, names, human = _clips("smplx", "galaxea_r1pro", [3, 4])
    with pytest.raises(ValueError):
        dataset.retarget_clips("smplx", "galaxea_r1pro", clips, heights=heights, human_names=names)


def test_bvh_driver_end_to_end(tmp_path):
    """scripts/bvh_to_robot_dataset.py on a folder of *.bvh files: GPU loader -> batched retarget -> pickles with
    the reference's keys (scripts/bvh_to_robot_dataset.py:144-156)."""
    import pickle
    import subprocess
    import sys
    z = np.load(os.path.join(os.path.dirname(__file__), "golden", "bvh_lafan.npz"))
    text = str(z["bvh_text"])
    src, tgt = tmp_path / "src", tmp_path / "tgt"
    (src / "sub").mkdir(parents=True)
    (src / "walk.bvh").write_text(text)
    lines = text.rstrip("\n").split("\n")
    short = "\n".join(lines[:-10]).replace("Frames: 24", "Frames: 14") + "\n"
    (src / "sub" / "short.bvh").write_text(short)
    script = os.path.join(os.path.dirname(__file__), "..", "scripts", "bvh_to_robot_dataset.py")
    r = subprocess.run([sys.executable, script, "--src_folder", str(src), "--tgt_folder", str(tgt), "--robot", "unitree_g1"],
                       capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout + r.stderr
    for rel, T in (("walk.pkl", 24), ("sub/short.pkl", 14)):
        with open(tgt / rel, "rb") as f:
            m = pickle.load(f)
        assert set(m) == {"fps", "root_pos", "root_rot", "dof_pos", "local_body_pos", "link_body_list"}
        assert m["root_pos"].shape == (T, 3) and m["root_rot"].shape == (T, 4) and m["dof_pos"].shape == (T, 29)
        assert np.isfinite(m["dof_pos"]).all() and abs(np.linalg.norm(m["root_rot"], axis=1) - 1).max() < 1e-9
    # the first 14 frames are the same motion: same result up to the height guess, which looks at the LAST frame
    assert np.isfinite(m["local_body_pos"]).all()


def test_smplx_driver_end_to_end(tmp_path):
    """scripts/smplx_to_robot_dataset.py on a folder of body-model outputs: fps alignment + orientation chain on the
    GPU -> batched retarget -> pickles (HEIGHT_ADJUST and ROOT_ORIGIN_OFFSET on, as the reference driver)."""
    import pickle
    import subprocess
    import sys
    from gmr_b200 import model_pack, synth
    m, t = model_pack.build_tables("smplx", "unitree_g1")
    human, heights, names = synth.synth_human_clips(m, t, "smplx", B=1, T=48)          # plausible joint positions
    z = np.load(os.path.join(os.path.dirname(__file__), "golden", "smplx_prep_skip4.npz"))
    src, tgt = tmp_path / "src", tmp_path / "tgt"
    src.mkdir()
    T = 41
    joints = np.zeros((T, 127, 3), np.float32)
    joints[:, :55] = human[0, :T, :, :3]
    np.savez(src / "a.npz", full_pose=z["pose"].reshape(T, -1) * 0.3, joints=joints, mocap_frame_rate=np.array(120.0),
             betas=np.zeros(16))
    np.savez(src / "b_stagei.npz", full_pose=z["pose"], joints=joints, mocap_frame_rate=np.array(120.0), betas=np.zeros(16))
    script = os.path.join(os.path.dirname(__file__), "..", "scripts", "smplx_to_robot_dataset.py")
    r = subprocess.run([sys.executable, script, "--src_folder", str(src), "--tgt_folder", str(tgt), "--robot", "unitree_g1"],
                       capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout + r.stderr
    assert sorted(os.listdir(tgt)) == ["a.pkl"]
    with open(tgt / "a.pkl", "rb") as f:
        mo = pickle.load(f)
    assert mo["root_pos"].shape == (10, 3) and abs(mo["fps"] - 10 / 41 * 120.0) < 1e-12
    np.testing.assert_allclose(mo["root_pos"][0, :2], 0.0, atol=1e-12)                # ROOT_ORIGIN_OFFSET
    assert np.isfinite(mo["dof_pos"]).all() and mo["local_body_pos"].shape[0] == 10


def test_clip_diagnostics_match_the_per_frame_log(tmp_path):
    """dataset.clip_diagnostics (all frames of a clip at once) == what scripts/fbx_to_robot.py:1183-1212 logs after
    each retarget() call: error1(), error2(), key-body position errors; and the --errors_csv file format."""
    from gmr_b200 import dataset, model_pack, synth
    from gmr_b200.engine import DeviceModel
    from gmr_b200.motion_retarget import GeneralMotionRetargeting
    from oracle.gmr_oracle_np import OracleRetargeter
    T = 7
    m, t = model_pack.build_tables("bvh", "unitree_g1")
    human, heights, names = synth.synth_human_clips(m, t, "bvh", B=1, T=T, seed=4, dtype=np.float64)
    frames = [{n: (human[0, i, k, :3].tolist(), human[0, i, k, 3:].tolist()) for k, n in enumerate(names)} for i in range(T)]
    gmr = GeneralMotionRetargeting("bvh", "unitree_g1", actual_human_height=float(heights[0]))
    ref = OracleRetargeter(m, t, actual_human_height=float(heights[0]))
    idx = [names.index(b) for b in t.scaled_names]
    rows, qpos = [], []
    for i in range(T):
        qpos.append(gmr.retarget(frames[i]))
        ref.retarget(human[0, i][idx])
        row = [gmr.error1(), gmr.error2()]
        # the oracle of the same two calls
        np.testing.assert_allclose(row[0], ref.error(ref._tasks(0, ref.scaled_human_data)), atol=1e-7)
        for rb, hb in dataset.KEY_BODIES:
            bid = gmr.model.body(rb).id
            row.append(float(np.linalg.norm(gmr.configuration.data.xpos[bid] - gmr.scaled_human_data[hb][0])))
        rows.append(row)
    rows = np.array(rows)
    dm = DeviceModel.load("bvh", "unitree_g1", names)
    d = dataset.clip_diagnostics(dm, human[0], np.array(qpos), height=float(heights[0]))
    np.testing.assert_allclose(d["error1"], rows[:, 0], atol=1e-10)
    np.testing.assert_allclose(d["error2"], rows[:, 1], atol=1e-10)
    for k, (rb, _) in enumerate(dataset.KEY_BODIES):
        np.testing.assert_allclose(d[rb], rows[:, 2 + k], atol=1e-10)
    p = tmp_path / "errors.csv"
    dataset.write_errors_csv(str(p), d)
    dataset.write_errors_csv(str(p), d)                       # appends without a second header
    lines = p.read_text().strip().split("\n")
    assert lines[0] == "frame,error1,error2,pelvis_pos_err,left_hand_pos_err,right_hand_pos_err" and len(lines) == 1 + 2 * T
    assert lines[1].split(",")[0] == "0" and abs(float(lines[1].split(",")[1]) - rows[0, 0]) < 1e-10
    # a robot without those bodies logs NaN there
    m2, t2 = model_pack.build_tables("smplx", "booster_t1")
    h2, hg2, n2 = synth.synth_human_clips(m2, t2, "smplx", B=1, T=3, seed=1, dtype=np.float64)
    dm2 = DeviceModel.load("smplx", "booster_t1", n2)
    d2 = dataset.clip_diagnostics(dm2, h2[0], np.tile(m2.qpos0, (3, 1)), height=float(hg2[0]))
    assert np.isnan(d2["left_wrist_yaw_link"]).all() and np.isfinite(d2["error1"]).all()


def test_bvh_driver_two_ranks_write_disjoint_halves(tmp_path):
    """The driver under torchrun's environment (RANK / WORLD_SIZE / LOCAL_RANK): the ranks cut the file list between
    them, each writes its own pickles (and .pt files), no communication; together they equal the one-process run."""
    import pickle
    import subprocess
    import sys
    z = np.load(os.path.join(os.path.dirname(__file__), "golden", "bvh_lafan.npz"))
    lines = str(z["bvh_text"]).rstrip("\n").split("\n")
    src = tmp_path / "src"
    src.mkdir()
    for k, drop in enumerate((0, 3, 6, 9, 12)):
        text = "\n".join(lines[:len(lines) - drop]).replace("Frames: 24", f"Frames: {24 - drop}") + "\n"
        (src / f"clip{k}.bvh").write_text(text)
    script = os.path.join(os.path.dirname(__file__), "..", "scripts", "bvh_to_robot_dataset.py")

    def run(tgt, env_extra, extra=()):
        env = dict(os.environ, **env_extra)
        r = subprocess.run([sys.executable, script, "--src_folder", str(src), "--tgt_folder", str(tgt), "--robot", "unitree_g1", *extra],
                           capture_output=True, text=True, timeout=600, env=env)
        assert r.returncode == 0, r.stdout + r.stderr
        return r.stdout
    run(tmp_path / "one", {})
    outs = [run(tmp_path / "two", {"RANK": str(r), "WORLD_SIZE": "2", "LOCAL_RANK": "0"}, ("--pt",)) for r in range(2)]
    assert "[rank 0/2]" in outs[0] and "[rank 1/2]" in outs[1]
    one = sorted(os.listdir(tmp_path / "one"))
    assert one == [f"clip{k}.pkl" for k in range(5)]
    assert sorted(f for f in os.listdir(tmp_path / "two") if f.endswith(".pkl")) == one
    assert len([f for f in os.listdir(tmp_path / "two") if f.endswith(".pt")]) == 5
    for f in one:
        a = pickle.load(open(tmp_path / "one" / f, "rb"))
        b = pickle.load(open(tmp_path / "two" / f, "rb"))
        np.testing.assert_array_equal(a["dof_pos"], b["dof_pos"])
        np.testing.assert_array_equal(a["local_body_pos"], b["local_body_pos"])
