"""The reference-facing class (gmr_b200.motion_retarget.GeneralMotionRetargeting) against the
numpy oracle of the reference class, used the way the reference's scripts use it
(scripts/bvh_to_robot_dataset.py:84-102, scripts/fbx_to_robot.py:996-1041,1184-1203)."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _frames(src, robot, T, seed=21):
    from gmr_b200 import model_pack, synth
    m, t = model_pack.build_tables(src, robot)
    human, heights, names = synth.synth_human_clips(m, t, src, B=1, T=T, seed=seed, dtype=np.float64)
    frames = [{n: (human[0, i, k, :3].tolist(), human[0, i, k, 3:].tolist()) for k, n in enumerate(names)} for i in range(T)]
    return m, t, human[0], float(heights[0]), names, frames


@pytest.mark.parametrize("src,robot", [("bvh", "unitree_g1"), ("smplx", "galaxea_r1pro"), ("smplx", "hightorque_hi")])
def test_retarget_like_the_dataset_driver(src, robot):
    from gmr_b200.motion_retarget import GeneralMotionRetargeting
    from oracle.gmr_oracle_np import OracleRetargeter
    T = 8
    m, t, human, height, names, frames = _frames(src, robot, T)
    gmr = GeneralMotionRetargeting(src_human=src, tgt_robot=robot, actual_human_height=height)
    assert gmr.xml_file.endswith(".xml")
    ref = OracleRetargeter(m, t, actual_human_height=height)
    idx = [names.index(b) for b in t.scaled_names]
    for i in range(T):
        q = gmr.retarget(frames[i])
        qr = ref.retarget(human[i][idx])
        assert q.shape == (m.nq,) and q.dtype == np.float64
        np.testing.assert_allclose(q, qr, atol=1e-7)
        # .scaled_human_data: dict body -> [pos, quat] after scale + offset (:124)
        for k, b in enumerate(t.scaled_names):
            np.testing.assert_allclose(gmr.scaled_human_data[b][0], ref.scaled_human_data[k, :3], atol=1e-12)
            np.testing.assert_allclose(gmr.scaled_human_data[b][1], ref.scaled_human_data[k, 3:], atol=1e-12)
    # to_numpy side effect on the caller's dict (:203-206)
    assert isinstance(frames[0][names[0]][0], np.ndarray)
    # diagnostics the fbx script reads (:1184-1203)
    for stage, (err_fn, tasks) in enumerate(((gmr.error1, gmr.tasks1), (gmr.error2, gmr.tasks2))):
        if not t.use_table[stage]:
            continue
        np.testing.assert_allclose(err_fn(), ref.error(ref._tasks(stage, ref.scaled_human_data)), atol=1e-9)
        assert [tk.frame_name for tk in tasks] == t.task_frame_names[stage]
        e0 = tasks[0].compute_error(gmr.configuration)
        assert e0.shape == (6,)
    np.testing.assert_allclose(gmr.configuration.data.qpos, ref.q, atol=1e-7)
    bid = gmr.model.body(t.task_frame_names[0][0]).id
    np.testing.assert_allclose(gmr.configuration.data.xpos[bid], ref.kin.xpos[bid], atol=1e-7)


def test_mutating_scale_table_and_offsets_takes_effect():
    """scripts/fbx_to_robot.py:1003-1041 edits human_scale_table and pos_offsets1 after construction."""
    from gmr_b200.motion_retarget import GeneralMotionRetargeting
    from oracle.gmr_oracle_np import OracleRetargeter
    m, t, human, height, names, frames = _frames("bvh", "unitree_g1", 4)
    gmr = GeneralMotionRetargeting("bvh", "unitree_g1")
    ref = OracleRetargeter(m, t)
    idx = [names.index(b) for b in t.scaled_names]
    np.testing.assert_allclose(gmr.retarget(frames[0]), ref.retarget(human[0][idx]), atol=1e-7)
    body = t.scaled_names[3]
    gmr.human_scale_table[body] *= 1.1
    ref.scale[3] *= 1.1
    gmr.pos_offsets1[body] = gmr.pos_offsets1[body] + np.array([0.01, 0.0, -0.02])
    t.pos_off[3] += np.array([0.01, 0.0, -0.02])
    np.testing.assert_allclose(gmr.retarget(frames[1]), ref.retarget(human[1][idx]), atol=1e-7)


def test_errors_mirror_the_reference():
    from gmr_b200.motion_retarget import GeneralMotionRetargeting
    with pytest.raises(KeyError):
        GeneralMotionRetargeting("bvh", "no_such_robot")
    with pytest.raises(KeyError):
        GeneralMotionRetargeting("no_such_source", "unitree_g1")
    m, t, human, height, names, frames = _frames("bvh", "unitree_g1", 1)
    gmr = GeneralMotionRetargeting("bvh", "unitree_g1")
    del frames[0][t.scaled_names[2]]
    with pytest.raises(KeyError):
        gmr.retarget(frames[0])


def test_scale_and_offset_helpers():
    from gmr_b200.motion_retarget import GeneralMotionRetargeting
    m, t, human, height, names, frames = _frames("bvh", "unitree_g1", 1)
    gmr = GeneralMotionRetargeting("bvh", "unitree_g1", actual_human_height=1.62)
    d = gmr.to_numpy(frames[0])
    scaled = gmr.scale_human_data(d, gmr.human_root_name, gmr.human_scale_table)
    assert set(scaled) == set(t.scaled_names)
    root = d[gmr.human_root_name][0]
    for b in t.scaled_names[1:]:
        exp = (d[b][0] - root) * gmr.human_scale_table[b] + gmr.human_scale_table[gmr.human_root_name] * root
        np.testing.assert_allclose(scaled[b][0], exp, atol=1e-12)
    off = gmr.offset_human_data(scaled, gmr.pos_offsets1, gmr.rot_offsets1)
    gmr.update_targets(frames[0])
    for b in t.scaled_names:
        np.testing.assert_allclose(off[b][0], gmr.scaled_human_data[b][0], atol=1e-12)
        np.testing.assert_allclose(off[b][1], gmr.scaled_human_data[b][1], atol=1e-12)


def test_helpers_then_retarget_like_the_fbx_script():
    """scripts/fbx_to_robot.py:1003-1041: scale_human_data / offset_human_data (helper batches of another footprint)
    between retarget() calls; the helpers must not disturb the main batch's launches or its warm start."""
    from gmr_b200.motion_retarget import GeneralMotionRetargeting
    from oracle.gmr_oracle_np import OracleRetargeter
    m, t, human, height, names, frames = _frames("bvh", "unitree_g1", 5)
    gmr = GeneralMotionRetargeting("bvh", "unitree_g1", actual_human_height=height)
    ref = OracleRetargeter(m, t, actual_human_height=height)
    idx = [names.index(b) for b in t.scaled_names]
    for i in range(5):
        d = gmr.to_numpy(frames[i])
        scaled = gmr.scale_human_data(d, gmr.human_root_name, gmr.human_scale_table)
        gmr.offset_human_data(scaled, gmr.pos_offsets1, gmr.rot_offsets1)
        np.testing.assert_allclose(gmr.retarget(frames[i]), ref.retarget(human[i][idx]), atol=1e-7)
        np.testing.assert_allclose(gmr.error2(), ref.error(ref._tasks(1, ref.scaled_human_data)), atol=1e-9)


def test_retarget_rate_of_one_object():
    """The B = 1 shim keeps its buffers and uploads tables only when they change: thousands of frames/s per object
    (the reference's README quotes 60-70 FPS for this loop)."""
    import time
    from gmr_b200.motion_retarget import GeneralMotionRetargeting
    m, t, human, height, names, frames = _frames("bvh", "unitree_g1", 200)
    gmr = GeneralMotionRetargeting("bvh", "unitree_g1", actual_human_height=height)
    for f in frames[:20]:
        gmr.retarget(f)
    t0 = time.perf_counter()
    for f in frames[20:]:
        gmr.retarget(f)
    rate = 180 / (time.perf_counter() - t0)
    print(f"B=1 shim: {rate:.0f} frames/s")
    assert rate > 1000
