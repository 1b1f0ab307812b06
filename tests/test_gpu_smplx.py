"""SMPL-X stream preparation on the GPU through the C ABI (gmr_smplx_to_stream): against
the golden vectors of the reference's own get_smplx_data_offline_fast, and a ragged batch against the oracle."""
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

GOLD = os.path.join(os.path.dirname(__file__), "golden")
TOL = 1e-12        # float64; same formulas as scipy, CUDA libm instead of glibc, chains of <= 9 joints


@pytest.mark.parametrize("name", ["skip4", "native", "skip1"])
def test_stream_matches_reference(name):
    from gmr_b200 import smplx_prep as sp
    z = np.load(os.path.join(GOLD, f"smplx_prep_{name}.npz"))
    clip = {"full_pose": z["pose"], "joints": z["joints"], "fps": float(z["src_fps"])}
    streams, fps = sp.smplx_to_stream([clip], tgt_fps=int(z["tgt_fps"]), dtype=np.float64)
    got = streams[0].cpu().numpy()
    assert fps[0] == float(z["aligned_fps"])
    np.testing.assert_allclose(got[..., :3], z["pos"], rtol=0, atol=TOL)
    np.testing.assert_allclose(got[..., 3:], z["quat"], rtol=0, atol=TOL)
    s32, _ = sp.smplx_to_stream([clip], tgt_fps=int(z["tgt_fps"]), dtype=np.float32)
    np.testing.assert_array_equal(s32[0].cpu().numpy(), got.astype(np.float32))


def test_ragged_batch_matches_oracle():
    from gmr_b200 import smplx_prep as sp
    from oracle import smplx_oracle as so
    rng = np.random.default_rng(11)
    clips = []
    for T, fps in ((1, 30.0), (5, 120.0), (64, 60.0), (301, 120.0), (40, 30.0), (9, 100.0)):
        pose = (rng.normal(0, 0.8, (T, 55, 3)) * rng.uniform(0, 1, (1, 55, 1))).astype(np.float32)
        pose[:, 3] *= 1e-5
        clips.append({"full_pose": pose.reshape(T, -1), "joints": rng.normal(0, 1, (T, 127, 3)).astype(np.float32), "fps": fps})
    streams, fps = sp.smplx_to_stream(clips, tgt_fps=30, dtype=np.float64)
    for c, s, f in zip(clips, streams, fps):
        T = c["full_pose"].shape[0]
        tt, rf = sp.alignment(T, c["fps"], 30)
        pos, quat = so.smplx_to_stream(c["full_pose"].reshape(T, 55, 3), c["joints"], sp.SMPLX_PARENTS, tt)
        got = s.cpu().numpy()
        assert got.shape == (len(pos), 55, 7) and f == rf
        np.testing.assert_allclose(got[..., :3], pos, rtol=0, atol=TOL)
        np.testing.assert_allclose(got[..., 3:], quat, rtol=0, atol=TOL)


def test_smplx_stream_feeds_the_retargeter():
    from gmr_b200 import smplx_prep as sp
    from gmr_b200.engine import BatchedRetargeter, DeviceModel
    z = np.load(os.path.join(GOLD, "smplx_prep_native.npz"))
    # a standing pose with plausible joint positions is not needed for this check: finite in, finite out
    clip = {"full_pose": z["pose"] * 0.2, "joints": z["joints"], "fps": 30.0}
    streams, _ = sp.smplx_to_stream([clip], tgt_fps=30, dtype=np.float32)
    dm = DeviceModel.load("smplx", "unitree_g1", list(sp.SMPLX_JOINT_NAMES))
    br = BatchedRetargeter(dm, 1)
    q = br.run(streams[0][None]).cpu().numpy()
    assert q.shape[1] == streams[0].shape[0] and np.isfinite(q).all()
