"""BVH loader row on the GPU, through the C ABI (gmr_bvh_to_stream): against the golden
vectors of the reference's own load_lafan1_file and, on larger random batches, against the numpy oracle."""
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

GOLD = os.path.join(os.path.dirname(__file__), "golden")
# float64 everywhere; the kernels keep numpy's operand order and unfused products, so the only difference
# is CUDA's sincos vs glibc's (<= 1 ulp each) carried down a chain of <= ~12 bones
TOL = 1e-12


def _gold(name):
    from gmr_b200 import bvh
    z = np.load(os.path.join(GOLD, f"bvh_{name}.npz"))
    return z, bvh.parse_bvh(str(z["bvh_text"]), is_text=True)


@pytest.mark.parametrize("name", ["lafan", "odd", "nofeet"])
def test_stream_matches_reference_loader(name):
    from gmr_b200 import bvh
    z, c = _gold(name)
    streams, rows, hs = bvh.load_bvh_clips([c], dtype=np.float64)
    assert rows == [str(x) for x in z["rows"]]
    np.testing.assert_allclose(streams[0].cpu().numpy(), z["frames"], rtol=0, atol=TOL)
    assert abs(hs[0] - float(z["height"])) <= TOL
    s32, _, _ = bvh.load_bvh_clips([c], dtype=np.float32)
    np.testing.assert_array_equal(s32[0].cpu().numpy(), streams[0].cpu().numpy().astype(np.float32))


def test_load_lafan1_file_dropin(tmp_path):
    from gmr_b200 import bvh
    z, _ = _gold("lafan")
    p = tmp_path / "clip.bvh"
    p.write_text(str(z["bvh_text"]))
    frames, height = bvh.load_lafan1_file(str(p))
    assert len(frames) == z["frames"].shape[0] and list(frames[0].keys()) == [str(x) for x in z["rows"]]
    for t in (0, 7, len(frames) - 1):
        for k, n in enumerate(frames[t].keys()):
            np.testing.assert_allclose(frames[t][n][0], z["frames"][t, k, :3], rtol=0, atol=TOL)
            np.testing.assert_allclose(frames[t][n][1], z["frames"][t, k, 3:], rtol=0, atol=TOL)
    assert abs(height - float(z["height"])) <= TOL


def test_ragged_batch_matches_oracle():
    """Many clips of different lengths in one call (incl. a 1-frame clip), wrapped angles, against the oracle."""
    from gmr_b200 import bvh
    from oracle import bvh_oracle as bo
    _, c0 = _gold("lafan")
    rng = np.random.default_rng(5)
    J = len(c0.names)
    clips = []
    for T in (1, 2, 33, 300, 57, 128):
        t = np.arange(T)[:, None, None] / 30.0
        ang = rng.uniform(20, 170, (1, J, 3)) * np.sin(2 * np.pi * rng.uniform(0.2, 2.0, (1, J, 3)) * t + rng.uniform(0, 6.28, (1, J, 3)))
        ang += 300.0 * t * (rng.random((1, J, 3)) < 0.2)
        ang = (ang + 180.0) % 360.0 - 180.0
        pos = np.repeat(c0.offsets[None], T, axis=0)
        pos[:, 0] = np.stack([40 * np.sin(t[:, 0, 0]), 92 + 3 * np.cos(t[:, 0, 0]), 100 * t[:, 0, 0]], -1)
        clips.append(bvh.BvhClip(c0.names, c0.parents, c0.offsets, pos, ang, c0.order, 1 / 30))
    streams, rows, hs = bvh.load_bvh_clips(clips, dtype=np.float64)
    _, sp, sr = bvh.stream_layout(c0.names)
    head, feet = bvh.height_rows(rows)
    flips = 0
    for c, s, hg in zip(clips, streams, hs):
        ref, rh = bo.bvh_to_stream(c.positions, c.eulers, c.parents, c.order, sp, sr, head, feet)
        np.testing.assert_allclose(s.cpu().numpy(), ref, rtol=0, atol=TOL)
        assert abs(hg - rh) <= TOL
        raw = bo.euler_to_quat(np.radians(c.eulers), c.order)
        flips += int((np.sum(raw[1:] * raw[:-1], axis=-1) < 0).sum())
    assert flips > 10


def test_bvh_stream_feeds_the_retargeter():
    """loader -> retargeter on the device: same qpos as feeding the oracle's stream."""
    import torch
    from gmr_b200 import bvh
    from gmr_b200.engine import BatchedRetargeter, DeviceModel
    from oracle import bvh_oracle as bo
    _, c = _gold("lafan")
    streams, rows, hs = bvh.load_bvh_clips([c], dtype=np.float64)
    _, sp, sr = bvh.stream_layout(c.names)
    head, feet = bvh.height_rows(rows)
    ref, rh = bo.bvh_to_stream(c.positions, c.eulers, c.parents, c.order, sp, sr, head, feet)
    dm = DeviceModel.load("bvh", "unitree_g1", rows)
    out = []
    for stream, hgt in ((streams[0][None], hs[0]), (torch.from_numpy(ref)[None].cuda(), rh)):
        br = BatchedRetargeter(dm, 1)
        br.set_heights([hgt])
        out.append(br.run(stream).cpu().numpy())
        br.close()
    np.testing.assert_allclose(out[0], out[1], rtol=0, atol=1e-7)
    assert np.isfinite(out[0]).all()
