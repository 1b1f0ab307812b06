"""gmr_bvh_to_stream_static (include/gmr_b200.h): the BVH -> frame-stream entry point that allocates nothing - its
scratch is a fixed array in the library's device image.  Through the C ABI, against gmr_bvh_to_stream on the same
inputs, two calls in a row (the second sees the first call's leftovers in the shared array), plus the size limit.

Collected last on purpose; passes alone in a fresh process and after the others.
"""
import ctypes as C
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

GOLD = os.path.join(os.path.dirname(__file__), "golden")
TOL = 1e-12          # the tolerance of tests/test_gpu_bvh.py; same kernels on the same inputs, so equality is expected
MAX_ELEMS = 1 << 19  # GMR_BVH_STATIC_MAX_ELEMS


def _clips(seed, lengths):
    from gmr_b200 import bvh
    z = np.load(os.path.join(GOLD, "bvh_lafan.npz"))
    c0 = bvh.parse_bvh(str(z["bvh_text"]), is_text=True)
    rng = np.random.default_rng(seed)
    J = len(c0.names)
    clips = []
    for T in lengths:
        t = np.arange(T)[:, None, None] / 30.0
        ang = rng.uniform(20, 170, (1, J, 3)) * np.sin(2 * np.pi * rng.uniform(0.2, 2.0, (1, J, 3)) * t + rng.uniform(0, 6.28, (1, J, 3)))
        ang += 300.0 * t * (rng.random((1, J, 3)) < 0.2)                     # wrapped angles: sign flips of the quaternions
        ang = (ang + 180.0) % 360.0 - 180.0
        pos = np.repeat(c0.offsets[None], T, axis=0)
        pos[:, 0] = np.stack([40 * np.sin(t[:, 0, 0]), 92 + 3 * np.cos(t[:, 0, 0]), 100 * t[:, 0, 0]], -1)
        clips.append(bvh.BvhClip(c0.names, c0.parents, c0.offsets, pos, ang, c0.order, 1 / 30))
    return c0, clips


def _call(entry, c0, clips, dtype):
    """One call of `entry` (gmr_bvh_to_stream or gmr_bvh_to_stream_static) through ctypes -> (stream, heights, rc)."""
    import torch
    from gmr_b200 import _lib, bvh
    from gmr_b200.abi import F32, F64
    L = _lib.lib()
    J = len(c0.names)
    rows, src_pos, src_rot = bvh.stream_layout(c0.names)
    head, feet = bvh.height_rows(rows)
    start = np.concatenate([[0], np.cumsum([c.positions.shape[0] for c in clips])]).astype(np.int32)
    N, h, Cn = int(start[-1]), len(rows), len(clips)
    dev = torch.device("cuda", 0)
    lpos = torch.from_numpy(np.concatenate([c.positions for c in clips], 0).reshape(N, J, 3)).to(dev)
    eul = torch.from_numpy(np.concatenate([c.eulers for c in clips], 0).reshape(N, J, 3)).to(dev)
    tdt = torch.float32 if dtype == np.float32 else torch.float64
    out = torch.full((N, h, 7), float("nan"), dtype=tdt, device=dev)
    heights = torch.full((Cn,), float("nan"), dtype=torch.float64, device=dev)
    i32 = lambda a: torch.from_numpy(np.ascontiguousarray(a, np.int32)).to(dev)      # noqa: E731
    d_start, d_par, d_sp, d_sr = i32(start), i32(c0.parents), i32(src_pos), i32(src_rot)
    ptr = lambda t: C.c_void_p(t.data_ptr())                                           # noqa: E731
    feet4 = (list(feet) + [-1] * 4)[:4]
    s = torch.cuda.current_stream(dev).cuda_stream
    rc = getattr(L, entry)(0, C.c_void_p(s), ptr(lpos), ptr(eul), ptr(d_start), Cn, N, J, ptr(d_par),
                           (C.c_int32 * 3)(*[int(a) for a in c0.order]), ptr(d_sp), ptr(d_sr), h,
                           F32 if dtype == np.float32 else F64, ptr(out), int(head), (C.c_int32 * 4)(*feet4), ptr(heights))
    torch.cuda.synchronize()
    return out.cpu().numpy(), heights.cpu().numpy(), rc


def test_static_scratch_entry_point_matches_the_allocating_one_twice_in_a_row():
    from gmr_b200 import _lib
    L = _lib.lib()
    # two different batches back to back: the second is SHORTER than the first and has other motions, so whatever the
    # first call left in the shared array (signs, global transforms) lies under and behind the second call's data
    for seed, lengths, dtype in ((11, (300, 1, 57, 128, 33), np.float64), (12, (2, 90, 17), np.float32)):
        c0, clips = _clips(seed, lengths)
        ref, ref_h, rc = _call("gmr_bvh_to_stream", c0, clips, dtype)
        assert rc == 0, L.gmr_last_error()
        got, got_h, rc = _call("gmr_bvh_to_stream_static", c0, clips, dtype)
        assert rc == 0, L.gmr_last_error()
        assert np.isfinite(got).all() and np.isfinite(got_h).all()
        np.testing.assert_allclose(got, ref, rtol=0, atol=TOL)
        np.testing.assert_allclose(got_h, ref_h, rtol=0, atol=TOL)
    # and the first batch again after the second: same answer as the first time
    c0, clips = _clips(11, (300, 1, 57, 128, 33))
    ref, ref_h, _ = _call("gmr_bvh_to_stream", c0, clips, np.float64)
    got, got_h, rc = _call("gmr_bvh_to_stream_static", c0, clips, np.float64)
    assert rc == 0
    np.testing.assert_array_equal(got, ref)
    np.testing.assert_array_equal(got_h, ref_h)


def test_static_scratch_entry_point_rejects_sizes_beyond_its_array():
    from gmr_b200 import _lib
    L = _lib.lib()
    c0, _ = _clips(1, (1,))
    J = len(c0.names)
    T = MAX_ELEMS // J + 1                      # one frame too many
    from gmr_b200 import bvh
    big = bvh.BvhClip(c0.names, c0.parents, c0.offsets, np.repeat(c0.offsets[None], T, axis=0), np.zeros((T, J, 3)), c0.order, 1 / 30)
    out, heights, rc = _call("gmr_bvh_to_stream_static", c0, [big], np.float32)
    assert rc == -4
    msg = L.gmr_last_error().decode()
    assert "GMR_BVH_STATIC_MAX_ELEMS" in msg and str(T * J) in msg
    assert np.isnan(out).all() and np.isnan(heights).all()          # nothing was launched
    ok = bvh.BvhClip(c0.names, c0.parents, c0.offsets, big.positions[:T - 1], big.eulers[:T - 1], c0.order, 1 / 30)
    out, heights, rc = _call("gmr_bvh_to_stream_static", c0, [ok], np.float32)          # the largest call that fits
    assert rc == 0, L.gmr_last_error()
    ref, ref_h, _ = _call("gmr_bvh_to_stream", c0, [ok], np.float32)
    np.testing.assert_array_equal(out, ref)
    np.testing.assert_array_equal(heights, ref_h)
