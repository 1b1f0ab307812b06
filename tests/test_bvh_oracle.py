"""BVH loader row on the CPU: the text parser and the numpy restatement against the golden
vectors produced by the reference's own load_lafan1_file / read_bvh (tools/make_golden_bvh.py)."""
import os

import numpy as np
import pytest

from gmr_b200 import bvh
from oracle import bvh_oracle as bo

GOLD = os.path.join(os.path.dirname(__file__), "golden")
CASES = ["lafan", "odd", "nofeet"]


def _gold(name):
    z = np.load(os.path.join(GOLD, f"bvh_{name}.npz"))
    return z, bvh.parse_bvh(str(z["bvh_text"]), is_text=True)


@pytest.mark.parametrize("name", CASES)
def test_parser_matches_reference_read_bvh(name):
    z, c = _gold(name)
    assert list(c.names) == [str(x) for x in z["bones"]]
    np.testing.assert_array_equal(c.parents, z["parents"])
    np.testing.assert_array_equal(c.positions, z["local_pos"])            # bit-exact: same float() of the same text
    lq = bo.remove_quat_discontinuities(bo.euler_to_quat(np.radians(c.eulers), c.order))
    np.testing.assert_array_equal(lq, z["local_quats"])


@pytest.mark.parametrize("name", CASES)
def test_oracle_matches_reference_loader(name):
    z, c = _gold(name)
    rows, sp, sr = bvh.stream_layout(c.names)
    assert rows == [str(x) for x in z["rows"]]
    head, feet = bvh.height_rows(rows)
    out, hgt = bo.bvh_to_stream(c.positions, c.eulers, c.parents, c.order, sp, sr, head, feet)
    np.testing.assert_array_equal(out, z["frames"])                        # numpy restatement: bit-exact
    assert hgt == float(z["height"])


def test_golden_cases_cover_the_branches():
    z, c = _gold("lafan")
    rows = [str(x) for x in z["rows"]]
    assert rows[-2:] == ["LeftFootMod", "RightFootMod"] and "Head" in rows       # lafan1.py:35-39, :45-53
    k_mod, k_foot, k_toe = rows.index("LeftFootMod"), rows.index("LeftFoot"), rows.index("LeftToe")
    np.testing.assert_array_equal(z["frames"][:, k_mod, :3], z["frames"][:, k_foot, :3])
    np.testing.assert_array_equal(z["frames"][:, k_mod, 3:], z["frames"][:, k_toe, 3:])
    raw = bo.euler_to_quat(np.radians(c.eulers), c.order)
    assert (np.sum(raw[1:] * raw[:-1], axis=-1) < 0).any()                        # sign continuity is exercised
    z2, c2 = _gold("odd")
    assert "Head" not in [str(x) for x in z2["rows"]] and len(z2["rows"]) == len(set(c2.names)) < len(c2.names)
    z3, _ = _gold("nofeet")
    assert float(z3["height"]) == 1.75                                            # clamp (:67-69)


def test_parser_rejects_garbage():
    with pytest.raises(ValueError):
        bvh.parse_bvh("HIERARCHY\nROOT a\n{\n}\nMOTION\n", is_text=True)
    z, c = _gold("lafan")
    text = str(z["bvh_text"])
    with pytest.raises(ValueError):
        bvh.parse_bvh(text[:text.rindex("\n", 0, len(text) - 2)].replace("Frames: 24", "Frames: 30"), is_text=True)


@pytest.mark.skipif(not os.path.exists("/root/reference/out/test_canonical.bvh"), reason="reference tree not on this machine")
def test_parser_and_oracle_on_the_reference_bvh():
    """The one real BVH the reference ships (101 bones, 250 frames), through the reference's own loader."""
    import sys
    import types
    import warnings
    pkg = types.ModuleType("general_motion_retargeting")
    pkg.__path__ = ["/root/reference/general_motion_retargeting"]
    saved = sys.modules.get("general_motion_retargeting")
    sys.modules["general_motion_retargeting"] = pkg
    try:
        with warnings.catch_warnings():
            warnings.simplefilter("ignore")
            from general_motion_retargeting.utils.lafan1 import load_lafan1_file
            frames, height = load_lafan1_file("/root/reference/out/test_canonical.bvh")
    finally:
        for k in [k for k in sys.modules if k.startswith("general_motion_retargeting")]:
            del sys.modules[k]
        if saved is not None:
            sys.modules["general_motion_retargeting"] = saved
    c = bvh.parse_bvh("/root/reference/out/test_canonical.bvh")
    rows, sp, sr = bvh.stream_layout(c.names)
    assert rows == list(frames[0].keys())
    head, feet = bvh.height_rows(rows)
    out, hgt = bo.bvh_to_stream(c.positions, c.eulers, c.parents, c.order, sp, sr, head, feet)
    assert hgt == height
    for t in (0, 1, 17, len(frames) - 1):
        for k, n in enumerate(rows):
            np.testing.assert_array_equal(out[t, k, :3], frames[t][n][0])
            np.testing.assert_array_equal(out[t, k, 3:], frames[t][n][1])
