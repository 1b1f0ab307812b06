"""SMPL-X stream preparation on the CPU: the numpy restatement against the golden vectors
produced by the reference's own get_smplx_data_offline_fast (tools/make_golden_smplx.py)."""
import os

import numpy as np
import pytest

from gmr_b200 import smplx_prep as sp
from oracle import smplx_oracle as so

GOLD = os.path.join(os.path.dirname(__file__), "golden")
CASES = ["skip4", "native", "skip1"]


@pytest.mark.parametrize("name", CASES)
def test_oracle_matches_reference(name):
    z = np.load(os.path.join(GOLD, f"smplx_prep_{name}.npz"))
    T = z["pose"].shape[0]
    tt, fps = sp.alignment(T, float(z["src_fps"]), int(z["tgt_fps"]))
    assert fps == float(z["aligned_fps"])
    assert (tt is None) == (name == "native")
    pos, quat = so.smplx_to_stream(z["pose"], z["joints"], sp.SMPLX_PARENTS, tt)
    assert pos.shape == z["pos"].shape
    # same formulas as scipy 1.18, different summation order inside its compiled norm / product: a few ulp
    np.testing.assert_allclose(pos, z["pos"], rtol=0, atol=1e-15)
    np.testing.assert_allclose(quat, z["quat"], rtol=0, atol=2e-15)


def test_golden_cases_cover_the_branches():
    z = np.load(os.path.join(GOLD, "smplx_prep_skip4.npz"))
    pose = z["pose"].astype(np.float64)
    ang = np.linalg.norm(pose, axis=-1)
    assert (ang <= 1e-3).any() and (ang > 3.2).any()                  # small-angle series; beyond pi (canonical sign)
    q = so.from_rotvec(pose)
    d = np.abs(np.sum(so.normalize(q[:-1]) * so.normalize(q[1:]), axis=-1))
    assert (d > 0.9995).any() and (d < 0.9995).any()                  # linear and spherical branches of slerp
    assert np.sum(q[:-1] * q[1:], axis=-1).min() < 0                  # hemisphere flip
    assert z["pos"].shape[0] == 41 // 4 and str(z["pos_dtype"]) == "float64"
    assert str(np.load(os.path.join(GOLD, "smplx_prep_native.npz"))["pos_dtype"]) == "float32"


def test_names_and_tree():
    assert len(sp.SMPLX_JOINT_NAMES) == len(sp.SMPLX_PARENTS) == 55
    assert all(p < i for i, p in enumerate(sp.SMPLX_PARENTS))
    from gmr_b200 import synth
    assert list(synth.SMPLX_BODIES) == list(sp.SMPLX_JOINT_NAMES)
