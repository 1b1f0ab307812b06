"""The bundled model pack (gmr_b200/data/model_pack.json) against a fresh parse of a GMR checkout.

The pack is what the GPU box uses (it has no checkout); here, where /root/reference (or $GMR_ROOT) exists, every
robot's flattened tree and every constructible pair's IK config is re-derived from the MJCF / JSON files and must
equal the bundled copy.  Reference inputs: assets/<robot>/*.xml (params.py:7-19), ik_configs/*.json (params.py:21-48).
"""
import json
import os

import numpy as np
import pytest

from gmr_b200 import params
from gmr_b200.ik_config import load_ik_config
from gmr_b200.mjcf import load_mjcf

ROOT = os.environ.get("GMR_ROOT", "/root/reference")
needs_checkout = pytest.mark.skipif(not os.path.isdir(os.path.join(ROOT, "assets")), reason="no GMR checkout on this box")


def _pack():
    with open(params.PACK_PATH) as f:
        return json.load(f)


def _same(a, b, path=""):
    if isinstance(a, dict):
        assert isinstance(b, dict) and sorted(a) == sorted(b), path
        for k in a:
            _same(a[k], b[k], f"{path}/{k}")
    elif isinstance(a, (list, tuple)):
        if a and isinstance(a[0], (int, float)) and not isinstance(a[0], bool):
            np.testing.assert_array_equal(np.asarray(a, float), np.asarray(b, float), err_msg=path)
        else:
            assert len(a) == len(b), path
            for i, (x, y) in enumerate(zip(a, b)):
                _same(x, y, f"{path}[{i}]")
    else:
        assert a == b, path


def test_pack_is_well_formed():
    p = _pack()
    assert p["format"] == 1
    assert {"unitree_g1", "unitree_g1_with_hands", "galaxea_r1pro"} <= set(p["robots"])
    for src, robot in p["pairs"]:
        assert robot in p["robots"]
        assert f"{src}_to_" in "".join(k for k in p["ik_configs"] if k.startswith(src))


@needs_checkout
def test_pack_matches_a_fresh_parse():
    prev = params.get_gmr_root()
    params.set_gmr_root(ROOT)
    try:
        p = _pack()
        pairs = params.constructible_pairs(ROOT)
        assert sorted(map(tuple, p["pairs"])) == sorted(pairs)
        robots = {r for _, r in pairs}
        assert robots <= set(p["robots"])
        for robot in sorted(p["robots"]):
            fresh = load_mjcf(os.fspath(params.ROBOT_XML_DICT[robot]), name=robot).to_dict()
            _same(json.loads(json.dumps(fresh)), p["robots"][robot], robot)
        for src, robot in pairs:
            path = params.IK_CONFIG_DICT[src][robot]
            _same(json.loads(json.dumps(load_ik_config(path))), p["ik_configs"][path.name], path.name)
    finally:
        params.set_gmr_root(prev)
