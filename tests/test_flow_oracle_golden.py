"""CPU: the flow-consistency oracle against the fixtures produced by the live gmflow reference."""
import glob
import os

import numpy as np
import pytest

from oracle import flow_oracle as fo  # checker only

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
FLOW = sorted(os.path.basename(p)[:-4] for p in glob.glob(os.path.join(GOLDEN, "flow_fb_*.npz")))
BAND = 1e-3          # |diff - threshold| below which a flag may legitimately differ (float32 on flows of tens of px)


def test_flow_fixture_set_present():
    assert len(FLOW) == 4


@pytest.mark.parametrize("name", FLOW)
def test_flow_fb_oracle_vs_golden(golden, name):
    g = golden(name)
    o = fo.forward_backward_consistency_check(g["fwd_flow"], g["bwd_flow"], float(g["alpha"]), float(g["beta"]))
    assert np.max(np.abs(o["threshold"] - g["aux_threshold"])) <= 1e-6
    for d in ("fwd", "bwd"):
        assert np.max(np.abs(o["diff_" + d] - g["aux_diff_" + d])) <= BAND
        near = np.abs(g["aux_diff_" + d] - g["aux_threshold"]) <= BAND
        assert ((o[d + "_occ"] == g["out_" + d + "_occ"]) | near).all()
        assert o[d + "_occ"].dtype == np.float32 and 0.05 < g["out_" + d + "_occ"].mean() < 0.95   # both classes present


def test_flow_warp_of_zero_flow_is_identity():
    rng = np.random.default_rng(5)
    f = rng.normal(0, 3, (2, 2, 17, 23)).astype(np.float32)
    w = fo.flow_warp(f, np.zeros_like(f))
    assert np.max(np.abs(w - f)) <= 1e-4          # the normalise / unnormalise round trip costs a few ulps of the coordinate
