"""GPU: forward-backward flow consistency (csrc/flow.cuh) against the live-reference fixtures and the oracle."""
import glob
import os

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

from oracle import flow_oracle as fo  # noqa: E402  (checker only)

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
FLOW = sorted(os.path.basename(p)[:-4] for p in glob.glob(os.path.join(GOLDEN, "flow_fb_*.npz")))
BAND = 1e-3


def _run(fwd, bwd, alpha, beta):
    from syn3r_b200 import flow
    o = flow.forward_backward_consistency_check(torch.from_numpy(fwd).cuda(), torch.from_numpy(bwd).cuda(), alpha, beta,
                                                return_diff=True)
    torch.cuda.synchronize()
    return [t.cpu().numpy() for t in o]


@pytest.mark.parametrize("name", FLOW)
def test_flow_fb_golden(golden, name):
    g = golden(name)
    fo_, bo_, df, db = _run(g["fwd_flow"], g["bwd_flow"], float(g["alpha"]), float(g["beta"]))
    for occ, diff, d in ((fo_, df, "fwd"), (bo_, db, "bwd")):
        assert occ.dtype == np.float32 and occ.shape == g["out_" + d + "_occ"].shape
        assert np.max(np.abs(diff - g["aux_diff_" + d])) <= BAND
        near = np.abs(g["aux_diff_" + d] - g["aux_threshold"]) <= BAND
        assert ((occ == g["out_" + d + "_occ"]) | near).all()


@pytest.mark.parametrize("B,H,W,amp", [(2, 64, 96, 4.0), (1, 37, 53, 40.0), (1, 2, 2, 1.0), (1, 576, 1024, 8.0)])
def test_flow_fb_vs_oracle(B, H, W, amp):
    """Same arithmetic on both sides (reciprocal multiply, fused taps): the norms agree to float32 rounding of the
    double-rounded oracle fma, the flags everywhere outside a 1e-5 band.  Includes flows leaving the frame, NaN and inf."""
    rng = np.random.default_rng(B * 100 + H)
    fwd = rng.normal(0, amp, (B, 2, H, W)).astype(np.float32)
    bwd = (-fwd + rng.normal(0, 0.4, (B, 2, H, W))).astype(np.float32)
    if H > 8:
        fwd[0, 0, 3, 5] = np.nan
        bwd[0, 1, 7, 2] = np.inf
        fwd[0, :, 5, 5] = 0.0
    r = fo.forward_backward_consistency_check(fwd, bwd, 0.01, 0.5)
    fo_, bo_, df, db = _run(fwd, bwd, 0.01, 0.5)
    for occ, diff, d in ((fo_, df, "fwd"), (bo_, db, "bwd")):
        ref = r["diff_" + d]
        fin = np.isfinite(ref)
        assert np.array_equal(np.isnan(diff), np.isnan(ref))
        assert np.max(np.abs(diff[fin] - ref[fin]), initial=0) <= 1e-5 * max(1.0, amp)
        with np.errstate(invalid="ignore"):
            near = np.abs(ref - r["threshold"]) <= 1e-5 * max(1.0, amp)
        assert ((occ == r[d + "_occ"]) | near).all()


def test_flow_fb_rejects_degenerate_sizes_and_empty_batch():
    from syn3r_b200 import _native as nat
    L = nat.lib()
    assert L.syn3r_flow_fb_consistency_batched(None, None, 0, 8, 8, 0.01, 0.5, None, None, None, None, None) == 0
    z = torch.zeros((1, 2, 1, 8), device="cuda")
    o = torch.zeros((1, 1, 8), device="cuda")
    assert L.syn3r_flow_fb_consistency_batched(nat.ptr(z), nat.ptr(z), 1, 1, 8, 0.01, 0.5, nat.ptr(o), None, None, None, None) != 0
