"""GPU: guided latent replacement (csrc/latent.cuh) against the fixtures and the oracle, bit-exact."""
import glob
import os

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

from oracle import latent_oracle as lo  # noqa: E402  (checker only)

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
LATENT = sorted(os.path.basename(p)[:-4] for p in glob.glob(os.path.join(GOLDEN, "latent_[0-9].npz")))


def _run(pred, tcl, mask, lam, step, **kw):
    from syn3r_b200 import latent
    up = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()  # noqa: E731
    p = up(pred.copy())
    out, ks, cv = latent.guided_replace_(p, up(tcl), up(mask), up(lam), step, return_cutoffs=True, **kw)
    torch.cuda.synchronize()
    assert out.data_ptr() == p.data_ptr()                                          # in place, like the reference
    return out.cpu().numpy(), ks.cpu().numpy(), cv.cpu().numpy()


@pytest.mark.parametrize("name", LATENT)
def test_guided_replace_golden(golden, name):
    g = golden(name)
    out, ks, cv = _run(g["pred"], g["temp_cond_latents"], g["mask"], g["lambda_ts"], int(g["step_index"]))
    assert np.array_equal(ks, g["aux_cutoff_index"])
    assert np.array_equal(cv, g["aux_cutoff_value"], equal_nan=True)
    assert np.array_equal(out, g["out"], equal_nan=True)


@pytest.mark.parametrize("b,F,c,h,w,bm,cm", [(1, 25, 4, 72, 128, 1, 1), (2, 7, 4, 20, 36, 2, 3), (1, 5, 1, 3, 5, 1, 1),
                                             (2, 6, 4, 16, 16, 1, 4), (2, 4, 4, 72, 128, 1, 1)])
def test_guided_replace_vs_oracle(b, F, c, h, w, bm, cm):
    """Production size (25 frames of 4 x 72 x 128) and odd ones; ties in the differences (quantised latents) so that
    the k-th smallest sits inside a run of equal keys.  The last case (2 x 4 x 72 x 128 keys = 288 KB) does not fit
    the shared-memory key cache and takes the recompute path."""
    rng = np.random.default_rng(b * 1000 + F * 10 + cm)
    pred = rng.normal(0, 1, (b, F, c, h, w)).astype(np.float32)
    tcl = rng.normal(0, 1, (2, F, c, h, w)).astype(np.float32)
    tcl[1] = pred[0] + np.round(rng.normal(0, 0.3, (F, c, h, w)) * 16).astype(np.float32) / 16     # many equal |diff|
    mask = (rng.random((bm, F - 1, cm, h, w)) < 0.3).astype(np.float32)
    lam = (rng.random((3, F)) * 1.3).astype(np.float32)
    r = lo.guided_replace(pred, tcl, mask, lam, 1)
    out, ks, cv = _run(pred, tcl, mask, lam, 1)
    assert np.array_equal(ks, r["cutoff_index"])
    assert np.array_equal(cv, r["cutoff_value"], equal_nan=True)
    assert np.array_equal(out, r["out"], equal_nan=True)


def test_guided_replace_sub_range_and_bad_arguments():
    from syn3r_b200 import _native as nat
    rng = np.random.default_rng(9)
    pred = rng.normal(0, 1, (1, 8, 4, 8, 8)).astype(np.float32)
    tcl = rng.normal(0, 1, (2, 8, 4, 8, 8)).astype(np.float32)
    mask = (rng.random((1, 7, 1, 8, 8)) < 0.3).astype(np.float32)
    lam = rng.random((2, 8)).astype(np.float32)
    r = lo.guided_replace(pred, tcl, mask, lam, 0, index_list=[2, 3, 4])
    out, _, _ = _run(pred, tcl, mask, lam, 0, index_list=[2, 3, 4])
    assert np.array_equal(out, r["out"], equal_nan=True)
    z = torch.zeros(8, device="cuda")
    assert nat.lib().syn3r_latent_guided_replace(nat.ptr(z), nat.ptr(z), nat.ptr(z), nat.ptr(z), 1, 8, 1, 1, 1, 1, 7, 1, 0, 3,
                                                 None, None, None) != 0            # tau0 = 0 is outside 1 .. F - 1


TOP = sorted(os.path.basename(p)[:-4] for p in glob.glob(os.path.join(GOLDEN, "latent_top_*.npz")))


@pytest.mark.parametrize("name", TOP)
def test_top_masks_golden(golden, name):
    """Selection-only entry point (step_single :548-589): same kernel, blend off, weight_clamp as the lower clamp."""
    from syn3r_b200 import latent
    g = golden(name)
    up = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()  # noqa: E731
    p = up(g["pred"])
    tm, ks, cv = latent.top_masks(p, up(g["temp_cond_latents"]), up(g["mask"]), up(g["lambda_ts"]), int(g["step_index"]),
                                  float(g["weight_clamp"]), return_cutoffs=True)
    torch.cuda.synchronize()
    assert tm.dtype == torch.bool and np.array_equal(tm.cpu().numpy(), g["out_top_masks"])
    assert np.array_equal(ks.cpu().numpy(), g["aux_cutoff_index"])
    assert np.array_equal(cv.cpu().numpy(), g["aux_cutoff_value"], equal_nan=True)
    assert np.array_equal(p.cpu().numpy(), g["pred"], equal_nan=True)                 # pred untouched


INTERP = sorted(os.path.basename(p)[:-4] for p in glob.glob(os.path.join(GOLDEN, "latent_interp_*.npz")))


@pytest.mark.parametrize("name", INTERP)
def test_step_interp_top_masks_golden(golden, name):
    """`step_interp` (:735-778) of the live scheduler: 23 mask frames, clamp 0.4, frames of ones on both ends."""
    from syn3r_b200 import latent
    g = golden(name)
    up = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()  # noqa: E731
    tm, ks, cv = latent.top_masks(up(g["pred"]), up(g["temp_cond_latents"]), up(g["mask"]), up(g["lambda_ts"]),
                                  int(g["step_index"]), float(g["weight_clamp"]), trailing_ones=True, return_cutoffs=True)
    torch.cuda.synchronize()
    assert tm.shape == g["out_top_masks"].shape and np.array_equal(tm.cpu().numpy(), g["out_top_masks"])
    assert np.array_equal(ks.cpu().numpy(), g["aux_cutoff_index"])
    assert np.array_equal(cv.cpu().numpy(), g["aux_cutoff_value"], equal_nan=True)
