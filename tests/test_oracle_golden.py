"""Pins the CPU oracle (oracle/) against outputs of the LIVE reference (tests/golden/*.npz, made by
tests/golden/make_golden.py from /root/reference @ 402f4e9).  CPU only."""
import numpy as np
import pytest
import torch

from oracle import cpu_oracle as orc

FW_CASES = ["fw_generic_0", "fw_generic_1", "fw_generic_2", "fw_mid", "fw_u8_f64pose", "fw_mask_K2",
            "fw_identity", "fw_integer_row", "fw_integer_shift", "fw_behind", "fw_all_out", "fw_all_behind",
            "fw_nan_depth"]


@pytest.mark.parametrize("name", FW_CASES)
def test_forward_warp_matches_reference(golden, name):
    g = golden(name)
    with np.errstate(all="ignore"):
        warped, mask2, flow = orc.forward_warp(g["frame"], g.get("mask1"), g["depth"], g["T1"], g["T2"], g["K"],
                                               g.get("K2"))
    assert warped.dtype == np.uint8 and mask2.dtype == np.bool_ and flow.dtype == np.float64
    # flow12 is fp64 and must be bit-identical (same op order as the host BLAS used for the goldens)
    assert np.array_equal(flow, g["flow12"], equal_nan=True)
    assert np.array_equal(mask2, g["mask2"])
    assert np.array_equal(warped, g["warped"])


def test_compute_transformed_points_dense(golden):
    for name in ("ctp_dense", "ctp_f64"):
        g = golden(name)
        q, X = orc.compute_transformed_points(g["depth"], g["T1"], g["T2"], g["K"], g.get("K2"))
        assert q.shape == g["q"].shape and X.shape == g["X"].shape
        assert np.array_equal(X, g["X"])
        assert np.array_equal(q, g["q"])


def test_bilinear_splatting_generic(golden):
    g = golden("bs_generic")
    out, mask = orc.bilinear_splatting(g["frame"], g["mask1"], g["depth"], g["flow12"], g["flow_mask"], False)
    assert np.array_equal(mask, g["mask"])
    np.testing.assert_allclose(out, g["out"], rtol=1e-12, atol=1e-13)


def test_create_grid(golden):
    g = golden("grid")["grid"]
    o = orc.create_grid(5, 7)
    assert o.dtype == g.dtype and np.array_equal(o, g)


def _band_ok(a, b, val, thr, rel=1e-3):
    """masks may differ from the live torch result only where `val` is within rel*thr of `thr`."""
    diff = a != b
    return bool(np.all(np.abs(val[diff] - thr) <= rel * abs(thr))), int(diff.sum())


def test_consistency_matches_reference(golden):
    g = golden("cc_generic")
    t = torch.from_numpy
    pts = orc.get_points_from_depth(t(g["depth1"]), t(g["K1"])).numpy()
    np.testing.assert_array_equal(pts, g["pts"])
    pts2 = orc.transform_points(t(g["pts"]), t(g["pose1"]), t(g["pose2"])).numpy()
    np.testing.assert_array_equal(pts2, g["pts2"])
    err = orc.consistency_check_with_depth(t(g["depth1"]), t(g["pose1"]), t(g["K1"]), t(g["depth2"]),
                                           t(g["pose2"]), t(g["K2"])).numpy()
    frac_exact = np.mean(err == g["err"])
    assert frac_exact > 0.99, frac_exact
    np.testing.assert_allclose(err, g["err"], rtol=2e-4, atol=2e-4)


@pytest.mark.parametrize("name", ["iw_generic_0", "iw_generic_1", "iw_generic_2"])
def test_inverse_warp_matches_reference(golden, name):
    g = golden(name)
    t = torch.from_numpy
    bg = t(g["bg_mask"]) if "bg_mask" in g else None
    bw = float(g["bandwidth"])
    with np.errstate(all="ignore"):
        o = orc.inverse_warp(t(g["img"]), t(g["depth"]), t(g["depth_pseudo"]), t(g["pose1"]), t(g["pose2"]), t(g["K"]),
                             bg_mask=bg, bandwidth=bw, return_err=True)
    for k in ("warped_img", "warped_depth", "mask_warp", "warped_masked_img"):
        assert np.array_equal(o[k].numpy(), g["out_" + k], equal_nan=True), k
    if bg is None:
        assert o["warped_bg_mask"] is None and g["out_warped_bg_mask"].size == 0
    else:
        assert np.array_equal(o["warped_bg_mask"].numpy(), g["out_warped_bg_mask"])
    for k in ("mask_depth", "mask_depth_strict", "mask", "mask_inv"):
        assert np.array_equal(o[k].numpy(), g["out_" + k]), k
    err = o["reproj_err"].numpy()
    ok, n = _band_ok(o["mask_reproj"].numpy(), g["out_mask_reproj"], err, bw)
    assert ok and n <= 2, n
    soft, gs = o["soft_mask_reproj"].numpy(), g["out_soft_mask_reproj"]
    both = np.isfinite(soft) & np.isfinite(gs)
    assert np.array_equal(np.isnan(soft), np.isnan(gs))
    assert np.max(np.abs(soft[both] - gs[both])) <= 2e-6
