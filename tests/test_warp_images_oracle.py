"""CPU: the oracle chain (host cv2.resize as the reference does + cpu_oracle + post_oracle) against the fixtures made by
exec-ing the reference's WHOLE methods `warp_images` / `warp_images_bw` (diffusionGS.py:1512-1605, :1367-1510) at their
own 576x1024 size (tests/golden/make_golden_warp_images.py)."""
import cv2
import numpy as np
import pytest
import torch

from oracle import cpu_oracle as orc  # checker only
from oracle import post_oracle as po

H, W = 576, 1024


def _bits(g, n):
    return np.unpackbits(g["out_mask_png_bits"])[:n * H * W].reshape(n, H, W).astype(bool)


def _sources(g, linear):
    interp = cv2.INTER_LINEAR if linear else cv2.INTER_NEAREST
    two = "image_r" in g
    imgs = [cv2.resize(g["image_l"], (W, H), interpolation=interp)] + ([cv2.resize(g["image_r"], (W, H), interpolation=interp)] if two else [])
    deps = [cv2.resize(g["depth_l"], (W, H), interpolation=cv2.INTER_NEAREST)] + \
           ([cv2.resize(g["depth_r"], (W, H), interpolation=cv2.INTER_NEAREST)] if two else [])
    poses = [g["poses"][0]] + ([g["poses"][-1]] if two else [])
    return two, imgs, deps, poses


@pytest.mark.parametrize("name,frames", [("warp_images_fw_two_anchors", (0, 11, 12, 22)), ("warp_images_fw_single_anchor", (0, 23))])
def test_forward_method_fixture_vs_oracle(golden, name, frames):
    g = golden(name)
    assert "model/diffusionGS.py:1512-1605 sha256=" in str(g["ref_source"])
    n = int(g["n_frames"])
    two, imgs, deps, poses = _sources(g, linear=True)
    assert np.array_equal((imgs[0] / 255.)[::37, ::41].astype(np.float32), g["out_image_o_probe"])
    png = _bits(g, n)
    full = {int(f): k for k, f in enumerate(g["full_frames"])}
    for i in frames:
        s = 0 if (not two or i < 12) else 1
        with np.errstate(all="ignore"):
            r = orc.forward_warp_ext([imgs[s]], [None], [deps[s]], [poses[s]], g["poses"][i + 1], [g["K"]], None)
        u8 = np.round(np.clip(r["out"], 0, 255)).astype(np.uint8)
        o = po.post_forward(u8, r["mask"])
        assert np.array_equal(o["mask_png"][..., 0] > 0, png[i]), i
        assert np.array_equal(o["latent_mask"], g["out_masks"][i].astype(np.float64)), i
        cu8 = np.round(o["cond"] * 255).astype(np.uint8)
        assert int(cu8.sum(dtype=np.int64)) == int(g["out_cond_u8_sum"][i]), i
        if i in full:
            assert np.array_equal(cu8, g["out_cond_u8_full"][full[i]]), i


def test_backward_method_fixture_vs_oracle(golden):
    g = golden("warp_images_bw_two_anchors")
    assert "model/diffusionGS.py:1367-1510 sha256=" in str(g["ref_source"])
    n = int(g["n_frames"])
    two, imgs, deps, poses = _sources(g, linear=False)
    png = _bits(g, n)
    full = {int(f): k for k, f in enumerate(g["full_frames"])}
    t_ = torch.from_numpy
    Kf = t_(g["K"].astype(np.float32))
    for i in (0, 12, 22):
        s = 0 if i < 12 else 1
        td = cv2.resize(g["tgt_depth"][i + 1], (W, H), interpolation=cv2.INTER_NEAREST)
        with np.errstate(all="ignore"):
            iw = orc.inverse_warp(t_(np.ascontiguousarray(np.moveaxis(imgs[s], -1, 0))), t_(deps[s][None]), t_(td[None]),
                                  t_(poses[s]), t_(g["poses"][i + 1]), Kf)
        o = po.post_backward(np.moveaxis(iw["warped_img"].numpy(), 0, -1), iw["mask_reproj"].numpy(), iw["soft_mask_reproj"].numpy())
        assert np.array_equal(o["mask_png"][..., 0] > 0, png[i]), i
        assert np.array_equal(o["latent_mask"], g["out_masks"][i].astype(np.float64)), i
        assert np.max(np.abs(o["soft_pooled"] - g["out_soft_pooled"][i])) <= 1e-6, i
        k = full[i]
        assert np.array_equal(np.round(o["cond"] * 255).astype(np.uint8), g["out_cond_u8_full"][k]), i
        assert np.array_equal(np.round(o["cond_ori"] * 255).astype(np.uint8), g["out_cond_ori_u8_full"][k]), i
        assert np.max(np.abs(o["soft_ori"][3::24] - g["out_soft_ori_rows"][k])) <= 1e-6, i
