"""Golden fixtures of the uncertainty fusion: tests/golden/fusion_*.npz.

Re-executes the reference's statements of /root/reference/model/diffusionGS.py:821-854 (intensity confidence,
product with the geometric confidence, 8x8 pooled mask, `np.where(unc > 0.5, gs, warp)`) and :876 (`np.clip`) with
NumPy itself on seeded inputs; the callers cannot be imported (SURVEY §8c caveat 3).  Only generalisation: the latent
grid `reshape(72,8,128,8)` is written `reshape(H//8,8,W//8,8)`.  float32 exp / pow are SIMD library calls whose last
bit depends on the host: consumers compare with a tolerance (1e-6) and exclude the |unc - 0.5| <= 1e-6 band of
the `np.where`.   Run:  python tests/golden/make_golden_fusion.py
"""
import os

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))


def get_intensity_confidence(warpped_image, pseudo_image, uncertainty_mask, sigma=0.1):      # :821-828
    intensity_conf = np.exp(-((np.linalg.norm(warpped_image - pseudo_image, axis=-1, keepdims=True)) / sigma) ** 3) * (1 - uncertainty_mask)
    return intensity_conf


def reference_fusion(cond_images_ori, gs_images, soft_masks_reproj_ori):
    """:831-854 + :876; cond_images_ori (N,H,W,3) f32 = warped / 255., gs_images (N,H,W,3) f32,
    soft_masks_reproj_ori (N,H,W) f32 = 1 - soft_mask_reproj."""
    N, H, W, _ = cond_images_ori.shape
    cond_image_ori = np.stack(list(cond_images_ori))
    intensity_conf = get_intensity_confidence(cond_image_ori, gs_images, 1 - (cond_image_ori.sum(axis=-1, keepdims=True) > 0), sigma=0.5)
    geo_inten_conf = (intensity_conf) * (1 - soft_masks_reproj_ori[..., None])
    geo_inten_uncertainty = 1 - geo_inten_conf
    mask_buf = []
    for mask_erosion_ in geo_inten_uncertainty:
        mask_erosion = np.mean(mask_erosion_, axis=-1)
        mask_erosion = mask_erosion.reshape(H // 8, 8, W // 8, 8).transpose(0, 2, 1, 3).reshape(H // 8, W // 8, 64)
        mask_erosion = np.mean(mask_erosion, axis=2)
        mask_buf.append(mask_erosion)
    masks = np.stack(mask_buf)
    masks = torch.from_numpy(masks).float()
    cond_image = np.where(geo_inten_uncertainty > 0.5, gs_images, np.stack(list(cond_images_ori)))
    cond_image = [np.clip(im, 0, 1) for im in cond_image]
    return dict(masks=masks.numpy(), cond_image=np.stack(cond_image), geo_inten_uncertainty=geo_inten_uncertainty)


def main():
    rng = np.random.default_rng(99)
    vers = np.array([f"numpy {np.__version__}", "statements of DecaYale/SYN3R@402f4e9 model/diffusionGS.py:821-854, :876"])
    for i, (N, H, W) in enumerate([(2, 48, 64), (3, 32, 40), (1, 72, 128)]):
        gs = rng.random((N, H, W, 3), dtype=np.float32)
        cond = (gs + np.float32(0.25) * rng.standard_normal((N, H, W, 3)).astype(np.float32) *
                (rng.random((N, H, W, 1)) > 0.5)).astype(np.float32)
        cond = np.clip(cond, 0, 1.2).astype(np.float32)
        cond[:, 5:9, 7:15] = 0                                 # blanked pixels: uncertainty_mask = 1
        soft_ori = (1 - np.exp(-(rng.random((N, H, W), dtype=np.float32) * 2) ** 3)).astype(np.float32)
        out = reference_fusion(cond.copy(), gs.copy(), soft_ori.copy())
        path = os.path.join(HERE, f"fusion_{i}.npz")
        np.savez_compressed(path, versions=vers, in_cond_ori=cond, in_gs=gs, in_soft_ori=soft_ori,
                            **{"out_" + k: v for k, v in out.items()})
        print(os.path.basename(path), os.path.getsize(path) // 1024, "KiB", {k: (v.dtype, v.shape) for k, v in out.items()})


if __name__ == "__main__":
    main()
