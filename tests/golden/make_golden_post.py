"""Golden fixtures of the caller-side post-processing: tests/golden/post_*.npz.

The callers (`model/diffusionGS.py`) cannot be imported here (FSGS, open3d, diffusers are absent, SURVEY §8c
caveat 3), so the fixture generator RE-EXECUTES the reference's own statements for this block with the same
third-party calls it makes (numpy, cv2.dilate, PIL round trips), line by line:

  forward branch   /root/reference/model/diffusionGS.py:1572-1597  (`warp_images`)
  backward branch  /root/reference/model/diffusionGS.py:1452-1491  (`warp_images_bw`)

The only generalisation: the hard-coded latent grid `reshape(72,8,128,8)` of a 576x1024 frame is written as
`reshape(H//8,8,W//8,8)` so that small fixtures can be stored.  Run in the build container:

    python tests/golden/make_golden_post.py
"""
import os
import sys

import cv2
import numpy as np
import PIL.Image

HERE = os.path.dirname(os.path.abspath(__file__))


def reference_post_forward(warped_frame2, mask2):
    """diffusionGS.py:1572-1597, statements in the reference's order."""
    H, W = mask2.shape
    mask = 1 - mask2
    mask[mask < 0.5] = 0
    mask[mask >= 0.5] = 1
    mask = np.repeat(mask[:, :, np.newaxis] * 255., repeats=3, axis=2)
    kernel = np.ones((5, 5), np.uint8)
    mask_erosion = cv2.dilate(np.array(mask), kernel, iterations=1)
    mask_erosion = PIL.Image.fromarray(np.uint8(mask_erosion))
    mask_png = np.array(mask_erosion)                                   # what `.save(..._mask.png)` stores
    mask_erosion_ = np.array(mask_erosion) / 255.
    mask_erosion_[mask_erosion_ < 0.5] = 0
    mask_erosion_[mask_erosion_ >= 0.5] = 1
    warped_frame2 = PIL.Image.fromarray(np.uint8(warped_frame2))
    warped_frame2 = PIL.Image.fromarray(np.uint8(warped_frame2 * (1 - mask_erosion_)))
    cond = np.asarray(warped_frame2, dtype=np.float32) / 255.
    mask_erosion = np.mean(mask_erosion_, axis=-1)
    mask_erosion = mask_erosion.reshape(H // 8, 8, W // 8, 8).transpose(0, 2, 1, 3).reshape(H // 8, W // 8, 64)
    mask_erosion = np.mean(mask_erosion, axis=2)
    mask_erosion[mask_erosion < 0.2] = 0
    mask_erosion[mask_erosion >= 0.2] = 1
    return dict(mask_png=mask_png, cond=cond, latent_mask=mask_erosion, masks_ero=mask_erosion_.astype(np.uint8))


def reference_post_backward(warped_frame2, mask2, soft):
    """diffusionGS.py:1452-1491: `mask2` = warp_dict['mask_reproj'] (bool), `warped_frame2` = warped_img as HWC
    float32, `soft` = warp_dict['soft_mask_reproj'] (H,W) float32."""
    H, W = mask2.shape
    mask = 1 - mask2
    mask[mask < 0.5] = 0
    mask[mask >= 0.5] = 1
    mask = np.repeat(mask[:, :, np.newaxis] * 255., repeats=3, axis=2)
    kernel = np.ones((5, 5), np.uint8)
    mask_erosion = cv2.dilate(np.array(mask), kernel, iterations=1)
    mask_erosion = PIL.Image.fromarray(np.uint8(mask_erosion))
    mask_png = np.array(mask_erosion)
    mask_erosion_ = np.array(mask_erosion) / 255.
    mask_erosion_[mask_erosion_ < 0.5] = 0
    mask_erosion_[mask_erosion_ >= 0.5] = 1
    cond_ori = warped_frame2 / 255.
    warped_frame2 = PIL.Image.fromarray(np.uint8(warped_frame2 * (1 - mask_erosion_)))
    cond = np.asarray(warped_frame2, dtype=np.float32) / 255.
    mask_erosion = np.mean(mask_erosion_, axis=-1)
    mask_erosion = mask_erosion.reshape(H // 8, 8, W // 8, 8).transpose(0, 2, 1, 3).reshape(H // 8, W // 8, 64)
    mask_erosion = np.mean(mask_erosion, axis=2)
    mask_erosion[mask_erosion < 0.2] = 0
    mask_erosion[mask_erosion >= 0.2] = 1
    soft_ori = 1 - soft
    soft_mask_reproj = 1 - soft
    soft_mask_reproj = soft_mask_reproj.reshape(H // 8, 8, W // 8, 8).transpose(0, 2, 1, 3).reshape(H // 8, W // 8, 64)
    soft_mask_reproj = np.mean(soft_mask_reproj, axis=2)
    return dict(mask_png=mask_png, cond=cond, cond_ori=cond_ori.astype(np.float32), latent_mask=mask_erosion,
                masks_ero=mask_erosion_.astype(np.uint8), soft_ori=soft_ori.astype(np.float32),
                soft_pooled=soft_mask_reproj.astype(np.float32))


def holes(rng, H, W, kind):
    """valid masks with the structures that matter to a 5x5 dilation + 8x8 pooling: blobs, single pixels, borders."""
    m = np.ones((H, W), bool)
    if kind == "blobs":
        for _ in range(6):
            y, x = rng.integers(0, H), rng.integers(0, W)
            m[max(0, y - rng.integers(1, 9)):y + rng.integers(1, 9), max(0, x - rng.integers(1, 13)):x + rng.integers(1, 13)] = False
        m &= rng.random((H, W)) > 0.01                          # isolated single-pixel holes
    elif kind == "border":
        m[0, :] = False; m[:, W - 1] = False; m[H - 1, ::7] = False; m[5::11, 0] = False
    elif kind == "sparse":
        m = rng.random((H, W)) > 0.004                          # block counts around the 13-of-64 threshold
    elif kind == "none":
        pass
    elif kind == "all":
        m[:] = False
    return m


def main():
    vers = np.array([f"numpy {np.__version__}", f"cv2 {cv2.__version__}", f"PIL {PIL.__version__}",
                     "statements of DecaYale/SYN3R@402f4e9 model/diffusionGS.py:1572-1597, :1452-1491"])
    rng = np.random.default_rng(2024)
    for i, (kind, H, W) in enumerate([("blobs", 64, 96), ("border", 40, 64), ("sparse", 64, 64), ("none", 16, 24),
                                      ("all", 16, 16), ("blobs", 72, 128)]):
        mask2 = holes(rng, H, W, kind)
        warped = rng.integers(0, 256, (H, W, 3), dtype=np.uint8)
        out = reference_post_forward(warped.copy(), mask2.copy())
        path = os.path.join(HERE, f"post_fw_{i}_{kind}.npz")
        np.savez_compressed(path, versions=vers, in_warped=warped, in_mask2=mask2, **{"out_" + k: v for k, v in out.items()})
        print(os.path.basename(path), os.path.getsize(path) // 1024, "KiB")
    for i, (kind, H, W) in enumerate([("blobs", 64, 96), ("sparse", 48, 64), ("border", 40, 64)]):
        mask2 = holes(rng, H, W, kind)
        warped = (rng.random((H, W, 3), dtype=np.float32) * np.float32(255.0)).astype(np.float32)
        soft = np.exp(-(rng.random((H, W), dtype=np.float32) * 3) ** 3).astype(np.float32)
        out = reference_post_backward(warped.copy(), mask2.copy(), soft.copy())
        path = os.path.join(HERE, f"post_bw_{i}_{kind}.npz")
        np.savez_compressed(path, versions=vers, in_warped=warped, in_mask2=mask2, in_soft=soft,
                            **{"out_" + k: v for k, v in out.items()})
        print(os.path.basename(path), os.path.getsize(path) // 1024, "KiB")


if __name__ == "__main__":
    sys.exit(main())
