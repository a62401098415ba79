"""Golden fixtures of the caller-side post-processing: tests/golden/post_*.npz.

The callers (`model/diffusionGS.py`) cannot be imported here (FSGS, open3d, diffusers are absent, SURVEY §8c
caveat 3), so the generator EXECUTES THE REFERENCE'S OWN SOURCE TEXT for this block: the loop bodies

  forward branch   /root/reference/model/diffusionGS.py:1572-1597  (`warp_images`)
  backward branch  /root/reference/model/diffusionGS.py:1452-1491  (`warp_images_bw`)

are sliced from the file at generation time, dedented and `exec`-ed (tests/golden/ref_exec.py) in a namespace that
supplies their free variables (`mask2`, `warped_frame2`, `warp_dict`, `save_path` = a temporary directory, the
result lists, `np`, `cv2`, `PIL`, `torch`, `os`).  The mask png is read back from the FILE the reference's
`mask_erosion.save(...)` wrote.  The only patch: the hard-coded latent grid `reshape(72,8,128,8)` /
`reshape(72,128,64)` of a 576x1024 frame becomes `reshape(H8,8,W8,8)` / `reshape(H8,W8,64)` so that small frames can
be stored; each fixture records the slice's SHA-256 and the patch (`versions`).  The same statements run UNPATCHED at
576x1024 inside the whole-method fixtures of make_golden_warp_images.py.

    python tests/golden/make_golden_post.py
"""
import os
import sys
import tempfile

import cv2
import numpy as np
import PIL.Image
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
import ref_exec  # noqa: E402

SRC = "model/diffusionGS.py"
GRID_PATCH = (("reshape(72,8,128,8)", "reshape(H8,8,W8,8)"), ("reshape(72,128,64)", "reshape(H8,W8,64)"))


def _namespace(H, W, tmp):
    return dict(np=np, cv2=cv2, PIL=PIL, torch=torch, os=os, H8=H // 8, W8=W // 8, save_path=tmp,
                save_file_name_prefix="fx", i=0, cond_image=[], masks=[], masks_ero=[], cond_images_ori=[],
                soft_masks_reproj=[], soft_masks_reproj_ori=[])


def reference_post_forward(warped_frame2, mask2):
    """diffusionGS.py:1572-1597 executed from the reference file."""
    H, W = mask2.shape
    with tempfile.TemporaryDirectory() as tmp:
        ns = _namespace(H, W, tmp)
        ns.update(warped_frame2=warped_frame2, mask2=mask2)
        tag, desc = ref_exec.exec_block(SRC, 1572, 1597, ns, GRID_PATCH)
        mask_png = np.array(PIL.Image.open(os.path.join(tmp, "fx_0000_mask.png")))
        frame_png = np.array(PIL.Image.open(os.path.join(tmp, "fx_0000.png")))
    cond = ns["cond_image"][0]
    assert np.array_equal(frame_png, np.asarray(ns["warped_frame2"]))
    return dict(mask_png=mask_png, cond=cond, latent_mask=ns["masks"][0][0].numpy(),
                masks_ero=ns["mask_erosion_"].astype(np.uint8)), tag, desc


class _T:                       # stands in for the CUDA tensors of warp_dict: `.cpu().numpy()`
    def __init__(self, a):
        self.a = a

    def cpu(self):
        return self

    def numpy(self):
        return self.a


def reference_post_backward(warped_frame2, mask2, soft):
    """diffusionGS.py:1452-1491 executed from the reference file: `mask2` = warp_dict['mask_reproj'] (bool),
    `warped_frame2` = warped_img as HWC float32, `soft` = warp_dict['soft_mask_reproj'] (H,W) float32."""
    H, W = mask2.shape
    with tempfile.TemporaryDirectory() as tmp:
        ns = _namespace(H, W, tmp)
        ns.update(warped_frame2=warped_frame2, mask2=mask2, warp_dict={"soft_mask_reproj": _T(soft)})
        tag, desc = ref_exec.exec_block(SRC, 1452, 1491, ns, GRID_PATCH)
        mask_png = np.array(PIL.Image.open(os.path.join(tmp, "fx_0000_mask.png")))
    return dict(mask_png=mask_png, cond=ns["cond_image"][0], cond_ori=ns["cond_images_ori"][0].astype(np.float32),
                latent_mask=ns["masks"][0][0].numpy(), masks_ero=ns["masks_ero"][0].astype(np.uint8),
                soft_ori=ns["soft_masks_reproj_ori"][0].astype(np.float32),
                soft_pooled=ns["soft_masks_reproj"][0][0].numpy().astype(np.float32)), tag, desc


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
    rng = np.random.default_rng(2024)
    for i, (kind, H, W) in enumerate([("blobs", 64, 96), ("border", 40, 64), ("sparse", 64, 64), ("none", 16, 24),
                                      ("all", 16, 16), ("blobs", 72, 128)]):
        mask2 = holes(rng, H, W, kind)
        warped = rng.integers(0, 256, (H, W, 3), dtype=np.uint8)
        out, tag, desc = reference_post_forward(warped.copy(), mask2.copy())
        path = os.path.join(HERE, f"post_fw_{i}_{kind}.npz")
        np.savez_compressed(path, versions=ref_exec.versions("exec of the reference source", tag, "patches: " + "; ".join(desc)),
                            in_warped=warped, in_mask2=mask2, **{"out_" + k: v for k, v in out.items()})
        print(os.path.basename(path), os.path.getsize(path) // 1024, "KiB")
    for i, (kind, H, W) in enumerate([("blobs", 64, 96), ("sparse", 48, 64), ("border", 40, 64)]):
        mask2 = holes(rng, H, W, kind)
        warped = (rng.random((H, W, 3), dtype=np.float32) * np.float32(255.0)).astype(np.float32)
        soft = np.exp(-(rng.random((H, W), dtype=np.float32) * 3) ** 3).astype(np.float32)
        out, tag, desc = reference_post_backward(warped.copy(), mask2.copy(), soft.copy())
        path = os.path.join(HERE, f"post_bw_{i}_{kind}.npz")
        np.savez_compressed(path, versions=ref_exec.versions("exec of the reference source", tag, "patches: " + "; ".join(desc)),
                            in_warped=warped, in_mask2=mask2, in_soft=soft, **{"out_" + k: v for k, v in out.items()})
        print(os.path.basename(path), os.path.getsize(path) // 1024, "KiB")


if __name__ == "__main__":
    sys.exit(main())
