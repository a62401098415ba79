"""Golden fixtures of the uncertainty fusion: tests/golden/fusion_*.npz.

EXECUTES THE REFERENCE'S OWN SOURCE TEXT (tests/golden/ref_exec.py): the statements of
/root/reference/model/diffusionGS.py:821-828 (`get_intensity_confidence`), :831-862 (the body of
`if 'Prob' in self.diffusion_type:` up to the `np.where(unc > 0.5, gs, warp)`) and :889 (`np.clip`) are sliced from the
file, dedented and exec-ed with `aux`, `pseudo_images`, `np`, `torch` in the namespace; the callers cannot be imported
(SURVEY §8c caveat 3).  Only patch: the latent grid `reshape(72,8,128,8)` -> `reshape(H8,8,W8,8)` for the small cases;
`fusion_3` runs UNPATCHED at 576x1024.  float32 exp / pow are SIMD library calls whose last bit depends on the host:
consumers compare with a tolerance (1e-6) and exclude the |unc - 0.5| <= 1e-6 band of the `np.where`.

    python tests/golden/make_golden_fusion.py
"""
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
import ref_exec  # noqa: E402

SRC = "model/diffusionGS.py"
GRID_PATCH = (("reshape(72,8,128,8)", "reshape(H8,8,W8,8)"), ("reshape(72,128,64)", "reshape(H8,W8,64)"))


def reference_fusion(cond_images_ori, gs_images, soft_masks_reproj_ori, patch=True):
    """cond_images_ori (N,H,W,3) f32 = warped / 255., gs_images (N,H,W,3) f32, soft_masks_reproj_ori (N,H,W) f32 =
    1 - soft_mask_reproj, handed over exactly as `_interpolate_between_gs_v3` holds them: `aux` of warp_images_bw
    (:1497-1504) and `pseudo_images` = [left anchor] + interpolated frames + [right anchor] (:800-805)."""
    N, H, W, _ = cond_images_ori.shape
    ns = dict(np=np, torch=torch, H8=H // 8, W8=W // 8,
              aux={"soft_masks_reproj": None, "soft_masks_reproj_ori": soft_masks_reproj_ori,
                   "cond_images_ori": list(cond_images_ori)},
              pseudo_images=[None] + list(gs_images) + [None])
    ns["pseudo_images"][0] = ns["pseudo_images"][-1] = gs_images[0]      # the anchors are sliced away by [1:-1] (:841)
    tags = [ref_exec.exec_block(SRC, 821, 828, ns)[0]]
    tag, desc = ref_exec.exec_block(SRC, 831, 862, ns, GRID_PATCH if patch else ())
    tags.append(tag)
    tags.append(ref_exec.exec_block(SRC, 889, 889, ns)[0])
    out = dict(masks=ns["masks"].numpy(), cond_image=np.stack(ns["cond_image"]), geo_inten_uncertainty=ns["geo_inten_uncertainty"])
    return out, tags, desc


def inputs(rng, N, H, W):
    gs = rng.random((N, H, W, 3), dtype=np.float32)
    cond = (gs + np.float32(0.25) * rng.standard_normal((N, H, W, 3)).astype(np.float32) *
            (rng.random((N, H, W, 1)) > 0.5)).astype(np.float32)
    cond = np.clip(cond, 0, 1.2).astype(np.float32)
    cond[:, 5:9, 7:15] = 0                                 # blanked pixels: uncertainty_mask = 1
    soft_ori = (1 - np.exp(-(rng.random((N, H, W), dtype=np.float32) * 2) ** 3)).astype(np.float32)
    return cond, gs, soft_ori


def main():
    rng = np.random.default_rng(99)
    for i, (N, H, W) in enumerate([(2, 48, 64), (3, 32, 40), (1, 72, 128)]):
        cond, gs, soft_ori = inputs(rng, N, H, W)
        out, tags, desc = reference_fusion(cond.copy(), gs.copy(), soft_ori.copy())
        path = os.path.join(HERE, f"fusion_{i}.npz")
        np.savez_compressed(path, versions=ref_exec.versions("exec of the reference source", *tags, "patches: " + "; ".join(desc)),
                            in_cond_ori=cond, in_gs=gs, in_soft_ori=soft_ori, **{"out_" + k: v for k, v in out.items()})
        print(os.path.basename(path), os.path.getsize(path) // 1024, "KiB", {k: (v.dtype, v.shape) for k, v in out.items()})
    # the statements as they stand (72 x 128 latent grid), one 576 x 1024 frame; inputs on an 8-bit lattice so that the
    # archive stays small (the outputs are reconstructed by the consumer: cond_image = where(sel, gs, cond))
    H, W = 576, 1024
    yy, xx = np.mgrid[0:H // 4, 0:W // 4].astype(np.float32) * 4
    gs8 = np.stack([(127 + 100 * np.sin(xx / 41) * np.cos(yy / 29)), (127 + 90 * np.cos(xx / 53 + yy / 31)), 0.2 * xx + 0.1 * yy], -1)
    gs8 = np.clip(gs8, 0, 255).astype(np.uint8).repeat(4, 0).repeat(4, 1)[None]        # 4x4 blocks: small archive
    d8 = (rng.integers(-40, 41, (1, H // 4, W // 4, 3)) * (rng.random((1, H // 4, W // 4, 1)) > 0.6)).astype(np.int16)
    d8 = d8.repeat(4, 1).repeat(4, 2)                      # blocky differences: the archive compresses
    cond8 = np.clip(gs8.astype(np.int16) + d8, 0, 255).astype(np.uint8)
    cond8[:, 100:140, 300:420] = 0
    soft8 = rng.integers(0, 256, (1, H // 4, W // 4), dtype=np.uint8).repeat(4, 1).repeat(4, 2)
    cond, gs, soft_ori = (cond8 / np.float32(255)).astype(np.float32), (gs8 / np.float32(255)).astype(np.float32), \
        (soft8 / np.float32(255)).astype(np.float32)
    out, tags, desc = reference_fusion(cond.copy(), gs.copy(), soft_ori.copy(), patch=False)
    assert not desc
    sel = out["geo_inten_uncertainty"] > 0.5
    assert np.array_equal(out["cond_image"], np.clip(np.where(sel, gs, cond), 0, 1))
    path = os.path.join(HERE, "fusion_3.npz")
    np.savez_compressed(path, versions=ref_exec.versions("exec of the reference source, unpatched, 576x1024", *tags),
                        in_cond_u8=cond8, in_gs_u8=gs8, in_soft_u8=soft8, out_masks=out["masks"],
                        out_sel_bits=np.packbits(sel[..., 0]),
                        out_geo_inten_uncertainty_rows9=out["geo_inten_uncertainty"][:, 1::36])    # every 36th row, float64
    print(os.path.basename(path), os.path.getsize(path) // 1024, "KiB")


if __name__ == "__main__":
    main()
