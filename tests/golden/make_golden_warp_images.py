"""Golden fixtures of the callers' whole methods: tests/golden/warp_images_*.npz.

`DiffusionGS.warp_images` (/root/reference/model/diffusionGS.py:1512-1605) and `DiffusionGS.warp_images_bw`
(:1367-1510) are sliced from the file AS WHOLE METHODS and exec-ed unpatched (tests/golden/ref_exec.py) around the LIVE
`forward_warp` / `inverse_warp` of the reference: the `cv2.resize` to 1024 x 576, the `i < 12` source switch, the
per-frame loop, the dilation / blanking / pooling block, the PNG files and the return tuple are all the reference's
own code running at its own (asserted) size.  `self` is a stub carrying `diffusion_width/height` and a `render_GS`
that hands back stored depth maps (the 3DGS renderer is outside the hot path); `torch` maps the hard-coded
device='cuda' of :1438-1441 to 'cpu'.

A 23-frame result is ~160 MB of float32, so the fixture keeps: the latent masks (all frames), the dilated masks as
bits (all frames), per-frame sums of the uint8 conditioning frames (all frames), four conditioning frames in full,
probes of the resized anchors, and for the backward variant the pooled soft masks (all frames) plus rows of the
full-resolution ones for the four full frames.  The PNG files the method writes are read back and must equal the returned arrays.

    python tests/golden/make_golden_warp_images.py        (~1 min)
"""
import os
import sys
import tempfile
import types

import cv2
import numpy as np
import PIL.Image
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
import make_golden as mg  # noqa: E402  (device shim, live reference modules, synth)
import ref_exec  # noqa: E402

SRC = "model/diffusionGS.py"
FULL = (0, 11, 12, 22)                 # frames stored in full (both sides of the source switch at i = 12)
ROWS = slice(3, None, 24)              # rows of the full-resolution float planes that are stored


def scene(seed, h0=144, w0=256):
    """Two anchors + 25 w2c poses at the diffusion resolution's intrinsics; anchors rendered at GS resolution h0 x w0
    (the reference resizes them, :1531-1549).  Integer-valued colours: the warped frames stay on the 8-bit lattice."""
    H, W = 576, 1024
    tr = mg.synth.make_trajectory(2, 25, h0, w0, seed=seed, smooth_rgb=True, pose_step=2.0)
    K = mg.synth.make_intrinsics(H, W)
    poses = [p for p in tr.pose_tgt]                      # poses[0] = left anchor, poses[24] = right anchor
    img = np.floor(tr.rgb).astype(np.float32)
    # pseudo depth the GS would render at each interpolated pose (half the anchors' resolution: the reference resizes
    # every map to 1024 x 576 on its own, :1434), consistent with the anchors' depth
    rng = np.random.default_rng(seed + 7)
    tgt_depth = np.stack([mg.synth.depth_field(h0 // 2, w0 // 2, rng, phase=0.3 * t / 24.0) * np.float32(1.003) for t in range(25)])
    return K, poses, img[0], img[1], tr.depth[0], tr.depth[1], tgt_depth


def run_forward(name, seed, single_anchor):
    fw, _ = ref_exec.live_warp_modules()
    K, poses, il, ir, dl, dr, _ = scene(seed)
    fn, tag, desc = ref_exec.exec_function(SRC, 1512, 1605, "warp_images",
                                           dict(np=np, cv2=cv2, PIL=PIL, torch=torch, os=os, forward_warp=fw.forward_warp))
    assert not desc
    self = types.SimpleNamespace(diffusion_width=1024, diffusion_height=576)
    with tempfile.TemporaryDirectory() as tmp:
        if single_anchor:
            image_o, image_o2, masks, cond = fn(self, K, poses, il, None, dl, None, save_path=tmp, save_file_name_prefix="7")
        else:
            image_o, image_o2, masks, cond = fn(self, K, poses, il, ir, dl, dr, save_path=tmp, save_file_name_prefix="7")
        n = len(cond)
        png_mask = np.stack([np.array(PIL.Image.open(os.path.join(tmp, f"7_{i:04d}_mask.png"))) for i in range(n)])
        png_frame = np.stack([np.array(PIL.Image.open(os.path.join(tmp, f"7_{i:04d}.png"))) for i in range(n)])
    cond_u8 = np.stack([np.round(c * 255).astype(np.uint8) for c in cond])
    assert all(np.array_equal(cond_u8[i].astype(np.float32) / 255., cond[i]) for i in range(n))
    assert np.array_equal(png_frame, cond_u8) and (png_mask[..., 0] == png_mask[..., 1]).all()
    assert masks.dtype == torch.float64 and tuple(masks.shape) == (n, 72, 128)
    arrs = dict(K=K, poses=np.stack(poses), image_l=il, depth_l=dl, n_frames=np.int64(n), ref_source=np.array(tag),
                out_image_o_probe=image_o[::37, ::41].astype(np.float32), out_masks=masks.numpy().astype(np.uint8),
                out_mask_png_bits=np.packbits(png_mask[..., 0] > 0), out_cond_u8_sum=cond_u8.reshape(n, -1).sum(1, dtype=np.int64),
                out_cond_u8_full=cond_u8[[f for f in FULL if f < n]], full_frames=np.array([f for f in FULL if f < n]))
    if not single_anchor:
        arrs.update(image_r=ir, depth_r=dr, out_image_o2_probe=image_o2[::37, ::41].astype(np.float32))
    else:
        assert image_o2 is None
    mg.save(name, **arrs)


def run_backward(name, seed):
    fw, _ = ref_exec.live_warp_modules()
    K, poses, il, ir, dl, dr, tgt_depth = scene(seed)
    calls = []

    def render_GS(pose=None):          # the reference asks for the pseudo view at interpolated_poses[i + 1] (:1431-1432)
        t = next(k for k, p in enumerate(poses) if p is pose)
        calls.append(t)
        return None, None, tgt_depth[t]

    fn, tag, desc = ref_exec.exec_function(SRC, 1367, 1510, "warp_images_bw",
                                           dict(np=np, cv2=cv2, PIL=PIL, torch=ref_exec.TorchOnCpu(), os=os,
                                                inverse_warp=fw.inverse_warp))
    assert not desc
    self = types.SimpleNamespace(diffusion_width=1024, diffusion_height=576, render_GS=render_GS)
    with tempfile.TemporaryDirectory() as tmp:
        image_o, image_o2, masks, cond, aux = fn(self, K, poses, il, ir, dl, dr, save_path=tmp, save_file_name_prefix="7",
                                                 return_ero_mask=True)
        n = len(cond)
        png_mask = np.stack([np.array(PIL.Image.open(os.path.join(tmp, f"7_{i:04d}_mask.png"))) for i in range(n)])
        png_frame = np.stack([np.array(PIL.Image.open(os.path.join(tmp, f"7_{i:04d}.png"))) for i in range(n)])
        png_depth = np.stack([np.array(PIL.Image.open(os.path.join(tmp, f"7_{i:04d}_depth.png"))) for i in range(n)])
    assert calls == list(range(1, n + 1))
    cond_u8 = np.stack([np.round(c * 255).astype(np.uint8) for c in cond])
    assert all(np.array_equal(cond_u8[i].astype(np.float32) / 255., cond[i]) for i in range(n))
    assert np.array_equal(png_frame, cond_u8)
    assert np.array_equal(aux["masks_ero"][..., 0] > 0, png_mask[..., 0] > 0) and aux["masks_ero"].dtype == np.uint8
    ori = np.stack(aux["cond_images_ori"])                                   # warped_img / 255., float32 (:1467)
    ori_u8 = np.round(ori * 255).astype(np.uint8)
    assert ori.dtype == np.float32 and np.array_equal((ori_u8 / np.float32(255.)).astype(np.float32), ori)   # 8-bit lattice
    assert aux["soft_masks_reproj"].dtype == torch.float32 and aux["soft_masks_reproj_ori"].dtype == np.float32
    mg.save(name, K=K, poses=np.stack(poses), image_l=il, depth_l=dl, image_r=ir, depth_r=dr, tgt_depth=tgt_depth,
            n_frames=np.int64(n), ref_source=np.array(tag), out_image_o_probe=image_o[::37, ::41].astype(np.float32),
            out_image_o2_probe=image_o2[::37, ::41].astype(np.float32), out_masks=masks.numpy().astype(np.uint8),
            out_mask_png_bits=np.packbits(png_mask[..., 0] > 0), out_cond_u8_sum=cond_u8.reshape(n, -1).sum(1, dtype=np.int64),
            out_cond_u8_full=cond_u8[list(FULL)], out_cond_ori_u8_full=ori_u8[list(FULL)], full_frames=np.array(FULL),
            out_cond_ori_u8_sum=ori_u8.reshape(n, -1).sum(1, dtype=np.int64),
            out_depth_png_probe=png_depth[:, ::37, ::41], out_soft_pooled=aux["soft_masks_reproj"].numpy(),
            out_soft_ori_rows=aux["soft_masks_reproj_ori"][list(FULL)][:, ROWS])


if __name__ == "__main__":
    run_forward("warp_images_fw_two_anchors", 31, single_anchor=False)
    run_forward("warp_images_fw_single_anchor", 32, single_anchor=True)
    run_backward("warp_images_bw_two_anchors", 33)
