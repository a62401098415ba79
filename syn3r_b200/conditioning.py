"""Caller-side post-processing of warped frames on the device.

The reference does this per frame on the host with cv2 / PIL between `forward_warp` / `inverse_warp` and the SVD
pipeline (/root/reference/model/diffusionGS.py:1572-1597 `warp_images`, :1452-1491 `warp_images_bw`): hole mask ->
5x5 dilation -> blanked conditioning frame -> 8x8 block mean thresholded at 0.2 (the latent-resolution mask) and the
8x8 pooled reprojection uncertainty.  Here it is ONE launch over all frames of a trajectory
(`syn3r_condition_post_batched`, csrc/postprocess.cuh); there is no CPU path.

  condition_post_batched(...)   device tensors in, device tensors out (what TrajectoryWarp uses)
  post_forward / post_backward  host arrays in the reference's own shapes and dtypes, for drop-in use and tests
"""
from __future__ import annotations

from typing import Dict, Optional, Sequence

import numpy as np
import torch

from . import _native as nat

POST_KEYS = ("mask_png", "cond_u8", "cond_f32", "cond_ori", "latent", "soft_pooled")


def condition_post_batched(valid: torch.Tensor, img_u8: Optional[torch.Tensor] = None,
                           img_chw: Optional[torch.Tensor] = None, soft: Optional[torch.Tensor] = None, *,
                           want: Sequence[str] = ("mask_png", "cond_u8", "latent"),
                           out: Optional[Dict[str, torch.Tensor]] = None) -> Dict[str, torch.Tensor]:
    """valid [P,H,W] u8/bool (1 = covered), img_u8 [P,H,W,3] u8 or img_chw [P,3,H,W] f32, soft [P,H,W] f32: CUDA
    tensors.  `want` picks from POST_KEYS; returns key -> CUDA tensor (mask_png [P,H,W] u8 0/255, cond_u8 [P,H,W,3],
    cond_f32 / cond_ori [P,H,W,3] f32, latent [P,H/8,W/8] u8 0/1, soft_pooled [P,H/8,W/8] f32)."""
    L = nat.lib()
    assert set(want) <= set(POST_KEYS), f"unknown output in {want}"
    if valid.dtype == torch.bool:
        valid = valid.view(torch.uint8)
    assert valid.dtype == torch.uint8 and valid.dim() == 3
    P, H, W = valid.shape
    assert H % 8 == 0 and W % 8 == 0, "H and W must be multiples of 8 (diffusionGS.py:1593 reshapes to 8x8 blocks)"
    dev = valid.device
    if img_u8 is not None:
        assert img_u8.shape == (P, H, W, 3) and img_u8.dtype == torch.uint8
    if img_chw is not None:
        assert img_chw.shape == (P, 3, H, W) and img_chw.dtype == torch.float32
    if soft is not None:
        assert soft.shape == (P, H, W) and soft.dtype == torch.float32
    shapes = {"mask_png": ((P, H, W), torch.uint8), "cond_u8": ((P, H, W, 3), torch.uint8),
              "cond_f32": ((P, H, W, 3), torch.float32), "cond_ori": ((P, H, W, 3), torch.float32),
              "latent": ((P, H // 8, W // 8), torch.uint8), "soft_pooled": ((P, H // 8, W // 8), torch.float32)}
    o = dict(out or {})
    for k in want:
        shp, dt = shapes[k]
        t = o.get(k)
        if t is None:
            t = o[k] = torch.empty(shp, dtype=dt, device=dev)
        assert t.shape == torch.Size(shp) and t.dtype == dt and t.is_contiguous()
    ptr = lambda k: nat.ptr(o[k]) if k in want else None  # noqa: E731
    nat.check(L.syn3r_condition_post_batched(nat.ptr(valid), nat.ptr(img_u8), nat.ptr(img_chw), nat.ptr(soft), P, H, W,
                                             ptr("mask_png"), ptr("cond_u8"), ptr("cond_f32"), ptr("cond_ori"),
                                             ptr("latent"), ptr("soft_pooled"),
                                             torch.cuda.current_stream(dev).cuda_stream),
              "syn3r_condition_post_batched")
    return {k: o[k] for k in want}


def pack_mask_bits(mask: torch.Tensor, out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """np.packbits(mask != 0) on the device (MSB first): mask [..., n] u8 / bool CUDA, contiguous -> uint8 [(numel + 7) // 8]
    (flat).  The hole mask of forward_warp and mask_reproj of inverse_warp are 0/1 bytes; packed they cross PCIe / NVLink
    at one bit per pixel.  `unpack_mask_bits` restores them on the host."""
    if mask.dtype == torch.bool:
        mask = mask.view(torch.uint8)
    assert mask.dtype == torch.uint8 and mask.is_cuda and mask.is_contiguous()
    n = mask.numel()
    if out is None:
        out = torch.empty(((n + 7) // 8,), dtype=torch.uint8, device=mask.device)
    assert out.dtype == torch.uint8 and out.is_contiguous() and out.numel() == (n + 7) // 8 and out.device == mask.device
    with torch.cuda.device(mask.device):
        nat.check(nat.lib().syn3r_pack_mask_bits(nat.ptr(mask), n, nat.ptr(out),
                                                 torch.cuda.current_stream(mask.device).cuda_stream), "syn3r_pack_mask_bits")
    return out


def unpack_mask_bits(bits, shape) -> np.ndarray:
    """Host side of pack_mask_bits: uint8 bit array (NumPy or CPU tensor) -> bool array of `shape`."""
    b = bits.numpy() if isinstance(bits, torch.Tensor) else np.asarray(bits)
    n = int(np.prod(shape))
    return np.unpackbits(b.reshape(-1), count=n).astype(bool).reshape(shape)


def _dev():
    if not torch.cuda.is_available():
        raise nat.Syn3rNativeError("syn3r_b200.conditioning needs a CUDA device; there is no CPU path")
    return torch.device("cuda", torch.cuda.current_device())


def post_forward(warped_frame2: np.ndarray, mask2: np.ndarray) -> dict:
    """The block of diffusionGS.py:1572-1597 for ONE frame, host arrays in and out, reference dtypes:
    mask_png (H,W,3) uint8, cond (H,W,3) float32, latent_mask (H/8,W/8) float64."""
    d = _dev()
    o = condition_post_batched(torch.from_numpy(np.ascontiguousarray(mask2)).to(d)[None],
                               img_u8=torch.from_numpy(np.ascontiguousarray(warped_frame2, dtype=np.uint8)).to(d)[None],
                               want=("mask_png", "cond_f32", "latent"))
    return dict(mask_png=np.repeat(o["mask_png"][0].cpu().numpy()[..., None], 3, axis=2),
                cond=o["cond_f32"][0].cpu().numpy(), latent_mask=o["latent"][0].cpu().numpy().astype(np.float64))


def post_backward(warped_frame2: np.ndarray, mask2: np.ndarray, soft_mask_reproj: np.ndarray) -> dict:
    """The block of diffusionGS.py:1452-1491 for ONE frame: warped_frame2 (H,W,3) float32 (warped_img transposed to
    HWC as at :1444), mask2 = mask_reproj (H,W) bool, soft_mask_reproj (H,W) float32."""
    d = _dev()
    chw = torch.from_numpy(np.ascontiguousarray(np.moveaxis(np.asarray(warped_frame2, np.float32), -1, 0))).to(d)[None]
    o = condition_post_batched(torch.from_numpy(np.ascontiguousarray(mask2)).to(d)[None], img_chw=chw,
                               soft=torch.from_numpy(np.ascontiguousarray(soft_mask_reproj, dtype=np.float32)).to(d)[None],
                               want=("mask_png", "cond_f32", "cond_ori", "latent", "soft_pooled"))
    return dict(mask_png=np.repeat(o["mask_png"][0].cpu().numpy()[..., None], 3, axis=2),
                cond=o["cond_f32"][0].cpu().numpy(), cond_ori=o["cond_ori"][0].cpu().numpy(),
                latent_mask=o["latent"][0].cpu().numpy().astype(np.float64),
                soft_pooled=o["soft_pooled"][0].cpu().numpy())


def fuse_uncertainty_batched(cond_ori: torch.Tensor, gs: torch.Tensor, soft_ori: torch.Tensor, *,
                             want: Sequence[str] = ("cond_image", "masks"),
                             out: Optional[Dict[str, torch.Tensor]] = None) -> Dict[str, torch.Tensor]:
    """Uncertainty fusion of diffusionGS.py:821-854 + the clip of :876 on the device.
    cond_ori, gs [N,H,W,3] f32; soft_ori [N,H,W] f32 (= 1 - soft_mask_reproj): CUDA tensors.  `want` picks from
    cond_image [N,H,W,3] f32, masks [N,H/8,W/8] f32, geo_inten_uncertainty [N,H,W] f64."""
    L = nat.lib()
    assert set(want) <= {"cond_image", "masks", "geo_inten_uncertainty"} and len(want) > 0
    N, H, W, three = cond_ori.shape
    assert three == 3 and gs.shape == cond_ori.shape and soft_ori.shape == (N, H, W)
    assert cond_ori.dtype == gs.dtype == soft_ori.dtype == torch.float32
    assert H % 8 == 0 and W % 8 == 0, "H and W must be multiples of 8 (diffusionGS.py:843 reshapes to 8x8 blocks)"
    shapes = {"cond_image": ((N, H, W, 3), torch.float32), "masks": ((N, H // 8, W // 8), torch.float32),
              "geo_inten_uncertainty": ((N, H, W), torch.float64)}
    o = dict(out or {})
    for k in want:
        shp, dt = shapes[k]
        if o.get(k) is None:
            o[k] = torch.empty(shp, dtype=dt, device=cond_ori.device)
        assert o[k].shape == torch.Size(shp) and o[k].dtype == dt and o[k].is_contiguous()
    ptr = lambda k: nat.ptr(o[k]) if k in want else None  # noqa: E731
    nat.check(L.syn3r_fuse_uncertainty_batched(nat.ptr(cond_ori), nat.ptr(gs), nat.ptr(soft_ori), N, H, W, ptr("cond_image"),
                                               ptr("masks"), ptr("geo_inten_uncertainty"),
                                               torch.cuda.current_stream(cond_ori.device).cuda_stream),
              "syn3r_fuse_uncertainty_batched")
    return {k: o[k] for k in want}


def nearby_pairs(n_frames: int, window_radius: int = 1):
    """(pair_src, pair_tgt, pair_first): the (ref_i -> cur_i) warps of diffusionGS.py:1322-1326 in the reference's loop
    order, grouped by target; pair_first[n] .. pair_first[n+1] are frame n's pairs.  The fusion kernel adds a frame's
    neighbours in list order, which is torch's mean over a stack of up to 4 tensors (its reduce keeps 4 running
    accumulators): window_radius <= 2.  The reference uses 1 (diffusionGS.py:1313)."""
    assert 1 <= window_radius <= 2, "k_nearby_uncertainty follows torch's reduction order for at most 4 neighbours"
    ps, pt, first = [], [], [0]
    for cur in range(n_frames):
        for ref in range(cur - window_radius, cur + window_radius + 1):
            if ref == cur or ref < 0 or ref >= n_frames:
                continue
            ps.append(ref)
            pt.append(cur)
        first.append(len(ps))
    return np.asarray(ps, np.int32), np.asarray(pt, np.int32), np.asarray(first, np.int32)


def nearby_uncertainty_batched(soft: torch.Tensor, warped: torch.Tensor, images_chw: torch.Tensor,
                               pair_first: torch.Tensor, *, sigma: float = 0.1,
                               out: Optional[Dict[str, torch.Tensor]] = None) -> Dict[str, torch.Tensor]:
    """The tail of consistency_check_from_nearby_images_bw (diffusionGS.py:1347-1355) for all frames in one launch.
    soft [P,H,W], warped [P,3,H,W] (inverse_warp_batched outputs), images_chw [N,3,H,W] f32, pair_first [N+1] i32: CUDA
    tensors.  Returns uncertainty [N,H,W] f32 (= 1 - mean soft) and intensity_uncertainty [N,H,W] f32."""
    L = nat.lib()
    N, three, H, W = images_chw.shape
    P = soft.shape[0]
    assert three == 3 and soft.shape == (P, H, W) and warped.shape == (P, 3, H, W) and pair_first.shape == (N + 1,)
    assert soft.dtype == warped.dtype == images_chw.dtype == torch.float32 and pair_first.dtype == torch.int32
    assert soft.is_contiguous() and warped.is_contiguous() and images_chw.is_contiguous()
    o = dict(out or {})
    for k in ("uncertainty", "intensity_uncertainty"):
        if o.get(k) is None:
            o[k] = torch.empty((N, H, W), dtype=torch.float32, device=soft.device)
        assert o[k].shape == (N, H, W) and o[k].dtype == torch.float32 and o[k].is_contiguous()
    inv_sigma = float(np.float32(1.0) / np.float32(sigma))     # tensor / python scalar on CUDA: * (1.0f / 0.1f)
    nat.check(L.syn3r_nearby_uncertainty_batched(nat.ptr(soft), nat.ptr(warped), nat.ptr(images_chw), nat.ptr(pair_first),
                                                 N, H, W, inv_sigma, nat.ptr(o["uncertainty"]),
                                                 nat.ptr(o["intensity_uncertainty"]),
                                                 torch.cuda.current_stream(soft.device).cuda_stream),
              "syn3r_nearby_uncertainty_batched")
    return {k: o[k] for k in ("uncertainty", "intensity_uncertainty")}


def consistency_check_from_nearby_images_bw(intrinsics, interpolated_poses, images, depths, *, window_radius: int = 1,
                                            bandwidth: float = 10.0, device=None):
    """Drop-in for diffusionGS.py:1300-1361: every frame is warped backward from its neighbours within the window
    (inverse_warp, bandwidth 10) and the soft reprojection masks / warped colours are averaged.  2 (N-1) warps and
    the fusion run as two launches instead of the reference's per-pair Python loop.
    interpolated_poses: N w2c 4x4; images: N (H,W,3) in [0,255]; depths: N (H,W).  Returns the reference's two lists
    as tensors: uncertainty_masks [N,H,W] f32, intensity_uncertainty_masks [N,H,W] f32 (CUDA)."""
    from . import warp as w
    dev = torch.device(device or "cuda")
    poses = np.stack([np.asarray(p, np.float32) for p in interpolated_poses])
    n = poses.shape[0]
    ps, pt, first = nearby_pairs(n, window_radius)
    if np.any(first[1:] == first[:-1]):
        raise RuntimeError("a frame without a neighbour: the reference's torch.stack(masks) raises on an empty list "
                           "(diffusionGS.py:1347)")
    up = lambda a, dt: torch.as_tensor(np.ascontiguousarray(a), dtype=dt).to(dev)  # noqa: E731
    img = up(np.stack([np.asarray(i, np.float32) for i in images]), torch.float32).permute(0, 3, 1, 2).contiguous()
    dep = up(np.stack([np.asarray(d, np.float32) for d in depths]), torch.float32)
    o = w.inverse_warp_batched(img, dep, dep, np.asarray(intrinsics, np.float32), up(w.inverse_pair_params(poses, poses, ps, pt),
                                                                                   torch.float32),
                               up(ps, torch.int32), up(pt, torch.int32), bandwidth=bandwidth,
                               want=("warped_img", "soft_mask_reproj"))
    f = nearby_uncertainty_batched(o["soft_mask_reproj"], o["warped_img"], img, up(first, torch.int32))
    return f["uncertainty"], f["intensity_uncertainty"]


# ---------------------------------------------------------------------------------------------------------
# Drop-in `DiffusionGS.warp_images` / `DiffusionGS.warp_images_bw`
# ---------------------------------------------------------------------------------------------------------

_plan_cache: dict = {}


def _diffusion_size(self):
    new_width = getattr(self, "diffusion_width", 1024)
    new_height = getattr(self, "diffusion_height", 576)
    assert new_width == 1024 and new_height == 576                     # diffusionGS.py:1525, :1382
    return new_width, new_height


def _save_pngs(jobs):
    """[(path, uint8 array)] -> PNG files, encoded on a small thread pool (zlib releases the GIL)."""
    import PIL.Image
    from concurrent.futures import ThreadPoolExecutor
    with ThreadPoolExecutor(max_workers=8) as ex:
        list(ex.map(lambda j: PIL.Image.fromarray(j[1]).save(j[0]), jobs))


def warp_images(self, intrinsics, interpolated_poses, image_l, image_r=None, depth_l=None, depth_r=None, save_path='',
                save_file_name_prefix=''):
    """Drop-in for `DiffusionGS.warp_images` (/root/reference/model/diffusionGS.py:1512-1605): same arguments, same
    return tuple `(image_o / 255., image_o2 / 255. | None, masks (23|24,72,128) float64 tensor, cond_image list of
    (576,1024,3) float32)`, same two PNG files per frame under `save_path`.  Bind it as the method
    (`DiffusionGS.warp_images = syn3r_b200.conditioning.warp_images`) or call it with any object carrying
    `diffusion_width` / `diffusion_height` (None: 1024 x 576).

    The host side does what the reference does on the host (`cv2.resize` LINEAR / NEAREST of :1531-1549); the 23 or 24
    `forward_warp` calls of the loop (:1554-1599) with its `i < 12` source switch and the per-frame cv2 / PIL block
    (:1572-1597) are ONE batched forward warp + ONE post-processing launch (`TrajectoryWarp(post=True)`)."""
    import os
    import cv2
    from . import trajectory, warp as w
    new_width, new_height = _diffusion_size(self)
    pose_s_l = interpolated_poses[0]
    image_l = cv2.resize(image_l, (new_width, new_height), interpolation=cv2.INTER_LINEAR)
    image_o = image_l
    depth_l = cv2.resize(depth_l, (new_width, new_height), interpolation=cv2.INTER_NEAREST)
    pose_s_r = interpolated_poses[-1]
    interp_num = len(interpolated_poses) - 1
    assert interp_num == 24                                             # :1544
    image_o2 = None
    if image_r is not None:
        interp_num = interp_num - 1
        image_r = cv2.resize(image_r, (new_width, new_height), interpolation=cv2.INTER_LINEAR)
        image_o2 = image_r
        depth_r = cv2.resize(depth_r, (new_width, new_height), interpolation=cv2.INTER_NEAREST)
    H, W = new_height, new_width
    two = image_r is not None
    S = 2 if two else 1
    pair_src = np.array([0 if (not two or i < 12) else 1 for i in range(interp_num)], np.int32)        # :1555-1567
    pair_tgt = np.arange(interp_num, dtype=np.int32)
    dev = _dev()
    key = ("fw", S, interp_num, H, W, str(dev))
    tw = _plan_cache.get(key)
    if tw is None:
        tw = _plan_cache[key] = trajectory.TrajectoryWarp(S, interp_num, H, W, pair_src, pair_tgt, device=dev, consistency=False,
                                                          want_f32=False, want_u8=True, want_winner=False, post=True)
    pose_src = np.stack([np.asarray(pose_s_l)] + ([np.asarray(pose_s_r)] if two else []))
    pose_tgt = np.stack([np.asarray(interpolated_poses[i + 1]) for i in range(interp_num)])
    tw.set_cameras(pose_src, pose_tgt, np.asarray(intrinsics))
    rgb = np.stack([image_l] + ([image_r] if two else []))
    dep = np.stack([depth_l] + ([depth_r] if two else []))
    tw.upload(torch.from_numpy(w.pack_rgbd(rgb, dep)))
    tw.run()
    o = tw.out_post
    png = o["post_mask_png"].cpu().numpy()
    cond_u8 = o["post_cond_u8"].cpu().numpy()
    latent = o["post_latent"].cpu().numpy()
    if int(tw.out_fwd["status"].max().item()) != 0:
        raise AssertionError("warped colour outside [0, 256] (forward_warp.py:121-125)")
    jobs = []
    for i in range(interp_num):
        stem = os.path.join(save_path, save_file_name_prefix + '_' + str(i).zfill(4))
        jobs.append((stem + "_mask.png", np.repeat(png[i][:, :, None], 3, axis=2)))                        # :1580
        jobs.append((stem + ".png", cond_u8[i]))                                                            # :1587
    _save_pngs(jobs)
    cond_image = [np.asarray(cond_u8[i], dtype=np.float32) / 255. for i in range(interp_num)]              # :1590
    masks = torch.from_numpy(latent.astype(np.float64))                                                     # :1592-1600
    return image_o / 255., (image_o2 / 255. if two else None), masks, cond_image


def warp_images_bw(self, intrinsics, interpolated_poses, image_l, image_r=None, depth_l=None, depth_r=None, save_path=None,
                   save_file_name_prefix='', return_ero_mask=False):
    """Drop-in for `DiffusionGS.warp_images_bw` (/root/reference/model/diffusionGS.py:1367-1510): same arguments and
    return values, incl. the `aux` dict when `return_ero_mask` (:1497-1504) and the three PNG files per frame.
    `self.render_GS(pose=...)` is the caller's 3DGS renderer (outside the hot path) and is called once per frame in
    the reference's order; the 23 or 24 `inverse_warp` calls (:1438-1441) and the per-frame host block (:1452-1491)
    are ONE batched backward warp + ONE post-processing launch.  Files are written when `save_path` is not None (the
    reference raises on `os.path.join(None, ...)`)."""
    import os
    import cv2
    from . import warp as w
    new_width, new_height = _diffusion_size(self)
    pose_s_l = interpolated_poses[0]
    image_l = cv2.resize(image_l, (new_width, new_height), interpolation=cv2.INTER_NEAREST)
    image_o = image_l
    depth_l = cv2.resize(depth_l, (new_width, new_height), interpolation=cv2.INTER_NEAREST)
    interp_num = len(interpolated_poses) - 1
    assert interp_num == 24                                             # :1395
    image_o2, pose_s_r = None, None
    if image_r is not None:
        interp_num = interp_num - 1
        pose_s_r = interpolated_poses[-1]
        image_r = cv2.resize(image_r, (new_width, new_height), interpolation=cv2.INTER_NEAREST)
        image_o2 = image_r
        depth_r = cv2.resize(depth_r, (new_width, new_height), interpolation=cv2.INTER_NEAREST)
    two = depth_r is not None                                           # the source switch keys on depth_r here (:1416)
    H, W = new_height, new_width
    tgt = []
    for i in range(interp_num):                                         # the caller's renderer, in the reference's order
        _, _rendered_t, rendered_depth_t = self.render_GS(pose=interpolated_poses[i + 1])
        tgt.append(cv2.resize(rendered_depth_t, (new_width, new_height), interpolation=cv2.INTER_NEAREST))      # :1434
    dev = _dev()
    pair_src = np.array([0 if (not two or i < 12) else 1 for i in range(interp_num)], np.int32)
    pair_tgt = np.arange(interp_num, dtype=np.int32)
    pose_src = np.stack([np.asarray(pose_s_l, np.float32)] + ([np.asarray(pose_s_r, np.float32)] if two else []))
    pose_tgt = np.stack([np.asarray(interpolated_poses[i + 1], np.float32) for i in range(interp_num)])
    up = lambda a, dt: torch.as_tensor(np.ascontiguousarray(a), dtype=dt).to(dev)  # noqa: E731
    img = up(np.stack([image_l] + ([image_r] if two else [])), torch.float32).permute(0, 3, 1, 2).contiguous()
    dep = up(np.stack([depth_l] + ([depth_r] if two else [])), torch.float32)
    tgd = up(np.stack(tgt), torch.float32)
    iw = w.inverse_warp_batched(img, dep, tgd, np.asarray(intrinsics, np.float32),
                                up(w.inverse_pair_params(pose_src, pose_tgt, pair_src, pair_tgt), torch.float32),
                                up(pair_src, torch.int32), up(pair_tgt, torch.int32),
                                want=("warped_img", "mask_reproj", "soft_mask_reproj"))
    o = condition_post_batched(iw["mask_reproj"], img_chw=iw["warped_img"], soft=iw["soft_mask_reproj"],
                               want=("mask_png", "cond_u8", "cond_ori", "latent", "soft_pooled"))
    png = o["mask_png"].cpu().numpy()
    cond_u8 = o["cond_u8"].cpu().numpy()
    latent = o["latent"].cpu().numpy()
    if save_path is not None:
        jobs = []
        for i in range(interp_num):
            stem = os.path.join(save_path, save_file_name_prefix + '_' + str(i).zfill(4))
            jobs.append((stem + "_depth.png", np.uint8(tgt[i])))                                            # :1447
            jobs.append((stem + "_mask.png", np.repeat(png[i][:, :, None], 3, axis=2)))                    # :1460
            jobs.append((stem + ".png", cond_u8[i]))                                                        # :1471
        _save_pngs(jobs)
    cond_image = [np.asarray(cond_u8[i], dtype=np.float32) / 255. for i in range(interp_num)]              # :1475
    masks = torch.from_numpy(latent.astype(np.float64))                                                     # :1477-1482, :1494
    ret = (image_o / 255., (image_o2 / 255. if image_r is not None else None), masks, cond_image)
    if not return_ero_mask:
        return ret
    cond_ori = o["cond_ori"].cpu().numpy()
    aux = {"masks_ero": np.repeat((png // 255)[..., None], 3, axis=3).astype(np.uint8),                    # :1463-1465, :1497
           "soft_masks_reproj": o["soft_pooled"].cpu().to(dtype=torch.float32),                             # :1486-1496
           "soft_masks_reproj_ori": (1 - iw["soft_mask_reproj"].cpu().numpy()).astype(np.float32),          # :1485, :1498
           "cond_images_ori": [cond_ori[i] for i in range(interp_num)]}                                     # :1467
    return ret + (aux,)
