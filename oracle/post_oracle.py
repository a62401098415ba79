"""CPU restatement of the caller-side post-processing of a warped frame.

TEST INFRASTRUCTURE ONLY: imported by tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg as the
checker, never by the product (syn3r_b200/ has no CPU path).

Follows /root/reference/model/diffusionGS.py:1572-1597 (`warp_images`, forward branch) and :1452-1491
(`warp_images_bw`, backward branch) in plain NumPy, without cv2 / PIL:

  hole      = 1 - mask2, binarised                                   (:1573-1575 / :1453-1455)
  dilated   = cv2.dilate(hole * 255, ones(5,5))                      (:1578-1579): OR over the 5x5 window centred on
              the pixel; cv2's default border for a dilation is "ignore what is outside the image"
  mask_png  = uint8(dilated), three identical channels               (:1580-1581)
  cond      = float32(uint8(warped * (1 - dilated))) / 255.          (:1586-1590): uint8() truncates toward zero;
              float32 array / python float stays float32 (one correctly rounded division)
  latent    = mean over each 8x8 block of dilated (exact in fp64: count / 64), >= 0.2 -> 1 else 0, i.e.
              count >= 13                                            (:1592-1596)
  soft      = mean over each 8x8 block of float32(1 - soft_mask_reproj)   (:1484-1487): np.mean of a contiguous
              float32 run of 64 uses NumPy's unrolled pairwise sum: eight running sums r[j] += a[8 i + j], then
              ((r0+r1)+(r2+r3))+((r4+r5)+(r6+r7)), divided by 64; after the reshape/transpose of :1485 element
              8 i + j of a block is (row i, column j), so r[j] is the column sum, rows in order.

Parity pinned by tests/golden/post_*.npz, which re-execute the reference's statements with cv2 / PIL themselves
(tests/golden/make_golden_post.py).
"""
import numpy as np


def dilate5(hole: np.ndarray) -> np.ndarray:
    """OR over [y-2, y+2] x [x-2, x+2], pixels outside the image ignored.  hole: bool (H, W)."""
    H, W = hole.shape
    pad = np.zeros((H + 4, W + 4), bool)
    pad[2:-2, 2:-2] = hole
    out = np.zeros((H, W), bool)
    for dy in range(5):
        for dx in range(5):
            out |= pad[dy:dy + H, dx:dx + W]
    return out


def latent_mask(dil: np.ndarray) -> np.ndarray:
    H, W = dil.shape
    cnt = dil.reshape(H // 8, 8, W // 8, 8).sum(axis=(1, 3))
    return (cnt >= 13).astype(np.float64)           # count / 64 >= 0.2


def pool_soft(soft: np.ndarray) -> np.ndarray:
    """8x8 block mean of float32(1 - soft) in NumPy's own summation order (see the module docstring)."""
    H, W = soft.shape
    u = (np.float32(1) - soft.astype(np.float32)).reshape(H // 8, 8, W // 8, 8)      # [by, i, bx, j]
    r = u[:, 0].copy()                                                                  # [by, bx, j]
    for i in range(1, 8):
        r = (r + u[:, i]).astype(np.float32)
    s = ((r[..., 0] + r[..., 1]) + (r[..., 2] + r[..., 3])) + ((r[..., 4] + r[..., 5]) + (r[..., 6] + r[..., 7]))
    return (s.astype(np.float32) / np.float32(64)).astype(np.float32)


def post_forward(warped_u8: np.ndarray, mask2: np.ndarray) -> dict:
    """diffusionGS.py:1572-1597.  warped_u8 (H,W,3) uint8, mask2 (H,W) bool (True = covered)."""
    dil = dilate5(~mask2.astype(bool))
    kept = np.where(dil[..., None], np.uint8(0), warped_u8).astype(np.uint8)
    return dict(mask_png=np.repeat((dil * np.uint8(255)).astype(np.uint8)[..., None], 3, axis=2),
                cond=(kept.astype(np.float32) / np.float32(255.0)).astype(np.float32),
                cond_u8=kept, latent_mask=latent_mask(dil), masks_ero=np.repeat(dil.astype(np.uint8)[..., None], 3, axis=2))


def post_backward(warped_f32: np.ndarray, mask2: np.ndarray, soft: np.ndarray) -> dict:
    """diffusionGS.py:1452-1491.  warped_f32 (H,W,3) float32 in [0, 256), mask2 = mask_reproj, soft (H,W) f32."""
    dil = dilate5(~mask2.astype(bool))
    trunc = np.trunc(warped_f32.astype(np.float64)).astype(np.uint8)      # np.uint8(x * 1.0) for x in [0, 256)
    kept = np.where(dil[..., None], np.uint8(0), trunc).astype(np.uint8)
    return dict(mask_png=np.repeat((dil * np.uint8(255)).astype(np.uint8)[..., None], 3, axis=2),
                cond=(kept.astype(np.float32) / np.float32(255.0)).astype(np.float32), cond_u8=kept,
                cond_ori=(warped_f32 / np.float32(255.0)).astype(np.float32), latent_mask=latent_mask(dil),
                masks_ero=np.repeat(dil.astype(np.uint8)[..., None], 3, axis=2),
                soft_ori=(np.float32(1) - soft).astype(np.float32), soft_pooled=pool_soft(soft))


def fuse_uncertainty(cond_ori: np.ndarray, gs: np.ndarray, soft_ori: np.ndarray) -> dict:
    """Uncertainty fusion of /root/reference/model/diffusionGS.py:821-854 and the clip of :876,
    every intermediate in the dtype NumPy gives it there:
      s      = (c0 + c1) + c2                      float32      (:833 `cond_image_ori.sum(axis=-1)`)
      nrm    = sqrt(((d0 d0 + d1 d1) + d2 d2))     float32      (np.linalg.norm over the last axis, d = cond - gs)
      conf   = exp(-(nrm / 0.5) ** 3)              float32, then * (1 - uncertainty_mask) [int64] -> float64
      geo    = conf * (1 - soft_ori)               float64 * float32 -> float64          (:836)
      unc    = 1 - geo                             float64                               (:838)
      masks  = 8x8 block mean of unc (float64, NumPy pairwise order), cast float32      (:841-848)
      cond   = clip(where(unc > 0.5, gs, cond_ori), 0, 1)                                (:854, :876)
    cond_ori, gs (N,H,W,3) float32; soft_ori (N,H,W) float32 (= 1 - soft_mask_reproj)."""
    c, g = cond_ori.astype(np.float32), gs.astype(np.float32)
    s = ((c[..., 0] + c[..., 1]) + c[..., 2]).astype(np.float32)
    d = (c - g).astype(np.float32)
    sq = (d * d).astype(np.float32)
    nrm = np.sqrt(((sq[..., 0] + sq[..., 1]) + sq[..., 2]).astype(np.float32)).astype(np.float32)
    t = (nrm / np.float32(0.5)).astype(np.float32)
    conf32 = np.exp(-((t * t).astype(np.float32) * t).astype(np.float32)).astype(np.float32)
    conf = conf32.astype(np.float64) * (s > 0).astype(np.float64)
    geo = conf * (np.float32(1) - soft_ori.astype(np.float32)).astype(np.float64)
    unc = 1.0 - geo
    N, H, W = unc.shape
    u = unc.reshape(N, H // 8, 8, W // 8, 8)
    r = u[:, :, 0].copy()
    for i in range(1, 8):
        r = r + u[:, :, i]
    tot = ((r[..., 0] + r[..., 1]) + (r[..., 2] + r[..., 3])) + ((r[..., 4] + r[..., 5]) + (r[..., 6] + r[..., 7]))
    masks = (tot / 64.0).astype(np.float32)
    cond = np.clip(np.where(unc[..., None] > 0.5, g, c), 0, 1).astype(np.float32)
    return dict(masks=masks, cond_image=cond, geo_inten_uncertainty=unc[..., None])


def nearby_uncertainty(soft: np.ndarray, warped: np.ndarray, images_chw: np.ndarray, first: np.ndarray,
                       sigma: float = 0.1) -> dict:
    """The tail of consistency_check_from_nearby_images_bw (/root/reference/model/diffusionGS.py:1347-1355) in float32,
    op by op as torch evaluates it on the CUDA device the reference runs it on:
      conf = stack(masks).mean(0)            sum in list order, times float32(1 / cnt)                     (:1347)
      unc  = 1 - conf                                                                                    (:1348)
      wimg = stack(warps).mean(0)                                                                        (:1350)
      nrm  = sqrt((d0 d0 + d1 d1) + d2 d2),  d = wimg - image_t                (torch.norm over dim 0)  (:1351)
      iunc = 1 - exp(-((nrm * float32(1 / 0.1)) ** 3))    tensor / python scalar = multiply by the reciprocal; x ** 3 =
             x * x * x                                                                                   (:1351-1353)
    (The CPU build of torch divides instead of multiplying: a last-place difference, inside the test tolerance; the
    fixtures tests/golden/nearby_*.npz were produced on the CPU.)
    soft [P,H,W], warped [P,3,H,W], images_chw [N,3,H,W] float32; first [N+1]: frame n owns pairs first[n]:first[n+1]."""
    f32 = np.float32
    N = images_chw.shape[0]
    unc = np.empty(images_chw.shape[:1] + images_chw.shape[2:], f32)
    iunc = np.empty_like(unc)
    inv_sigma = f32(1.0) / f32(sigma)
    for n in range(N):
        k0, k1 = int(first[n]), int(first[n + 1])
        if k1 == k0:
            raise RuntimeError("stack expects a non-empty TensorList")
        factor = f32(1.0) / f32(k1 - k0)
        cs, ws = soft[k0].astype(f32), warped[k0].astype(f32)
        for k in range(k0 + 1, k1):
            cs = (cs + soft[k]).astype(f32)
            ws = (ws + warped[k]).astype(f32)
        unc[n] = f32(1) - (cs * factor).astype(f32)
        d = ((ws * factor).astype(f32) - images_chw[n]).astype(f32)
        sq = (d * d).astype(f32)
        nrm = np.sqrt(((sq[0] + sq[1]).astype(f32) + sq[2]).astype(f32)).astype(f32)
        t = (nrm * inv_sigma).astype(f32)
        iunc[n] = f32(1) - np.exp(-((t * t).astype(f32) * t).astype(f32)).astype(f32)
    return dict(uncertainty=unc, intensity_uncertainty=iunc)


def nearby_consistency(K, poses, images, depths, window_radius: int = 1, bandwidth: float = 10.0) -> dict:
    """consistency_check_from_nearby_images_bw (diffusionGS.py:1300-1361) end to end on the CPU: the oracle's
    inverse_warp (oracle/cpu_oracle.py) per (neighbour -> frame) pair, then nearby_uncertainty."""
    import torch
    from . import cpu_oracle as orc
    n = len(poses)
    soft, warped, first = [], [], [0]
    t_ = lambda a: torch.from_numpy(np.ascontiguousarray(a, np.float32))  # noqa: E731
    for cur in range(n):
        for ref in range(cur - window_radius, cur + window_radius + 1):
            if ref == cur or ref < 0 or ref >= n:
                continue
            r = orc.inverse_warp(t_(np.moveaxis(images[ref], -1, 0)), t_(depths[ref][None]), t_(depths[cur][None]),
                                 t_(poses[ref]), t_(poses[cur]), t_(np.asarray(K, np.float32)), bandwidth=bandwidth)
            soft.append(r["soft_mask_reproj"].numpy())
            warped.append(r["warped_img"].numpy())
        first.append(len(soft))
    if first[-1] == 0:
        raise RuntimeError("stack expects a non-empty TensorList")
    img = np.ascontiguousarray(np.moveaxis(np.asarray(images, np.float32), -1, 1))
    return nearby_uncertainty(np.stack(soft), np.stack(warped), img, np.asarray(first))
