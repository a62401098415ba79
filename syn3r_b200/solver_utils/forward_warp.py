"""Drop-in mirror of /root/reference/solver_utils/forward_warp.py (DecaYale/SYN3R @ 402f4e9).

Same function names, positional order, defaults, return types and numerical quirks; underneath every
function launches the sm_100a kernels of libsyn3r_warp.so through `syn3r_b200.warp` with one pair.
There is no CPU path: without the CUDA library (or a GPU) these functions raise.

Differences a caller can observe (documented in DESIGN.md):
  * the colour-range failure of forward_warp.py:121-125 raises AssertionError instead of opening pdb;
  * colours are accumulated in integer block floating point with 21-22 significant bits per cell (DESIGN.md 4.1;
    the reference accumulates in float64): uint8 outputs can differ by one level at rounding ties
    (about 3e-5 of values), float outputs agree to < 1e-5 on the [0,1] scale; masks are bit-exact;
  * float64 images are rounded to float32 on upload (depth, poses and intrinsics keep full precision).
"""
from __future__ import annotations

from typing import Optional, Tuple

import numpy as np
import torch

from .. import _native as nat
from .. import warp as _w
from .consistency import consistency_check_with_depth  # noqa: F401  (the reference imports it here, :4)


def _device():
    if not torch.cuda.is_available():
        raise nat.Syn3rNativeError("syn3r_b200 needs a CUDA device (sm_100a); there is no CPU fallback")
    return torch.device("cuda", torch.cuda.current_device())


def _up(a, dtype=None):
    t = torch.from_numpy(np.ascontiguousarray(a))
    if dtype is not None and t.dtype != dtype:
        t = t.to(dtype)
    return t.to(_device(), non_blocking=False)


def create_grid(h, w):
    """forward_warp.py:133-139 — integer pixel grid [h,w,2] ([...,0]=x, [...,1]=y).  Host-side index
    bookkeeping only (the kernels derive x,y from the thread index)."""
    g = np.empty((h, w, 2), dtype=np.int64)
    g[..., 0] = np.arange(0, w)[None]
    g[..., 1] = np.arange(0, h)[:, None]
    return g


def _pair_matrices(transformation1, transformation2, intrinsic1, intrinsic2):
    Kinv, par = _w.forward_pair_params(np.asarray(transformation1)[None], np.asarray(transformation2)[None],
                                       np.asarray(intrinsic1)[None],
                                       None if intrinsic2 is None else np.asarray(intrinsic2), [0], [0])
    return Kinv, par


def compute_transformed_points(depth1: np.ndarray, transformation1: np.ndarray, transformation2: np.ndarray,
                               intrinsic1: np.ndarray, intrinsic2: Optional[np.ndarray]):
    """forward_warp.py:7-37 -> (trans_norm_points [h,w,3,1], world_points [h,w,3,1]) float64."""
    h, w = depth1.shape
    Kinv, par = _pair_matrices(transformation1, transformation2, intrinsic1, intrinsic2)
    d = _up(np.asarray(depth1), torch.float64)
    q = torch.empty((h, w, 3), dtype=torch.float64, device=d.device)
    X = torch.empty_like(q)
    kinv_d, par_d = _up(Kinv), _up(par)        # keep alive until the launch is enqueued
    rc = nat.lib().syn3r_compute_transformed_points(nat.ptr(d), nat.ptr(kinv_d), nat.ptr(par_d), h, w,
                                                    nat.ptr(q), nat.ptr(X), torch.cuda.current_stream().cuda_stream)
    nat.check(rc, "syn3r_compute_transformed_points")
    return q.cpu().numpy()[..., None], X.cpu().numpy()[..., None]


def bilinear_splatting(frame1: np.ndarray, mask1: Optional[np.ndarray], depth1: np.ndarray, flow12: np.ndarray,
                       flow12_mask: Optional[np.ndarray], is_image: bool = False) -> Tuple[np.ndarray, np.ndarray]:
    """forward_warp.py:41-129 -> (warped_frame2 [h,w,c] (uint8 if is_image else float64), mask2 bool [h,w])."""
    h, w, c = frame1.shape
    L = nat.lib()
    dev = _device()
    z = _up(np.asarray(depth1), torch.float64)
    fl = _up(np.asarray(flow12), torch.float64)
    m1 = None if mask1 is None else _up(np.asarray(mask1).astype(np.uint8))
    fm = None if flow12_mask is None else _up(np.asarray(flow12_mask).astype(np.uint8))
    flags = nat.FWD_IS_IMAGE if is_image else 0
    ws = _w._workspace(L.syn3r_fwd_workspace_bytes(h, w, 1, 1, 1, flags | nat.FWD_LEGACY_GLOBAL), dev, "fwd")
    outs, mask = [], None
    frame_f32 = np.asarray(frame1).astype(np.float32, copy=False)
    for c0 in range(0, c, 3):                       # accumulators hold 3 channels + weight
        cc = min(3, c - c0)
        fr = _up(np.ascontiguousarray(frame_f32[..., c0:c0 + cc]))
        out = torch.empty((h, w, cc), dtype=torch.uint8 if is_image else torch.float32, device=dev)
        mk = torch.empty((h, w), dtype=torch.uint8, device=dev)
        st = torch.zeros((1,), dtype=torch.int32, device=dev)
        rc = L.syn3r_bilinear_splat(nat.ptr(fr), nat.ptr(m1), nat.ptr(z), nat.ptr(fl), nat.ptr(fm), h, w, cc, flags,
                                    None if is_image else nat.ptr(out), nat.ptr(out) if is_image else None,
                                    nat.ptr(mk), nat.ptr(st), ws.data_ptr(), ws.numel(),
                                    torch.cuda.current_stream().cuda_stream)
        nat.check(rc, "syn3r_bilinear_splat")
        if is_image:
            assert int(st.item()) == 0, "warped colour outside [0, 256] (forward_warp.py:121-125)"
        outs.append(out)
        mask = mk
    res = torch.cat(outs, dim=2).cpu().numpy()
    if not is_image:
        res = res.astype(np.float64)
    return res, mask.cpu().numpy().astype(bool)


def forward_warp(frame1: np.ndarray, mask1: Optional[np.ndarray], depth1: np.ndarray, transformation1: np.ndarray,
                 transformation2: np.ndarray, intrinsic1: np.ndarray, intrinsic2: Optional[np.ndarray],
                 *, gate06: bool = False):
    """forward_warp.py:141-182 -> (warped_frame2 uint8 [h,w,3], mask2 bool [h,w], flow12 float64 [h,w,2]).
    `gate06=True` adds the `mask &= weights <= 0.6` gate of the demo copies (model/SVD_1pass.py:853-855)."""
    h, w = frame1.shape[:2]
    assert frame1.shape == (h, w, 3)
    assert mask1 is None or mask1.shape == (h, w)
    assert depth1.shape == (h, w)
    assert transformation1.shape == (4, 4)
    assert transformation2.shape == (4, 4)
    assert intrinsic1.shape == (3, 3)
    assert intrinsic2 is None or intrinsic2.shape == (3, 3)
    Kinv, par = _pair_matrices(transformation1, transformation2, intrinsic1, intrinsic2)
    depth1 = np.asarray(depth1)
    rgbd = _up(_w.pack_rgbd(np.asarray(frame1)[None], depth1[None].astype(np.float32)))
    d64 = None if depth1.dtype == np.float32 else _up(depth1[None], torch.float64)
    m1 = None if mask1 is None else _up(np.asarray(mask1).astype(np.uint8)[None])
    pair_src = torch.zeros((1,), dtype=torch.int32, device=rgbd.device)
    o = _w.forward_warp_batched(rgbd, _up(Kinv), _up(par), pair_src, depth64=d64, src_mask=m1, want_f32=False,
                                want_u8=True, want_flow=True, gate06=gate06, is_image=True)
    assert int(o["status"].item()) == 0, "warped colour outside [0, 256] (forward_warp.py:121-125)"
    return o["rgb_u8"][0].cpu().numpy(), o["mask"][0].cpu().numpy().astype(bool), o["flow"][0].cpu().numpy()


def inverse_warp(img, depth, depth_pseudo, pose1, pose2, K, bg_mask=None, bandwidth=20):
    """forward_warp.py:187-279 — backward nearest warp of the closest view into the pseudo view plus the
    warp / depth / reprojection masks.  img [3,H,W], depth, depth_pseudo [1,H,W], pose1 (closest view) and
    pose2 (pseudo view) w2c [4,4], K [3,3]; tensors on any device.  Returns the reference's 11-key dict of
    CUDA tensors."""
    dev = _device()
    _, H, W = img.shape
    f32 = lambda t: t.detach().to(dev, torch.float32).contiguous()  # noqa: E731
    pair_P = torch.from_numpy(_w.inverse_pair_params(pose1.detach().cpu()[None], pose2.detach().cpu()[None],
                                                     [0], [0])).to(dev)
    idx = torch.zeros((1,), dtype=torch.int32, device=dev)
    bg = None
    if bg_mask is not None:
        bg = (bg_mask.detach().to(dev).float() > 0.5).to(torch.uint8).reshape(1, H, W).contiguous()
    want = ["warped_img", "warped_depth", "mask_warp", "mask_depth", "mask", "warped_masked_img", "mask_inv",
            "mask_depth_strict", "mask_reproj", "soft_mask_reproj"] + (["warped_bg"] if bg is not None else [])
    o = _w.inverse_warp_batched(f32(img)[None], f32(depth).reshape(1, H, W), f32(depth_pseudo).reshape(1, H, W),
                                K.detach().cpu() if isinstance(K, torch.Tensor) else K, pair_P, idx, idx, src_bg=bg,
                                bandwidth=float(bandwidth), want=want)
    b = lambda k: o[k][0].bool()  # noqa: E731
    return {"warped_img": o["warped_img"][0], "warped_depth": o["warped_depth"],
            "mask_warp": b("mask_warp"), "mask_depth": b("mask_depth"), "mask": b("mask"),
            "warped_masked_img": o["warped_masked_img"][0], "mask_inv": b("mask_inv"),
            "mask_depth_strict": b("mask_depth_strict"),
            "warped_bg_mask": None if bg is None else o["warped_bg"].bool(),
            "mask_reproj": b("mask_reproj"),
            "soft_mask_reproj": o["soft_mask_reproj"][0],
            }
