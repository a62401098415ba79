"""ctypes front-end of the CPU oracle (oracle/syn3r_oracle.c) — TEST INFRASTRUCTURE ONLY.

May be imported only from tests/, `__graft_entry__.smoke()` and bench.py's `cpu_baseline` /
`--impl reference` legs.  The product package `syn3r_b200` never imports it.

The functions keep the reference's names, argument order and return types
(/root/reference/solver_utils/forward_warp.py, consistency.py) so parity tests read like calls to the
reference.  Tiny host-side matrix algebra (4x4 / 3x3 inverses and products) is done with the same
library calls the reference makes (np.linalg.inv / np.matmul in the caller's dtype for the forward
path, forward_warp.py:16,25; torch.inverse / @ in fp32 for the backward path, forward_warp.py:217,
consistency.py:21,39); the per-pixel arithmetic is the C restatement.
"""
from __future__ import annotations

import ctypes
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "_build", "libsyn3r_oracle.so")
_lib = None


def build(force: bool = False) -> str:
    """Compile the C restatement with the committed Makefile (gcc only)."""
    if force or not os.path.exists(_LIB_PATH) or \
            os.path.getmtime(_LIB_PATH) < os.path.getmtime(os.path.join(_HERE, "syn3r_oracle.c")):
        subprocess.check_call(["make", "-C", _HERE, "-s"] + (["-B"] if force else []))
    return _LIB_PATH


def lib():
    global _lib
    if _lib is None:
        build()
        _lib = ctypes.CDLL(_LIB_PATH)
    return _lib


def _p(a):
    return None if a is None else a.ctypes.data_as(ctypes.c_void_p)


def _c(a, dt):
    return None if a is None else np.ascontiguousarray(a, dtype=dt)


# ------------------------------------------------------------------------------------------------
# forward path (fp64)
# ------------------------------------------------------------------------------------------------

def create_grid(h, w):
    """forward_warp.py:133-139 — int64 [h,w,2], [...,0]=x, [...,1]=y."""
    g = np.empty((h, w, 2), dtype=np.int64)
    g[..., 0] = np.arange(w)[None, :]
    g[..., 1] = np.arange(h)[:, None]
    return g


def host_matrices(transformation1, transformation2, intrinsic1, intrinsic2):
    """Host-side small-matrix step of forward_warp.py:14-16,25 (same NumPy calls, caller's dtype)."""
    if intrinsic2 is None:
        intrinsic2 = np.copy(intrinsic1)
    M = np.matmul(transformation2, np.linalg.inv(transformation1))
    K1inv = np.linalg.inv(intrinsic1)
    return (np.ascontiguousarray(M, dtype=np.float64), np.ascontiguousarray(K1inv, dtype=np.float64),
            np.ascontiguousarray(intrinsic2, dtype=np.float64))


def compute_transformed_points(depth1, transformation1, transformation2, intrinsic1, intrinsic2):
    """forward_warp.py:7-37 -> (trans_norm_points [h,w,3,1], world_points [h,w,3,1]) f64."""
    h, w = depth1.shape
    M, K1inv, K2 = host_matrices(transformation1, transformation2, intrinsic1, intrinsic2)
    d = _c(depth1, np.float64)
    q = np.empty((h, w, 3), np.float64)
    X = np.empty((h, w, 3), np.float64)
    rc = lib().orc_project(_p(d), _p(M), _p(K1inv), _p(K2), h, w, _p(q), _p(X))
    assert rc == 0
    return q[..., None], X[..., None]


def _splat_resolve(frames, mask1s, zs, flows, flow_masks, is_image, gate06=False, want_winner=False):
    """One accumulator, len(frames) sources (1 = PAIR / reference; >1 = FUSED extension)."""
    h, w, c = frames[0].shape
    L = lib()
    acc_i = np.zeros((h + 2, w + 2, c), np.float64)
    acc_w = np.zeros((h + 2, w + 2), np.float64)
    key = np.full((h + 2, w + 2), np.iinfo(np.uint64).max, np.uint64) if want_winner else None
    for s, (fr, m1, z, fl, fm) in enumerate(zip(frames, mask1s, zs, flows, flow_masks)):
        rc = L.orc_splat(_p(_c(fr, np.float64)), _p(_c(m1, np.uint8)), _p(_c(z, np.float64)),
                         _p(_c(fl, np.float64)), _p(_c(fm, np.uint8)), h, w, c, s, _p(acc_i), _p(acc_w), _p(key))
        assert rc == 0
    out = np.empty((h, w, c), np.float64)
    out_u8 = np.empty((h, w, c), np.uint8) if is_image else None
    mask = np.empty((h, w), np.uint8)
    winner = np.empty((h, w), np.int32) if want_winner else None
    S = np.empty((h, w), np.float64)
    ok = ctypes.c_int(1)
    rc = L.orc_resolve(_p(acc_i), _p(acc_w), _p(key), h, w, c, int(is_image), int(gate06), _p(out), _p(out_u8),
                       _p(mask), _p(winner), _p(S), ctypes.byref(ok))
    assert rc == 0
    return dict(out=out, out_u8=out_u8, mask=mask.astype(bool), winner=winner, S=S,
                acc_img=acc_i[1:-1, 1:-1], range_ok=bool(ok.value))


def bilinear_splatting(frame1, mask1, depth1, flow12, flow12_mask, is_image=False):
    """forward_warp.py:41-129 -> (warped_frame2 [h,w,c] (uint8 if is_image else f64), mask bool [h,w])."""
    r = _splat_resolve([frame1], [mask1], [depth1], [flow12], [flow12_mask], is_image)
    if is_image:
        assert r["range_ok"], "colour range assert (forward_warp.py:121-125)"
        return r["out_u8"], r["mask"]
    return r["out"], r["mask"]


def _flow_z(depth1, transformation1, transformation2, intrinsic1, intrinsic2):
    h, w = depth1.shape
    q, _ = compute_transformed_points(depth1, transformation1, transformation2, intrinsic1, intrinsic2)
    q = np.ascontiguousarray(q[..., 0])
    flow = np.empty((h, w, 2), np.float64)
    z = np.empty((h, w), np.float64)
    rc = lib().orc_flow_from_q(_p(q), h, w, _p(flow), _p(z))
    assert rc == 0
    return flow, z


def forward_warp(frame1, mask1, depth1, transformation1, transformation2, intrinsic1, intrinsic2):
    """forward_warp.py:141-182 -> (uint8 [h,w,3], bool [h,w], f64 flow12 [h,w,2])."""
    h, w = frame1.shape[:2]
    assert frame1.shape == (h, w, 3)
    assert mask1 is None or mask1.shape == (h, w)
    assert depth1.shape == (h, w)
    assert transformation1.shape == (4, 4) and transformation2.shape == (4, 4)
    assert intrinsic1.shape == (3, 3) and (intrinsic2 is None or intrinsic2.shape == (3, 3))
    flow, z = _flow_z(depth1, transformation1, transformation2, intrinsic1, intrinsic2)
    warped, mask2 = bilinear_splatting(frame1, mask1, z, flow, None, is_image=True)
    return warped, mask2, flow


def forward_warp_ext(frames, mask1s, depths, transformations1, transformation2, intrinsics1, intrinsic2,
                     gate06=False):
    """Extension outputs (parity unpinned by the reference): S sources fused into one
    accumulator (S=1 -> PAIR), returns dict with f64 colour `out`, `out_u8`, `mask`, `winner`
    (src_id*H*W + y*W + x of the nearest contributing source pixel, -1 = hole), `S`, `flows`."""
    flows, zs = [], []
    for d, T1, K1 in zip(depths, transformations1, intrinsics1):
        f, z = _flow_z(d, T1, transformation2, K1, intrinsic2)
        flows.append(f)
        zs.append(z)
    r = _splat_resolve(frames, mask1s, zs, flows, [None] * len(frames), True, gate06, want_winner=True)
    r["flows"] = flows
    return r


# ------------------------------------------------------------------------------------------------
# backward path (fp32, torch tensors in / out like the reference)
# ------------------------------------------------------------------------------------------------

def _np32(t):
    import torch
    if isinstance(t, torch.Tensor):
        t = t.detach().cpu().numpy()
    return np.ascontiguousarray(t, dtype=np.float32)


def get_points_from_depth(depth, intrinsics):
    """consistency.py:6-23 -> [h,w,3] tensor."""
    import torch
    h, w = depth.shape
    Kinv = _np32(torch.inverse(intrinsics.cpu()))
    pts = np.empty((h, w, 3), np.float32)
    assert lib().orc_points_from_depth(_p(_np32(depth)), _p(Kinv), h, w, _p(pts)) == 0
    return torch.from_numpy(pts)


def transform_points(pts, pose1, pose2):
    """consistency.py:25-42 -> [h,w,3] tensor in frame 2."""
    import torch
    h, w, _ = pts.shape
    T = _np32(pose2.cpu() @ torch.inverse(pose1.cpu()))
    out = np.empty((h, w, 3), np.float32)
    assert lib().orc_transform_points(_p(_np32(pts)), _p(T), h, w, _p(out)) == 0
    return torch.from_numpy(out)


def consistency_check_with_depth(depth1, pose1, intrinsics1, depth2, pose2, intrinsics2):
    """consistency.py:44-91 -> reprojection error [h,w] (pixels)."""
    import torch
    h, w = depth1.shape
    pose1, pose2 = pose1.cpu().float(), pose2.cpu().float()
    K1, K2 = intrinsics1.cpu().float(), intrinsics2.cpu().float()
    K1inv = _np32(torch.inverse(K1))
    T12 = _np32(pose2 @ torch.inverse(pose1))
    T21 = _np32(pose1 @ torch.inverse(pose2))
    err = np.empty((h, w), np.float32)
    rc = lib().orc_consistency(_p(_np32(depth1)), _p(_np32(K1)), _p(K1inv), _p(T12), _p(_np32(depth2)),
                               _p(_np32(K2)), _p(T21), h, w, _p(err))
    assert rc == 0
    return torch.from_numpy(err)


_IW_KEYS = ("warped_img", "warped_depth", "mask_warp", "mask_depth", "mask", "warped_masked_img", "mask_inv",
            "mask_depth_strict", "warped_bg_mask", "mask_reproj", "soft_mask_reproj")


def inverse_warp(img, depth, depth_pseudo, pose1, pose2, K, bg_mask=None, bandwidth=20, return_err=False):
    """forward_warp.py:187-279 -> dict with the reference's 11 keys (CPU tensors)."""
    import torch
    _, H, W = img.shape
    pose1, pose2, Kt = pose1.cpu().float(), pose2.cpu().float(), K.cpu().float()
    P = _np32(pose1 @ torch.inverse(pose2))
    Pinv = _np32(pose2 @ torch.inverse(pose1))
    Kinv = _np32(torch.inverse(Kt))
    f32 = lambda *s: np.empty(s, np.float32)  # noqa: E731
    u8 = lambda *s: np.empty(s, np.uint8)  # noqa: E731
    o = dict(warped_img=f32(3, H, W), warped_depth=f32(1, H, W), mask_warp=u8(H, W), mask_depth=u8(H, W),
             mask=u8(H, W), warped_masked_img=f32(3, H, W), mask_inv=u8(H, W), mask_depth_strict=u8(H, W),
             warped_bg_mask=u8(1, H, W) if bg_mask is not None else None, mask_reproj=u8(H, W),
             soft_mask_reproj=f32(H, W))
    err = f32(H, W)
    bg = None if bg_mask is None else np.ascontiguousarray(bg_mask.cpu().numpy().astype(np.uint8))
    rc = lib().orc_inverse_warp(_p(_np32(img)), _p(_np32(depth)), _p(_np32(depth_pseudo)), _p(P), _p(Pinv),
                                _p(_np32(Kt)), _p(Kinv), _p(bg), ctypes.c_float(float(bandwidth)), H, W,
                                _p(o["warped_img"]), _p(o["warped_depth"]), _p(o["mask_warp"]), _p(o["mask_depth"]),
                                _p(o["mask"]), _p(o["warped_masked_img"]), _p(o["mask_inv"]),
                                _p(o["mask_depth_strict"]), _p(o["warped_bg_mask"]), _p(o["mask_reproj"]),
                                _p(o["soft_mask_reproj"]), _p(err))
    assert rc == 0
    out = {}
    for k in _IW_KEYS:
        v = o[k]
        if v is None:
            out[k] = None
        elif v.dtype == np.uint8:
            out[k] = torch.from_numpy(v.astype(bool))
        else:
            out[k] = torch.from_numpy(v)
    if return_err:
        out["reproj_err"] = torch.from_numpy(err)
    return out
