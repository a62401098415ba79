"""Drop-in mirror of /root/reference/solver_utils/consistency.py (DecaYale/SYN3R @ 402f4e9).

Same names and signatures; the arithmetic runs in the sm_100a kernels of libsyn3r_warp.so (fp32, op
order pinned by the oracle).  Inputs may live on any device; results are CUDA tensors.  No CPU path.
"""
from __future__ import annotations

import numpy as np
import torch

from .. import _native as nat
from .. import warp as _w


def _device():
    if not torch.cuda.is_available():
        raise nat.Syn3rNativeError("syn3r_b200 needs a CUDA device (sm_100a); there is no CPU fallback")
    return torch.device("cuda", torch.cuda.current_device())


def _f32(t, dev):
    return torch.as_tensor(t).detach().to(dev, torch.float32).contiguous()


def get_points_from_depth(depth, intrinsics):
    """consistency.py:6-23 — depth (h,w), intrinsics (3,3) -> points (h,w,3) in the camera frame."""
    dev = _device()
    h, w = depth.shape
    Kinv = torch.inverse(torch.as_tensor(intrinsics).detach().float().cpu())
    karr, kp = nat.host_f32(Kinv.numpy())      # karr keeps the host buffer alive across the call
    d = _f32(depth, dev)
    pts = torch.empty((h, w, 3), dtype=torch.float32, device=dev)
    rc = nat.lib().syn3r_points_from_depth(nat.ptr(d), kp, h, w, nat.ptr(pts),
                                           torch.cuda.current_stream().cuda_stream)
    nat.check(rc, "syn3r_points_from_depth")
    del karr
    return pts


def transform_points(pts, pose1, pose2):
    """consistency.py:25-42 — points (h,w,3) from frame 1 to frame 2; poses are w2c 4x4."""
    dev = _device()
    h, w, _ = pts.shape
    p1 = torch.as_tensor(pose1).detach().float().cpu()
    p2 = torch.as_tensor(pose2).detach().float().cpu()
    tarr, tp = nat.host_f32((p2 @ torch.inverse(p1)).numpy())
    src = _f32(pts, dev)
    out = torch.empty((h, w, 3), dtype=torch.float32, device=dev)
    rc = nat.lib().syn3r_transform_points(nat.ptr(src), tp, h, w, nat.ptr(out),
                                          torch.cuda.current_stream().cuda_stream)
    nat.check(rc, "syn3r_transform_points")
    del tarr
    return out


def consistency_check_with_depth(depth1, pose1, intrinsics1, depth2, pose2, intrinsics2):
    """consistency.py:44-91 — round-trip reprojection error (pixels) of frame-1 pixels through frame 2.
    depth1, depth2 (h,w); pose1, pose2 w2c (4,4); intrinsics (3,3).  Returns (h,w) float32 CUDA tensor."""
    dev = _device()
    p1 = torch.as_tensor(pose1).detach().cpu()[None]
    p2 = torch.as_tensor(pose2).detach().cpu()[None]
    # inverse_pair_params(src=A, tgt=B) gives [A @ inv(B), B @ inv(A)]; we need [pose2 @ inv(pose1), pose1 @ inv(pose2)]
    pair_T = torch.from_numpy(_w.inverse_pair_params(p2, p1, [0], [0])).to(dev)
    idx = torch.zeros((1,), dtype=torch.int32, device=dev)
    K1 = np.asarray(torch.as_tensor(intrinsics1).detach().cpu())
    K2 = np.asarray(torch.as_tensor(intrinsics2).detach().cpu())
    err = _w.consistency_batched(_f32(depth1, dev)[None], _f32(depth2, dev)[None], K1, K2, pair_T, idx, idx)
    return err[0]
