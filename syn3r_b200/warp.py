"""Batched device-side entry points: whole trajectories of (source, target) pair warps in one call.

These are what the per-pair reference-signature functions in `syn3r_b200.solver_utils` call with
P = 1, and what `bench.py` / the multi-GPU driver call with a whole trajectory.  Inputs and outputs
are CUDA tensors; all kernels run on `torch.cuda.current_stream()`; nothing synchronises.

Small-matrix preparation (pose products, inverses) is done on the host with the SAME library calls
the reference makes, so their rounding is the reference's:
  forward  — np.matmul(T2, np.linalg.inv(T1)) and np.linalg.inv(K1) in the caller's dtype
             (/root/reference/solver_utils/forward_warp.py:16,25), then promoted to f64;
  backward — pose1 @ torch.inverse(pose2), torch.inverse(K) in fp32 on the CPU
             (forward_warp.py:217, consistency.py:21,39).
"""
from __future__ import annotations

import ctypes

import numpy as np
import torch

from . import _native as nat

_ws_cache: dict = {}


def _workspace(nbytes: int, device, tag: str) -> torch.Tensor:
    """Caller-owned workspace, cached per (device, stream, tag) and grown on demand (torch's caching
    allocator returns 512-byte aligned blocks).  The workspace holds kernel state (work counters, max-z' keys,
    block boxes, the redo list), so calls on different streams must not share one; calls on ONE stream are
    ordered by the stream."""
    key = (str(device), torch.cuda.current_stream(device).cuda_stream, tag)
    t = _ws_cache.get(key)
    if t is None or t.numel() < nbytes:
        t = torch.empty(max(nbytes, 256), dtype=torch.uint8, device=device)
        _ws_cache[key] = t
    return t


_last_fwd: dict = {}


def redo_items(device) -> int:
    """Work items ((slot, tile) pairs) the integer tile kernel handed to the fp32 kernel in the LAST forward call on
    `device` (crowded cells, NaN-poisoned pairs, non-finite colours; normally 0).  Reads the redo counter of the call's
    workspace (abi.cu fwd_layout: counter block behind the per-pair max-z' keys and NaN flags); synchronises."""
    device = torch.device(device)
    if device.index is None:                               # "cuda" -> the current device, like torch itself
        device = torch.device("cuda", torch.cuda.current_device())
    ent = _last_fwd.get(str(device))
    if ent is None:
        return 0
    ws, P = ent
    up = lambda v: (v + 255) // 256 * 256  # noqa: E731
    off = up(P * 8) + up(P * 4) + 32 * 4
    return int(ws[off:off + 4].view(torch.int32).item())


def _stream_ptr(device=None) -> int:
    return torch.cuda.current_stream(device).cuda_stream


# --------------------------------------------------------------------------------------------
# forward path
# --------------------------------------------------------------------------------------------

def pack_rgbd(rgb, depth) -> np.ndarray:
    """[S,H,W,3] colour (any real dtype, 0..255 scale) + [S,H,W] depth -> [S,H,W,4] float32 (host)."""
    rgb = np.asarray(rgb)
    out = np.empty(rgb.shape[:-1] + (4,), np.float32)
    out[..., :3] = rgb
    out[..., 3] = depth
    return out


def forward_pair_params(T_src, T_tgt, K_src, K_tgt, pair_src, pair_tgt):
    """Host-side matrices of forward_warp.py:14-16,25 for every pair.

    T_src [S,4,4], T_tgt [T,4,4] w2c in the CALLER's dtype (f32 in production, f64 in the demos);
    K_src [S,3,3] (or [3,3]); K_tgt None (-> copy of the pair's K_src, :14-15), [3,3] or [T,3,3].
    Returns (Kinv_src [S,9] f64, pair_par [P,24] f64)."""
    T_src, T_tgt = np.asarray(T_src), np.asarray(T_tgt)
    S = T_src.shape[0]
    K_src = np.asarray(K_src)
    if K_src.ndim == 2:
        K_src = np.broadcast_to(K_src, (S, 3, 3))
    Kinv = np.stack([np.linalg.inv(K_src[s]) for s in range(S)]).astype(np.float64).reshape(S, 9)
    P = len(pair_src)
    par = np.zeros((P, 24), np.float64)
    inv_cache = {}
    for p, (s, t) in enumerate(zip(pair_src, pair_tgt)):
        s, t = int(s), int(t)
        if s not in inv_cache:
            inv_cache[s] = np.linalg.inv(T_src[s])
        M = np.matmul(T_tgt[t], inv_cache[s])                    # caller's dtype, then promoted
        par[p, :12] = np.asarray(M, dtype=np.float64)[:3].ravel()
        if K_tgt is None:
            K2 = K_src[s]
        else:
            K2 = np.asarray(K_tgt)
            K2 = K2 if K2.ndim == 2 else K2[t]
        par[p, 12:21] = np.asarray(K2, dtype=np.float64).ravel()
    return np.ascontiguousarray(Kinv), par


def forward_warp_batched(rgbd, Kinv_src, pair_par, pair_src, *, pair_slot=None, n_slots=None, depth64=None,
                         src_mask=None, want_f32=True, want_u8=True, want_winner=False, want_flow=False,
                         gate06=False, is_image=True, out=None, phases=0, exact_weight=False, legacy=False,
                         fp32_tile=False, workspace=None):
    """Launch the forward warp for P pairs.  All array arguments are CUDA tensors:
    rgbd [S,H,W,4] f32, Kinv_src [S,9] f64, pair_par [P,24] f64, pair_src [P] i32,
    pair_slot [P] i32 or None (PAIR mode).  Returns a dict of CUDA tensors
    (rgb_f32 [n,H,W,3], rgb_u8 [n,H,W,3], mask [n,H,W] u8, winner [n,H,W] i32, flow [P,H,W,2] f64,
    status [n] i32).  `out` may carry preallocated tensors under the same keys.
    `exact_weight`: IEEE log/exp/div for the soft depth weight (default: fast fp64 table path, rel. err < 3e-7);
    `legacy`: first-generation global-accumulator kernels, `fp32_tile`: second-generation per-tile fp32
    accumulators (A/B measurements; default is the integer block-floating-point tile kernel).
    `workspace`: a caller-owned uint8 CUDA tensor of at least `fwd_workspace_bytes` (a holder: `workspace[0]` is
    replaced when it has to grow); default: one cached per (device, stream).  A caller that runs the phases of one
    call on two streams (TrajectoryWarp) must own it."""
    L = nat.lib()
    S, H, W, four = rgbd.shape
    assert four == 4 and rgbd.dtype == torch.float32
    P = pair_src.numel()
    n = P if pair_slot is None else int(n_slots)
    dev = rgbd.device
    flags = (nat.FWD_GATE_0P6 if gate06 else 0) | (nat.FWD_WINNER if want_winner else 0) | \
            (nat.FWD_IS_IMAGE if is_image else 0) | int(phases) | (nat.FWD_EXACT_WEIGHT if exact_weight else 0) | \
            (nat.FWD_LEGACY_GLOBAL if legacy else 0) | (nat.FWD_FP32_TILE if fp32_tile else 0)
    o = dict(out or {})

    def get(key, want, shape, dtype):
        if not want:
            return None
        t = o.get(key)
        if t is None:
            t = torch.empty(shape, dtype=dtype, device=dev)
            o[key] = t
        assert t.shape == torch.Size(shape) and t.dtype == dtype and t.is_contiguous()
        return t

    rgb_f32 = get("rgb_f32", want_f32, (n, H, W, 3), torch.float32)
    rgb_u8 = get("rgb_u8", want_u8 and is_image, (n, H, W, 3), torch.uint8)
    mask = get("mask", True, (n, H, W), torch.uint8)
    winner = get("winner", want_winner, (n, H, W), torch.int32)
    flow = get("flow", want_flow, (P, H, W, 2), torch.float64)
    status = get("status", True, (n,), torch.int32)
    nbytes = L.syn3r_fwd_workspace_bytes(H, W, S, P, n, flags)
    if workspace is not None:
        if workspace[0] is None or workspace[0].numel() < nbytes:
            workspace[0] = torch.empty(max(nbytes, 256), dtype=torch.uint8, device=dev)
        ws = workspace[0]
    else:
        ws = _workspace(nbytes, dev, "fwd")
    _last_fwd[str(dev)] = (ws, P)
    with torch.cuda.device(dev):        # the C side sizes its grid from, and launches on, the CURRENT device
        rc = L.syn3r_forward_warp_batched(
            nat.ptr(rgbd), nat.ptr(depth64), nat.ptr(src_mask), nat.ptr(Kinv_src), nat.ptr(pair_par), nat.ptr(pair_src),
            nat.ptr(pair_slot), S, P, n, H, W, flags, nat.ptr(rgb_f32), nat.ptr(rgb_u8), nat.ptr(mask), nat.ptr(winner),
            nat.ptr(flow), nat.ptr(status), ws.data_ptr(), ws.numel(), _stream_ptr(dev))
    nat.check(rc, "syn3r_forward_warp_batched")
    return o


# --------------------------------------------------------------------------------------------
# backward / consistency path
# --------------------------------------------------------------------------------------------

INV_KEYS_F32 = ("warped_img", "warped_depth", "warped_masked_img", "soft_mask_reproj", "reproj_err")
INV_KEYS_U8 = ("mask_warp", "mask_depth", "mask", "mask_inv", "mask_depth_strict", "warped_bg", "mask_reproj")


def inverse_pair_params(pose_src, pose_tgt, pair_src, pair_tgt) -> np.ndarray:
    """[P,2,16] f32: pose1 @ inverse(pose2) then pose2 @ inverse(pose1) (pose1 = source w2c,
    pose2 = target w2c), with torch CPU fp32 like forward_warp.py:217 / consistency.py:39."""
    ps = torch.as_tensor(np.asarray(pose_src)).float().cpu()
    pt = torch.as_tensor(np.asarray(pose_tgt)).float().cpu()
    inv_s = [torch.inverse(ps[i]) for i in range(ps.shape[0])]
    inv_t = [torch.inverse(pt[i]) for i in range(pt.shape[0])]
    out = np.empty((len(pair_src), 2, 16), np.float32)
    for p, (s, t) in enumerate(zip(pair_src, pair_tgt)):
        s, t = int(s), int(t)
        out[p, 0] = (ps[s] @ inv_t[t]).numpy().ravel()
        out[p, 1] = (pt[t] @ inv_s[s]).numpy().ravel()
    return out


def inverse_warp_batched(src_img, src_depth, tgt_depth, K, pair_P, pair_src, pair_tgt, *, src_bg=None,
                         rgbd=None,
                         bandwidth=20.0, thr_depth=0.3, thr_depth_strict=0.1,
                         want=("warped_img", "mask_reproj", "soft_mask_reproj"), out=None):
    """Launch the backward warp + consistency masks for P pairs.  CUDA tensors: src_img [S,3,H,W] f32,
    src_depth [S,H,W] f32, tgt_depth [T,H,W] f32, pair_P [P,2,16] f32, pair_src/pair_tgt [P] i32;
    K is a HOST 3x3 (numpy / CPU tensor).  `want` selects outputs (keys of the reference's dict plus
    'reproj_err'); returns dict key -> CUDA tensor with a leading pair axis.
    Instead of `src_depth`, `rgbd` [S,H,W,4] f32 (the forward path's input) may be given: its depth lane
    is read in place.  `src_img` may be None when no colour output is wanted."""
    L = nat.lib()
    if rgbd is not None:
        assert src_depth is None and rgbd.dtype == torch.float32 and rgbd.is_contiguous()
        S, H, W, _ = rgbd.shape
        depth_ptr, depth_stride, dev = rgbd.data_ptr() + 12, 4, rgbd.device
    else:
        S, H, W = src_depth.shape
        assert src_depth.dtype == torch.float32
        depth_ptr, depth_stride, dev = nat.ptr(src_depth), 1, src_depth.device
    if src_img is not None:
        assert src_img.shape == (S, 3, H, W) and src_img.dtype == torch.float32
    P = pair_src.numel()
    Kt = torch.as_tensor(np.asarray(K.cpu() if isinstance(K, torch.Tensor) else K)).float().cpu()
    Kinv = torch.inverse(Kt)
    k_arr, k_ptr = nat.host_f32(Kt.numpy())
    ki_arr, ki_ptr = nat.host_f32(Kinv.numpy())
    o = dict(out or {})
    outs = nat.InvOutputs()
    for key in want:
        if key in INV_KEYS_F32:
            shape = (P, 3, H, W) if key in ("warped_img", "warped_masked_img") else (P, H, W)
            dtype = torch.float32
        elif key in INV_KEYS_U8:
            shape, dtype = (P, H, W), torch.uint8
        else:
            raise KeyError(key)
        t = o.get(key)
        if t is None:
            t = torch.empty(shape, dtype=dtype, device=dev)
            o[key] = t
        assert t.shape == torch.Size(shape) and t.dtype == dtype and t.is_contiguous()
        setattr(outs, key, t.data_ptr())
    nbytes = L.syn3r_inv_workspace_bytes(P)
    ws = _workspace(nbytes, dev, "inv")
    with torch.cuda.device(dev):
        rc = L.syn3r_inverse_warp_batched(
            nat.ptr(src_img), depth_ptr, depth_stride, nat.ptr(src_bg), nat.ptr(tgt_depth), k_ptr, ki_ptr, nat.ptr(pair_P),
            nat.ptr(pair_src), nat.ptr(pair_tgt), P, H, W, float(bandwidth), float(thr_depth), float(thr_depth_strict),
            ctypes.byref(outs), ws.data_ptr(), ws.numel(), _stream_ptr(dev))
    nat.check(rc, "syn3r_inverse_warp_batched")
    return o


def consistency_batched(depth1, depth2, K1, K2, pair_T, pair_1, pair_2, out=None):
    """consistency_check_with_depth for P pairs: depth1 [A,H,W], depth2 [B,H,W] f32 CUDA, K1/K2 HOST 3x3,
    pair_T [P,2,16] f32 CUDA (pose2 @ inv(pose1), pose1 @ inv(pose2)).  Returns err [P,H,W] f32."""
    L = nat.lib()
    _, H, W = depth1.shape
    P = pair_1.numel()
    K1t = torch.as_tensor(np.asarray(K1)).float().cpu()
    K2t = torch.as_tensor(np.asarray(K2)).float().cpu()
    a1, p1 = nat.host_f32(K1t.numpy())
    a1i, p1i = nat.host_f32(torch.inverse(K1t).numpy())
    a2, p2 = nat.host_f32(K2t.numpy())
    err = out if out is not None else torch.empty((P, H, W), dtype=torch.float32, device=depth1.device)
    with torch.cuda.device(depth1.device):
        rc = L.syn3r_consistency_batched(nat.ptr(depth1), nat.ptr(depth2), p1, p1i, p2, nat.ptr(pair_T), nat.ptr(pair_1),
                                         nat.ptr(pair_2), P, H, W, nat.ptr(err), _stream_ptr(depth1.device))
    nat.check(rc, "syn3r_consistency_batched")
    return err
