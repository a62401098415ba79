"""CPU restatement of gmflow's forward-backward flow consistency check.

TEST INFRASTRUCTURE ONLY: imported by tests/ as the checker, never by the product (syn3r_b200/ has no CPU path).

Follows /root/reference/thirdparty/gmflow/gmflow/geometry.py in float32 NumPy, op by op:
  coords_grid                       :5-21    x = column, y = row, as float32
  bilinear_sample                   :41-62   x_grid = 2 sx / (w - 1) - 1 (a tensor divided by a Python scalar is a multiply
                                             by float32(1 / (w - 1)) on the CUDA device gmflow runs on), then
                                             F.grid_sample(bilinear, zeros, align_corners=True): ix = ((x_grid + 1) / 2) (w - 1),
                                             corner weights (ix_se - ix)(iy_se - iy) ..., out = sum of in-bounds value * weight
                                             accumulated nw, ne, sw, se with fused multiply-adds (ATen GridSampler.cu)
  flow_warp                         :65-72   sample `feature` at coords + flow
  forward_backward_consistency_check :75-96  |fwd + warp(bwd, fwd)| > alpha (|fwd| + |bwd|) + beta, and the mirror image

Parity pinned by tests/golden/flow_fb_*.npz, produced by importing and calling the live reference on the CPU
(tests/golden/make_golden_flow.py).  torch's CPU build divides by (w - 1) and sums the four taps without fused
multiply-adds, so intermediates agree to a few float32 ulps (1e-4 absolute on flows of tens of pixels) and the binary
maps agree wherever |diff - threshold| exceeds that.
"""
import numpy as np

f32 = np.float32


def _fma(a, b, c):
    """float32 fused multiply-add: the product is exact in float64, one rounding of the sum to float64 and one to
    float32 (double rounding differs from a true fma only on exact float64 ties)."""
    return (a.astype(np.float64) * b.astype(np.float64) + c.astype(np.float64)).astype(f32)


def flow_warp(feature: np.ndarray, flow: np.ndarray) -> np.ndarray:
    """geometry.py:65-72 / :41-62.  feature, flow: [B, 2, H, W] float32."""
    with np.errstate(invalid="ignore", over="ignore"):        # inf / NaN flows propagate as in torch
        return _flow_warp(feature, flow)


def _flow_warp(feature, flow):
    B, C, H, W = feature.shape
    yy, xx = np.meshgrid(np.arange(H, dtype=f32), np.arange(W, dtype=f32), indexing="ij")
    sx = (xx[None] + flow[:, 0]).astype(f32)
    sy = (yy[None] + flow[:, 1]).astype(f32)
    xg = ((f32(2) * sx).astype(f32) * (f32(1) / f32(W - 1))).astype(f32) - f32(1)
    yg = ((f32(2) * sy).astype(f32) * (f32(1) / f32(H - 1))).astype(f32) - f32(1)
    ix = (((xg + f32(1)).astype(f32) * f32(0.5)).astype(f32) * f32(W - 1)).astype(f32)
    iy = (((yg + f32(1)).astype(f32) * f32(0.5)).astype(f32) * f32(H - 1)).astype(f32)
    x0, y0 = np.floor(ix), np.floor(iy)
    x1, y1 = x0 + f32(1), y0 + f32(1)
    wts = [((x1 - ix).astype(f32) * (y1 - iy).astype(f32)).astype(f32),      # nw
           ((ix - x0).astype(f32) * (y1 - iy).astype(f32)).astype(f32),      # ne
           ((x1 - ix).astype(f32) * (iy - y0).astype(f32)).astype(f32),      # sw
           ((ix - x0).astype(f32) * (iy - y0).astype(f32)).astype(f32)]      # se
    taps = [(x0, y0), (x1, y0), (x0, y1), (x1, y1)]
    out = np.zeros_like(feature)
    bidx = np.arange(B)[:, None, None]
    with np.errstate(invalid="ignore"):
        for (tx, ty), wt in zip(taps, wts):
            ok = (tx >= 0) & (tx <= W - 1) & (ty >= 0) & (ty <= H - 1)       # NaN coordinates fail every test
            xi = np.where(ok, tx, 0).astype(np.int64)
            yi = np.where(ok, ty, 0).astype(np.int64)
            for c in range(C):
                v = feature[bidx, c, yi, xi]
                out[:, c] = np.where(ok, _fma(v, wt, out[:, c]), out[:, c])
    return out


def _norm2(a: np.ndarray) -> np.ndarray:
    return np.sqrt(((a[:, 0] * a[:, 0]).astype(f32) + (a[:, 1] * a[:, 1]).astype(f32)).astype(f32)).astype(f32)


def forward_backward_consistency_check(fwd_flow: np.ndarray, bwd_flow: np.ndarray, alpha: float = 0.01, beta: float = 0.5) -> dict:
    """geometry.py:75-96.  Returns the two occlusion maps (float32 0 / 1) and the intermediates."""
    fwd_flow, bwd_flow = fwd_flow.astype(f32), bwd_flow.astype(f32)
    assert fwd_flow.ndim == 4 and bwd_flow.ndim == 4 and fwd_flow.shape[1] == 2 and bwd_flow.shape[1] == 2
    mag = (_norm2(fwd_flow) + _norm2(bwd_flow)).astype(f32)
    diff_fwd = _norm2((fwd_flow + flow_warp(bwd_flow, fwd_flow)).astype(f32))
    diff_bwd = _norm2((bwd_flow + flow_warp(fwd_flow, bwd_flow)).astype(f32))
    thr = ((f32(alpha) * mag).astype(f32) + f32(beta)).astype(f32)
    return dict(fwd_occ=(diff_fwd > thr).astype(f32), bwd_occ=(diff_bwd > thr).astype(f32), diff_fwd=diff_fwd,
                diff_bwd=diff_bwd, threshold=thr)
