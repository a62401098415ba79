"""Forward-backward optical-flow consistency on the device (gmflow half).

Mirror of /root/reference/thirdparty/gmflow/gmflow/geometry.py:75-96 `forward_backward_consistency_check` (same
name, arguments and return value): the reference's ~40 elementwise / grid_sample passes become one launch of
`k_flow_fb_consistency` (csrc/flow.cuh) through `syn3r_flow_fb_consistency_batched`.  There is no CPU path.
"""
from __future__ import annotations

from typing import Tuple

import torch

from . import _native as nat


def forward_backward_consistency_check(fwd_flow: torch.Tensor, bwd_flow: torch.Tensor, alpha: float = 0.01,
                                       beta: float = 0.5, *, return_diff: bool = False) -> Tuple[torch.Tensor, ...]:
    """fwd_flow, bwd_flow: [B, 2, H, W] float32 CUDA tensors.  Returns (fwd_occ, bwd_occ), float32 [B, H, W] of 0 / 1
    (geometry.py:93-96); with return_diff also the two norms that were thresholded (extension)."""
    assert fwd_flow.dim() == 4 and bwd_flow.dim() == 4                      # geometry.py:81-82
    assert fwd_flow.size(1) == 2 and bwd_flow.size(1) == 2
    assert fwd_flow.shape == bwd_flow.shape and fwd_flow.dtype == bwd_flow.dtype == torch.float32
    assert fwd_flow.is_cuda and bwd_flow.device == fwd_flow.device
    fwd_flow, bwd_flow = fwd_flow.contiguous(), bwd_flow.contiguous()
    B, _, H, W = fwd_flow.shape
    mk = lambda: torch.empty((B, H, W), dtype=torch.float32, device=fwd_flow.device)  # noqa: E731
    fwd_occ, bwd_occ = mk(), mk()
    d_f, d_b = (mk(), mk()) if return_diff else (None, None)
    nat.check(nat.lib().syn3r_flow_fb_consistency_batched(nat.ptr(fwd_flow), nat.ptr(bwd_flow), B, H, W, float(alpha), float(beta),
                                                          nat.ptr(fwd_occ), nat.ptr(bwd_occ),
                                                          nat.ptr(d_f) if return_diff else None,
                                                          nat.ptr(d_b) if return_diff else None,
                                                          torch.cuda.current_stream(fwd_flow.device).cuda_stream),
              "syn3r_flow_fb_consistency_batched")
    return (fwd_occ, bwd_occ, d_f, d_b) if return_diff else (fwd_occ, bwd_occ)
