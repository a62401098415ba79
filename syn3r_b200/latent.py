"""Guided latent replacement of the patched Euler scheduler on the device (scheduler half).

The reference (/root/reference/thirdparty/diffusers/src/diffusers/schedulers/scheduling_euler_discrete.py:1442-1499,
`step_interp_prob_uncertain`; `step_single` :548-589 and `step_interp` :735-778 keep the selection only, `top_masks`) loops over the 23
conditioning frames with a `torch.sort` and ~25 small launches each, 100 steps x 2 passes per clip.  Here the whole block
is ONE launch (`k_latent_guided_replace`, csrc/latent.cuh: a radix select per frame instead of the sort) through
`syn3r_latent_guided_replace`; results are bit-identical.  There is no CPU path.
"""
from __future__ import annotations

from typing import Optional, Sequence, Tuple

import torch

from . import _native as nat


def guided_replace_(pred_original_sample: torch.Tensor, temp_cond_latents: torch.Tensor, mask: torch.Tensor,
                    lambda_ts: torch.Tensor, step_index: int, index_list: Optional[Sequence[int]] = None, *,
                    return_cutoffs: bool = False):
    """In place on pred_original_sample [b,F,c,h,w] (float32, contiguous, CUDA), like the reference's block.
    temp_cond_latents [2,F,c,h,w] (batch item 1 is used, :1454), mask [bm,F-1,cm,h,w] with 1 = hole as the pipeline
    passes it, lambda_ts [steps,F], index_list: the frames to blend (default 1 .. F-2; the reference hard-codes
    1 .. 23 for F = 25) - a contiguous range.  Returns pred_original_sample (and the cutoff indices / values)."""
    p = pred_original_sample
    assert p.is_cuda and p.dtype == torch.float32 and p.is_contiguous() and p.dim() == 5
    b, F, c, h, w = p.shape
    assert temp_cond_latents.shape[1:] == (F, c, h, w) and temp_cond_latents.shape[0] >= 2
    cond = temp_cond_latents[1].to(torch.float32).contiguous()
    m = mask.to(torch.float32).contiguous()
    assert m.dim() == 5 and m.shape[3:] == (h, w)
    idx = list(range(1, F - 1)) if index_list is None else [int(i) for i in index_list]
    assert idx == list(range(idx[0], idx[0] + len(idx))) if idx else True, "index_list must be a contiguous range"
    assert not idx or m.shape[1] >= idx[-1], "frame tau reads mask[:, tau - 1] (the scheduler prepends a frame of ones, :1449)"
    row = lambda_ts[step_index].to(device=p.device, dtype=torch.float32).contiguous()
    assert row.shape == (F,)
    n_tau = len(idx)
    ks = torch.empty(max(n_tau, 1), dtype=torch.int32, device=p.device) if return_cutoffs else None
    cv = torch.empty(max(n_tau, 1), dtype=torch.float32, device=p.device) if return_cutoffs else None
    nat.check(nat.lib().syn3r_latent_guided_replace(nat.ptr(p), nat.ptr(cond), nat.ptr(m), nat.ptr(row), b, F, c, h, w,
                                                    m.shape[0], m.shape[1], m.shape[2], idx[0] if idx else 1, n_tau,
                                                    nat.ptr(ks) if return_cutoffs else None,
                                                    nat.ptr(cv) if return_cutoffs else None,
                                                    torch.cuda.current_stream(p.device).cuda_stream),
              "syn3r_latent_guided_replace")
    return (p, ks[:n_tau], cv[:n_tau]) if return_cutoffs else p


def top_masks(pred_original_sample: torch.Tensor, temp_cond_latents: torch.Tensor, mask: torch.Tensor,
              lambda_ts: torch.Tensor, step_index: int, weight_clamp: float, index_list: Optional[Sequence[int]] = None, *,
              trailing_ones: bool = False, return_cutoffs: bool = False):
    """The selection-only form of `step_single` (:548-589) / `step_interp` (:735-778): same cutoff as guided_replace_
    with `weight_clamp` as the lower clamp, nothing written to pred.  Returns the reference's top_masks, bool
    [1, 1 + len(index_list) * b, c, h, w] (frame of ones first).  index_list defaults to 1 .. F-1 (step_single).
    `trailing_ones=True` is `step_interp`: index_list defaults to 1 .. F-2 and a second frame of ones closes the
    stack (:775-777), so the result covers all F frames.
    dtype contract: pred / cond latents float32; the mask is thresholded as float32 `(1 - mask) > 0.5` (:743),
    which equals the float64 comparison of the non-Prob pipeline for every 0/1-valued or pooled 1/64-step mask."""
    p = pred_original_sample
    assert p.is_cuda and p.dtype == torch.float32 and p.is_contiguous() and p.dim() == 5
    b, F, c, h, w = p.shape
    assert temp_cond_latents.shape[1:] == (F, c, h, w) and temp_cond_latents.shape[0] >= 2
    cond = temp_cond_latents[1].to(torch.float32).contiguous()
    m = mask.to(torch.float32).contiguous()
    assert m.dim() == 5 and m.shape[3:] == (h, w)
    idx = list(range(1, F - 1 if trailing_ones else F)) if index_list is None else [int(i) for i in index_list]
    assert idx and idx == list(range(idx[0], idx[0] + len(idx))), "index_list must be a contiguous range"
    assert m.shape[1] >= idx[-1], "frame tau reads mask[:, tau - 1] (the scheduler prepends a frame of ones, :556)"
    row = lambda_ts[step_index].to(device=p.device, dtype=torch.float32).contiguous()
    n_tau = len(idx)
    top = torch.empty((n_tau * b, c, h, w), dtype=torch.uint8, device=p.device)
    ks = torch.empty(n_tau, dtype=torch.int32, device=p.device)
    cv = torch.empty(n_tau, dtype=torch.float32, device=p.device)
    nat.check(nat.lib().syn3r_latent_top_masks(nat.ptr(p), nat.ptr(cond), nat.ptr(m), nat.ptr(row), b, F, c, h, w, m.shape[0],
                                               m.shape[1], m.shape[2], idx[0], n_tau, float(weight_clamp), nat.ptr(top), nat.ptr(ks),
                                               nat.ptr(cv), torch.cuda.current_stream(p.device).cuda_stream),
              "syn3r_latent_top_masks")
    ones = torch.ones((1, 1, c, h, w), dtype=torch.bool, device=p.device)
    out = torch.cat([ones, top.bool()[None]] + ([ones] if trailing_ones else []), 1)                         # :586-589, :775-777
    return (out, ks, cv) if return_cutoffs else out
