"""Batched scoring of perturbed pose candidates.

The reference scores every perturbed pose of every anchor with one `inverse_warp` call and a mean
(/root/reference/model/diffusionGS.py:716-766 `_perturb_and_select_interp_poses`: nearest reference by translation
distance :718-726, `inverse_warp(ref_img, ref_depth, depth_pseudo, ref_pose, perturbed_pose, K, bandwidth=20)` :758,
`uncertainty = (1 - soft_mask_reproj).mean()` :759-761, `argmax` per anchor group :763): 25 anchors x 6 candidates =
150 calls per view pair.  Here all candidates are one batched consistency launch (`k_reproj_masks`) plus one
deterministic reduction (`syn3r_mean_uncertainty_batched`); the rendering of the candidates' pseudo depth
(`render_GS`) stays with the caller.  No CPU path.
"""
from __future__ import annotations

from typing import List, Sequence

import numpy as np
import torch

from . import _native as nat
from . import warp as _w


def nearest_ref_index(pose: np.ndarray, ref_poses: Sequence[np.ndarray]) -> int:
    """diffusionGS.py:718-726: index of the reference pose whose translation column is nearest."""
    ref_t = np.array([p[:3, 3] for p in ref_poses])
    return int(np.argmin(np.linalg.norm(ref_t - pose[:3, 3], axis=1)))


def mean_uncertainty(soft: torch.Tensor) -> torch.Tensor:
    """soft [P,H,W] f32 CUDA -> [P] f32: mean over the frame of 1 - soft (diffusionGS.py:759-761)."""
    L = nat.lib()
    assert soft.is_cuda and soft.dtype == torch.float32 and soft.dim() == 3 and soft.is_contiguous()
    P, H, W = soft.shape
    out = torch.empty((P,), dtype=torch.float32, device=soft.device)
    ws = torch.empty(max(int(L.syn3r_mean_uncertainty_workspace_bytes(P)), 8), dtype=torch.uint8, device=soft.device)
    nat.check(L.syn3r_mean_uncertainty_batched(nat.ptr(soft), P, H, W, nat.ptr(out), nat.ptr(ws), ws.numel(),
                                               torch.cuda.current_stream(soft.device).cuda_stream),
              "syn3r_mean_uncertainty_batched")
    return out


def score_pose_candidates(ref_depths: torch.Tensor, ref_poses: Sequence[np.ndarray], cand_depths: torch.Tensor,
                          cand_poses: Sequence[np.ndarray], K, bandwidth: float = 20.0) -> torch.Tensor:
    """Uncertainty of every candidate: ref_depths [R,H,W] f32 CUDA (GS depth renders of the reference poses),
    cand_depths [C,H,W] f32 CUDA (the candidates' pseudo depth), poses w2c 4x4, K 3x3.  Returns [C] f32 CUDA.
    Candidate c is warped against its nearest reference (`nearest_ref_index`), exactly the pair the reference forms."""
    ref_poses = np.asarray(ref_poses, np.float32)
    cand_poses = np.asarray(cand_poses, np.float32)
    C = len(cand_poses)
    assert cand_depths.shape[0] == C and ref_depths.shape[0] == len(ref_poses)
    src = np.array([nearest_ref_index(p, ref_poses) for p in cand_poses], np.int32)
    tgt = np.arange(C, dtype=np.int32)
    dev = ref_depths.device
    pair_P = torch.from_numpy(_w.inverse_pair_params(ref_poses, cand_poses, src, tgt)).to(dev)
    o = _w.inverse_warp_batched(None, ref_depths, cand_depths, K, pair_P, torch.from_numpy(src).to(dev),
                                torch.from_numpy(tgt).to(dev), bandwidth=bandwidth, want=("soft_mask_reproj",))
    return mean_uncertainty(o["soft_mask_reproj"])


def select_most_uncertain(groups: Sequence[Sequence[np.ndarray]], uncertainties: np.ndarray) -> List[np.ndarray]:
    """diffusionGS.py:763-765: per anchor group the candidate with the highest uncertainty (first one on ties).
    `uncertainties` follows the groups flattened in order."""
    out, k = [], 0
    for g in groups:
        u = np.asarray(uncertainties[k:k + len(g)])
        out.append(g[int(np.argmax(u))])
        k += len(g)
    return out
