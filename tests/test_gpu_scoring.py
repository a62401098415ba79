"""GPU: batched pose-candidate scoring against the per-candidate form the reference uses
(model/diffusionGS.py:742-765): one inverse_warp per candidate, `(1 - soft_mask_reproj).mean()`, argmax per group."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

from oracle import cpu_oracle as orc  # noqa: E402  (checker only)
from syn3r_b200 import synth  # noqa: E402


def test_mean_uncertainty_vs_float64_and_determinism():
    from syn3r_b200 import scoring
    g = torch.Generator(device="cuda").manual_seed(1)
    for (P, H, W) in [(7, 576, 1024), (3, 61, 36), (1, 5, 7), (150, 96, 128)]:
        if (H * W) % 4 and P > 1:
            continue
        soft = torch.rand((P, H, W), device="cuda", generator=g)
        a, b = scoring.mean_uncertainty(soft), scoring.mean_uncertainty(soft)
        assert torch.equal(a, b)                                           # fixed-order reduction
        ref = (1.0 - soft.double()).mean(dim=(1, 2))
        assert float((a.double() - ref).abs().max()) <= 2e-7
        t = (1 - soft).mean(dim=(1, 2))                                    # what the reference evaluates (torch, CUDA)
        assert float((a - t).abs().max()) <= 1e-6


def test_score_pose_candidates_matches_per_candidate_reference_form():
    from syn3r_b200 import scoring
    S, T, H, W = 3, 4, 72, 96
    tr = synth.make_trajectory(S, T, H, W, seed=17, pose_step=1.5)
    rng = np.random.default_rng(3)
    groups = []
    for a in range(T):                                                     # anchors = the trajectory's targets
        grp = [tr.pose_tgt[a].copy()]
        for _ in range(3):
            p = tr.pose_tgt[a].copy()
            p[:3, 3] += rng.normal(0, 0.01, 3).astype(np.float32)
            grp.append(p)
        groups.append(grp)
    cand = [p for g in groups for p in g]
    cand_depth = np.stack([tr.depth_tgt[i // 4] * np.float32(1 + 0.002 * (i % 4)) for i in range(len(cand))])
    up = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()  # noqa: E731
    u = scoring.score_pose_candidates(up(tr.depth), list(tr.pose_src), up(cand_depth), cand, tr.K.astype(np.float32))
    torch.cuda.synchronize()
    u = u.cpu().numpy()
    t_ = torch.from_numpy
    img0 = np.zeros((3, H, W), np.float32)
    ref_u = []
    for i, pose in enumerate(cand):                                        # the reference's loop, through the oracle
        r = scoring.nearest_ref_index(pose, list(tr.pose_src))
        iw = orc.inverse_warp(t_(img0), t_(tr.depth[r][None]), t_(cand_depth[i][None]), t_(tr.pose_src[r]), t_(pose),
                              t_(tr.K.astype(np.float32)), bandwidth=20)
        ref_u.append(float((1 - iw["soft_mask_reproj"].double()).mean()))
    ref_u = np.array(ref_u)
    assert np.max(np.abs(u - ref_u)) <= 2e-6
    sel = scoring.select_most_uncertain(groups, u)
    for gi, g in enumerate(groups):
        assert sel[gi] is g[int(np.argmax(ref_u[4 * gi:4 * gi + 4]))]
