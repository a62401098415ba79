"""Golden fixtures of the nearby-frame consistency check (second half): tests/golden/nearby_*.npz.

`consistency_check_from_nearby_images_bw` (/root/reference/model/diffusionGS.py:1300-1361) is a method of a class that
cannot be imported here (FSGS, open3d, diffusers are absent, SURVEY §8c caveat 3).  Its body is a loop around the
LIVE reference `inverse_warp` (solver_utils/forward_warp.py:185, imported unmodified through make_golden.py's
'cuda'->'cpu' shim); the generator re-executes the loop's own statements around that call, with device='cpu' and
without the hard-coded 1024x576 assert so that small frames can be stored.  Run in the build container:

    python tests/golden/make_golden_nearby.py
"""
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
import make_golden as mg  # noqa: E402  (device shim + the live reference modules)

inverse_warp = mg.ref_fw.inverse_warp
synth = mg.synth


def reference_nearby(intrinsics, interpolated_poses, images, depths, window_radius=1):
    """diffusionGS.py:1318-1357, statements in the reference's order (device='cuda' -> 'cpu')."""
    uncertainty_masks = []
    intensity_uncertainty_masks = []
    for cur_i in range(len(interpolated_poses)):
        masks = []
        warps = []
        for ref_i in range(cur_i - window_radius, cur_i + window_radius + 1):
            if ref_i == cur_i or ref_i < 0 or ref_i >= len(interpolated_poses):
                continue
            pose_s = interpolated_poses[ref_i]
            pose_t = interpolated_poses[cur_i]
            depth_s = depths[ref_i]
            depth_t = depths[cur_i]
            image_s = images[ref_i]
            image_t = images[cur_i]
            warp_dict = inverse_warp(
                torch.tensor(image_s, device='cpu', dtype=torch.float32).permute([2, 0, 1]),
                torch.tensor(depth_s[None], device='cpu'),
                depth_pseudo=torch.tensor(depth_t[None], device='cpu'),
                pose1=torch.tensor(pose_s, device='cpu', dtype=torch.float32),
                pose2=torch.tensor(pose_t, device='cpu', dtype=torch.float32),
                K=torch.tensor(intrinsics, device='cpu', dtype=torch.float32),
                bandwidth=10)
            masks.append(warp_dict['soft_mask_reproj'])
            warps.append(warp_dict['warped_img'])
        conf_mask = torch.stack(masks).mean(dim=0)
        uncertainty_masks.append(1 - conf_mask)
        warped_image = torch.stack(warps).mean(dim=0)
        intensity_conf = torch.exp(-((torch.norm(
            warped_image - torch.tensor(image_t, device='cpu', dtype=torch.float32).permute([2, 0, 1]), dim=0)) / 0.1) ** 3)
        intensity_uncertainty_masks.append(1 - intensity_conf)
    return uncertainty_masks, intensity_uncertainty_masks


def low_contrast_images(rng, N, H, W):
    """frames whose neighbours differ by a fraction of a grey level: ||warp - image|| spreads over the steep part of
    exp(-(d / 0.1)^3) instead of saturating at 0."""
    yy, xx = np.mgrid[0:H, 0:W].astype(np.float32)
    img = np.empty((N, H, W, 3), np.float32)
    for i in range(N):
        for c in range(3):
            img[i, ..., c] = (100.0 + 20.0 * c + 0.08 * np.sin(0.11 * xx + 0.3 * c + 0.05 * i) * np.cos(0.07 * yy)
                              + 0.03 * i + 0.02 * rng.random((H, W), dtype=np.float32))
    return img


def main():
    for k, (N, H, W, seed, low) in enumerate([(4, 48, 64, 61, True), (3, 40, 56, 62, False), (1, 16, 24, 63, True)]):
        tr = synth.make_trajectory(N, 1, H, W, seed=seed, pose_step=0.6)
        rng = np.random.default_rng(seed + 100)
        images = low_contrast_images(rng, N, H, W) if low else tr.rgb.astype(np.float32)
        if N == 1:                       # a single frame has no neighbour: torch.stack([]) raises in the reference
            try:
                reference_nearby(tr.K, list(tr.pose_src), list(images), list(tr.depth))
                raised = False
            except RuntimeError:
                raised = True
            mg.save(f"nearby_{k}", K=tr.K, poses=tr.pose_src, images=images, depths=tr.depth, raised=np.array(raised))
            continue
        unc, iunc = reference_nearby(tr.K, list(tr.pose_src), list(images), list(tr.depth))
        mg.save(f"nearby_{k}", K=tr.K, poses=tr.pose_src, images=images, depths=tr.depth,
                out_uncertainty=np.stack([u.numpy() for u in unc]), out_intensity=np.stack([u.numpy() for u in iunc]))


if __name__ == "__main__":
    sys.exit(main())
