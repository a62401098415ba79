"""Golden fixtures of the nearby-frame consistency check (second half): tests/golden/nearby_*.npz.

`consistency_check_from_nearby_images_bw` (/root/reference/model/diffusionGS.py:1300-1361) is a method of a class that
cannot be imported here (FSGS, open3d, diffusers are absent, SURVEY §8c caveat 3).  The generator slices the METHOD'S
OWN SOURCE TEXT from the file and exec-s it (tests/golden/ref_exec.py) around the LIVE reference `inverse_warp`
(solver_utils/forward_warp.py:187, imported unmodified through make_golden.py's 'cuda'->'cpu' shim); no line of it is
retyped or patched, and each fixture records the slice's SHA-256.  Run in the build container:

    python tests/golden/make_golden_nearby.py
"""
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
import make_golden as mg  # noqa: E402  (device shim + the live reference modules)
import ref_exec  # noqa: E402

inverse_warp = mg.ref_fw.inverse_warp
synth = mg.synth


def reference_nearby(intrinsics, interpolated_poses, images, depths):
    """The reference's method itself: `consistency_check_from_nearby_images_bw` (diffusionGS.py:1300-1361) sliced from
    the file and exec-ed (ref_exec.py) with the live `inverse_warp`, a stub `self` carrying the asserted 1024 x 576
    diffusion size (the method never uses it: frames of any size run) and a `torch` whose factory calls map the
    hard-coded device='cuda' to 'cpu'.  No textual patch."""
    import types
    fn, tag, desc = ref_exec.exec_function("model/diffusionGS.py", 1300, 1361, "consistency_check_from_nearby_images_bw",
                                           dict(torch=ref_exec.TorchOnCpu(), np=np, inverse_warp=inverse_warp))
    assert not desc
    self = types.SimpleNamespace(diffusion_width=1024, diffusion_height=576)
    unc, iunc = fn(self, intrinsics, interpolated_poses, images, depths)
    return unc, iunc, tag


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
            mg.save(f"nearby_{k}", K=tr.K, poses=tr.pose_src, images=images, depths=tr.depth, raised=np.array(raised),
                    ref_source=np.array(ref_exec.source("model/diffusionGS.py", 1300, 1361)[1]))
            continue
        unc, iunc, tag = reference_nearby(tr.K, list(tr.pose_src), list(images), list(tr.depth))
        mg.save(f"nearby_{k}", K=tr.K, poses=tr.pose_src, images=images, depths=tr.depth, ref_source=np.array(tag),
                out_uncertainty=np.stack([u.numpy() for u in unc]), out_intensity=np.stack([u.numpy() for u in iunc]))


if __name__ == "__main__":
    sys.exit(main())
