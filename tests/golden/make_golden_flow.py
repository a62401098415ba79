"""Golden fixtures of the forward-backward flow consistency check (gmflow half):
tests/golden/flow_fb_*.npz.

RUNS THE LIVE REFERENCE: /root/reference/thirdparty/gmflow/gmflow/geometry.py is imported unmodified (it needs
only torch) and `forward_backward_consistency_check` (:75-96) is called on seeded flow pairs, CPU tensors.  Stored
besides the two occlusion maps: the reference's own intermediates recomputed with its helpers (`flow_warp`, :65-72),
so that the test can tell a pixel that sits on the threshold from a wrong one.  Run in the build container:

    python tests/golden/make_golden_flow.py
"""
import importlib.util
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
REF = os.environ.get("SYN3R_REFERENCE", "/root/reference")
spec = importlib.util.spec_from_file_location("ref_gmflow_geometry", os.path.join(REF, "thirdparty/gmflow/gmflow/geometry.py"))
geo = importlib.util.module_from_spec(spec)
spec.loader.exec_module(geo)


def flows(rng, B, H, W, kind):
    """a smooth forward flow and a backward flow that undoes it up to noise, plus regions that disagree (occlusions),
    flows leaving the frame, and exactly-integer displacements."""
    yy, xx = np.mgrid[0:H, 0:W].astype(np.float32)
    fwd = np.empty((B, 2, H, W), np.float32)
    bwd = np.empty((B, 2, H, W), np.float32)
    for b in range(B):
        a = 0.6 + 0.4 * b                                # slowly varying: bwd(x + fwd(x)) ~ -fwd(x) up to the noise
        fwd[b, 0] = 2.5 + a * np.sin(0.09 * yy + b) + 0.01 * xx
        fwd[b, 1] = -1.75 + a * np.cos(0.07 * xx) - 0.01 * yy
        bwd[b] = -fwd[b] + rng.normal(0, 0.3 if kind != "integer" else 0.0, (2, H, W)).astype(np.float32)
        if kind == "occluded":
            bwd[b, :, H // 4:H // 2, W // 3:W // 2] += 4.0
            fwd[b, 0, :, -6:] += 30.0                    # leaves the frame on the right
            bwd[b, 1, :5, :] -= 25.0                     # leaves the frame at the top
        if kind == "integer":
            fwd[b] = np.round(fwd[b])
            bwd[b] = -fwd[b]
            bwd[b, 0, ::3, ::4] += 1.0
    return fwd, bwd


def main():
    vers = np.array([f"numpy {np.__version__}", f"torch {torch.__version__}",
                     "live DecaYale/SYN3R@402f4e9 thirdparty/gmflow/gmflow/geometry.py:75-96"])
    rng = np.random.default_rng(404)
    for i, (kind, B, H, W, alpha, beta) in enumerate([("smooth", 2, 40, 56, 0.01, 0.5), ("occluded", 2, 48, 64, 0.01, 0.5),
                                                       ("integer", 1, 24, 40, 0.01, 0.5), ("smooth", 1, 33, 47, 0.05, 0.1)]):
        fwd, bwd = flows(rng, B, H, W, kind)
        tf, tb = torch.from_numpy(fwd), torch.from_numpy(bwd)
        fwd_occ, bwd_occ = geo.forward_backward_consistency_check(tf, tb, alpha=alpha, beta=beta)
        diff_fwd = torch.norm(tf + geo.flow_warp(tb, tf), dim=1)
        diff_bwd = torch.norm(tb + geo.flow_warp(tf, tb), dim=1)
        thr = alpha * (torch.norm(tf, dim=1) + torch.norm(tb, dim=1)) + beta
        path = os.path.join(HERE, f"flow_fb_{i}_{kind}.npz")
        np.savez_compressed(path, versions=vers, fwd_flow=fwd, bwd_flow=bwd, alpha=np.float64(alpha), beta=np.float64(beta),
                            out_fwd_occ=fwd_occ.numpy(), out_bwd_occ=bwd_occ.numpy(), aux_diff_fwd=diff_fwd.numpy(),
                            aux_diff_bwd=diff_bwd.numpy(), aux_threshold=thr.numpy())
        print(os.path.basename(path), os.path.getsize(path) // 1024, "KiB", "occluded fraction",
              float(fwd_occ.mean()), float(bwd_occ.mean()))


if __name__ == "__main__":
    sys.exit(main())
