"""Full-size golden fixtures from the LIVE reference at the benchmark sizes: tests/golden/fullsize_{B,C,D}.npz.

For each of the single-GPU configurations of BASELINE.json (B 576x1024, C 300x400, D 540x960) two (source, target)
pairs of the benchmark's own synthetic trajectory (`synth.make_trajectory(S, T, H, W, seed=0)`, regenerated from the
seed by the consumer and guarded by a checksum) go through the unmodified `forward_warp` and `inverse_warp` of
/root/reference/solver_utils/forward_warp.py.  Stored per pair: the hole mask and `mask_reproj` as bits, the uint8
frame (first pair in full, second as per-row sums), rows of `soft_mask_reproj`, probes of `flow12`.  `-m gpu` tests
on the box then compare the CUDA path with REFERENCE outputs at the benchmark size, not only with the port.

    python tests/golden/make_golden_fullsize.py        (~20 s)
"""
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
import make_golden as mg  # noqa: E402

ROWS = slice(5, None, 32)
PAIRS = {"B_svd": [(1, 12), (2, 24)], "C_dtu": [(0, 40), (2, 7)], "D_dl3dv": [(1, 3), (0, 24)]}


def checksum(tr):
    """Order-sensitive digest of the regenerated inputs (float64 dot with a fixed ramp would drown in rounding; exact
    integer sums of the raw bit patterns do not)."""
    parts = [tr.rgb, tr.depth, tr.depth_tgt, tr.pose_src, tr.pose_tgt]
    return np.array([int(np.ascontiguousarray(a).view(np.uint32).astype(np.uint64).sum()) for a in parts], np.uint64)


def main():
    t_ = torch.from_numpy
    for cfg, pairs in PAIRS.items():
        S, T, H, W = mg.synth.CONFIGS[cfg]
        tr = mg.synth.make_trajectory(S, T, H, W, seed=0)
        arrs = dict(pairs=np.array(pairs, np.int32), input_checksum=checksum(tr), rows=np.array([ROWS.start, ROWS.step]))
        for k, (s, t) in enumerate(pairs):
            warped, mask2, flow = mg.ref_fw.forward_warp(tr.rgb[s], None, tr.depth[s], tr.pose_src[s], tr.pose_tgt[t], tr.K, None)
            iw = mg.ref_fw.inverse_warp(t_(np.ascontiguousarray(np.moveaxis(tr.rgb[s], -1, 0))), t_(tr.depth[s][None]),
                                        t_(tr.depth_tgt[t][None]), t_(tr.pose_src[s]), t_(tr.pose_tgt[t]),
                                        t_(tr.K.astype(np.float32)))
            arrs[f"p{k}_mask_bits"] = np.packbits(mask2)
            if k == 0:
                arrs[f"p{k}_warped_u8"] = warped
            arrs[f"p{k}_warped_rowsum"] = warped.reshape(H, -1).sum(1, dtype=np.int64)
            arrs[f"p{k}_flow_probe"] = flow[::29, ::31]
            arrs[f"p{k}_mask_reproj_bits"] = np.packbits(iw["mask_reproj"].numpy())
            arrs[f"p{k}_mask_warp_bits"] = np.packbits(iw["mask_warp"].numpy())
            arrs[f"p{k}_soft_rows"] = iw["soft_mask_reproj"].numpy()[ROWS]
            # distance of the reprojection error to the bandwidth where the thresholded mask could flip
            err = mg.ref_cc.consistency_check_with_depth(t_(tr.depth_tgt[t]), t_(tr.pose_tgt[t]), t_(tr.K.astype(np.float32)),
                                                         t_(tr.depth[s]), t_(tr.pose_src[s]), t_(tr.K.astype(np.float32))).numpy()
            arrs[f"p{k}_band_bits"] = np.packbits(np.abs(err - 20.0) <= 1e-3 * 20.0)
        mg.save("fullsize_" + cfg[0], **arrs)


if __name__ == "__main__":
    main()
