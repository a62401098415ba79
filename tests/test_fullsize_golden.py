"""CPU: the oracle against LIVE-REFERENCE outputs at the benchmark sizes (576x1024, 300x400, 540x960; two pairs
each, tests/golden/fullsize_*.npz).  The small fixtures stop at 120x160; this is the committed evidence that the
restatement the GPU is checked against is the reference at the sizes that are timed."""
import numpy as np
import pytest
import torch

from oracle import cpu_oracle as orc  # checker only
from _fullsize_util import bits, load


@pytest.mark.parametrize("letter", ["B", "C", "D"])
def test_oracle_matches_live_reference_at_benchmark_size(golden, letter):
    tr, g, pairs, rows = load(golden, letter)
    H, W = tr.depth.shape[1:]
    t_ = torch.from_numpy
    for k, (s, t) in enumerate(pairs):
        with np.errstate(all="ignore"):
            r = orc.forward_warp_ext([tr.rgb[s]], [None], [tr.depth[s]], [tr.pose_src[s]], tr.pose_tgt[t], [tr.K], None)
            iw = orc.inverse_warp(t_(np.ascontiguousarray(np.moveaxis(tr.rgb[s], -1, 0))), t_(tr.depth[s][None]),
                                  t_(tr.depth_tgt[t][None]), t_(tr.pose_src[s]), t_(tr.pose_tgt[t]), t_(tr.K.astype(np.float32)))
        assert np.array_equal(r["mask"], bits(g, f"p{k}_mask_bits", H, W)), (letter, k, "hole mask")
        u8 = np.round(np.clip(r["out"], 0, 255)).astype(np.uint8)
        assert np.array_equal(u8.reshape(H, -1).sum(1, dtype=np.int64), g[f"p{k}_warped_rowsum"]), (letter, k, "uint8 row sums")
        if k == 0:
            assert np.array_equal(u8, g["p0_warped_u8"]), (letter, "uint8 frame")
        assert np.array_equal(iw["mask_warp"].numpy(), bits(g, f"p{k}_mask_warp_bits", H, W))
        # fp32 threshold mask: the oracle follows torch's CPU op order, so it is exact here; the band is recorded for the GPU
        assert np.array_equal(iw["mask_reproj"].numpy(), bits(g, f"p{k}_mask_reproj_bits", H, W)), (letter, k, "mask_reproj")
        a, b = iw["soft_mask_reproj"].numpy()[rows], g[f"p{k}_soft_rows"]
        fin = np.isfinite(a) & np.isfinite(b)
        assert np.array_equal(np.isnan(a), np.isnan(b)) and np.max(np.abs(a[fin] - b[fin]), initial=0) <= 1e-6
