"""Shared by the CPU and GPU full-size tests: the benchmark trajectory regenerated from its seed, guarded by the
checksum stored beside the live-reference outputs (tests/golden/fullsize_*.npz, make_golden_fullsize.py)."""
import numpy as np

from syn3r_b200 import synth

CFG = {"B": "B_svd", "C": "C_dtu", "D": "D_dl3dv"}


def checksum(tr):
    parts = [tr.rgb, tr.depth, tr.depth_tgt, tr.pose_src, tr.pose_tgt]
    return np.array([int(np.ascontiguousarray(a).view(np.uint32).astype(np.uint64).sum()) for a in parts], np.uint64)


def load(golden, letter):
    """(trajectory, fixture dict, [(s, t)], row slice of the stored soft-mask rows)."""
    g = golden("fullsize_" + letter)
    S, T, H, W = synth.CONFIGS[CFG[letter]]
    tr = synth.make_trajectory(S, T, H, W, seed=0)
    assert np.array_equal(checksum(tr), g["input_checksum"]), \
        "the regenerated synthetic trajectory differs from the one the live reference was run on (numpy / scipy drift?)"
    r0, rs = (int(v) for v in g["rows"])
    return tr, g, [tuple(int(v) for v in p) for p in g["pairs"]], slice(r0, None, rs)


def bits(g, key, H, W):
    return np.unpackbits(g[key])[:H * W].reshape(H, W).astype(bool)
