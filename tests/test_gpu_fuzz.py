"""GPU fuzz of the forward tile kernel against the oracle: random poses (large rotations, zoom in / out), random
intrinsics, depth fields with zeros, negatives, NaN, inf, huge values (z' > 5000: the upper clip of forward_warp.py:82)
and tiny scenes (50 / Lmax > 500: sweep 1 falls back to the exact weight, forward_fx.cuh fx_weight_upper).
Masks and winner indices bit-exact, colours within 1e-5 on the [0,1] scale, for every case."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

from oracle import cpu_oracle as orc  # noqa: E402  (checker only)
from syn3r_b200 import synth  # noqa: E402


def _rand_pose(rng, rot_deg, trans):
    T = np.eye(4)
    T[:3, :3] = synth.euler_deg_to_R(*(rng.uniform(-rot_deg, rot_deg, 3)))
    T[:3, 3] = rng.uniform(-trans, trans, 3)
    return T.astype(np.float32)


def _case(rng, kind):
    H, W = int(rng.integers(6, 70)), int(rng.integers(6, 90))
    S, T = int(rng.integers(1, 3)), int(rng.integers(1, 3))
    scale = {"tiny": 0.02, "huge": 3000.0}.get(kind, 1.0)
    depth = (scale * (1.0 + 2.0 * rng.random((S, H, W)))).astype(np.float32)
    if kind == "special":
        for s in range(S):
            for val in (0.0, -1.3, np.nan, np.inf, 1e4):
                y, x = rng.integers(0, H), rng.integers(0, W)
                depth[s, y:y + 3, x:x + 4] = val
    if kind == "huge":
        depth[:, ::5, ::3] = 9000.0
    rgb = (rng.random((S, H, W, 3), dtype=np.float32) * np.float32(255)).astype(np.float32)
    f = rng.uniform(0.4, 2.5) * W
    K = np.array([[f, 0, W / 2 + rng.uniform(-3, 3)], [0, f * rng.uniform(0.8, 1.2), H / 2 + rng.uniform(-3, 3)], [0, 0, 1]], np.float64)
    rot, tr = {"zoom": (3, 0.1), "wild": (35, 1.5 * scale)}.get(kind, (8, 0.4 * scale))
    pose_src = np.stack([_rand_pose(rng, rot, tr) for _ in range(S)])
    pose_tgt = np.stack([_rand_pose(rng, rot, tr) for _ in range(T)])
    if kind == "zoom":                                       # strong minification / magnification along the axis
        pose_tgt[:, 2, 3] += rng.choice([-0.8, 2.5, 12.0]) * scale
    return synth.Trajectory(rgb, depth, K, pose_src, pose_tgt, np.ones((T, H, W), np.float32))


@pytest.mark.parametrize("kind,seed", [(k, s) for k in ("plain", "wild", "zoom", "special", "tiny", "huge") for s in range(8)])
def test_forward_fuzz_vs_oracle(kind, seed):
    from syn3r_b200 import warp as w
    rng = np.random.default_rng(1000 * sum(map(ord, kind)) + seed)
    tr = _case(rng, kind)
    S, T, H, W = tr.shape
    ps, pt = synth.all_pairs(S, T)
    up = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()  # noqa: E731
    Kinv, par = w.forward_pair_params(tr.pose_src, tr.pose_tgt, tr.K, None, ps, pt)
    o = w.forward_warp_batched(up(w.pack_rgbd(tr.rgb, tr.depth)), up(Kinv), up(par), up(ps), want_winner=True,
                               want_u8=False, want_f32=True)
    torch.cuda.synchronize()
    for p, (s, t) in enumerate(zip(ps, pt)):
        with np.errstate(all="ignore"):
            r = orc.forward_warp_ext([tr.rgb[s]], [None], [tr.depth[s]], [tr.pose_src[s]], tr.pose_tgt[t], [tr.K], None)
        assert np.array_equal(o["mask"][p].cpu().numpy().astype(bool), r["mask"]), (kind, seed, p)
        win = np.where(r["winner"] >= 0, r["winner"] + s * H * W, -1)
        assert np.array_equal(o["winner"][p].cpu().numpy(), win), (kind, seed, p)
        d = np.abs(o["rgb_f32"][p].cpu().numpy() - r["out"])
        assert np.max(d, initial=0) <= 255e-5, (kind, seed, p, float(np.max(d)))


@pytest.mark.parametrize("kind,seed", [(k, s) for k in ("plain", "wild", "zoom", "special") for s in range(5)])
def test_consistency_fuzz_vs_oracle(kind, seed):
    """The trajectory consistency kernel (lean path) on the same random cases: mask_reproj bit-exact, soft <= 2e-6."""
    from syn3r_b200 import warp as w
    rng = np.random.default_rng(77 * sum(map(ord, kind)) + seed)
    tr = _case(rng, kind)
    S, T, H, W = tr.shape
    tr.depth_tgt = (tr.depth[rng.integers(0, S, T)] * np.float32(1.0 + 0.02 * rng.standard_normal())).astype(np.float32)
    ps, pt = synth.all_pairs(S, T)
    up = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()  # noqa: E731
    bw = float(rng.choice([20.0, 10.0, 3.0]))
    o = w.inverse_warp_batched(None, up(tr.depth), up(tr.depth_tgt), tr.K.astype(np.float32),
                               up(w.inverse_pair_params(tr.pose_src, tr.pose_tgt, ps, pt)), up(ps), up(pt), bandwidth=bw,
                               want=("mask_reproj", "soft_mask_reproj", "mask_warp"))
    torch.cuda.synchronize()
    t_ = torch.from_numpy
    img0 = np.zeros((3, H, W), np.float32)
    for p, (s, t) in enumerate(zip(ps, pt)):
        r = orc.inverse_warp(t_(img0), t_(tr.depth[s][None]), t_(tr.depth_tgt[t][None]), t_(tr.pose_src[s]),
                             t_(tr.pose_tgt[t]), t_(tr.K.astype(np.float32)), bandwidth=bw)
        assert np.array_equal(o["mask_warp"][p].cpu().numpy().astype(bool), r["mask_warp"].numpy()), (kind, seed, p)
        assert np.array_equal(o["mask_reproj"][p].cpu().numpy().astype(bool), r["mask_reproj"].numpy()), (kind, seed, p)
        a, b = o["soft_mask_reproj"][p].cpu().numpy(), r["soft_mask_reproj"].numpy()
        fin = np.isfinite(a) & np.isfinite(b)
        assert np.array_equal(np.isnan(a), np.isnan(b)), (kind, seed, p)
        assert np.max(np.abs(a[fin] - b[fin]), initial=0) <= 2e-6, (kind, seed, p)


# ---- §8f rank 4 kernels: random shapes and value ranges against their oracles ------------------------------------

@pytest.mark.parametrize("seed", range(12))
def test_fuzz_latent_guided_replace(seed):
    """Random batch / frame / channel / latent sizes, mask layouts, hole fractions from none to all, lambda rows
    outside the clamp range, duplicated and special values: bit-exact against oracle/latent_oracle.py."""
    from oracle import latent_oracle as lo
    from syn3r_b200 import latent
    rng = np.random.default_rng(9000 + seed)
    b, F, c = int(rng.integers(1, 3)), int(rng.integers(3, 9)), int(rng.integers(1, 5))
    h, w = int(rng.integers(1, 24)), int(rng.integers(1, 40))
    bm = b if rng.random() < 0.5 else 1
    cm = int(rng.choice([1, 2, 3, 4]))
    pred = rng.normal(0, rng.choice([1e-3, 1.0, 50.0]), (b, F, c, h, w)).astype(np.float32)
    tcl = rng.normal(0, 1, (2, F, c, h, w)).astype(np.float32)
    q = float(rng.choice([0.0, 1 / 8, 1 / 64]))
    noise = rng.normal(0, 0.3, (F, c, h, w))
    tcl[1] = (pred[0] + (np.round(noise / q) * q if q else noise)).astype(np.float32)
    mask = (rng.random((bm, F - 1, cm, h, w)) < rng.choice([0.0, 0.2, 0.7, 1.0])).astype(np.float32)
    if seed % 4 == 1:
        pred.ravel()[rng.integers(0, pred.size, 3)] = [np.nan, np.inf, -np.inf]
    lam = (rng.random((2, F)) * 1.5).astype(np.float32)
    r = lo.guided_replace(pred, tcl, mask, lam, 1)
    up = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()  # noqa: E731
    out, ks, cv = latent.guided_replace_(up(pred.copy()), up(tcl), up(mask), up(lam), 1, return_cutoffs=True)
    torch.cuda.synchronize()
    assert np.array_equal(ks.cpu().numpy(), r["cutoff_index"])
    assert np.array_equal(cv.cpu().numpy(), r["cutoff_value"], equal_nan=True)
    assert np.array_equal(out.cpu().numpy(), r["out"], equal_nan=True)


@pytest.mark.parametrize("seed", range(12))
def test_fuzz_flow_fb_consistency(seed):
    """Random sizes (odd widths, tiny frames), flow magnitudes from sub-pixel to larger than the frame, thresholds:
    norms to float32 rounding, flags outside a narrow band around the threshold."""
    from oracle import flow_oracle as fo
    from syn3r_b200 import flow
    rng = np.random.default_rng(9100 + seed)
    B, H, W = int(rng.integers(1, 4)), int(rng.integers(2, 70)), int(rng.integers(2, 90))
    amp = float(rng.choice([0.2, 2.0, 15.0, 200.0]))
    fwd = rng.normal(0, amp, (B, 2, H, W)).astype(np.float32)
    bwd = (-fwd + rng.normal(0, 0.3, (B, 2, H, W))).astype(np.float32)
    if seed % 3 == 0:
        fwd = np.round(fwd).astype(np.float32)            # taps exactly on pixel centres and on the frame border
    alpha, beta = float(rng.choice([0.01, 0.05, 0.0])), float(rng.choice([0.5, 0.1, 1.0]))
    r = fo.forward_backward_consistency_check(fwd, bwd, alpha, beta)
    o = flow.forward_backward_consistency_check(torch.from_numpy(fwd).cuda(), torch.from_numpy(bwd).cuda(), alpha, beta,
                                                return_diff=True)
    torch.cuda.synchronize()
    tol = 2e-6 * max(1.0, amp * 4)
    for occ, diff, d in ((o[0], o[2], "fwd"), (o[1], o[3], "bwd")):
        assert np.max(np.abs(diff.cpu().numpy() - r["diff_" + d])) <= tol
        near = np.abs(r["diff_" + d] - r["threshold"]) <= tol
        assert ((occ.cpu().numpy() == r[d + "_occ"]) | near).all()
