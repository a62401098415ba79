"""Full-size runs of the BASELINE.json configurations on a B200 (`-m gpu`): oracle spot checks on a few pairs plus
size-independent properties (bitwise reproducibility, fused mask == OR of pair masks, winner
consistency, geometric sanity of the winner) where the oracle would take minutes."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

from oracle import cpu_oracle as orc  # noqa: E402  (checker only)
from syn3r_b200 import synth, trajectory, warp as w  # noqa: E402

COLOUR_TOL_255 = 1e-5 * 255.0


def _plan(cfg, T=None, **kw):
    S, T0, H, W = synth.CONFIGS[cfg]
    T = T or T0
    tr = synth.make_trajectory(S, T, H, W, seed=3)
    tw = trajectory.from_synth(tr, **kw)
    tw.upload(torch.from_numpy(w.pack_rgbd(tr.rgb, tr.depth)), torch.from_numpy(tr.depth_tgt))
    return tr, tw


def _oracle_fwd(tr, s, t):
    r = orc.forward_warp_ext([tr.rgb[s]], [None], [tr.depth[s]], [tr.pose_src[s]], tr.pose_tgt[t], [tr.K], None)
    H, W = tr.depth.shape[1:]
    r["winner"] = np.where(r["winner"] >= 0, r["winner"] + s * H * W, -1).astype(np.int32)
    return r


def _spot_check(tr, tw, out, pairs):
    t_ = torch.from_numpy
    for p in pairs:
        s, t = int(tw.pair_src_h[p]), int(tw.pair_tgt_h[p])
        r = _oracle_fwd(tr, s, t)
        assert np.array_equal(out.mask[p].cpu().numpy().astype(bool), r["mask"]), p
        assert np.array_equal(out.winner[p].cpu().numpy(), r["winner"]), p
        assert np.max(np.abs(out.rgb_f32[p].cpu().numpy() - r["out"])) <= COLOUR_TOL_255, p
        iw = orc.inverse_warp(t_(np.ascontiguousarray(np.moveaxis(tr.rgb[s], -1, 0))), t_(tr.depth[s][None]),
                              t_(tr.depth_tgt[t][None]), t_(tr.pose_src[s]), t_(tr.pose_tgt[t]),
                              t_(tr.K.astype(np.float32)))
        assert np.array_equal(out.mask_reproj[p].cpu().numpy().astype(bool), iw["mask_reproj"].numpy()), p
        a, b = out.soft_mask_reproj[p].cpu().numpy(), iw["soft_mask_reproj"].numpy()
        fin = np.isfinite(a) & np.isfinite(b)
        assert np.array_equal(np.isnan(a), np.isnan(b)) and np.max(np.abs(a[fin] - b[fin]), initial=0) <= 2e-6, p


def test_config_B_full_size():
    """3 sources -> 25 targets at 576x1024 (the headline workload): oracle on 3 of the 75 pairs + properties."""
    tr, tw = _plan("B_svd", want_u8=True)
    out = tw.run()
    torch.cuda.synchronize()
    assert int(out.status.max().item()) == 0
    # pair 0 is (source 0 -> target 0) with target 0 at source 0's pose up to rounding: nearly every pixel lands
    # within 1e-12 of an integer, i.e. the whole pair runs through the bit-faithful fallback chain
    _spot_check(tr, tw, out, [0, 36, 74])
    # winner >= 0 exactly where the hole mask is set
    assert torch.equal(out.winner >= 0, out.mask.bool())
    # integer accumulation: a second run is bitwise identical, colours included
    snap = {k: v.clone() for k, v in {**tw.out_fwd, **tw.out_inv}.items()}
    tw.run()
    torch.cuda.synchronize()
    for k, v in {**tw.out_fwd, **tw.out_inv}.items():
        assert torch.equal(v, snap[k]) or (v.dtype.is_floating_point and torch.equal(torch.nan_to_num(v), torch.nan_to_num(snap[k]))), k


def test_winner_lands_next_to_its_cell():
    """The winning source pixel of every covered cell projects into the 2x2 neighbourhood of that cell (fp64 NumPy)."""
    tr, tw = _plan("D_dl3dv", T=3, consistency=False)
    out = tw.run()
    torch.cuda.synchronize()
    S, T, H, W = tr.shape
    Kinv, par = w.forward_pair_params(tr.pose_src, tr.pose_tgt, tr.K, None, tw.pair_src_h, tw.pair_tgt_h)
    for p in (1, 4):
        s = int(tw.pair_src_h[p])
        win = out.winner[p].cpu().numpy()
        ys, xs = np.nonzero(win >= 0)
        src = win[ys, xs] - s * H * W
        assert src.min() >= 0 and src.max() < H * W
        sy, sx = np.divmod(src, W)
        X = (np.stack([sx, sy, np.ones_like(sx)], -1).astype(np.float64) @ Kinv[s].reshape(3, 3).T) * \
            tr.depth[s][sy, sx].astype(np.float64)[:, None]
        M, K2 = par[p, :12].reshape(3, 4), par[p, 12:21].reshape(3, 3)
        q = (X @ M[:, :3].T + M[:, 3]) @ K2.T
        cx, cy = q[:, 0] / q[:, 2], q[:, 1] / q[:, 2]
        assert np.all(np.abs(cx - xs) < 1.0 + 1e-9) and np.all(np.abs(cy - ys) < 1.0 + 1e-9)


@pytest.mark.parametrize("cfg,pairs", [("C_dtu", [0, 100, 191]), ("D_dl3dv", [3, 30, 49])])
def test_configs_C_and_D_spot_checks(cfg, pairs):
    tr, tw = _plan(cfg)
    out = tw.run()
    torch.cuda.synchronize()
    assert int(out.status.max().item()) == 0
    _spot_check(tr, tw, out, pairs)


def test_config_E_shape_fused_heavy_collision():
    """4K sources (more super-blocks than one cull list holds), FUSED mode: 4 sources into each of 2 target
    frames.  fused mask == OR of the pair masks, bitwise reproducible, winner consistent."""
    S, T, H, W = 4, 2, 2160, 3840
    tr = synth.make_trajectory(S, T, H, W, seed=5)
    ps, pt = synth.all_pairs(S, T)
    dev = torch.device("cuda")
    Kinv, par = w.forward_pair_params(tr.pose_src, tr.pose_tgt, tr.K, None, ps, pt)
    up = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)  # noqa: E731
    rgbd = up(w.pack_rgbd(tr.rgb, tr.depth))
    args = (rgbd, up(Kinv), up(par), up(ps))
    fused = w.forward_warp_batched(*args, pair_slot=up(pt), n_slots=T, want_winner=True, want_u8=True, want_f32=False)
    torch.cuda.synchronize()
    assert int(fused["status"].max().item()) == 0
    snap = {k: v.clone() for k, v in fused.items()}
    union = torch.zeros((T, H, W), dtype=torch.bool, device=dev)
    for s in range(S):                                   # one source at a time keeps the pair outputs small
        sel = np.nonzero(ps == s)[0]
        pair = w.forward_warp_batched(rgbd, args[1], up(par[sel]), up(ps[sel]), want_winner=False, want_u8=False,
                                      want_f32=False)
        union |= pair["mask"].bool()
    assert torch.equal(fused["mask"].bool(), union)
    assert torch.equal(fused["winner"] >= 0, fused["mask"].bool())
    again = w.forward_warp_batched(*args, pair_slot=up(pt), n_slots=T, want_winner=True, want_u8=True, want_f32=False)
    torch.cuda.synchronize()
    for k in ("mask", "winner", "rgb_u8"):
        assert torch.equal(again[k], snap[k]), k
    # the nearest surface wins: the winner of a cell covered by the foreground occluder of some source is a
    # foreground pixel (depth 0.9) of that source, never a background one
    win = fused["winner"][0].cpu().numpy()
    depth_flat = tr.depth.reshape(-1)
    wd = depth_flat[win[win >= 0]]
    assert wd.min() == np.float32(0.9)


@pytest.mark.parametrize("letter", ["B", "C", "D"])
def test_gpu_matches_live_reference_at_benchmark_size(golden, letter):
    """The CUDA path against outputs of the LIVE reference (forward_warp + inverse_warp of /root/reference, stored by
    tests/golden/make_golden_fullsize.py) on two pairs of the benchmark trajectory of configs B, C, D.  Bars: hole
    mask bit-exact; uint8 frame identical except single levels at rounding ties (<= 2e-4 of values);
    mask_reproj may differ from torch's result only inside |err - bw| <= 1e-3 bw; soft mask <= 2e-6."""
    from _fullsize_util import bits, load
    tr, g, pairs, rows = load(golden, letter)
    S, T, H, W = tr.shape
    ps = np.array([p[0] for p in pairs], np.int32)
    pt = np.array([p[1] for p in pairs], np.int32)
    tw = trajectory.from_synth(tr, pairs=(ps, pt), want_u8=True, want_f32=False, want_winner=False)
    tw.upload(torch.from_numpy(w.pack_rgbd(tr.rgb, tr.depth)), torch.from_numpy(tr.depth_tgt))
    out = tw.run()
    torch.cuda.synchronize()
    assert int(out.status.max().item()) == 0
    for k in range(len(pairs)):
        assert np.array_equal(out.mask[k].cpu().numpy().astype(bool), bits(g, f"p{k}_mask_bits", H, W)), (letter, k)
        u8 = out.rgb_u8[k].cpu().numpy()
        rs = u8.reshape(H, -1).sum(1, dtype=np.int64)
        assert np.abs(rs - g[f"p{k}_warped_rowsum"]).max() <= 3, (letter, k)
        if k == 0:
            d = np.abs(u8.astype(np.int16) - g["p0_warped_u8"].astype(np.int16))
            assert d.max() <= 1 and np.count_nonzero(d) <= 2e-4 * d.size, (letter, int(d.max()), np.count_nonzero(d))
        xor = out.mask_reproj[k].cpu().numpy().astype(bool) ^ bits(g, f"p{k}_mask_reproj_bits", H, W)
        assert not (xor & ~bits(g, f"p{k}_band_bits", H, W)).any() and np.count_nonzero(xor) <= 8, (letter, k, np.count_nonzero(xor))
        a, b = out.soft_mask_reproj[k].cpu().numpy()[rows], g[f"p{k}_soft_rows"]
        fin = np.isfinite(a) & np.isfinite(b)
        assert np.array_equal(np.isnan(a), np.isnan(b)) and np.max(np.abs(a[fin] - b[fin]), initial=0) <= 2e-6, (letter, k)
