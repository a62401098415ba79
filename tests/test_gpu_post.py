"""GPU parity of the warped-frame post-processing through the C ABI: against the golden
fixtures (reference statements re-executed with cv2 / PIL) and against the NumPy oracle at full size."""
import glob
import os

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

from oracle import post_oracle as po  # noqa: E402  (checker only)

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
FW = sorted(os.path.basename(p)[:-4] for p in glob.glob(os.path.join(GOLDEN, "post_fw_*.npz")))
BW = sorted(os.path.basename(p)[:-4] for p in glob.glob(os.path.join(GOLDEN, "post_bw_*.npz")))


@pytest.mark.parametrize("name", FW)
def test_post_forward_golden(golden, name):
    from syn3r_b200 import conditioning as c
    g = golden(name)
    o = c.post_forward(g["in_warped"], g["in_mask2"])
    for k in ("mask_png", "cond", "latent_mask"):
        assert o[k].dtype == g["out_" + k].dtype and np.array_equal(o[k], g["out_" + k]), k


@pytest.mark.parametrize("name", BW)
def test_post_backward_golden(golden, name):
    from syn3r_b200 import conditioning as c
    g = golden(name)
    o = c.post_backward(g["in_warped"], g["in_mask2"], g["in_soft"])
    for k in ("mask_png", "cond", "cond_ori", "latent_mask", "soft_pooled"):
        assert o[k].dtype == g["out_" + k].dtype and np.array_equal(o[k], g["out_" + k]), k


def test_post_batched_full_size_vs_oracle():
    """5 frames at 576x1024 (the reference's own 72x128 latent grid), ragged hole structure, every output."""
    from syn3r_b200 import conditioning as c
    rng = np.random.default_rng(77)
    P, H, W = 5, 576, 1024
    valid = rng.random((P, H, W)) > 0.003
    valid[0, 100:160, 300:420] = False
    valid[1, :, :3] = False
    valid[2] = True
    valid[3] = False
    img = rng.integers(0, 256, (P, H, W, 3), dtype=np.uint8)
    soft = np.exp(-(rng.random((P, H, W), dtype=np.float32) * 3) ** 3).astype(np.float32)
    up = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()  # noqa: E731
    o = c.condition_post_batched(up(valid), img_u8=up(img), soft=up(soft), want=c.POST_KEYS)
    torch.cuda.synchronize()
    for p in range(P):
        r = po.post_forward(img[p], valid[p])
        assert np.array_equal(o["mask_png"][p].cpu().numpy(), r["mask_png"][..., 0]), p
        assert np.array_equal(o["cond_u8"][p].cpu().numpy(), r["cond_u8"]), p
        assert np.array_equal(o["cond_f32"][p].cpu().numpy(), r["cond"]), p
        assert np.array_equal(o["cond_ori"][p].cpu().numpy(), img[p].astype(np.float32) / np.float32(255)), p
        assert np.array_equal(o["latent"][p].cpu().numpy().astype(np.float64), r["latent_mask"]), p
        assert np.array_equal(o["soft_pooled"][p].cpu().numpy(), po.pool_soft(soft[p])), p
    # properties: dilation only grows holes; an all-covered frame passes through; an all-hole frame is blank
    assert bool((o["mask_png"].cpu().numpy() >= np.where(valid, 0, 255)).all())
    assert np.array_equal(o["cond_u8"][2].cpu().numpy(), img[2]) and int(o["latent"][2].sum()) == 0
    assert int(o["cond_u8"][3].max()) == 0 and int(o["latent"][3].min()) == 1


def test_post_rejects_sizes_that_do_not_divide_by_8():
    from syn3r_b200 import conditioning as c
    with pytest.raises(AssertionError):
        c.condition_post_batched(torch.ones((1, 20, 24), dtype=torch.uint8, device="cuda"))


def test_post_float_frame_truncates_like_np_uint8():
    """backward branch: np.uint8(float32 frame * (1 - m)) truncates toward zero (diffusionGS.py:1468)."""
    from syn3r_b200 import conditioning as c
    rng = np.random.default_rng(3)
    H, W = 32, 48
    f = (rng.random((H, W, 3), dtype=np.float32) * np.float32(255.9)).astype(np.float32)
    f[0, 0] = [0.0, 254.99998, 255.5]
    m = np.ones((H, W), bool)
    m[10:12, 20:23] = False
    s = rng.random((H, W), dtype=np.float32)
    o = c.post_backward(f, m, s)
    r = po.post_backward(f, m, s)
    for k in ("mask_png", "cond", "cond_ori", "latent_mask", "soft_pooled"):
        assert np.array_equal(o[k], r[k]), k


def test_trajectory_run_post_matches_oracle_on_the_plan_outputs():
    """TrajectoryWarp.run() -> run_post(): the reference's caller block applied to every pair's warped frame, hole
    mask, reprojection mask and soft mask, checked against the oracle fed with the plan's own outputs."""
    from syn3r_b200 import synth, trajectory, warp as w
    S, T, H, W = 2, 3, 64, 96
    tr = synth.make_trajectory(S, T, H, W, seed=91)
    tw = trajectory.from_synth(tr, want_u8=True, want_f32=False)
    tw.upload(torch.from_numpy(w.pack_rgbd(tr.rgb, tr.depth)), torch.from_numpy(tr.depth_tgt))
    out = tw.run()
    post = tw.run_post()
    torch.cuda.synchronize()
    for p in range(tw.P):
        r = po.post_forward(out.rgb_u8[p].cpu().numpy(), out.mask[p].cpu().numpy().astype(bool))
        assert np.array_equal(post["fw"]["mask_png"][p].cpu().numpy(), r["mask_png"][..., 0])
        assert np.array_equal(post["fw"]["cond_u8"][p].cpu().numpy(), r["cond_u8"])
        assert np.array_equal(post["fw"]["latent"][p].cpu().numpy().astype(np.float64), r["latent_mask"])
        dil = po.dilate5(~out.mask_reproj[p].cpu().numpy().astype(bool))
        assert np.array_equal(post["bw"]["latent"][p].cpu().numpy().astype(np.float64), po.latent_mask(dil))
        assert np.array_equal(post["bw"]["soft_pooled"][p].cpu().numpy(), po.pool_soft(out.soft_mask_reproj[p].cpu().numpy()))


FUSION = sorted(os.path.basename(p)[:-4] for p in glob.glob(os.path.join(GOLDEN, "fusion_*.npz")))


def _check_fusion(o, ref_unc, ref_masks, ref_cond, rows=slice(None)):
    unc = o["geo_inten_uncertainty"].cpu().numpy()
    assert np.max(np.abs(unc[:, rows] - ref_unc)) <= 1e-6
    assert np.max(np.abs(o["masks"].cpu().numpy() - ref_masks)) <= 1e-6
    band = (np.abs(unc - 0.5) <= 2e-6)[..., None]               # the np.where(unc > 0.5) choice is free inside the band
    assert ((o["cond_image"].cpu().numpy() == ref_cond) | band).all()


@pytest.mark.parametrize("name", FUSION)
def test_fuse_uncertainty_golden(golden, name):
    """SURVEY §8f rank 2 against the reference's statements re-executed with NumPy (float32 exp / pow: 1e-6)."""
    from syn3r_b200 import conditioning as c
    from conftest import expand_fusion
    cond, gs, soft, ref_masks, ref_cond, ref_unc, rows = expand_fusion(golden(name))
    up = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()  # noqa: E731
    o = c.fuse_uncertainty_batched(up(cond), up(gs), up(soft), want=("cond_image", "masks", "geo_inten_uncertainty"))
    torch.cuda.synchronize()
    assert o["masks"].dtype == torch.float32 and o["geo_inten_uncertainty"].dtype == torch.float64
    _check_fusion(o, ref_unc, ref_masks, ref_cond, rows)


def test_fuse_uncertainty_full_size_vs_oracle():
    from syn3r_b200 import conditioning as c
    rng = np.random.default_rng(8)
    N, H, W = 4, 576, 1024
    gs = rng.random((N, H, W, 3), dtype=np.float32)
    cond = np.clip(gs + np.float32(0.3) * rng.standard_normal((N, H, W, 3)).astype(np.float32) * (rng.random((N, H, W, 1)) > 0.6),
                   -0.1, 1.3).astype(np.float32)
    cond[:, 100:140, 200:330] = 0
    soft_ori = rng.random((N, H, W), dtype=np.float32)
    up = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()  # noqa: E731
    o = c.fuse_uncertainty_batched(up(cond), up(gs), up(soft_ori), want=("cond_image", "masks", "geo_inten_uncertainty"))
    torch.cuda.synchronize()
    r = po.fuse_uncertainty(cond, gs, soft_ori)
    _check_fusion(o, r["geo_inten_uncertainty"][..., 0], r["masks"], r["cond_image"])
    ci = o["cond_image"].cpu().numpy()
    assert ci.min() >= 0.0 and ci.max() <= 1.0                  # :876
    assert np.all(o["geo_inten_uncertainty"].cpu().numpy()[:, 100:140, 200:330] == 1.0)   # blanked pixels: confidence 0


# ---- nearby-frame consistency (diffusionGS.py:1300-1361) ------------------------------------------

@pytest.mark.parametrize("name", ["nearby_0", "nearby_1"])
def test_nearby_consistency_golden(golden, name):
    """Drop-in call (two launches: batched inverse warp + fusion) against the reference loop's outputs."""
    from syn3r_b200 import conditioning as c
    g = golden(name)
    unc, iunc = c.consistency_check_from_nearby_images_bw(g["K"], list(g["poses"]), list(g["images"]), list(g["depths"]))
    assert unc.dtype == torch.float32 and tuple(unc.shape) == g["out_uncertainty"].shape
    assert np.max(np.abs(unc.cpu().numpy() - g["out_uncertainty"])) <= 2e-6      # soft mask bar (test_gpu_parity)
    assert np.max(np.abs(iunc.cpu().numpy() - g["out_intensity"])) <= 2e-6


def test_nearby_consistency_single_frame_raises(golden):
    from syn3r_b200 import conditioning as c
    g = golden("nearby_2")
    with pytest.raises(RuntimeError):
        c.consistency_check_from_nearby_images_bw(g["K"], list(g["poses"]), list(g["images"]), list(g["depths"]))


@pytest.mark.parametrize("N,H,W,radius", [(5, 40, 64, 1), (4, 33, 47, 2), (2, 576, 1024, 1)])
def test_nearby_uncertainty_kernel_vs_oracle(N, H, W, radius):
    """The fusion kernel alone on the oracle's own inputs: 1..2*radius neighbours per frame, odd sizes (scalar path),
    NaN soft masks (exp of a NaN error) and differences across the whole range of exp(-(d / 0.1)^3)."""
    from syn3r_b200 import conditioning as c
    rng = np.random.default_rng(N * 1000 + H)
    ps, pt, first = c.nearby_pairs(N, radius)
    P = len(ps)
    images = (rng.random((N, 3, H, W), dtype=np.float32) * 255).astype(np.float32)
    warped = (images[pt] + (rng.random((P, 3, H, W), dtype=np.float32) - 0.5) * rng.choice(
        np.float32([0.0, 0.05, 0.2, 1.0, 50.0]), size=(P, 1, H, W))).astype(np.float32)
    soft = rng.random((P, H, W), dtype=np.float32)
    soft[0, ::7, ::5] = np.nan
    r = po.nearby_uncertainty(soft, warped, images, first)
    up = lambda a: torch.from_numpy(a).cuda()  # noqa: E731
    o = c.nearby_uncertainty_batched(up(soft), up(warped), up(images), up(first))
    torch.cuda.synchronize()
    assert np.array_equal(o["uncertainty"].cpu().numpy(), r["uncertainty"], equal_nan=True)     # adds and one multiply
    assert np.max(np.abs(o["intensity_uncertainty"].cpu().numpy() - r["intensity_uncertainty"])) <= 1e-6


def test_nearby_uncertainty_empty_is_a_no_op():
    from syn3r_b200 import _native as nat
    assert nat.lib().syn3r_nearby_uncertainty_batched(None, None, None, None, 0, 8, 8, 10.0, None, None, None) == 0


@pytest.mark.parametrize("n", [16, 17, 31, 4096, 4096 + 5, 576 * 1024, 300 * 400 * 3 + 7])
def test_pack_mask_bits_is_np_packbits(n):
    """syn3r_pack_mask_bits == np.packbits(mask != 0) (MSB first, zero-padded tail), aligned and unaligned starts."""
    from syn3r_b200 import conditioning as c
    rng = np.random.default_rng(n)
    m = (rng.random(n + 3) < 0.4).astype(np.uint8) * rng.integers(1, 256, n + 3).astype(np.uint8)   # any non-zero byte is True
    for off in (0, 3):                                   # off = 3: a 16-byte-misaligned start takes the byte-wise kernel
        t = torch.from_numpy(m).cuda()[off:off + n].contiguous() if off == 0 else torch.from_numpy(m).cuda()[off:off + n]
        bits = c.pack_mask_bits(t if t.is_contiguous() else t.contiguous())
        ref = np.packbits(m[off:off + n] != 0)
        assert np.array_equal(bits.cpu().numpy(), ref)
        assert np.array_equal(c.unpack_mask_bits(bits.cpu(), (n,)), m[off:off + n] != 0)


def test_trajectory_packed_masks_and_step_host_match_run():
    """pack_masks=True: the bit planes are np.packbits of the byte masks per slot; step_host (upload stream, slot ranges,
    D2H on a side stream) delivers exactly what run() leaves on the device."""
    from syn3r_b200 import synth, trajectory, warp as w
    S, T, H, W = 2, 5, 64, 96
    tr = synth.make_trajectory(S, T, H, W, seed=11)
    ps, pt = synth.target_major(*synth.all_pairs(S, T))
    keys = ("rgb_u8", "mask_bits", "mask_reproj_bits", "soft_mask_reproj")
    tw = trajectory.from_synth(tr, pairs=(ps, pt), want_f32=False, want_u8=True, want_winner=True, pack_masks=True,
                               pinned_host=True, host_outputs=keys)
    rgbd_h = torch.from_numpy(w.pack_rgbd(tr.rgb, tr.depth)).pin_memory()
    tgt_h = torch.from_numpy(tr.depth_tgt).pin_memory()
    tw.upload(rgbd_h, tgt_h)
    o = tw.run()
    torch.cuda.synchronize()
    mb = tw.out_fwd["mask_bits"].cpu().numpy()
    rb = tw.out_inv["mask_reproj_bits"].cpu().numpy()
    for p in range(tw.P):
        assert np.array_equal(mb[p], np.packbits(o.mask[p].cpu().numpy().reshape(-1) != 0))
        assert np.array_equal(rb[p], np.packbits(o.mask_reproj[p].cpu().numpy().reshape(-1) != 0))
    ref = {k: v.clone() for k, v in {**tw.out_fwd, **tw.out_inv}.items()}
    for d in (tw.out_fwd, tw.out_inv):
        for v in d.values():
            v.zero_()
    tw.rgbd.zero_()
    tw.tgt_depth.zero_()
    pinned = tw.step_host(rgbd_h, tgt_h, chunks=3)
    torch.cuda.synchronize()
    for k in keys:
        assert torch.equal(pinned["out_" + k], ref[k].cpu()), k
