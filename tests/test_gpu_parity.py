"""GPU parity tests proper: the CUDA path (through the C ABI) against the golden outputs of the live
reference (tests/golden) and against the CPU oracle on seeded inputs.  Run with `-m gpu` on a B200."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

from oracle import cpu_oracle as orc  # noqa: E402  (checker only)
from syn3r_b200 import synth  # noqa: E402

FW_CASES = ["fw_generic_0", "fw_generic_1", "fw_generic_2", "fw_mid", "fw_u8_f64pose", "fw_mask_K2",
            "fw_identity", "fw_integer_row", "fw_integer_shift", "fw_behind", "fw_all_out", "fw_all_behind",
            "fw_nan_depth"]

COLOUR_TOL_01 = 1e-5          # north_star: max-abs on the [0,1] scale
COLOUR_TOL_255 = COLOUR_TOL_01 * 255.0


@pytest.fixture(scope="module")
def fw():
    import syn3r_b200.solver_utils.forward_warp as m
    return m


@pytest.fixture(scope="module")
def cc():
    import syn3r_b200.solver_utils.consistency as m
    return m


def _u8_close(a, b, max_frac=2e-4):
    """uint8 colours: identical up to one level at rounding ties (fp32 accumulation)."""
    d = np.abs(a.astype(np.int16) - b.astype(np.int16))
    assert d.max(initial=0) <= 1, int(d.max())
    assert (d > 0).mean() <= max_frac, float((d > 0).mean())


@pytest.mark.parametrize("name", FW_CASES)
def test_forward_warp_golden(golden, fw, name):
    g = golden(name)
    warped, mask2, flow = fw.forward_warp(g["frame"], g.get("mask1"), g["depth"], g["T1"], g["T2"], g["K"],
                                          g.get("K2"))
    assert warped.dtype == np.uint8 and mask2.dtype == np.bool_ and flow.dtype == np.float64
    assert warped.shape == g["warped"].shape and mask2.shape == g["mask2"].shape
    assert np.array_equal(flow, g["flow12"], equal_nan=True)      # fp64 projection bit-exact
    assert np.array_equal(mask2, g["mask2"])                      # hole mask bit-exact
    _u8_close(warped, g["warped"])


def test_compute_transformed_points_golden(golden, fw):
    for name in ("ctp_dense", "ctp_f64"):
        g = golden(name)
        q, X = fw.compute_transformed_points(g["depth"], g["T1"], g["T2"], g["K"], g.get("K2"))
        assert q.shape == g["q"].shape and X.shape == g["X"].shape and q.dtype == np.float64
        assert np.array_equal(X, g["X"])
        assert np.array_equal(q, g["q"])


def test_bilinear_splatting_golden(golden, fw):
    g = golden("bs_generic")
    out, mask = fw.bilinear_splatting(g["frame"], g["mask1"], g["depth"], g["flow12"], g["flow_mask"], False)
    assert out.dtype == np.float64 and mask.dtype == np.bool_
    assert np.array_equal(mask, g["mask"])
    scale = max(1.0, float(np.abs(g["out"]).max()))
    assert np.max(np.abs(out - g["out"])) <= 1e-5 * scale


def test_create_grid_golden(golden, fw):
    g = golden("grid")["grid"]
    o = fw.create_grid(5, 7)
    assert o.dtype == g.dtype and np.array_equal(o, g)


def test_consistency_golden(golden, cc):
    g = golden("cc_generic")
    t = torch.from_numpy
    pts = cc.get_points_from_depth(t(g["depth1"]), t(g["K1"])).cpu().numpy()
    np.testing.assert_array_equal(pts, g["pts"])
    pts2 = cc.transform_points(t(g["pts"]), t(g["pose1"]), t(g["pose2"])).cpu().numpy()
    np.testing.assert_array_equal(pts2, g["pts2"])
    err = cc.consistency_check_with_depth(t(g["depth1"]), t(g["pose1"]), t(g["K1"]), t(g["depth2"]), t(g["pose2"]),
                                          t(g["K2"]))
    assert err.is_cuda and err.dtype == torch.float32
    err = err.cpu().numpy()
    # bit-exact against the oracle (same op order) ...
    ref = orc.consistency_check_with_depth(t(g["depth1"]), t(g["pose1"]), t(g["K1"]), t(g["depth2"]), t(g["pose2"]),
                                           t(g["K2"])).numpy()
    assert np.array_equal(err, ref, equal_nan=True)
    # ... and within the documented band of the live torch result
    assert np.mean(err == g["err"]) > 0.99
    np.testing.assert_allclose(err, g["err"], rtol=2e-4, atol=2e-4)


def _band_ok(a, b, val, thr, rel=1e-3):
    diff = a != b
    return bool(np.all(np.abs(val[diff] - thr) <= rel * abs(thr))), int(diff.sum())


@pytest.mark.parametrize("name", ["iw_generic_0", "iw_generic_1", "iw_generic_2"])
def test_inverse_warp_golden(golden, fw, name):
    g = golden(name)
    t = torch.from_numpy
    bg = t(g["bg_mask"]) if "bg_mask" in g else None
    bw = float(g["bandwidth"])
    args = (t(g["img"]), t(g["depth"]), t(g["depth_pseudo"]), t(g["pose1"]), t(g["pose2"]), t(g["K"]))
    o = fw.inverse_warp(*args, bg_mask=bg, bandwidth=bw)
    assert set(o.keys()) == {"warped_img", "warped_depth", "mask_warp", "mask_depth", "mask", "warped_masked_img",
                             "mask_inv", "mask_depth_strict", "warped_bg_mask", "mask_reproj", "soft_mask_reproj"}
    ref = orc.inverse_warp(*args, bg_mask=bg, bandwidth=bw, return_err=True)
    for k, v in o.items():
        if v is None:
            assert ref[k] is None and g["out_" + k].size == 0
            continue
        assert v.is_cuda, k
        a = v.cpu().numpy()
        assert a.shape == g["out_" + k].shape and a.dtype == g["out_" + k].dtype, k
        if k == "soft_mask_reproj":
            r = ref[k].numpy()
            assert np.array_equal(np.isnan(a), np.isnan(r))
            fin = np.isfinite(a) & np.isfinite(r)
            assert np.max(np.abs(a[fin] - r[fin]), initial=0) <= 2e-6
            both = fin & np.isfinite(g["out_" + k])
            assert np.max(np.abs(a[both] - g["out_" + k][both]), initial=0) <= 2e-6
        else:
            assert np.array_equal(a, ref[k].numpy(), equal_nan=True), k        # vs oracle: bit-exact
    for k in ("warped_img", "warped_depth", "mask_warp", "warped_masked_img", "mask_depth", "mask_depth_strict",
              "mask", "mask_inv"):
        assert np.array_equal(o[k].cpu().numpy(), g["out_" + k], equal_nan=True), k   # vs live reference
    ok, n = _band_ok(o["mask_reproj"].cpu().numpy(), g["out_mask_reproj"], ref["reproj_err"].numpy(), bw)
    assert ok and n <= 2, n


# ---------------------------------------------------------------------------------------------
# seeded synthetic trajectories against the oracle (PAIR and FUSED, winner index)
# ---------------------------------------------------------------------------------------------

def _run_pairs(tr, ps, pt, fused=False, **kw):
    from syn3r_b200 import warp as w
    dev = torch.device("cuda")
    S, T, H, W = tr.shape
    Kinv, par = w.forward_pair_params(tr.pose_src, tr.pose_tgt, tr.K, None, ps, pt)
    rgbd = torch.from_numpy(w.pack_rgbd(tr.rgb, tr.depth)).to(dev)
    up = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)  # noqa: E731
    slot = up(np.asarray(pt, np.int32)) if fused else None
    o = w.forward_warp_batched(rgbd, up(Kinv), up(par), up(np.asarray(ps, np.int32)), pair_slot=slot,
                               n_slots=T if fused else None, want_winner=True, want_u8=True, want_f32=True, **kw)
    torch.cuda.synchronize()
    return {k: v.cpu().numpy() for k, v in o.items()}


def _oracle_pair(tr, s_list, t):
    return orc.forward_warp_ext([tr.rgb[s] for s in s_list], [None] * len(s_list), [tr.depth[s] for s in s_list],
                                [tr.pose_src[s] for s in s_list], tr.pose_tgt[t], [tr.K] * len(s_list), None)


def _check_slot(o, i, r, hw_offsets=None):
    assert np.array_equal(o["mask"][i].astype(bool), r["mask"])
    assert np.max(np.abs(o["rgb_f32"][i] - r["out"])) <= COLOUR_TOL_255
    _u8_close(o["rgb_u8"][i], r["out_u8"])
    assert np.array_equal(o["winner"][i], r["winner"])
    assert o["status"][i] == 0


def test_shared_reciprocal_division_is_ieee():
    """ddiv2_rn (forward_tile.cuh) == div.rn.f64 bit for bit on 2 x 128M random operand triples per mode."""
    from syn3r_b200 import _native as nat
    cnt = torch.zeros(1, dtype=torch.int64, device="cuda")
    for mode in (0, 1):
        nat.check(nat.lib().syn3r_selftest_ddiv2(1 << 27, 12345 + mode, mode, cnt.data_ptr(),
                                                 torch.cuda.current_stream().cuda_stream))
    torch.cuda.synchronize()
    assert int(cnt.item()) == 0


@pytest.mark.parametrize("size", [1024, 576, 960, 540, 400, 300, 504, 378, 3840, 2160, 128, 96, 131, 97])
def test_const_division_exhaustive(size):
    """x / ((size - 1) / 2) of consistency.py:71-73 through the rounded reciprocal (div_const_rn, inverse_kernels.cuh)
    has the bits of the IEEE division for EVERY finite float x with a normal quotient (all 2^32 patterns tried)."""
    import ctypes
    from syn3r_b200 import _native as nat
    c = np.float32((size - 1) / 2.0)
    cnt = torch.zeros(2, dtype=torch.int64, device="cuda")      # [mismatches, bit pattern of one failing x]
    rc = ctypes.c_float(0.0)
    nat.check(nat.lib().syn3r_selftest_div_const(float(c), 0, 1 << 32, 1, cnt.data_ptr(), ctypes.byref(rc),
                                                 torch.cuda.current_stream().cuda_stream))
    torch.cuda.synchronize()
    assert rc.value != 0.0 and abs(rc.value * float(c) - 1.0) < 1e-7
    assert int(cnt[0].item()) == 0, hex(int(cnt[1].item()))


@pytest.mark.parametrize("variant", [{}, {"fp32_tile": True}, {"exact_weight": True}, {"legacy": True}])
@pytest.mark.parametrize("shape,seed", [((3, 4, 60, 80), 21), ((2, 3, 97, 131), 22), ((1, 2, 16, 8), 23),
                                        ((2, 2, 150, 300), 24)])
def test_pair_mode_vs_oracle(shape, seed, variant):
    S, T, H, W = shape
    tr = synth.make_trajectory(S, T, H, W, seed=seed, pose_step=1.5)
    ps, pt = synth.all_pairs(S, T)
    o = _run_pairs(tr, ps, pt, **variant)
    for p, (s, t) in enumerate(zip(ps, pt)):
        r = _oracle_pair(tr, [s], t)
        # oracle winner indices are relative to its own source list: shift to the global source id
        r["winner"] = np.where(r["winner"] >= 0, r["winner"] + s * H * W, -1).astype(np.int32)
        _check_slot(o, p, r)


@pytest.mark.parametrize("variant", [{}, {"legacy": True}])
def test_fused_mode_vs_oracle(variant):
    S, T, H, W = 4, 3, 72, 96
    tr = synth.make_trajectory(S, T, H, W, seed=31, pose_step=1.0)
    ps, pt = synth.all_pairs(S, T)
    o = _run_pairs(tr, ps, pt, fused=True, **variant)
    pair = _run_pairs(tr, ps, pt)
    for t in range(T):
        r = _oracle_pair(tr, list(range(S)), t)
        _check_slot(o, t, r)
        # mask_fused == OR(mask_pair): derivable from the live reference
        union = np.zeros((H, W), bool)
        for p in np.nonzero(pt == t)[0]:
            union |= pair["mask"][p].astype(bool)
        assert np.array_equal(o["mask"][t].astype(bool), union)


def test_crowded_cells_take_redo_path():
    """A target camera pulled far back: the whole 96x128 source lands on a handful of cells, > 512 corner updates
    each, which the integer tile kernel hands to the fp32 tile kernel through the redo list (forward_fx.cuh, FX_MAX_N).
    Neighbouring tiles stay on the integer path; both must agree with the oracle."""
    tr = synth.make_trajectory(1, 2, 96, 128, seed=51, pose_step=1.0)
    tr.pose_tgt[0, 2, 3] += 60.0                       # w2c translation along the optical axis: ~30x minification
    ps, pt = synth.all_pairs(1, 2)
    o = _run_pairs(tr, ps, pt)
    for p, (s, t) in enumerate(zip(ps, pt)):
        r = _oracle_pair(tr, [s], t)
        if p == 0:
            assert r["mask"].sum() < 64                # really crowded
        _check_slot(o, p, r)


@pytest.mark.parametrize("pull_back,expect_redo", [(1.5, 0), (2.0, 3), (5.5, None)])
def test_many_fused_sources_under_minification(pull_back, expect_redo):
    """Config-E style crowding: 16 sources FUSED into a target camera pulled back along its optical axis, so that a
    cell receives ~ 4 S m^2 corner updates (m = minification).  Past FX_MAX_N = 512 updates a work item goes through
    the redo list to the fp32 tile kernel; below it the integer kernel handles all 16 sources.  pull_back 1.5: up to
    420 updates per cell, no item redone; 2.0: 3 of the 6 tiles hold a cell past 512 (565 / 567 / 599 updates, counted
    on the host with the reference's floor / ceil corners); 5.5: every tile that is touched.  All must agree with the
    oracle; the number of redone items is printed."""
    from syn3r_b200 import warp as w
    S, T, H, W = 16, 1, 144, 256
    tr = synth.make_trajectory(S, T, H, W, seed=61, pose_step=0.25)
    tr.pose_tgt[:, 2, 3] += pull_back                  # w2c translation along the optical axis
    ps, pt = synth.all_pairs(S, T)
    o = _run_pairs(tr, ps, pt, fused=True)
    redo = w.redo_items(torch.device("cuda"))
    n_items = T * ((W + 127) // 128) * ((H + 47) // 48)
    print(f"pull_back {pull_back}: {redo} of {n_items} work items took the fp32 redo kernel")
    r = _oracle_pair(tr, list(range(S)), 0)
    assert r["mask"].sum() > 500
    _check_slot(o, 0, r)
    if expect_redo is not None:
        assert redo == expect_redo
    else:
        assert redo > 0


def test_large_frame_takes_two_level_cull():
    """1088 x 576 = 2880 source-block slots > SYN3R_FX_FLAT (2304): the integer tile kernel is launched without the
    one-pass cull (abi.cu launch_tile_fx) and goes through the super-block / block lists; pair and fused mode."""
    tr = synth.make_trajectory(2, 1, 576, 1088, seed=57)
    ps, pt = synth.all_pairs(2, 1)
    o = _run_pairs(tr, ps, pt)
    for p, (s, t) in enumerate(zip(ps, pt)):
        r = _oracle_pair(tr, [s], t)
        r["winner"] = np.where(r["winner"] >= 0, r["winner"] + s * 576 * 1088, -1).astype(np.int32)
        _check_slot(o, p, r)
    o = _run_pairs(tr, ps, pt, fused=True)
    _check_slot(o, 0, _oracle_pair(tr, list(ps), 0))


def test_gate06_and_smooth_rgb():
    tr = synth.make_trajectory(2, 2, 64, 64, seed=41, smooth_rgb=True)
    ps, pt = synth.all_pairs(2, 2)
    o = _run_pairs(tr, ps, pt, gate06=True)
    for p, (s, t) in enumerate(zip(ps, pt)):
        r = orc.forward_warp_ext([tr.rgb[s]], [None], [tr.depth[s]], [tr.pose_src[s]], tr.pose_tgt[t], [tr.K], None,
                                 gate06=True)
        band = np.abs(r["S"] - 0.6) < 1e-6          # fp32 sum vs fp64 sum at the 0.6 gate
        m = o["mask"][p].astype(bool)
        assert np.array_equal(m[~band], r["mask"][~band])


def test_trajectory_plan_matches_per_pair_api(fw):
    """The batched TrajectoryWarp (what bench.py times) == the reference-signature per-pair calls."""
    from syn3r_b200 import trajectory
    S, T, H, W = 2, 3, 48, 64
    tr = synth.make_trajectory(S, T, H, W, seed=51)
    tw = trajectory.from_synth(tr, want_u8=True)
    from syn3r_b200 import warp as w
    tw.upload(torch.from_numpy(w.pack_rgbd(tr.rgb, tr.depth)), torch.from_numpy(tr.depth_tgt))
    out = tw.run()
    torch.cuda.synchronize()
    t_ = torch.from_numpy
    for p, (s, t) in enumerate(zip(tw.pair_src_h, tw.pair_tgt_h)):
        warped, mask2, _ = fw.forward_warp(tr.rgb[s], None, tr.depth[s], tr.pose_src[s], tr.pose_tgt[t], tr.K, None)
        assert np.array_equal(out.mask[p].cpu().numpy().astype(bool), mask2)
        _u8_close(out.rgb_u8[p].cpu().numpy(), warped, max_frac=1e-3)   # atomics order differs run to run
        iw = fw.inverse_warp(t_(np.moveaxis(tr.rgb[s], -1, 0).copy()), t_(tr.depth[s][None]),
                             t_(tr.depth_tgt[t][None]), t_(tr.pose_src[s]), t_(tr.pose_tgt[t]),
                             t_(tr.K.astype(np.float32)))
        assert torch.equal(out.mask_reproj[p].bool(), iw["mask_reproj"])
        # the trajectory kernel forms the soft mask with sqrt.approx / ex2.approx (inverse_kernels.cuh): <= 2e-6
        a, b = out.soft_mask_reproj[p].cpu().numpy(), iw["soft_mask_reproj"].cpu().numpy()
        fin = np.isfinite(a) & np.isfinite(b)
        assert np.array_equal(np.isnan(a), np.isnan(b)) and np.max(np.abs(a[fin] - b[fin]), initial=0) <= 2e-6


def test_inverse_vs_oracle_seeded():
    from syn3r_b200 import warp as w
    S, T, H, W = 2, 3, 75, 100
    tr = synth.make_trajectory(S, T, H, W, seed=61, pose_step=2.0)
    ps, pt = synth.all_pairs(S, T)
    dev = torch.device("cuda")
    up = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)  # noqa: E731
    img = np.moveaxis(tr.rgb, -1, 1).copy()
    want = ["warped_img", "warped_depth", "mask_warp", "mask_depth", "mask", "warped_masked_img", "mask_inv",
            "mask_depth_strict", "mask_reproj", "soft_mask_reproj", "reproj_err"]
    o = w.inverse_warp_batched(up(img), up(tr.depth), up(tr.depth_tgt), tr.K.astype(np.float32),
                               up(w.inverse_pair_params(tr.pose_src, tr.pose_tgt, ps, pt)), up(ps), up(pt),
                               bandwidth=10.0, want=want)
    t_ = torch.from_numpy
    for p, (s, t) in enumerate(zip(ps, pt)):
        r = orc.inverse_warp(t_(img[s]), t_(tr.depth[s][None]), t_(tr.depth_tgt[t][None]), t_(tr.pose_src[s]),
                             t_(tr.pose_tgt[t]), t_(tr.K.astype(np.float32)), bandwidth=10.0, return_err=True)
        for k in want:
            a = o[k][p].cpu().numpy()
            b = r[k].numpy().reshape(a.shape)
            if k == "soft_mask_reproj":
                fin = np.isfinite(a) & np.isfinite(b)
                assert np.array_equal(np.isnan(a), np.isnan(b)) and np.max(np.abs(a[fin] - b[fin]), initial=0) <= 2e-6
            else:
                assert np.array_equal(a.astype(b.dtype), b, equal_nan=True), (k, p)


@pytest.mark.parametrize("bw", [20.0, 10.0, 0.3, 7.77, 1e-3, 1e-30, 0.0, -1.0, float("inf"), 3e19])
def test_lean_reproj_mask_equals_err_below_bandwidth(bw):
    """The trajectory kernel never forms sqrt(e2): mask_reproj = e2 < threshold(bw) & mask_warp must equal
    (sqrt_rn(e2) < bw) & mask_warp of the kernel that does (same launch path, reproj_err requested)."""
    from syn3r_b200 import warp as w
    S, T, H, W = 2, 2, 61, 97
    tr = synth.make_trajectory(S, T, H, W, seed=63, pose_step=2.0)
    ps, pt = synth.all_pairs(S, T)
    up = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()  # noqa: E731
    args = (None, up(tr.depth), up(tr.depth_tgt), tr.K.astype(np.float32),
            up(w.inverse_pair_params(tr.pose_src, tr.pose_tgt, ps, pt)), up(ps), up(pt))
    lean = w.inverse_warp_batched(*args, bandwidth=bw, want=("mask_reproj", "soft_mask_reproj", "mask_warp"))
    full = w.inverse_warp_batched(*args, bandwidth=bw, want=("mask_reproj", "soft_mask_reproj", "mask_warp", "reproj_err"))
    torch.cuda.synchronize()
    assert torch.equal(lean["mask_warp"], full["mask_warp"])
    assert torch.equal(lean["mask_reproj"], full["mask_reproj"])
    assert torch.equal(full["mask_reproj"].bool(), (full["reproj_err"] < bw) & full["mask_warp"].bool())
    if 1e-3 <= bw <= 1e3:
        a, b = lean["soft_mask_reproj"].cpu().numpy(), full["soft_mask_reproj"].cpu().numpy()
        fin = np.isfinite(a) & np.isfinite(b)
        assert np.array_equal(np.isnan(a), np.isnan(b)) and np.max(np.abs(a[fin] - b[fin]), initial=0) <= 2e-6


def test_empty_batch_is_a_no_op():
    """Zero pairs: the reference's per-frame loop does not run; the batched entry points return empty outputs."""
    from syn3r_b200 import conditioning, warp as w
    tr = synth.make_trajectory(1, 1, 8, 16, seed=1)
    e = np.zeros((0,), np.int32)
    up = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()  # noqa: E731
    Kinv, par = w.forward_pair_params(tr.pose_src, tr.pose_tgt, tr.K, None, e, e)
    o = w.forward_warp_batched(up(w.pack_rgbd(tr.rgb, tr.depth)), up(Kinv), up(par.reshape(0, 24)), up(e), want_winner=True)
    assert o["mask"].shape == (0, 8, 16) and o["winner"].shape == (0, 8, 16)
    i = w.inverse_warp_batched(None, up(tr.depth), up(tr.depth_tgt), tr.K.astype(np.float32),
                               up(np.zeros((0, 2, 16), np.float32)), up(e), up(e), want=("mask_reproj", "soft_mask_reproj"))
    assert i["mask_reproj"].shape == (0, 8, 16)
    c = conditioning.condition_post_batched(torch.zeros((0, 8, 16), dtype=torch.uint8, device="cuda"))
    assert c["latent"].shape == (0, 1, 2)
    torch.cuda.synchronize()


def test_step_host_matches_run():
    """The pipelined end-to-end step (H2D, kernels in pair ranges, overlapped D2H) == run() + download()."""
    from syn3r_b200 import trajectory, warp as w
    S, T, H, W = 2, 4, 48, 160
    tr = synth.make_trajectory(S, T, H, W, seed=71)
    tw = trajectory.from_synth(tr, want_u8=True, pinned_host=True, post=True)
    rgbd_h, tgt_h = torch.from_numpy(w.pack_rgbd(tr.rgb, tr.depth)), torch.from_numpy(tr.depth_tgt)
    tw.upload(rgbd_h, tgt_h)
    tw.run()
    tw.download()
    torch.cuda.synchronize()
    ref = {k: v.clone() for k, v in tw.pinned.items() if k.startswith("out_")}
    for k in ref:
        tw.pinned[k].zero_()
    got = tw.step_host(rgbd_h, tgt_h, chunks=3)
    torch.cuda.synchronize()
    for k, v in got.items():
        if k.startswith("out_"):
            assert torch.equal(torch.nan_to_num(v.float()), torch.nan_to_num(ref[k].float())), k


@pytest.mark.parametrize("case", ["generic", "hostile", "identity", "projective", "tiny_scene"])
@pytest.mark.parametrize("shape", [(61, 97), (64, 256), (1, 40)])
def test_packed_reproj_matches_scalar(case, shape):
    """k_reproj_masks2 (two rows per thread, packed fp32 instructions, range-tested division fast path with the scalar
    arithmetic as its slow path) against k_reproj_masks (one row per thread, IEEE divisions): the same BITS on every
    output, for both depth layouts, with odd heights, depths that are zero / negative / NaN / inf / denormal / huge, an
    identity pose (exact zeros everywhere), a pose matrix with a projective bottom row and a scene of size 1e-15."""
    from syn3r_b200 import _native as nat, warp as w
    H, W = shape
    S, T = 2, 3
    tr = synth.make_trajectory(S, T, H, W, seed=71 + H, pose_step=2.0)
    depth, depth_tgt = tr.depth.copy(), tr.depth_tgt.copy()
    pose_src, pose_tgt = tr.pose_src.copy(), tr.pose_tgt.copy()
    rng = np.random.default_rng(5)
    if case == "hostile":
        for d in (depth, depth_tgt):
            flat = d.reshape(-1)
            idx = rng.choice(flat.size, size=max(8, flat.size // 6), replace=False)
            flat[idx] = rng.choice(np.array([0.0, -0.0, -1.5, np.nan, np.inf, -np.inf, 1e-42, 1e-30, 1e30, 3e38, 1e-12],
                                            np.float32), size=idx.size)
    elif case == "identity":
        pose_tgt[0] = pose_src[0]
        depth_tgt[0] = depth[0]
    elif case == "tiny_scene":
        depth *= 1e-15
        depth_tgt *= 1e-15
        pose_src[:, :3, 3] *= 1e-15
        pose_tgt[:, :3, 3] *= 1e-15
    ps, pt = synth.all_pairs(S, T)
    par = w.inverse_pair_params(pose_src, pose_tgt, ps, pt)
    if case == "projective":
        par = par.copy()
        par[:, :, 12:16] += np.array([1e-4, -2e-4, 3e-4, 0.01], np.float32)     # w != 1: the scalar path decides
    up = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()  # noqa: E731
    want = ("mask_reproj", "soft_mask_reproj", "mask_warp")
    rgbd = up(w.pack_rgbd(tr.rgb, depth))
    L = nat.lib()
    res = {}
    for packed in (1, 0):
        was = L.syn3r_set_reproj_packed(packed)
        try:
            planar = w.inverse_warp_batched(None, up(depth), up(depth_tgt), tr.K.astype(np.float32), up(par), up(ps), up(pt),
                                            bandwidth=3.0, want=want)
            lane = w.inverse_warp_batched(None, None, up(depth_tgt), tr.K.astype(np.float32), up(par), up(ps), up(pt), rgbd=rgbd,
                                          bandwidth=3.0, want=want)
            torch.cuda.synchronize()
        finally:
            L.syn3r_set_reproj_packed(was)
        res[packed] = (planar, lane)
    for k in want:
        for layout in (0, 1):
            a, b = res[1][layout][k].cpu().numpy(), res[0][layout][k].cpu().numpy()
            if a.dtype == np.float32:
                assert np.array_equal(a.view(np.uint32), b.view(np.uint32)), (case, shape, k, layout)
            else:
                assert np.array_equal(a, b), (case, shape, k, layout)
        a, b = res[1][0][k].cpu().numpy(), res[1][1][k].cpu().numpy()                      # planar == RGB-D lane
        assert np.array_equal(a.view(np.uint32), b.view(np.uint32)) if a.dtype == np.float32 else np.array_equal(a, b)
    if case == "generic":
        assert 0 < int(res[1][0]["mask_reproj"].sum()) < res[1][0]["mask_reproj"].numel() or H == 1


@pytest.mark.parametrize("shape", [(3, 5, 96, 160), (2, 3, 61, 97), (2, 2, 200, 300), (3, 4, 576, 1024)])
def test_overlapped_tile_kernel_matches(shape):
    """k_fwd_tile_fx_ov (resolve step split, its second phase overlapped with the next work item's first sweep, three
    barriers per item) against k_fwd_tile_fx: the same BITS on every forward output, with and without the winner index,
    for frames with ragged tiles (non-vector stores), several items per CTA and more items than CTAs."""
    from syn3r_b200 import _native as nat, trajectory, warp as w
    S, T, H, W = shape
    tr = synth.make_trajectory(S, T, H, W, seed=90 + H, pose_step=2.0)
    L = nat.lib()
    res = {}
    for winner in (True, False):
        for ov in (1, 0):
            was = L.syn3r_set_fx_overlap(ov)
            try:
                tw = trajectory.from_synth(tr, want_f32=True, want_u8=True, want_winner=winner, consistency=False)
                tw.upload(torch.from_numpy(w.pack_rgbd(tr.rgb, tr.depth)))
                o = tw.run()
                o2 = tw.run()                      # a second run on the same workspace: nothing stale in the planes
                torch.cuda.synchronize()
            finally:
                L.syn3r_set_fx_overlap(was)
            assert int(o.status.max().item()) == 0
            res[(winner, ov)] = {k: getattr(o2, k).cpu().numpy().copy() for k in ("rgb_f32", "rgb_u8", "mask") +
                                 (("winner",) if winner else ())}
        for k, a in res[(winner, 1)].items():
            b = res[(winner, 0)][k]
            same = np.array_equal(a.view(np.uint32), b.view(np.uint32)) if a.dtype == np.float32 else np.array_equal(a, b)
            assert same, (shape, winner, k, int((a != b).sum()))
    assert res[(True, 1)]["mask"].any()
