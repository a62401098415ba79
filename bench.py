#!/usr/bin/env python
"""bench.py — headline benchmark of the geometric-conditioning hot path (BASELINE.json `metric`).

One "step" = one pass of the hot path over one trajectory batch: config B, 3 source RGB-D views ->
25 target frames at 576x1024, PAIR mode (75 pair warps): forward warp (colour f32 + hole mask + winner
index) + reprojection-consistency masks per pair.  metric = target-view Mpix/s = T*H*W / time.

  python bench.py [--gpus N --steps K --warmup W]            our CUDA path (torchrun for N > 1)
  python bench.py --impl reference [...]                     the CPU restatement of the reference on host cores

Prints ONE JSON line (rank 0).  See DESIGN.md "Measurement" for every field.
"""
import argparse
import json
import os
import statistics
import subprocess
import sys
import tempfile
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

from syn3r_b200 import synth  # noqa: E402

METRIC = "target_view_mpix_per_s_warped_masked"
UNIT = "Mpix/s"


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--config", default="B_svd", choices=list(synth.CONFIGS))
    ap.add_argument("--pairs", default="all", choices=["all", "reference_policy"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--cpu-sample-seconds", type=float, default=12.0)
    ap.add_argument("--e2e-chunks", type=int, default=10, help="pair ranges whose D2H overlaps the next range's kernels")
    return ap.parse_args()


def workload_desc(args, S, T, H, W, P):
    return {"workload": f"{args.config}: {S} src RGB-D f32 -> {T} target poses at {H}x{W}, PAIR mode ({P} pair warps), "
                        "forward soft-z splat (rgb f32 + hole mask + winner idx) + reprojection consistency "
                        "(mask_reproj u8 + soft_mask f32) per pair",
            "S": S, "T": T, "H": H, "W": W, "pairs": P, "pair_policy": args.pairs,
            "l2_policy": "256 MiB buffer written between timed iterations (L2 flush); outputs 973 MB >> 126 MB L2"}


def pair_lists(args, S, T):
    return synth.all_pairs(S, T) if args.pairs == "all" else synth.reference_policy_pairs(T, S)


# -------------------------------------------------------------------------------------------------
# CPU legs (oracle port = C restatement of the reference; /root/reference cannot travel to the GPU box)
# -------------------------------------------------------------------------------------------------

def _cpu_one_pair(job):
    """forward_warp (+ winner) and inverse_warp-style consistency of ONE pair through the oracle port."""
    import torch
    from oracle import cpu_oracle as orc
    tr, s, t = job
    t_ = torch.from_numpy
    with np.errstate(all="ignore"):
        orc.forward_warp_ext([tr.rgb[s]], [None], [tr.depth[s]], [tr.pose_src[s]], tr.pose_tgt[t], [tr.K], None)
        orc.inverse_warp(t_(np.ascontiguousarray(np.moveaxis(tr.rgb[s], -1, 0))), t_(tr.depth[s][None]),
                         t_(tr.depth_tgt[t][None]), t_(tr.pose_src[s]), t_(tr.pose_tgt[t]),
                         t_(tr.K.astype(np.float32)))
    return 1


_G = {}


def _pool_job(i):
    tr, ps, pt = _G["tr"], _G["ps"], _G["pt"]
    return _cpu_one_pair((tr, int(ps[i % len(ps)]), int(pt[i % len(pt)])))


def cpu_baseline_sample(tr, ps, pt, seconds):
    """Scalar oracle port on ONE host core over a bounded sample of the same pairs."""
    from oracle import cpu_oracle as orc
    orc.lib()
    S, T, H, W = tr.shape
    _cpu_one_pair((tr, int(ps[0]), int(pt[0])))           # warm (page in, build)
    n, t0 = 0, time.perf_counter()
    while True:
        _cpu_one_pair((tr, int(ps[n % len(ps)]), int(pt[n % len(pt)])))
        n += 1
        el = time.perf_counter() - t0
        if el >= seconds or n >= len(ps):
            break
    pairs_per_view = len(ps) / T
    val = (n / pairs_per_view) * H * W / el / 1e6
    return {"value": val, "unit": UNIT, "cores": 1, "kind": "port",
            "sample": f"{n} of {len(ps)} pair warps (forward+winner and inverse/consistency) at {H}x{W} in {el:.1f} s, "
                      f"oracle/syn3r_oracle.c single thread; target-view rate = pair rate / {pairs_per_view:g}"}


def parity_gates(tr, ps, pt, tw, n_pairs=3):
    """SURVEY 8d 'parity gates reported with every timing': the plan's device outputs of the timed workload against
    the oracle on a few pairs at full size (first, middle, last)."""
    import torch
    from oracle import cpu_oracle as orc
    S, T, H, W = tr.shape
    t_ = torch.from_numpy
    sel = sorted({0, len(ps) // 2, len(ps) - 1})[:n_pairs]
    g = {"pairs_checked": [int(p) for p in sel], "hole_mask_xor": 0, "winner_mismatch": 0, "colour_max_abs_01": 0.0,
         "mask_reproj_xor": 0, "soft_mask_max_abs": 0.0}
    for p in sel:
        s, t = int(ps[p]), int(pt[p])
        with np.errstate(all="ignore"):
            r = orc.forward_warp_ext([tr.rgb[s]], [None], [tr.depth[s]], [tr.pose_src[s]], tr.pose_tgt[t], [tr.K], None)
            iw = orc.inverse_warp(t_(np.ascontiguousarray(np.moveaxis(tr.rgb[s], -1, 0))), t_(tr.depth[s][None]),
                                  t_(tr.depth_tgt[t][None]), t_(tr.pose_src[s]), t_(tr.pose_tgt[t]),
                                  t_(tr.K.astype(np.float32)))
        g["hole_mask_xor"] += int(np.count_nonzero(tw.out_fwd["mask"][p].cpu().numpy().astype(bool) ^ r["mask"]))
        win = np.where(r["winner"] >= 0, r["winner"] + s * H * W, -1)
        g["winner_mismatch"] += int(np.count_nonzero(tw.out_fwd["winner"][p].cpu().numpy() != win))
        g["colour_max_abs_01"] = max(g["colour_max_abs_01"],
                                     float(np.max(np.abs(tw.out_fwd["rgb_f32"][p].cpu().numpy() - r["out"]))) / 255.0)
        g["mask_reproj_xor"] += int(np.count_nonzero(tw.out_inv["mask_reproj"][p].cpu().numpy().astype(bool) ^
                                                     iw["mask_reproj"].numpy()))
        a, b = tw.out_inv["soft_mask_reproj"][p].cpu().numpy(), iw["soft_mask_reproj"].numpy()
        fin = np.isfinite(a) & np.isfinite(b)
        g["soft_mask_max_abs"] = max(g["soft_mask_max_abs"], float(np.max(np.abs(a[fin] - b[fin]), initial=0)))
    g["bars"] = "masks and winner: 0 mismatches; colour <= 1e-5 on [0,1]; soft mask <= 2e-6"
    return g


def run_reference(args):
    """--impl reference: the reference's CPU algorithm (oracle port) on all host cores, frames farmed over a
    process pool (frames are independent, SURVEY §8d 'fair multi-core')."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    import multiprocessing as mp
    from oracle import cpu_oracle as orc
    orc.build()
    S, T, H, W = synth.CONFIGS[args.config]
    tr = synth.make_trajectory(S, T, H, W, seed=0)
    ps, pt = pair_lists(args, S, T)
    cores = os.cpu_count() or 1
    workers = max(1, min(cores, 64))
    per_step = min(len(ps), max(workers, 8))                # bounded sample of the 75-pair step
    _G.update(tr=tr, ps=ps, pt=pt)
    ctx = mp.get_context("fork")
    with ctx.Pool(workers) as pool:
        for _ in range(max(args.warmup, 1)):
            pool.map(_pool_job, range(workers), chunksize=1)
        t0 = time.perf_counter()
        for _ in range(args.steps):
            pool.map(_pool_job, range(per_step), chunksize=1)
        el = time.perf_counter() - t0
    pairs_per_view = len(ps) / T
    val = (args.steps * per_step / pairs_per_view) * H * W / el / 1e6
    sample = (f"{per_step} of {len(ps)} pair warps per step at {H}x{W}, {workers} worker processes "
              f"(os.cpu_count()={cores}), oracle/syn3r_oracle.c")
    line = {"impl": "reference", "metric": METRIC, "value": val, "unit": UNIT, "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": el / args.steps * 1e3,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64 forward / f32 backward",
            "data": "synthetic", "config": workload_desc(args, S, T, H, W, len(ps)),
            "cpu_baseline": {"value": val, "unit": UNIT, "cores": workers, "kind": "port", "sample": sample},
            "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line))


# -------------------------------------------------------------------------------------------------
# GPU arm
# -------------------------------------------------------------------------------------------------

class ClockSampler:
    """nvidia-smi clocks/throttle reasons sampled every 200 ms DURING the timed region."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.proc, self.path = index, None, None

    def __enter__(self):
        try:
            f = tempfile.NamedTemporaryFile("w", suffix=".csv", delete=False)
            self.path = f.name
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "200"], stdout=f,
                                         stderr=subprocess.DEVNULL)
        except Exception:
            self.proc = None
        return self

    def __exit__(self, *a):
        if self.proc is not None:
            time.sleep(0.25)
            self.proc.terminate()
            try:
                self.proc.wait(timeout=5)
            except Exception:
                self.proc.kill()

    def summary(self):
        out = {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        try:
            rows = [r.split(",") for r in open(self.path).read().strip().splitlines() if r.strip()]
            sm = [float(r[0]) for r in rows]
            out["samples"] = len(sm)
            if sm:
                out["sm_mhz"] = statistics.median(sm)
                out["sm_max_mhz"] = float(rows[0][1])
                out["power_w_max"] = max(float(r[2]) for r in rows)
                names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
                out["reasons"] = [n for i, n in enumerate(names) if any("Active" == r[3 + i].strip() for r in rows)]
            os.unlink(self.path)
        except Exception as e:  # noqa: BLE001
            out["error"] = str(e)
        return out


def run_ours(args):
    import torch
    import torch.distributed as dist
    from syn3r_b200 import _native as nat, sharding, trajectory, warp as w

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    assert torch.cuda.is_available(), "bench.py (impl=ours) needs a GPU; there is no CPU fallback"
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        if os.environ.get("NCCL_DEBUG", "VERSION").upper() == "VERSION":
            os.environ["NCCL_DEBUG"] = "WARN"              # keep stdout to the one JSON line
        dist.init_process_group("nccl", device_id=dev)
    nat.lib()

    S, T, H, W = synth.CONFIGS[args.config]
    tr = synth.make_trajectory(S, T, H, W, seed=rank)        # weak scaling: one trajectory per rank
    ps, pt = synth.target_major(*pair_lists(args, S, T))    # same pairs; a pair range then needs few target maps
    tw = trajectory.from_synth(tr, pairs=(ps, pt), want_f32=True, want_u8=False, want_winner=True,
                               pinned_host=not args.no_e2e)
    rgbd_h = torch.from_numpy(w.pack_rgbd(tr.rgb, tr.depth))
    tgtd_h = torch.from_numpy(tr.depth_tgt)
    if not args.no_e2e:
        tw.pinned["rgbd"].copy_(rgbd_h)
        tw.pinned["tgt_depth"].copy_(tgtd_h)
        rgbd_h, tgtd_h = tw.pinned["rgbd"], tw.pinned["tgt_depth"]
    tw.upload(rgbd_h, tgtd_h)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)

    def barrier():
        if world > 1:
            dist.barrier()

    def timed(fn, steps, between=None):
        evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
        torch.cuda.synchronize()
        barrier()
        for e0, e1 in evs:
            if between is not None:
                between()
            e0.record()
            fn()
            e1.record()
        torch.cuda.synchronize()
        barrier()
        return [e0.elapsed_time(e1) for e0, e1 in evs]

    for _ in range(max(args.warmup, 3)):
        tw.run()
    torch.cuda.synchronize()
    assert int(tw.out_fwd["status"].max().item()) == 0, "colour range check failed"

    # ---- device-resident throughput (`value`) ------------------------------------------------------
    with ClockSampler(local) as cs:
        ms = timed(tw.run, args.steps, between=flush.zero_)
    clocks = cs.summary()
    total_ms = sum(ms)
    if world > 1:
        tt = torch.tensor([total_ms], device=dev, dtype=torch.float64)
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        total_ms = float(tt.item())
    mpix_step = T * H * W / 1e6
    value = world * args.steps * mpix_step / (total_ms / 1e3)

    # ---- per-kernel durations (CUDA events on the launching stream, phases launched one at a time) ---
    def phase(bits):
        return lambda: w.forward_warp_batched(tw.rgbd, tw.Kinv_src, tw.pair_par, tw.pair_src, want_f32=True,
                                              want_u8=False, want_winner=True, out=tw.out_fwd, phases=bits)

    def inv():
        return w.inverse_warp_batched(None, None, tw.tgt_depth, tw.K_host, tw.pair_P, tw.pair_src, tw.pair_tgt,
                                      rgbd=tw.rgbd, bandwidth=tw.bandwidth, want=("mask_reproj", "soft_mask_reproj"),
                                      out=tw.out_inv)

    kern = {}
    n_k = max(3, min(args.steps, 10))
    for name, fn in (("fwd_prep(stats+bounds+max z')", phase(nat.FWD_PHASE_ZMAX)),
                     ("fwd_tile(splat+resolve)", phase(nat.FWD_PHASE_SPLAT)), ("inverse_consistency", inv)):
        kern[name] = statistics.mean(timed(fn, n_k, between=flush.zero_))
    ksum = sum(kern.values())
    dominant = max(kern, key=kern.get)

    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    peak = float(peaks.get("hbm_gbs", 6650.0))
    hw = H * W
    b_step = tw.algorithmic_bytes()
    # algorithmic bytes of ONE launch of each kernel (inputs once + outputs once, accumulators zero)
    b_kernel = {"fwd_prep(stats+bounds+max z')": 2 * S * hw * 16,                      # sources read by stats and by max z'
                "fwd_tile(splat+resolve)": S * hw * 16 + len(ps) * hw * 17,
                "inverse_consistency": S * hw * 4 + T * hw * 4 + len(ps) * hw * 5}
    traffic = None
    try:    # dram__bytes_read.sum + dram__bytes_write.sum per launch, from the committed ncu launch list
        traffic = json.load(open(os.path.join(ROOT, "profiles", "ncu_traffic.json"))).get(dominant)
    except Exception:
        pass
    ach_k = b_kernel[dominant] / (kern[dominant] / 1e3) / 1e9
    ach_s = b_step / (ksum / 1e3) / 1e9
    roofline = {"bound": "hbm", "kernel": dominant, "achieved": ach_k, "peak": peak, "unit": "GB/s",
                "frac": ach_k / peak, "traffic": traffic,
                "peak_source": "MEASURED_PEAKS.json hbm_gbs (burst copy), of measured" if peaks else "fallback 6650 GB/s, of fallback",
                "algorithmic_bytes_per_launch": b_kernel[dominant],
                "step": {"achieved": ach_s, "frac": ach_s / peak, "algorithmic_bytes_per_step": b_step,
                         "scope": "all kernels of one step: SURVEY 8d B_alg / their summed CUDA-event time"},
                "kernel_ms": kern, "kernel_share": {k: v / ksum for k, v in kern.items()},
                "note": "instruction-issue bound, not HBM bound: see DESIGN.md 'Roofline' (fp64 projection + "
                        "soft-z weight + 16 integer shared-memory atomics per pair-pixel)"}

    # ---- the callers' per-frame post-processing as one launch per branch (HBM-bound byte work) ----
    post = None
    if H % 8 == 0 and W % 8 == 0:
        from syn3r_b200 import conditioning
        P_ = len(ps)
        m_fw, m_bw = tw.out_fwd["mask"], tw.out_inv["mask_reproj"]
        img8 = torch.randint(0, 256, (P_, H, W, 3), dtype=torch.uint8, device=dev)
        o_fw, o_bw = {}, {}

        def post_fw():
            o_fw.update(conditioning.condition_post_batched(m_fw, img_u8=img8, want=("mask_png", "cond_u8", "latent"), out=o_fw))

        def post_bw():
            o_bw.update(conditioning.condition_post_batched(m_bw, soft=tw.out_inv["soft_mask_reproj"],
                                                            want=("latent", "soft_pooled"), out=o_bw))
        post_fw(); post_bw()
        t_fw = statistics.mean(timed(post_fw, n_k, between=flush.zero_))
        t_bw = statistics.mean(timed(post_bw, n_k, between=flush.zero_))
        b_fw = P_ * hw * 8 + P_ * (hw // 64)                  # valid 1 + frame 3 in; png 1 + blanked frame 3 + latent out
        b_bw = P_ * hw * 5 + P_ * (hw // 64) * 5              # valid 1 + soft 4 in; latent 1 + pooled 4 per 8x8 block out
        # uncertainty fusion of the T - 2 interior frames (diffusionGS.py:821-854), 28 B in + 12 B out
        Nf = max(T - 2, 1)
        c_ori = torch.rand((Nf, H, W, 3), device=dev)
        g_img = torch.rand((Nf, H, W, 3), device=dev)
        s_ori = torch.rand((Nf, H, W), device=dev)
        o_fu = {}

        def fuse():
            o_fu.update(conditioning.fuse_uncertainty_batched(c_ori, g_img, s_ori, want=("cond_image", "masks"), out=o_fu))
        fuse()
        t_fu = statistics.mean(timed(fuse, n_k, between=flush.zero_))
        b_fu = Nf * hw * 40 + Nf * (hw // 64) * 4
        del c_ori, g_img, s_ori
        # second half of rank 2: nearby-frame consistency of the T interpolated frames (diffusionGS.py:1300-1361):
        # 2 (T - 1) neighbour warps (batched inverse warp: warped_img + soft mask) and their fusion
        Nn = max(T, 2)
        ps_n, pt_n, first_n = conditioning.nearby_pairs(Nn, 1)
        Pn = len(ps_n)
        soft_n = torch.rand((Pn, H, W), device=dev)
        warp_n = torch.rand((Pn, 3, H, W), device=dev) * 255
        img_n = torch.rand((Nn, 3, H, W), device=dev) * 255
        first_d = torch.from_numpy(first_n).to(dev)
        o_nb = {}

        def nearby():
            o_nb.update(conditioning.nearby_uncertainty_batched(soft_n, warp_n, img_n, first_d, out=o_nb))
        nearby()
        t_nb = statistics.mean(timed(nearby, n_k, between=flush.zero_))
        b_nb = Pn * hw * 16 + Nn * hw * 20                      # soft 4 + warp 12 per pair; image 12 in, 2 x 4 out per frame
        del soft_n, warp_n, img_n
        # rank 4, gmflow half: forward-backward flow consistency of T - 1 frame pairs (geometry.py:75-96)
        from syn3r_b200 import _native as nat
        Bf = max(T - 1, 1)
        yy_, xx_ = torch.meshgrid(torch.arange(H, device=dev, dtype=torch.float32),
                                  torch.arange(W, device=dev, dtype=torch.float32), indexing="ij")
        f_fl = torch.stack([2.5 + 3.0 * torch.sin(0.02 * yy_) + 0.01 * xx_, -1.75 + 3.0 * torch.cos(0.015 * xx_) - 0.01 * yy_])
        f_fl = (f_fl[None] + torch.randn((Bf, 2, H, W), device=dev) * 0.3).contiguous()     # smooth motion + sub-pixel noise
        b_fl = -f_fl + torch.randn((Bf, 2, H, W), device=dev) * 0.3
        o_fl = [torch.empty((Bf, H, W), device=dev) for _ in range(2)]

        def flow_fb():
            nat.check(nat.lib().syn3r_flow_fb_consistency_batched(nat.ptr(f_fl), nat.ptr(b_fl), Bf, H, W, 0.01, 0.5, nat.ptr(o_fl[0]),
                                                                  nat.ptr(o_fl[1]), None, None,
                                                                  torch.cuda.current_stream().cuda_stream), "flow_fb")
        flow_fb()
        t_fl = statistics.mean(timed(flow_fb, n_k, between=flush.zero_))
        b_fl_bytes = Bf * hw * 24                               # two 2-channel flows in, two flags out
        del f_fl, b_fl
        # rank 4, scheduler half: guided latent replacement of one diffusion step (scheduling_euler_discrete.py:1442-1499),
        # 25 frames of 4 x 72 x 128 latents; launch-latency bound (25 CTAs, ~11 MB), reported as time per step
        from syn3r_b200 import latent as latent_mod
        Fl, hl, wl = 25, H // 8, W // 8
        p_l = torch.randn((1, Fl, 4, hl, wl), device=dev)
        c_l = torch.randn((2, Fl, 4, hl, wl), device=dev)
        m_l = (torch.rand((1, Fl - 1, 1, hl, wl), device=dev) < 0.3).float()
        lam_l = torch.rand((100, Fl), device=dev)

        cond_l, row_l, REP = c_l[1].contiguous(), lam_l[7].contiguous(), 10

        def latent_step():      # the C-ABI call itself, REP diffusion steps back to back (a 50 us kernel: one Python-level
            for _ in range(REP):  # call between two events would time the host, not the device)
                nat.check(nat.lib().syn3r_latent_guided_replace(nat.ptr(p_l), nat.ptr(cond_l), nat.ptr(m_l), nat.ptr(row_l), 1, Fl, 4,
                                                                hl, wl, 1, Fl - 1, 1, 1, Fl - 2, None, None,
                                                                torch.cuda.current_stream().cuda_stream), "latent")
        latent_mod.guided_replace_(p_l, c_l, m_l, lam_l, 7)
        latent_step()
        t_lat = statistics.mean(timed(latent_step, n_k, between=flush.zero_)) / REP
        b_lat = (Fl - 2) * 4 * hl * wl * 4 * 3 + (Fl - 2) * hl * wl * 4 + 2 * 4 * hl * wl * 8
        post = {"latent_replace": {"kernel": "k_latent_guided_replace", "frames": Fl, "latent": [4, hl, wl], "ms": t_lat,
                                   "algorithmic_bytes": b_lat, "achieved_GBps": b_lat / t_lat / 1e6,
                                   "frac": b_lat / t_lat / 1e6 / peak,
                                   "note": "one CTA per frame (25 CTAs): latency bound by design; the reference issues a torch.sort "
                                           "and ~25 small launches per frame for the same step; mean of 10 back-to-back steps"},
                "flow_fb": {"kernel": "k_flow_fb_consistency", "frame_pairs": Bf, "ms": t_fl, "algorithmic_bytes": b_fl_bytes,
                            "achieved_GBps": b_fl_bytes / t_fl / 1e6, "frac": b_fl_bytes / t_fl / 1e6 / peak,
                            "input": "smooth synthetic motion of a few pixels + 0.3 px noise (taps of neighbouring pixels share lines)"},
                "nearby": {"kernel": "k_nearby_uncertainty", "frames": Nn, "pairs": Pn, "ms": t_nb, "algorithmic_bytes": b_nb,
                           "achieved_GBps": b_nb / t_nb / 1e6, "frac": b_nb / t_nb / 1e6 / peak},
                "kernel": "k_condition_post", "forward_branch_ms": t_fw, "backward_branch_ms": t_bw,
                "fusion": {"kernel": "k_fuse_uncertainty", "frames": Nf, "ms": t_fu, "algorithmic_bytes": b_fu,
                           "achieved_GBps": b_fu / t_fu / 1e6, "frac": b_fu / t_fu / 1e6 / peak},
                "forward_branch": {"algorithmic_bytes": b_fw, "achieved_GBps": b_fw / t_fw / 1e6, "frac": b_fw / t_fw / 1e6 / peak},
                "backward_branch": {"algorithmic_bytes": b_bw, "achieved_GBps": b_bw / t_bw / 1e6, "frac": b_bw / t_bw / 1e6 / peak},
                "scope": "diffusionGS.py:1572-1597 / :1452-1491 for all pairs; not part of `value` (SURVEY 8f rank 1)"}

    # ---- N > 1: the one exchange step of the path, the gather of finished frames on the diffusion rank -------
    gather, gather_info = None, None
    if world > 1:
        wire = sharding.compact_wire(tw.out_fwd, tw.out_inv)
        gather = sharding.FrameGather(wire, [tw.P] * world, dst=0)
        for _ in range(2):
            gather(wire)
        ms_g = timed(lambda: gather(wire), max(3, min(args.steps, 10)))
        tg = torch.tensor([statistics.mean(ms_g)], device=dev, dtype=torch.float64)
        dist.all_reduce(tg, op=dist.ReduceOp.MAX)
        nbytes = gather.bytes_received(wire)
        gather_info = {"ms": float(tg.item()), "bytes_into_rank0": nbytes, "GBps_into_rank0": nbytes / float(tg.item()) / 1e6,
                       "wire": sorted(wire.keys()), "backend": "nccl grouped send/recv (batch_isend_irecv)"}

    # ---- end to end through the plan with HOST buffers ------------------------------------------------
    # `e2e`: what the reference's functions hand back to their callers travels to the host — forward_warp's uint8
    # frame + hole mask (forward_warp.py:182) and the consumed inverse_warp keys mask_reproj + soft_mask_reproj
    # (diffusionGS.py:1346-1347); the winner index (extension) is computed and stays on the device.
    # `e2e_all_outputs`: the `value` plan itself with every output (fp32 colour + winner) copied back.
    e2e, e2e_all, e2e_post = None, None, None
    if not args.no_e2e:
        def measure(plan, how):
            g_wire = g_plan = None
            if world > 1:                                   # the plan's own wire: uint8 frames for the reference-facing plan
                g_wire = ({k: plan.out_post[k] for k in plan.host_outputs if k in plan.out_post} if plan.post
                          else sharding.compact_wire(plan.out_fwd, plan.out_inv))
                g_plan = gather if plan is tw else sharding.FrameGather(g_wire, [plan.P] * world, dst=0)

            def step_e2e():
                plan.step_host(rgbd_h, tgtd_h, chunks=args.e2e_chunks)
                if g_plan is not None:
                    g_plan(g_wire)
            for _ in range(2):
                step_e2e()
            ms_e = timed(step_e2e, args.steps)
            tot = sum(ms_e)
            if world > 1:
                tt = torch.tensor([tot], device=dev, dtype=torch.float64)
                dist.all_reduce(tt, op=dist.ReduceOp.MAX)
                tot = float(tt.item())
            return {"value": world * args.steps * mpix_step / (tot / 1e3), "unit": UNIT,
                    "h2d_bytes_per_step": plan.h2d_bytes(), "d2h_bytes_per_step": plan.d2h_bytes(),
                    "ms_per_step": tot / args.steps, "host_outputs": list(plan.host_outputs),
                    "gather_wire": sorted(g_wire.keys()) if g_wire else None, "how": how}
        e2e_all = measure(tw, f"TrajectoryWarp.step_host on the `value` plan: pinned H2D of RGB-D + target depth, kernels in "
                              f"{args.e2e_chunks} pair ranges, D2H of EVERY output (fp32 colour, winner) overlapping the next range")
        tw_ref = trajectory.from_synth(tr, pairs=(ps, pt), want_f32=False, want_u8=True, want_winner=True, pinned_host=True,
                                       host_outputs=("rgb_u8", "mask", "mask_reproj", "soft_mask_reproj"))
        e2e = measure(tw_ref, f"TrajectoryWarp.step_host: pinned H2D of RGB-D + target depth, kernels in {args.e2e_chunks} pair "
                              "ranges (uint8 frame + hole mask + winner + consistency masks), D2H of the outputs the "
                              "reference's forward_warp / inverse_warp return to their callers, overlapping the next "
                              "range (PCIe-bound)")
        assert int(tw_ref.out_fwd["status"].max().item()) == 0
        del tw_ref
        # the same step when the callers' post-processing also runs on the device: only what the SVD
        # pipeline consumes travels back - the blanked uint8 frame, the two latent masks, the pooled uncertainty
        if H % 8 == 0 and W % 8 == 0:
            tw_post = trajectory.from_synth(tr, pairs=(ps, pt), want_f32=False, want_u8=True, want_winner=True,
                                            pinned_host=True, post=True,
                                            host_outputs=("post_cond_u8", "post_latent", "post_latent_reproj", "post_soft_pooled"))
            e2e_post = measure(tw_post, "as `e2e`, plus k_condition_post on every pair range; D2H of the blanked uint8 frame, "
                                        "the latent masks of the hole and reprojection masks and the pooled uncertainty "
                                        "(diffusionGS.py:1572-1597, :1452-1491 on the device)")
            del tw_post

    cpu, parity = None, None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        cpu = cpu_baseline_sample(tr, ps, pt, args.cpu_sample_seconds)
        parity = parity_gates(tr, ps, pt, tw)

    if rank == 0:
        line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
                "warmup": max(args.warmup, 3), "ms_per_step": total_ms / args.steps, "higher_is_better": True,
                "scaling": "weak", "vs_baseline": None, "dtype": "f64 projection / block-floating-point i32 accumulate / f32 consistency", "data": "synthetic",
                "config": workload_desc(args, S, T, H, W, len(ps)), "clocks": clocks, "e2e": e2e,
                "e2e_all_outputs": e2e_all, "e2e_conditioning": e2e_post,
                "gpu_launches": args.steps * tw.launches_per_run(), "roofline": roofline, "cpu_baseline": cpu,
                "pair_mpix_per_s": world * args.steps * len(ps) * H * W / 1e6 / (total_ms / 1e3), "parity": parity}
        if gather_info is not None:
            line["gather"] = gather_info
        if post is not None:
            line["post"] = post
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    a = parse_args()
    if a.impl == "reference":
        run_reference(a)
    else:
        run_ours(a)
