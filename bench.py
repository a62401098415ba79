#!/usr/bin/env python
"""bench.py — headline benchmark of the geometric-conditioning hot path (BASELINE.json `metric`).

One "step" = one pass of the hot path over one workload of BASELINE.json (`--config`, default B: 3 source RGB-D views
-> 25 target frames at 576x1024, PAIR mode = 75 pair warps: forward warp (colour f32 + hole mask + winner index) +
reprojection-consistency masks per pair).  metric = target-view Mpix/s = frames * H * W / time.

  python bench.py [--gpus N --steps K --warmup W] [--config B_svd|C_dtu|D_dl3dv|E_stress] [--mode pair|fused]
        our CUDA path.  N > 1 (torchrun): the workload's target frames are PARTITIONED over the ranks (north_star's
        split of ONE trajectory; whole trajectories first for config D) and the gather of the finished frames on the
        diffusion rank (rank 0, NCCL over NVLink or direct peer stores) is INSIDE the timed region: `scaling: "strong"`.
  python bench.py --impl reference [...]
        the reference's CPU algorithm on the host cores: the live reference staged under oracle/_ref
        (tools/stage_reference.sh) when present, else the C port of it (oracle/syn3r_oracle.c).

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

from syn3r_b200 import synth, workload  # noqa: E402

METRIC = "target_view_mpix_per_s_warped_masked"
UNIT = "Mpix/s"


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--config", default="B_svd", choices=list(synth.CONFIGS))
    ap.add_argument("--mode", default=None, choices=["pair", "fused"], help="default: the config's own (E: fused, else pair)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-post", action="store_true", help="skip the SURVEY 8f kernels' roofline object")
    ap.add_argument("--cpu-sample-seconds", type=float, default=12.0)
    ap.add_argument("--e2e-chunks", type=int, default=4, help="slot ranges whose D2H overlaps the next range's kernels")
    ap.add_argument("--e2e-equal-ranges", action="store_true",
                    help="equal slot ranges instead of the doubling ramp (A/B: the first D2H then starts later)")
    ap.add_argument("--gather", default="auto", choices=["auto", "nccl", "peer"],
                    help="N > 1: how finished frames reach rank 0 (auto: direct peer stores when symmetric memory works)")
    ap.add_argument("--gather-chunks", type=int, default=4, help="N > 1, nccl: slot ranges whose send overlaps the next range")
    ap.add_argument("--wire", default="compact", choices=["compact", "bytes", "post"],
                    help="N > 1: what travels to rank 0: the reference-facing outputs (uint8 frame + masks + soft mask, 9 B/px) "
                         "or what the SVD pipeline consumes after the post-processing (3.1 B/px)")
    ap.add_argument("--weak", action="store_true", help="N > 1: one whole workload per rank, no collective (secondary figure)")
    return ap.parse_args()


def config_desc(args, wl, extra=None):
    S, T = wl.trajs[0]
    d = {"workload": wl.describe(), "S": S, "T": T, "trajectories": len(wl.trajs), "H": wl.H, "W": wl.W, "pairs": wl.pairs,
         "frames": wl.frames, "mode": wl.mode,
         "l2_policy": "256 MiB buffer written between timed iterations (L2 flush); outputs >> 126 MB L2"}
    if extra:
        d.update(extra)
    return d


# -------------------------------------------------------------------------------------------------
# CPU legs: the live reference (oracle/_ref, staged copy of the reference's two files) or its C port
# -------------------------------------------------------------------------------------------------

REF_DIR = os.path.join(ROOT, "oracle", "_ref")
_G = {}


def live_reference():
    """(forward_warp, inverse_warp) of the staged live reference, or None.  `inverse_warp` hard-codes .to('cuda')
    (forward_warp.py:225-256): for the CPU baseline a shim maps it to 'cpu' (SURVEY §8c caveat 1)."""
    if "live" in _G:
        return _G["live"]
    _G["live"] = None
    if os.path.exists(os.path.join(REF_DIR, "solver_utils", "forward_warp.py")):
        import torch
        orig_to = torch.Tensor.to

        def to_shim(self, *a, **k):
            a = tuple("cpu" if (isinstance(x, str) and x.startswith("cuda")) else x for x in a)
            if isinstance(k.get("device"), str) and k["device"].startswith("cuda"):
                k["device"] = "cpu"
            return orig_to(self, *a, **k)
        torch.Tensor.to = to_shim
        sys.path.insert(0, REF_DIR)
        import warnings
        warnings.filterwarnings("ignore")
        import solver_utils.forward_warp as fw
        _G["live"] = (fw.forward_warp, fw.inverse_warp)
    return _G["live"]


def _cpu_one_pair(job):
    """forward_warp (+ inverse_warp when the workload has the consistency pass) of ONE pair on the CPU."""
    import torch
    tr, s, t, consistency, kind = job
    t_ = torch.from_numpy
    with np.errstate(all="ignore"):
        if kind == "reference":
            fwd, inv = live_reference()
            fwd(tr.rgb[s], None, tr.depth[s], tr.pose_src[s], tr.pose_tgt[t], tr.K, None)
            if consistency:
                inv(t_(np.ascontiguousarray(np.moveaxis(tr.rgb[s], -1, 0))), t_(tr.depth[s][None]), t_(tr.depth_tgt[t][None]),
                    t_(tr.pose_src[s]), t_(tr.pose_tgt[t]), t_(tr.K.astype(np.float32)))
        else:
            from oracle import cpu_oracle as orc
            orc.forward_warp_ext([tr.rgb[s]], [None], [tr.depth[s]], [tr.pose_src[s]], tr.pose_tgt[t], [tr.K], None)
            if consistency:
                orc.inverse_warp(t_(np.ascontiguousarray(np.moveaxis(tr.rgb[s], -1, 0))), t_(tr.depth[s][None]),
                                 t_(tr.depth_tgt[t][None]), t_(tr.pose_src[s]), t_(tr.pose_tgt[t]),
                                 t_(tr.K.astype(np.float32)))
    return 1


def _pool_job(i):
    tr, pairs, cons, kind = _G["tr"], _G["pairs"], _G["cons"], _G["kind"]
    s, t = pairs[i % len(pairs)]
    import torch
    torch.set_num_threads(1)
    return _cpu_one_pair((tr, s, t, cons, kind))


def cpu_trajectory(wl):
    """One trajectory of the workload on the host (config E: two sources are enough for a bounded sample)."""
    S, T = wl.trajs[0]
    if wl.lazy_sources:
        S = 2
        K, ps, pt = workload.lazy_cameras(wl.trajs[0][0], T, wl.H, wl.W, wl.pose_step)
        src = [workload.lazy_source(wl.H, wl.W, s, 0) for s in range(S)]
        return synth.Trajectory(np.stack([a for a, _ in src]), np.stack([b for _, b in src]), K, ps[:S], pt,
                                np.zeros((0, wl.H, wl.W), np.float32))
    return synth.make_trajectory(S, T, wl.H, wl.W, seed=0, pose_step=wl.pose_step)


def cpu_baseline_sample(wl, seconds):
    """ONE host core over a bounded sample of the workload's own pair warps: the live reference when staged (`kind:
    "reference"`, the sequential loop of diffusionGS.py:1554), else the scalar C port (`kind: "port"`)."""
    import torch
    kind = "reference" if live_reference() else "port"
    if kind == "port":
        from oracle import cpu_oracle as orc
        orc.lib()
    tr = cpu_trajectory(wl)
    S, T, H, W = tr.depth.shape[0], tr.pose_tgt.shape[0], wl.H, wl.W
    pairs = [(s, t) for t in range(T) for s in range(S)]
    nt = torch.get_num_threads()
    torch.set_num_threads(1)
    _cpu_one_pair((tr, pairs[0][0], pairs[0][1], wl.consistency, kind))           # warm (page in, build)
    n, t0 = 0, time.perf_counter()
    while True:
        s, t = pairs[(n * 7) % len(pairs)]
        _cpu_one_pair((tr, s, t, wl.consistency, kind))
        n += 1
        el = time.perf_counter() - t0
        if el >= seconds or n >= len(pairs):
            break
    torch.set_num_threads(nt)
    pairs_per_view = wl.trajs[0][0]
    val = (n / pairs_per_view) * H * W / el / 1e6
    what = ("live reference forward_warp (NumPy f64)" + (" + inverse_warp (torch CPU, 1 thread)" if wl.consistency else "")
            if kind == "reference" else "oracle/syn3r_oracle.c (C port of the reference), single thread")
    return {"value": val, "unit": UNIT, "cores": 1, "kind": kind,
            "sample": f"{n} of {wl.pairs} pair warps ({'forward + inverse/consistency' if wl.consistency else 'forward only'}) at "
                      f"{H}x{W} in {el:.1f} s, {what}; target-view rate = pair rate / {pairs_per_view:g}"}


def run_reference(args):
    """--impl reference: the reference's CPU implementation of the path on all host cores, pair warps farmed over a
    process pool (frames are independent, SURVEY §8d 'fair multi-core').  Each step is a bounded sample of the workload."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    import multiprocessing as mp
    wl = workload.for_config(args.config, args.mode)
    kind = "reference" if live_reference() else "port"
    if kind == "port":
        from oracle import cpu_oracle as orc
        orc.build()
    tr = cpu_trajectory(wl)
    S, T, H, W = tr.depth.shape[0], tr.pose_tgt.shape[0], wl.H, wl.W
    pairs = [(s, t) for t in range(T) for s in range(S)]
    cores = os.cpu_count() or 1
    workers = max(1, min(cores, 64))
    per_step = min(len(pairs), max(workers, 8))             # bounded sample of the step's pair warps
    _G.update(tr=tr, pairs=pairs, cons=wl.consistency, kind=kind)
    ctx = mp.get_context("fork")
    with ctx.Pool(workers) as pool:
        for _ in range(max(min(args.warmup, 2), 1)):
            pool.map(_pool_job, range(workers), chunksize=1)
        t0 = time.perf_counter()
        for _ in range(args.steps):
            pool.map(_pool_job, range(per_step), chunksize=1)
        el = time.perf_counter() - t0
    pairs_per_view = wl.trajs[0][0]
    val = (args.steps * per_step / pairs_per_view) * H * W / el / 1e6
    ms_sample = el / args.steps * 1e3
    what = ("live reference (oracle/_ref: forward_warp NumPy f64" + (" + inverse_warp torch CPU)" if wl.consistency else ")")
            if kind == "reference" else "oracle/syn3r_oracle.c")
    sample = (f"{per_step} of {wl.pairs} pair warps per step at {H}x{W}, {workers} worker processes (os.cpu_count()={cores}), {what}")
    line = {"impl": "reference", "metric": METRIC, "value": val, "unit": UNIT, "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms_sample * wl.pairs / per_step,                # a WHOLE step of wl.pairs pair warps at this rate
            "ms_per_sample": ms_sample,
            "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64 forward / f32 backward",
            "data": "synthetic", "config": config_desc(args, wl),
            "cpu_baseline": {"value": val, "unit": UNIT, "cores": workers, "kind": kind, "sample": sample},
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


def parity_gates(wl, share, n_pairs=3):
    """SURVEY 8d 'parity gates reported with every timing': the plan's device outputs of the timed workload against
    the oracle (pinned to the live reference at these sizes, tests/test_fullsize_golden.py) on a few full-size pairs."""
    import torch
    from oracle import cpu_oracle as orc
    tw = share.plan
    H, W = wl.H, wl.W
    t_ = torch.from_numpy
    if wl.mode == "fused":
        return fused_gates(wl, share)
    P = tw.P
    sel = sorted({0, P // 2, P - 1})[:n_pairs]
    g = {"pairs_checked": [int(p) for p in sel], "hole_mask_xor": 0, "winner_mismatch": 0, "colour_max_abs_01": 0.0,
         "uint8_values": 0, "uint8_exact_match": 0, "uint8_max_abs": 0, "mask_reproj_xor": 0, "soft_mask_max_abs": 0.0}
    for p in sel:
        k, s, t = share.pair_global[p]
        tr = share.trajs[share.traj_ids.index(k)]
        with np.errstate(all="ignore"):
            r = orc.forward_warp_ext([tr.rgb[s]], [None], [tr.depth[s]], [tr.pose_src[s]], tr.pose_tgt[t], [tr.K], None)
        g["hole_mask_xor"] += int(np.count_nonzero(tw.out_fwd["mask"][p].cpu().numpy().astype(bool) ^ r["mask"]))
        s_loc = int(tw.pair_src_h[p])
        win = np.where(r["winner"] >= 0, r["winner"] + s_loc * H * W, -1)
        g["winner_mismatch"] += int(np.count_nonzero(tw.out_fwd["winner"][p].cpu().numpy() != win))
        f32 = tw.out_fwd["rgb_f32"][p].cpu().numpy()
        g["colour_max_abs_01"] = max(g["colour_max_abs_01"], float(np.max(np.abs(f32 - r["out"]))) / 255.0)
        # the reference's actual return value is the uint8 frame (forward_warp.py:127-128): exact-match count
        u_ref = np.round(np.clip(r["out"], 0, 255)).astype(np.uint8)
        u_gpu = np.rint(np.clip(f32, 0, 255)).astype(np.uint8)
        d = np.abs(u_ref.astype(np.int16) - u_gpu.astype(np.int16))
        g["uint8_values"] += int(d.size)
        g["uint8_exact_match"] += int(np.count_nonzero(d == 0))
        g["uint8_max_abs"] = max(g["uint8_max_abs"], int(d.max()))
        if wl.consistency:
            with np.errstate(all="ignore"):
                iw = orc.inverse_warp(t_(np.ascontiguousarray(np.moveaxis(tr.rgb[s], -1, 0))), t_(tr.depth[s][None]),
                                      t_(tr.depth_tgt[t][None]), t_(tr.pose_src[s]), t_(tr.pose_tgt[t]),
                                      t_(tr.K.astype(np.float32)))
            g["mask_reproj_xor"] += int(np.count_nonzero(tw.out_inv["mask_reproj"][p].cpu().numpy().astype(bool) ^
                                                         iw["mask_reproj"].numpy()))
            a, b = tw.out_inv["soft_mask_reproj"][p].cpu().numpy(), iw["soft_mask_reproj"].numpy()
            fin = np.isfinite(a) & np.isfinite(b)
            g["soft_mask_max_abs"] = max(g["soft_mask_max_abs"], float(np.max(np.abs(a[fin] - b[fin]), initial=0)))
    g["bars"] = ("masks and winner: 0 mismatches; colour <= 1e-5 on [0,1]; uint8 frame: single levels at rounding ties only "
                 "(SURVEY N5); soft mask <= 2e-6; mask_reproj follows torch's CPU op order (the reference runs it on CUDA: "
                 "flips only inside |err - bw| <= 1e-3 bw, SURVEY H4)")
    return g


def fused_gates(wl, share):
    """FUSED mode is an extension: what the live reference can pin is `mask_fused == OR(mask_pair)`.  Checked
    on the device for the first and last frame of the timed plan (PAIR-mode warps of the same sources, themselves pinned
    against the oracle / reference by the tests), plus `winner >= 0 == mask`."""
    import torch
    from syn3r_b200 import warp as w
    tw = share.plan
    g = {"frames_checked": [], "fused_mask_vs_or_of_pair_masks_xor": 0, "winner_vs_mask_xor": 0}
    for t in sorted({0, tw.T - 1}):
        p0, p1 = (int(v) for v in np.searchsorted(tw.pair_tgt_h, [t, t + 1]))
        o = w.forward_warp_batched(tw.rgbd, tw.Kinv_src, tw.pair_par[p0:p1].contiguous(), tw.pair_src[p0:p1].contiguous(),
                                   want_f32=False, want_u8=False, want_winner=False)
        union = o["mask"].bool().any(0)
        g["fused_mask_vs_or_of_pair_masks_xor"] += int((union ^ tw.out_fwd["mask"][t].bool()).sum().item())
        g["winner_vs_mask_xor"] += int(((tw.out_fwd["winner"][t] >= 0) ^ tw.out_fwd["mask"][t].bool()).sum().item())
        g["frames_checked"].append(int(t))
        del o
    torch.cuda.empty_cache()
    g["bars"] = "0 mismatches (the fused accumulator itself is oracle-defined: tests/test_gpu_parity.py, test_gpu_fullsize.py)"
    return g


def run_ours(args):
    import torch
    import torch.distributed as dist
    from syn3r_b200 import _native as nat, sharding, warp as w

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    assert torch.cuda.is_available(), "bench.py (impl=ours) needs a GPU; there is no CPU fallback"
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    nat.lib()

    wl = workload.for_config(args.config, args.mode)
    H, W, hw = wl.H, wl.W, wl.H * wl.W
    big = wl.lazy_sources                                         # config E: GBs of inputs / outputs per rank
    strong = world > 1 and not args.weak
    # strong scaling: the workload's output slots (pair warps in PAIR mode, frames in FUSED mode) in contiguous blocks, one
    # per rank; in PAIR mode the cut may fall inside a target frame (75 pair warps over 8 GPUs: 10,10,10,9,... instead of
    # 4,3,3,... frames).  (Measured, N = 8, config B: giving rank 0 - which also ingests every peer's stores - a LIGHT
    # share is slower, 0.503 vs 0.477 ms: a heavy rank's extra work is cheapest where its stores are local.)
    slot_parts = workload.partition_slots(wl, world) if strong else [(0, sum(workload.slots_per_frame(wl)))] * world
    parts, trims = [], []
    for q0, q1 in slot_parts:
        if q1 > q0:
            a, b, h, t = workload.frames_of_slots(wl, q0, q1)
        else:
            a, b, h, t = 0, 0, 0, 0
        parts.append((a, b))
        trims.append((h, t))
    f0, f1 = parts[rank]
    assert f1 > f0, "more ranks than output slots"
    packed = strong and args.wire == "compact" and hw % 16 == 0     # the 0/1 masks travel as bits (np.packbits order)
    share = workload.build_plan(wl, f0, f1, seed=(0 if (strong or world == 1) else rank), trim=trims[rank], want_f32=True,
                                want_u8=strong, want_winner=True, pack_masks=packed)
    tw = share.plan
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)

    def barrier():
        if world > 1:
            dist.barrier()

    def all_max(v):
        if world == 1:
            return float(v)
        t = torch.tensor([v], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

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

    # ---- N > 1: the gather of the finished frames on the diffusion rank is part of the step ---------------------
    gather_info, step, counts = None, tw.run, None
    if strong:
        counts = [q1 - q0 for q0, q1 in slot_parts]
        if args.wire == "post":
            tw._alloc_post()
            tw.post = True
            wire = {k: tw.out_post[k] for k in ("post_cond_u8", "post_latent", "post_latent_reproj", "post_soft_pooled")
                    if k in tw.out_post}
        else:
            wire = sharding.compact_wire(tw.out_fwd, tw.out_inv)
        per_row = sharding.wire_bytes(wire) // max(tw.N, 1)
        how = args.gather
        peer, peer_err = None, ""
        if how in ("auto", "peer"):
            ok = 1
            try:
                peer = sharding.PeerOutputs(wire, counts, dst=0)
            except Exception as e:  # noqa: BLE001
                ok, peer, peer_err = 0, None, f"{type(e).__name__}: {e}"
            t_ok = torch.tensor([ok], device=dev)
            dist.all_reduce(t_ok, op=dist.ReduceOp.MIN)
            if int(t_ok.item()) == 0:
                peer = None
                if how == "peer":
                    raise RuntimeError("symmetric memory unavailable: " + (peer_err or "on another rank"))
        local_outs = {}
        if peer is not None:
            # the plan's outputs that travel ARE rank 0's buffers: kernels store over NVLink while they compute
            for k, v in peer.local_views.items():
                for d in (tw.out_fwd, tw.out_inv, tw.out_post):
                    if k in d:
                        local_outs[k] = (d, d[k], v)
                        d[k] = v

            def step():
                tw.run()
                peer.barrier()
            gather_info = {"how": "direct peer stores: the resolve step of k_fwd_tile_fx and k_reproj_masks2 write into rank 0's "
                                  "symmetric buffers over NVLink (no send/recv, no staging copy); one barrier closes the step"}
            nbytes = peer.bytes_received(wire)
        else:
            pg = sharding.PipelinedGather(wire, counts, args.gather_chunks, dst=0)

            def step():
                pg.begin()
                for c in range(pg.n_chunks):
                    a, b = pg.bounds(c)
                    if b > a:
                        tw._run_range(a, b)
                    pg.push(c, wire)
                pg.end()
            gather_info = {"how": f"NCCL grouped send/recv (batch_isend_irecv) per slot range, {pg.n_chunks} ranges per rank, on a "
                                  "side stream behind an event recorded after the range's kernels",
                           "peer_store_unavailable": peer_err or None}
            nbytes = pg.bytes_received(wire)
        gather_info.update({"wire": sorted(wire.keys()), "wire_bytes_per_slot_pixel": per_row / hw, "bytes_into_rank0": nbytes})

    for _ in range(max(args.warmup, 3)):
        step()
    torch.cuda.synchronize()
    assert int(tw.out_fwd["status"].max().item()) == 0, "colour range check failed"

    # ---- device-resident throughput (`value`) ------------------------------------------------------
    with ClockSampler(local) as cs:
        ms = timed(step, args.steps, between=flush.zero_)
    clocks = cs.summary()
    total_ms = all_max(sum(ms))
    frames_per_step = wl.frames if (strong or world == 1) else world * wl.frames
    mpix_step = frames_per_step * hw / 1e6
    value = args.steps * mpix_step / (total_ms / 1e3)
    redo = w.redo_items(dev)
    if strong:
        # compute-only per rank (same plan, no gather / barrier): what the exchange adds is the difference
        for k, (d, loc, _) in local_outs.items():
            d[k] = loc                                               # peer mode: outputs in local memory for this leg
        comp = timed(tw.run, max(3, min(args.steps, 10)), between=flush.zero_)
        for k, (d, _, rem) in local_outs.items():
            d[k] = rem
        t_comp = torch.tensor([statistics.mean(comp)], device=dev, dtype=torch.float64)
        allc = [torch.zeros_like(t_comp) for _ in range(world)]
        dist.all_gather(allc, t_comp)
        gather_info.update({"compute_only_ms_per_rank": [float(t.item()) for t in allc],
                            "step_ms_with_gather": total_ms / args.steps,
                            "GBps_into_rank0_over_step": gather_info["bytes_into_rank0"] / (total_ms / args.steps) / 1e6})

    # ---- per-kernel durations (CUDA events on the launching stream, phases launched one at a time) ---
    def phase(bits):
        return lambda: tw._forward(phases=bits)

    kern = {}
    n_k = max(3, min(args.steps, 10))
    legs = [("fwd_prep(stats+bounds+max z')", phase(nat.FWD_PHASE_ZMAX)), ("fwd_tile(splat+resolve)", phase(nat.FWD_PHASE_SPLAT))]
    if wl.consistency:
        legs.append(("inverse_consistency", tw._inverse))
    for name, fn in legs:
        kern[name] = statistics.mean(timed(fn, n_k, between=flush.zero_))
    ksum = sum(kern.values())
    dominant = max(kern, key=kern.get)

    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    peak = float(peaks.get("hbm_gbs", 6650.0))
    b_step = tw.algorithmic_bytes()
    out_px = (12 if tw.want_f32 else 0) + (3 if tw.want_u8 else 0) + 1 + (4 if tw.want_winner else 0)
    # algorithmic bytes of ONE launch of each kernel on THIS rank (inputs once + outputs once, accumulators zero)
    b_kernel = {"fwd_prep(stats+bounds+max z')": 2 * tw.S * hw * 16,                  # sources read by stats and by max z'
                "fwd_tile(splat+resolve)": tw.S * hw * 16 + tw.N * hw * out_px,
                "inverse_consistency": tw.S * hw * 4 + tw.T * hw * 4 + tw.P * hw * 5}
    traffic = None
    try:    # dram__bytes_read.sum + dram__bytes_write.sum per launch, from the committed ncu launch list (config B, N = 1)
        if args.config == "B_svd" and world == 1 and wl.mode == "pair":
            traffic = json.load(open(os.path.join(ROOT, "profiles", "ncu_traffic.json"))).get(dominant)
    except Exception:
        pass
    ach_k = b_kernel[dominant] / (kern[dominant] / 1e3) / 1e9
    ach_s = b_step / (ksum / 1e3) / 1e9
    tiles = ((H + 47) // 48) * ((W + 127) // 128)
    roofline = {"bound": "hbm", "kernel": dominant, "achieved": ach_k, "peak": peak, "unit": "GB/s",
                "frac": ach_k / peak, "traffic": traffic,
                "peak_source": "MEASURED_PEAKS.json hbm_gbs (burst copy), of measured" if peaks else "fallback 6650 GB/s, of fallback",
                "algorithmic_bytes_per_launch": b_kernel[dominant],
                "step": {"achieved": ach_s, "frac": ach_s / peak, "algorithmic_bytes_per_step": b_step,
                         "scope": "all kernels of one step on rank 0's share: SURVEY 8d B_alg / their summed CUDA-event time"},
                "kernel_ms": kern, "kernel_share": {k: v / ksum for k, v in kern.items()},
                "redo_items": redo, "work_items": tw.N * tiles,
                "note": "instruction-issue bound, not HBM bound: see DESIGN.md 'Roofline'; redo_items = (slot, tile) work items "
                        "the integer tile kernel handed to the fp32 kernel (crowded cells), of work_items"}

    post = None
    if world == 1 and not args.no_post and not big and H % 8 == 0 and W % 8 == 0 and wl.consistency:
        post = post_kernels(args, tw, wl, timed, flush, peak, dev, n_k)

    # ---- end to end with HOST buffers ---------------------------------------------------------------
    e2e, e2e_all, e2e_post, e2e_dropin = None, None, None, None
    if not args.no_e2e:
        if world == 1:
            e2e, e2e_all, e2e_post, e2e_dropin = e2e_single(args, wl, share, timed, mpix_step, big)
        elif strong:
            e2e = e2e_strong(args, wl, share, parts, trims, slot_parts, timed, mpix_step, dev, rank, world, all_max)

    cpu, parity = None, None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        cpu = cpu_baseline_sample(wl, args.cpu_sample_seconds)
        tw.run()
        torch.cuda.synchronize()
        parity = parity_gates(wl, share)

    if rank == 0:
        extra = {}
        if world > 1:
            extra["partition"] = ({"slots_per_rank": counts, "frames_touched_per_rank": [b - a for a, b in parts]} if strong
                                  else "one whole workload per rank (seed = rank), no collective")
        line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
                "warmup": max(args.warmup, 3), "ms_per_step": total_ms / args.steps, "higher_is_better": True,
                "scaling": "strong" if (strong or world == 1) else "weak", "vs_baseline": None,
                "dtype": "f64 projection / block-floating-point i32 accumulate / f32 consistency", "data": "synthetic",
                "config": config_desc(args, wl, extra), "clocks": clocks, "e2e": e2e,
                "gpu_launches": args.steps * tw.launches_per_run(), "roofline": roofline, "cpu_baseline": cpu,
                "pair_mpix_per_s": args.steps * (wl.pairs if (strong or world == 1) else world * wl.pairs) * hw / 1e6 / (total_ms / 1e3),
                "parity": parity}
        for k, v in (("e2e_all_outputs", e2e_all), ("e2e_conditioning", e2e_post), ("e2e_dropin", e2e_dropin),
                     ("gather", gather_info), ("post", post)):
            if v is not None:
                line[k] = v
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


def e2e_single(args, wl, share, timed, mpix_step, big):
    """N = 1, HOST buffers, everything inside the timed region (pinned H2D of the step's inputs, kernels in slot ranges,
    D2H of each range overlapping the next range's kernels)."""
    import torch

    def measure(plan, rgbd_h, tgt_h, how):
        def step_e2e():
            plan.step_host(rgbd_h, tgt_h, chunks=args.e2e_chunks, ramp=False if args.e2e_equal_ranges else None)
        for _ in range(2):
            step_e2e()
        ms_e = timed(step_e2e, max(3, args.steps // (4 if big else 1)))
        tot = sum(ms_e) / len(ms_e)
        return {"value": mpix_step / (tot / 1e3), "unit": UNIT, "h2d_bytes_per_step": plan.h2d_bytes(),
                "d2h_bytes_per_step": plan.d2h_bytes(), "ms_per_step": tot, "host_outputs": list(plan.host_outputs), "how": how}

    def plan_like(**kw):
        sh = workload.build_plan(wl, 0, wl.frames, seed=0, pinned_host=True, **kw)
        return sh.plan, sh.host_rgbd, sh.host_tgt

    e2e_all = None
    packed = (wl.H * wl.W) % 16 == 0
    how = (f"TrajectoryWarp.step_host: pinned H2D of RGB-D (+ target depth) on an upload stream, kernels in {args.e2e_chunks} "
           + ("equal " if args.e2e_equal_ranges else "doubling (equal for the post=True plan) ") +
           "slot ranges (uint8 frame + hole mask + winner" + (" + consistency masks" if wl.consistency else "") + "), D2H of "
           "the outputs the reference's forward_warp / inverse_warp return to their callers, overlapping the next range")
    if packed:
        ref_keys = ("rgb_u8", "mask_bits") + (("mask_reproj_bits", "soft_mask_reproj") if wl.consistency else ())
        p_ref, r_h, t_h = plan_like(want_f32=False, want_u8=True, want_winner=True, host_outputs=ref_keys, pack_masks=True)
        e2e = measure(p_ref, r_h, t_h, how + "; the two 0/1 masks cross PCIe as bits (np.packbits order, "
                                             "conditioning.unpack_mask_bits on the host)")
        assert int(p_ref.out_fwd["status"].max().item()) == 0
        del p_ref, r_h, t_h
        torch.cuda.empty_cache()
    ref_keys = ("rgb_u8", "mask") + (("mask_reproj", "soft_mask_reproj") if wl.consistency else ())
    p_ref, r_h, t_h = plan_like(want_f32=False, want_u8=True, want_winner=True, host_outputs=ref_keys)
    e2e_bytes = measure(p_ref, r_h, t_h, how + "; masks as one byte per pixel")
    assert int(p_ref.out_fwd["status"].max().item()) == 0
    del p_ref, r_h, t_h
    torch.cuda.empty_cache()
    if packed:
        e2e["byte_masks"] = {k: e2e_bytes[k] for k in ("value", "ms_per_step", "d2h_bytes_per_step", "host_outputs")}
    else:
        e2e = e2e_bytes
    e2e_post, e2e_dropin = None, None
    if wl.H % 8 == 0 and wl.W % 8 == 0 and not big:
        keys = ("post_cond_u8", "post_latent") + (("post_latent_reproj", "post_soft_pooled") if wl.consistency else ())
        p_post, r_h, t_h = plan_like(want_f32=False, want_u8=True, want_winner=True, post=True, host_outputs=keys)
        e2e_post = measure(p_post, r_h, t_h, "as `e2e`, plus k_condition_post on every slot range; D2H of what the SVD pipeline "
                                              "consumes: the blanked uint8 frame, the latent masks, the pooled uncertainty "
                                              "(diffusionGS.py:1572-1597, :1452-1491 on the device)")
        del p_post, r_h, t_h
        torch.cuda.empty_cache()
    if not big and wl.mode == "pair":
        e2e_dropin = dropin_e2e(wl, share, mpix_step)
    return e2e, e2e_all, e2e_post, e2e_dropin


def dropin_e2e(wl, share, mpix_step, max_pairs=24):
    """The cost of NOT editing the caller (zero-edit switch): the per-pair mirror functions
    `solver_utils.forward_warp` (+ `inverse_warp`) called in a Python loop with NumPy / torch host arrays in and out,
    one pair per call like diffusionGS.py:1554, each call uploading pageable memory and synchronising.  Wall clock
    over a bounded number of pairs, scaled to the whole step."""
    import torch
    from syn3r_b200.solver_utils import forward_warp as fwm
    tr = share.trajs[0]
    pairs = [(s, t) for (k, s, t) in share.pair_global if k == share.traj_ids[0]][:max_pairs]
    t_ = torch.from_numpy
    Kf = t_(tr.K.astype(np.float32))

    def one(s, t):
        fwm.forward_warp(tr.rgb[s], None, tr.depth[s], tr.pose_src[s], tr.pose_tgt[t], tr.K, None)
        if wl.consistency:
            d = fwm.inverse_warp(t_(np.ascontiguousarray(np.moveaxis(tr.rgb[s], -1, 0))), t_(tr.depth[s][None]),
                                 t_(tr.depth_tgt[t][None]), t_(tr.pose_src[s]), t_(tr.pose_tgt[t]), Kf)
            d["mask_reproj"].cpu(), d["soft_mask_reproj"].cpu()                      # what the callers read back (:1444-1445)
    one(*pairs[0])
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for s, t in pairs:
        one(s, t)
    torch.cuda.synchronize()
    el = time.perf_counter() - t0
    ms_step = el / len(pairs) * wl.pairs * 1e3
    return {"value": mpix_step / (ms_step / 1e3), "unit": UNIT, "ms_per_step": ms_step, "pairs_timed": len(pairs),
            "how": "per-pair mirror solver_utils.forward_warp" + (" + inverse_warp" if wl.consistency else "") +
                   " in a Python loop, NumPy / torch HOST arrays in and out (pageable H2D, one synchronisation per call, the "
                   "flow12 / 11-key dict the reference returns): wall clock, scaled from the timed pairs to the whole step"}


def e2e_strong(args, wl, share, parts, trims, slot_parts, timed, mpix_step, dev, rank, world, all_max):
    """N > 1, ONE consumer: host inputs on rank 0 -> H2D -> broadcast over NVLink -> partitioned warps -> gather of the
    reference-facing outputs on rank 0 -> D2H on rank 0 only."""
    import torch
    import torch.distributed as dist
    from syn3r_b200 import sharding
    S_all = sum(s for s, _ in wl.trajs)
    if len(wl.trajs) > 1 or wl.lazy_sources:
        return {"value": None, "unit": UNIT, "how": "not measured for multi-trajectory / 4K workloads at N > 1 (the broadcast would "
                                                    "ship every trajectory's sources to every rank); see `value` and `gather`"}
    full_rgbd = torch.empty((S_all, wl.H, wl.W, 4), dtype=torch.float32, device=dev)
    full_tgt = torch.empty((wl.frames, wl.H, wl.W), dtype=torch.float32, device=dev) if wl.consistency else None
    host_rgbd = host_tgt = None
    if rank == 0:
        sh0 = workload.build_plan(wl, 0, wl.frames, seed=0, pinned_host=True, want_f32=False, want_u8=False, want_winner=False)
        host_rgbd, host_tgt = sh0.host_rgbd, sh0.host_tgt
        del sh0
    f0, f1 = parts[rank]
    # its own plan with local outputs (the `value` plan may be writing into rank 0's peer buffers)
    tw = workload.build_plan(wl, f0, f1, seed=0, trim=trims[rank], want_f32=False, want_u8=True, want_winner=True,
                             pack_masks=(wl.H * wl.W) % 16 == 0).plan
    local_wire = sharding.compact_wire(tw.out_fwd, tw.out_inv)
    counts = [q1 - q0 for q0, q1 in slot_parts]
    pg = sharding.PipelinedGather(local_wire, counts, args.gather_chunks, dst=0)
    pinned = None
    if rank == 0:
        pinned = {k: torch.empty(tuple(v.shape), dtype=v.dtype, pin_memory=True) for k, v in pg.buffers.items()}

    def step():
        if rank == 0:
            full_rgbd.copy_(host_rgbd, non_blocking=True)
            if full_tgt is not None:
                full_tgt.copy_(host_tgt, non_blocking=True)
        dist.broadcast(full_rgbd, 0)
        if full_tgt is not None:
            dist.broadcast(full_tgt, 0)
        tw.rgbd.copy_(full_rgbd)
        if full_tgt is not None:
            tw.tgt_depth.copy_(full_tgt[f0:f1])
        pg.begin()
        for c in range(pg.n_chunks):
            a, b = pg.bounds(c)
            if b > a:
                tw._run_range(a, b)
            pg.push(c, local_wire)
        bufs = pg.end()
        if rank == 0:
            for k, v in bufs.items():
                pinned[k].copy_(v, non_blocking=True)
    for _ in range(2):
        step()
    ms_e = timed(step, max(3, args.steps))
    tot = all_max(sum(ms_e) / len(ms_e))
    h2d = full_rgbd.numel() * 4 + (full_tgt.numel() * 4 if full_tgt is not None else 0)
    d2h = sum(v.numel() * v.element_size() for v in pg.buffers.values()) if rank == 0 else 0
    return {"value": mpix_step / (tot / 1e3), "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h, "ms_per_step": tot,
            "how": "ONE consumer: pinned H2D on rank 0, NCCL broadcast of the inputs, partitioned warps, gather of the "
                   "reference-facing outputs on rank 0, D2H on rank 0 only (bound by rank 0's PCIe link, not by the GPUs)"}


def post_kernels(args, tw, wl, timed, flush, peak, dev, n_k):
    """the callers' per-frame post-processing and its neighbours, one launch each (HBM-bound byte work)."""
    import torch
    from syn3r_b200 import _native as nat, conditioning, latent as latent_mod
    H, W, hw, T = wl.H, wl.W, wl.H * wl.W, wl.trajs[0][1]
    P_ = tw.P
    m_fw, m_bw = tw.out_fwd["mask"], tw.out_inv["mask_reproj"]
    img8 = torch.randint(0, 256, (P_, H, W, 3), dtype=torch.uint8, device=dev)
    o_fw, o_bw = {}, {}

    def post_fw():
        o_fw.update(conditioning.condition_post_batched(m_fw, img_u8=img8, want=("mask_png", "cond_u8", "latent"), out=o_fw))

    def post_bw():
        o_bw.update(conditioning.condition_post_batched(m_bw, soft=tw.out_inv["soft_mask_reproj"],
                                                        want=("latent", "soft_pooled"), out=o_bw))
    post_fw(), post_bw()
    t_fw = statistics.mean(timed(post_fw, n_k, between=flush.zero_))
    t_bw = statistics.mean(timed(post_bw, n_k, between=flush.zero_))
    b_fw = P_ * hw * 8 + P_ * (hw // 64)                  # valid 1 + frame 3 in; png 1 + blanked frame 3 + latent out
    b_bw = P_ * hw * 5 + P_ * (hw // 64) * 5              # valid 1 + soft 4 in; latent 1 + pooled 4 per 8x8 block out
    del img8
    # uncertainty fusion of the T - 2 interior frames (diffusionGS.py:821-862), 28 B in + 12 B out
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
    # second half of rank 2: nearby-frame consistency of the T interpolated frames (diffusionGS.py:1300-1361)
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
    # rank 4, scheduler half: guided latent replacement of one diffusion step (scheduling_euler_discrete.py:1442-1499)
    Fl, hl, wl_ = 25, H // 8, W // 8
    p_l = torch.randn((1, Fl, 4, hl, wl_), device=dev)
    c_l = torch.randn((2, Fl, 4, hl, wl_), device=dev)
    m_l = (torch.rand((1, Fl - 2, 1, hl, wl_), device=dev) < 0.3).float()
    lam_l = torch.rand((100, Fl), device=dev)
    cond_l, row_l, REP = c_l[1].contiguous(), lam_l[7].contiguous(), 10

    def latent_step():      # the C-ABI call itself, REP diffusion steps back to back (a 50 us kernel: one Python-level
        for _ in range(REP):  # call between two events would time the host, not the device)
            nat.check(nat.lib().syn3r_latent_guided_replace(nat.ptr(p_l), nat.ptr(cond_l), nat.ptr(m_l), nat.ptr(row_l), 1, Fl, 4,
                                                            hl, wl_, 1, Fl - 2, 1, 1, Fl - 2, None, None,
                                                            torch.cuda.current_stream().cuda_stream), "latent")
    latent_mod.guided_replace_(p_l, c_l, m_l, lam_l, 7)
    latent_step()
    t_lat = statistics.mean(timed(latent_step, n_k, between=flush.zero_)) / REP
    b_lat = (Fl - 2) * 4 * hl * wl_ * 4 * 3 + (Fl - 2) * hl * wl_ * 4 + 2 * 4 * hl * wl_ * 8

    def gb(b, t):
        return b / t / 1e6
    return {"latent_replace": {"kernel": "k_latent_guided_replace", "frames": Fl, "latent": [4, hl, wl_], "ms": t_lat,
                               "algorithmic_bytes": b_lat, "achieved_GBps": gb(b_lat, t_lat), "frac": gb(b_lat, t_lat) / peak,
                               "note": "one CTA per frame: latency bound by design; mean of 10 back-to-back steps"},
            "flow_fb": {"kernel": "k_flow_fb_consistency", "frame_pairs": Bf, "ms": t_fl, "algorithmic_bytes": b_fl_bytes,
                        "achieved_GBps": gb(b_fl_bytes, t_fl), "frac": gb(b_fl_bytes, t_fl) / peak},
            "nearby": {"kernel": "k_nearby_uncertainty", "frames": Nn, "pairs": Pn, "ms": t_nb, "algorithmic_bytes": b_nb,
                       "achieved_GBps": gb(b_nb, t_nb), "frac": gb(b_nb, t_nb) / peak},
            "fusion": {"kernel": "k_fuse_uncertainty", "frames": Nf, "ms": t_fu, "algorithmic_bytes": b_fu,
                       "achieved_GBps": gb(b_fu, t_fu), "frac": gb(b_fu, t_fu) / peak},
            "kernel": "k_condition_post", "forward_branch_ms": t_fw, "backward_branch_ms": t_bw,
            "forward_branch": {"algorithmic_bytes": b_fw, "achieved_GBps": gb(b_fw, t_fw), "frac": gb(b_fw, t_fw) / peak},
            "backward_branch": {"algorithmic_bytes": b_bw, "achieved_GBps": gb(b_bw, t_bw), "frac": gb(b_bw, t_bw) / peak},
            "scope": "diffusionGS.py:1572-1597 / :1452-1491 for all pairs; not part of `value` (SURVEY 8f)"}


if __name__ == "__main__":
    a = parse_args()
    if a.impl == "reference":
        run_reference(a)
    else:
        run_ours(a)
