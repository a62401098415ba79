#!/usr/bin/env python
"""A/B of the forward tile kernel: k_fwd_tile_fx_ov (syn3r_set_fx_overlap(1), default) against k_fwd_tile_fx (0), per-kernel
CUDA-event times with a flushed L2 and output checksums.  Usage: tools/ab_fx_overlap.py [config] [reps]"""
import json
import os
import statistics
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from syn3r_b200 import _native as nat, synth, trajectory, warp as w  # noqa: E402

cfg = sys.argv[1] if len(sys.argv) > 1 else "B_svd"
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 10
S, T, H, W = synth.CONFIGS[cfg]
tr = synth.make_trajectory(S, T, H, W, seed=0)
tw = trajectory.from_synth(tr, want_f32=True, want_u8=False, want_winner=True)
tw.upload(torch.from_numpy(w.pack_rgbd(tr.rgb, tr.depth)), torch.from_numpy(tr.depth_tgt))
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
L = nat.lib()


def timed(fn, n):
    out = []
    for _ in range(3):
        fn()
    for _ in range(n):
        flush.zero_()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        fn()
        b.record()
        torch.cuda.synchronize()
        out.append(a.elapsed_time(b))
    return round(statistics.mean(out), 4)


res = {"config": cfg}
for ov in (1, 0, 1):
    L.syn3r_set_fx_overlap(ov)
    t = timed(lambda: tw._forward(phases=nat.FWD_PHASE_SPLAT), reps)
    st = timed(tw.run, reps)
    o = tw.run()
    torch.cuda.synchronize()
    res.setdefault(f"overlap{ov}", []).append({"tile_ms": t, "step_ms": st, "check": [int(o.mask.sum().item()), float(o.rgb_f32.double().sum().item()),
                                                                                    int(o.winner.long().sum().item()), w.redo_items(o.mask.device)]})
L.syn3r_set_fx_overlap(1)
print(json.dumps(res))
