#!/usr/bin/env python
"""Per-kernel CUDA-event timings of one workload for A/B builds (SYN3R_WARP_LIB selects the library).
Usage: tools/time_kernels.py [config] [reps]   ->  one JSON line: prep / tile / consistency / step in ms (mean of reps,
256 MiB L2 flush between timed launches) and, with SYN3R_TIME_CHECK=1, a checksum of the outputs (to tell builds that
change results from builds that do not)."""
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
tw = trajectory.from_synth(tr, want_f32=True, want_u8=False, want_winner=os.environ.get("SYN3R_TIME_NO_WINNER") != "1")
tw.upload(torch.from_numpy(w.pack_rgbd(tr.rgb, tr.depth)), torch.from_numpy(tr.depth_tgt))
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")


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
    return statistics.mean(out), min(out)


res = {"lib": os.path.basename(nat.LIB_PATH), "config": cfg}
for name, fn in (("prep", lambda: tw._forward(phases=nat.FWD_PHASE_ZMAX)), ("tile", lambda: tw._forward(phases=nat.FWD_PHASE_SPLAT)),
                 ("consistency", tw._inverse), ("step", tw.run)):
    m, lo = timed(fn, reps)
    res[name + "_ms"] = round(m, 4)
    res[name + "_min_ms"] = round(lo, 4)
if os.environ.get("SYN3R_TIME_CHECK") == "1":
    o = tw.run()
    torch.cuda.synchronize()
    res["check"] = {"mask_sum": int(o.mask.sum().item()), "rgb_sum": float(o.rgb_f32.double().sum().item()),
                    "winner_sum": int(o.winner.long().sum().item()) if o.winner is not None else None, "reproj_sum": int(o.mask_reproj.sum().item()),
                    "soft_sum": float(o.soft_mask_reproj.double().sum().item()), "redo": w.redo_items(o.mask.device)}
print(json.dumps(res))
