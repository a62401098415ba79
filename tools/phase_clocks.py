#!/usr/bin/env python
"""Where a work item of k_fwd_tile_fx spends its cycles (A/B build with -DSYN3R_FX_PHASE_CLOCKS=1, selected through
SYN3R_WARP_LIB): claim / cull / sweep 1 / its tail at the barrier / sweep 2 / its tail / resolve, summed per CTA by
thread 0 with clock().  Usage: SYN3R_WARP_LIB=.../libsyn3r_warp_phase.so tools/phase_clocks.py [config]"""
import ctypes
import json
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from syn3r_b200 import _native as nat, synth, trajectory, warp as w  # noqa: E402

cfg = sys.argv[1] if len(sys.argv) > 1 else "B_svd"
S, T, H, W = synth.CONFIGS[cfg]
tr = synth.make_trajectory(S, T, H, W, seed=0)
tw = trajectory.from_synth(tr, want_f32=True, want_u8=False, want_winner=True)
tw.upload(torch.from_numpy(w.pack_rgbd(tr.rgb, tr.depth)), torch.from_numpy(tr.depth_tgt))
for _ in range(3):
    tw.run()
torch.cuda.synchronize()
n = nat.device_info()["sm_count"]
buf = np.zeros((n, 8), dtype=np.uint64)
fn = nat.lib().syn3r_debug_fx_phase
fn.restype, fn.argtypes = ctypes.c_int, [ctypes.c_void_p, ctypes.c_int]
nat.check(fn(buf.ctypes.data, n), "syn3r_debug_fx_phase")
names = ["claim", "cull", "sweep1", "sweep1_tail", "sweep2", "sweep2_tail", "resolve"]
tot = buf[:, :7].sum(axis=1).astype(np.float64)
items = buf[:, 7].astype(np.float64)
res = {"config": cfg, "ctas": int(n), "items": int(items.sum()), "cycles_per_cta": float(tot.mean()),
       "share": {k: round(float(buf[:, i].sum() / tot.sum()), 4) for i, k in enumerate(names)},
       "cycles_per_item": {k: round(float(buf[:, i].sum() / items.sum()), 1) for i, k in enumerate(names)}}
print(json.dumps(res))
