#!/usr/bin/env python
"""One profiled step of the headline workload for ncu (`--profile-from-start off`): two warm-up steps, then
cudaProfilerStart / one TrajectoryWarp.run() / cudaProfilerStop.  Usage: tools/profile_step.py [config] [pairs]"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from syn3r_b200 import synth, trajectory, warp as w  # noqa: E402

cfg = sys.argv[1] if len(sys.argv) > 1 else "B_svd"
S, T, H, W = synth.CONFIGS[cfg]
if len(sys.argv) > 2:
    T = int(sys.argv[2])
tr = synth.make_trajectory(S, T, H, W, seed=0)
post = os.environ.get("SYN3R_PROFILE_POST", "0") == "1"     # also the callers' post-processing (k_condition_post)
tw = trajectory.from_synth(tr, want_f32=True, want_u8=post, want_winner=True, post=post)
tw.upload(torch.from_numpy(w.pack_rgbd(tr.rgb, tr.depth)), torch.from_numpy(tr.depth_tgt))
for _ in range(2):
    tw.run()
torch.cuda.synchronize()
torch.cuda.profiler.start()
tw.run()
torch.cuda.synchronize()
torch.cuda.profiler.stop()
print("profiled one step:", cfg, "pairs", tw.P, "status", int(tw.out_fwd["status"].max().item()))
