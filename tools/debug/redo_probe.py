"""Debug aid: counters of the forward workspace after a crowded call (work counter, cmax bits, redo count)."""
import sys, os, numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, 'tests'))
from syn3r_b200 import synth, warp as w
import test_gpu_parity as tp
for pb, fused in ((9.0, True), (9.0, False)):
    S, T, H, W = 16, 1, 144, 256
    tr = synth.make_trajectory(S, T, H, W, seed=61, pose_step=0.25)
    tr.pose_tgt[:, 2, 3] += pb
    ps, pt = synth.all_pairs(S, T)
    o = tp._run_pairs(tr, ps, pt, fused=fused)
    ws, P = w._last_fwd[str(torch.device('cuda', 0))] if str(torch.device('cuda', 0)) in w._last_fwd else list(w._last_fwd.values())[0]
    print(pb, fused, 'keys', list(w._last_fwd.keys()), 'P', P, 'redo', w.redo_items(torch.device('cuda')))
    up = lambda v: (v + 255) // 256 * 256
    off = up(P * 8) + up(P * 4)
    print(ws[off:off + 256].view(torch.int32).cpu().numpy().reshape(4, 16))
