import sys, os, json, statistics, torch
sys.path.insert(0, '/root/repo')
from syn3r_b200 import _native as nat, synth, trajectory, warp as w
cfg = sys.argv[1] if len(sys.argv) > 1 else "B_svd"
S, T, H, W = synth.CONFIGS[cfg]
tr = synth.make_trajectory(S, T, H, W, seed=0)
tw = trajectory.from_synth(tr, want_f32=True, want_u8=False, want_winner=True)
tw.upload(torch.from_numpy(w.pack_rgbd(tr.rgb, tr.depth)), torch.from_numpy(tr.depth_tgt))
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
L = nat.lib()
def timed(fn, n=10):
    out = []
    for _ in range(3): fn()
    for _ in range(n):
        flush.zero_()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record(); torch.cuda.synchronize(); out.append(a.elapsed_time(b))
    return round(statistics.mean(out), 4)
res = {"config": cfg}
for packed in (1, 0):
    L.syn3r_set_reproj_packed(packed)
    res[f"consistency_ms_packed{packed}"] = timed(tw._inverse)
    res[f"step_ms_packed{packed}"] = timed(tw.run)
    o = tw.run(); torch.cuda.synchronize()
    res[f"check{packed}"] = [int(o.mask_reproj.sum().item()), float(o.soft_mask_reproj.double().sum().item())]
L.syn3r_set_reproj_packed(1)
print(json.dumps(res))
