import sys, numpy as np, torch
sys.path.insert(0, '/root/repo')
from syn3r_b200 import _native as nat, synth, warp as w
H, W, S, T = 61, 97, 2, 3
tr = synth.make_trajectory(S, T, H, W, seed=71 + H, pose_step=2.0)
depth, depth_tgt = tr.depth.copy(), tr.depth_tgt.copy()
rng = np.random.default_rng(5)
for d in (depth, depth_tgt):
    flat = d.reshape(-1)
    idx = rng.choice(flat.size, size=max(8, flat.size // 6), replace=False)
    flat[idx] = rng.choice(np.array([0.0, -0.0, -1.5, np.nan, np.inf, -np.inf, 1e-42, 1e-30, 1e30, 3e38, 1e-12], np.float32), size=idx.size)
ps, pt = synth.all_pairs(S, T)
par = w.inverse_pair_params(tr.pose_src, tr.pose_tgt, ps, pt)
up = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()
want = ("mask_reproj", "soft_mask_reproj", "mask_warp")
L = nat.lib(); res = {}
for packed in (1, 0):
    L.syn3r_set_reproj_packed(packed)
    res[packed] = {k: v.cpu().numpy() for k, v in w.inverse_warp_batched(None, up(depth), up(depth_tgt), tr.K.astype(np.float32), up(par), up(ps), up(pt), bandwidth=3.0, want=want).items()}
for k in want:
    a, b = res[1][k], res[0][k]
    if a.dtype == np.float32:
        bad = a.view(np.uint32) != b.view(np.uint32)
    else:
        bad = a != b
    print(k, int(bad.sum()))
    for (p, y, x) in np.argwhere(bad)[:12]:
        print("  pair", p, "src", ps[p], "tgt", pt[p], "y", y, "x", x, "packed", a[p, y, x], "scalar", b[p, y, x], "z", depth_tgt[pt[p], y, x],
              "z_other_row", depth_tgt[pt[p], y ^ 1 if (y ^ 1) < H else y, x])
