#!/usr/bin/env python
"""Time k_flow_fb_consistency alone at the bench shape (A/B of build variants through SYN3R_WARP_LIB)."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
from syn3r_b200 import _native as nat  # noqa: E402

B, H, W = 24, 576, 1024
dev = "cuda"
yy, xx = torch.meshgrid(torch.arange(H, device=dev, dtype=torch.float32), torch.arange(W, device=dev, dtype=torch.float32),
                        indexing="ij")
f = torch.stack([2.5 + 3.0 * torch.sin(0.02 * yy) + 0.01 * xx, -1.75 + 3.0 * torch.cos(0.015 * xx) - 0.01 * yy])
f = (f[None] + torch.randn((B, 2, H, W), device=dev) * 0.3).contiguous()
b = -f + torch.randn((B, 2, H, W), device=dev) * 0.3
o = [torch.empty((B, H, W), device=dev) for _ in range(2)]
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
L = nat.lib()
ts = []
for it in range(25):
    flush.zero_()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    nat.check(L.syn3r_flow_fb_consistency_batched(nat.ptr(f), nat.ptr(b), B, H, W, 0.01, 0.5, nat.ptr(o[0]), nat.ptr(o[1]), None, None,
                                                  torch.cuda.current_stream().cuda_stream), "flow")
    e1.record()
    torch.cuda.synchronize()
    if it >= 5:
        ts.append(e0.elapsed_time(e1))
print(os.environ.get("SYN3R_WARP_LIB", "default"), "ms", sum(ts) / len(ts), "min", min(ts))
