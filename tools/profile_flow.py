#!/usr/bin/env python
"""One profiled launch of the flow-consistency and nearby-uncertainty kernels at the bench shapes for ncu
(`--profile-from-start off`)."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from syn3r_b200 import conditioning as c, flow  # noqa: E402

B, H, W = 24, 576, 1024
dev = "cuda"
yy, xx = torch.meshgrid(torch.arange(H, device=dev, dtype=torch.float32), torch.arange(W, device=dev, dtype=torch.float32),
                        indexing="ij")
f = torch.stack([2.5 + 3.0 * torch.sin(0.02 * yy) + 0.01 * xx, -1.75 + 3.0 * torch.cos(0.015 * xx) - 0.01 * yy])
f = (f[None] + torch.randn((B, 2, H, W), device=dev) * 0.3).contiguous()
b = -f + torch.randn((B, 2, H, W), device=dev) * 0.3
N = 25
ps, pt, first = c.nearby_pairs(N, 1)
soft = torch.rand((len(ps), H, W), device=dev)
warp = torch.rand((len(ps), 3, H, W), device=dev) * 255
img = torch.rand((N, 3, H, W), device=dev) * 255
fd = torch.from_numpy(first).to(dev)
for _ in range(2):
    flow.forward_backward_consistency_check(f, b)
    c.nearby_uncertainty_batched(soft, warp, img, fd)
torch.cuda.synchronize()
torch.cuda.profiler.start()
flow.forward_backward_consistency_check(f, b)
c.nearby_uncertainty_batched(soft, warp, img, fd)
torch.cuda.synchronize()
torch.cuda.profiler.stop()
print("profiled k_flow_fb_consistency, k_nearby_uncertainty")
