#!/usr/bin/env python
"""One profiled launch of each branch of the post-processing kernel at the headline shape (75 x 576 x 1024) for ncu
(`--profile-from-start off`)."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from syn3r_b200 import conditioning as c  # noqa: E402

P, H, W = 75, 576, 1024
g = torch.Generator(device="cuda").manual_seed(0)
valid = (torch.rand((P, H, W), device="cuda", generator=g) > 0.1).to(torch.uint8)
img = torch.randint(0, 256, (P, H, W, 3), dtype=torch.uint8, device="cuda", generator=g)
soft = torch.rand((P, H, W), device="cuda", generator=g)
o1, o2 = {}, {}
for _ in range(2):
    o1 = c.condition_post_batched(valid, img_u8=img, want=("mask_png", "cond_u8", "latent"), out=o1)
    o2 = c.condition_post_batched(valid, soft=soft, want=("latent", "soft_pooled"), out=o2)
torch.cuda.synchronize()
torch.cuda.profiler.start()
c.condition_post_batched(valid, img_u8=img, want=("mask_png", "cond_u8", "latent"), out=o1)
c.condition_post_batched(valid, soft=soft, want=("latent", "soft_pooled"), out=o2)
torch.cuda.synchronize()
torch.cuda.profiler.stop()
print("profiled k_condition_post: forward branch, backward branch")
