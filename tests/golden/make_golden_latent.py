"""Golden fixtures of the scheduler's guided latent replacement (scheduler half):
tests/golden/latent_*.npz.

The patched scheduler (/root/reference/thirdparty/diffusers/src/diffusers/schedulers/scheduling_euler_discrete.py)
cannot be imported here: the vendored diffusers needs `huggingface_hub.cached_download`, which the installed hub no
longer has.  The fixture generator therefore RE-EXECUTES the statements of the replacement block (:1442-1499, the
soft-blend form of `step_interp_prob_uncertain`) and of the selection-only form of `step_single` (:548-589; `step_interp`
:735-778 differs only in its index list) with torch on the CPU, in the reference's order; `self.step_index` becomes an
argument and the hard-coded `index_list = [1..23]` becomes `range(1, F - 1)` (identical for the production F = 25)
so that short clips can be stored.  Run in the build container:

    python tests/golden/make_golden_latent.py
"""
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))


def reference_replace(pred_original_sample, temp_cond_latents, mask, lambda_ts, step_index):
    """scheduling_euler_discrete.py:1442-1499, statements in the reference's order."""
    b, cond_len, c, h, w = temp_cond_latents.shape
    index_list = list(range(1, pred_original_sample.shape[1] - 1))
    mask = (1 - mask) > 0.5
    mask_ones = torch.ones_like(mask[:, 0:1, :, :, :])
    mask = torch.cat((mask_ones, mask), 1)
    aux = []
    for ii, tau in enumerate(index_list):
        FEA = temp_cond_latents[1:2, tau, :, :, :]
        weight = lambda_ts[step_index, tau]
        mask_t_ = torch.mean(mask[:, tau, :, :, :].float(), 1, keepdim=True)
        mask_t = mask_t_ > 0.5
        mask_zero = ~mask_t
        num_zero = torch.sum(mask_zero)
        masked_tensor1 = pred_original_sample[:, tau, :, :, :] * mask_t
        masked_tensor2 = FEA * mask_t
        masked_diff = masked_tensor1 - masked_tensor2
        flat_diff = torch.abs(masked_diff.flatten())
        sorted_diff, indices = torch.sort(flat_diff)
        weight = torch.clamp(weight, 0.4, 1)
        cutoff_index_s = num_zero - 1  # noqa: F841
        cutoff_index_e = int(weight * (len(sorted_diff) - num_zero)) + num_zero
        cutoff_value = sorted_diff[cutoff_index_e - 1]
        top_mask = (torch.abs(masked_diff) <= cutoff_value) & mask_t  # noqa: F841
        weight_func = lambda x: x / (1 + x)  # noqa: E731
        w = weight_func(1 / (1.0 - mask_t_.to(torch.float32) + 1e-6))
        w = torch.where(w >= 0.51, w, torch.zeros_like(w))
        w = (torch.abs(masked_diff) <= cutoff_value).float() * w
        pred_original_sample[:, tau, :, :, :] = (1 - w) * pred_original_sample[:, tau, :, :, :] + w * FEA
        aux.append((int(cutoff_index_e), float(cutoff_value)))
    pred_original_sample[:, 0:1, :, :, :] = temp_cond_latents[1:2, 0:1, :, :, :]
    pred_original_sample[:, -1:, :, :, :] = temp_cond_latents[1:2, -1:, :, :, :]
    return pred_original_sample, aux


def reference_top_masks(pred_original_sample, temp_cond_latents, mask, lambda_ts, step_index, weight_clamp):
    """scheduling_euler_discrete.py:548-589 (`step_single`), statements in the reference's order; index_list
    [1..24] of the production F = 25 becomes range(1, F)."""
    b, cond_len, c, h, w = temp_cond_latents.shape
    index_list = list(range(1, pred_original_sample.shape[1]))
    mask = (1 - mask) > 0.5
    mask_ones = torch.ones_like(mask[:, 0:1, :, :, :])
    mask = torch.cat((mask_ones, mask), 1)
    top_masks = []
    aux = []
    for ii, tau in enumerate(index_list):
        FEA = temp_cond_latents[1:2, tau, :, :, :]
        weight = lambda_ts[step_index, tau]
        mask_t = torch.mean(mask[:, tau, :, :, :].float(), 1, keepdim=True)
        mask_t = mask_t > 0.5
        mask_zero = ~mask_t
        num_zero = torch.sum(mask_zero)
        masked_tensor1 = pred_original_sample[:, tau, :, :, :] * mask_t
        masked_tensor2 = FEA * mask_t
        masked_diff = masked_tensor1 - masked_tensor2
        flat_diff = torch.abs(masked_diff.flatten())
        sorted_diff, indices = torch.sort(flat_diff)
        weight = torch.clamp(weight, weight_clamp, 1)
        cutoff_index_e = int(weight * (len(sorted_diff) - num_zero)) + num_zero
        cutoff_value = sorted_diff[cutoff_index_e - 1]
        top_mask = (torch.abs(masked_diff) <= cutoff_value) & mask_t
        top_masks.append(top_mask)
        aux.append((int(cutoff_index_e), float(cutoff_value)))
    top_masks = torch.cat(top_masks, 0)
    top_masks = top_masks.unsqueeze(0)
    mask_ones = torch.ones_like(top_masks[:, 0:1, :, :, :])
    top_masks = torch.cat((mask_ones, top_masks), 1)
    return top_masks, aux


def main():
    vers = np.array([f"numpy {np.__version__}", f"torch {torch.__version__}",
                     "statements of DecaYale/SYN3R@402f4e9 scheduling_euler_discrete.py:1442-1499"])
    g = torch.Generator().manual_seed(77)
    cases = [  # b, F, c, h, w, mask batch, mask channels, hole fraction, step
        (1, 6, 4, 9, 16, 1, 1, 0.3, 0),
        (2, 5, 4, 8, 8, 2, 4, 0.5, 2),
        (1, 4, 4, 9, 16, 1, 4, 0.0, 1),       # nothing masked out
        (1, 4, 2, 6, 10, 1, 1, 1.0, 3),       # everything masked out
        (1, 25, 4, 9, 16, 1, 1, 0.25, 5),     # the production frame count
    ]
    for i, (b, F, c, h, w, bm, cm, frac, step) in enumerate(cases):
        pred = torch.randn((b, F, c, h, w), generator=g)
        tcl = torch.randn((2, F, c, h, w), generator=g)
        tcl[1] = pred[0] + 0.3 * torch.randn((F, c, h, w), generator=g)       # the conditioning latents are close
        mask = (torch.rand((bm, F - 1, cm, h, w), generator=g) < frac).float()  # 1 = hole (the pipeline's convention)
        if cm == 4:
            mask[:, :, 2:] = mask[:, :, :1]                                    # channels mostly agree, some 2-of-4 ties
        lambda_ts = torch.rand((6, F), generator=g) * 1.2                     # some below 0.4, some above 1
        if i == 0:
            pred[0, 2, 1, 3, 4] = float("nan")
            pred[0, 3, 0, 0, 0] = float("inf")
            tcl[1, 1, 0, 2, 2] = pred[0, 1, 0, 2, 2]                          # an exact zero difference
        out, aux = reference_replace(pred.clone(), tcl, mask.clone(), lambda_ts, step)
        path = os.path.join(HERE, f"latent_{i}.npz")
        np.savez_compressed(path, versions=vers, pred=pred.numpy(), temp_cond_latents=tcl.numpy(), mask=mask.numpy(),
                            lambda_ts=lambda_ts.numpy(), step_index=np.int64(step), out=out.numpy(),
                            aux_cutoff_index=np.array([a[0] for a in aux]), aux_cutoff_value=np.array([a[1] for a in aux], np.float32))
        print(os.path.basename(path), os.path.getsize(path) // 1024, "KiB", "replaced fraction",
              float((out[:, 1:-1] != pred[:, 1:-1]).float().mean()))
        if i in (0, 1, 4):
            wc = [0.5, 0.2, 0.8][(0, 1, 4).index(i)]
            tm, aux = reference_top_masks(pred.clone(), tcl, mask.clone(), lambda_ts, step, wc)
            path = os.path.join(HERE, f"latent_top_{i}.npz")
            np.savez_compressed(path, versions=vers, pred=pred.numpy(), temp_cond_latents=tcl.numpy(), mask=mask.numpy(),
                                lambda_ts=lambda_ts.numpy(), step_index=np.int64(step), weight_clamp=np.float64(wc),
                                out_top_masks=tm.numpy(), aux_cutoff_index=np.array([a[0] for a in aux]),
                                aux_cutoff_value=np.array([a[1] for a in aux], np.float32))
            print(os.path.basename(path), os.path.getsize(path) // 1024, "KiB", "selected fraction", float(tm.float().mean()))


if __name__ == "__main__":
    sys.exit(main())
