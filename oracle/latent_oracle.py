"""CPU restatement of the scheduler's guided latent replacement (scheduler half).

TEST INFRASTRUCTURE ONLY: imported by tests/ as the checker, never by the product (syn3r_b200/ has no CPU path).

Follows /root/reference/thirdparty/diffusers/src/diffusers/schedulers/scheduling_euler_discrete.py:1442-1499
(`step_interp_prob_uncertain`) in float32 NumPy; `top_masks` restates the selection-only form of `step_single` (:548-589)
and `step_interp` (:735-778), which clamp lambda at the caller's `weight_clamp` and keep the boolean mask:

  valid    = (1 - mask) > 0.5, a frame of ones prepended                                     (:1447-1451)
  per frame tau of index_list:
    m_     = mean over the mask's channels of valid (float32); m = m_ > 0.5; nz = count(~m)   (:1457-1461)
    d      = |pred[:, tau] * m - FEA * m|,  FEA = temp_cond_latents[1, tau]                   (:1464-1469)
    k      = int(clamp(lambda_ts[step, tau], 0.4, 1) * (d.size - nz)) + nz   (float32 product) (:1476-1479)
    cutoff = the k-th smallest of d (sorted[k - 1]; k = 0 wraps to the largest; NaN sorts last) (:1473, :1480)
    w      = x / (1 + x), x = 1 / (1 - m_ + 1e-6); w < 0.51 -> 0; w *= (d <= cutoff)          (:1487-1491)
    pred[:, tau] = (1 - w) * pred[:, tau] + w * FEA                                           (:1493)
  pred[:, 0] = temp_cond_latents[1, 0]; pred[:, -1] = temp_cond_latents[1, -1]                (:1496-1499)

A selection (np.partition semantics via a full sort) and exact float32 elementwise work: the bar is bit-exact.
Parity pinned by tests/golden/latent_*.npz, latent_top_*.npz, latent_interp_*.npz: outputs of the LIVE patched
`EulerDiscreteScheduler` of the reference (`step_interp_prob_uncertain`, `step_single`, `step_interp` called on seeded
inputs; tests/golden/make_golden_latent.py).
"""
import numpy as np

f32 = np.float32


def guided_replace(pred: np.ndarray, temp_cond_latents: np.ndarray, mask: np.ndarray, lambda_ts: np.ndarray,
                   step_index: int, index_list=None) -> dict:
    """pred [b,F,c,h,w] f32, temp_cond_latents [2,F,c,h,w] f32, mask [bm,F-1,cm,h,w] (1 = hole), lambda_ts [steps,F].
    Returns out (a new array) and the per-frame cutoff index / value."""
    pred = pred.astype(f32).copy()
    b, F, c, h, w = pred.shape
    index_list = list(range(1, F - 1)) if index_list is None else list(index_list)
    valid = (f32(1) - mask.astype(f32)) > f32(0.5)
    valid = np.concatenate([np.ones_like(valid[:, :1]), valid], axis=1)
    ks, cuts = [], []
    with np.errstate(invalid="ignore", over="ignore", divide="ignore"):
        for tau in index_list:
            fea = temp_cond_latents[1:2, tau].astype(f32)
            cm = valid.shape[2]
            m_ = (valid[:, tau].astype(f32).sum(axis=1, keepdims=True, dtype=f32) * (f32(1) / f32(cm))).astype(f32)
            m = m_ > f32(0.5)
            nz = int((~m).sum())
            mf = m.astype(f32)
            d = np.abs(((pred[:, tau] * mf).astype(f32) - (fea * mf).astype(f32)).astype(f32))
            flat = np.sort(d.ravel())                      # NaN last, like torch.sort
            wt = f32(min(max(f32(lambda_ts[step_index, tau]), f32(0.4)), f32(1)))
            k = int(f32(wt * f32(flat.size - nz))) + nz
            cutoff = flat[k - 1]
            x = (f32(1) / ((f32(1.0) - m_).astype(f32) + f32(1e-6)).astype(f32)).astype(f32)
            wgt = (x / (f32(1) + x).astype(f32)).astype(f32)
            wgt = np.where(wgt >= f32(0.51), wgt, f32(0)).astype(f32)
            wgt = ((d <= cutoff).astype(f32) * wgt).astype(f32)
            pred[:, tau] = (((f32(1) - wgt).astype(f32) * pred[:, tau]).astype(f32) + (wgt * fea).astype(f32)).astype(f32)
            ks.append(k)
            cuts.append(cutoff)
    pred[:, 0:1] = temp_cond_latents[1:2, 0:1]
    pred[:, -1:] = temp_cond_latents[1:2, -1:]
    return dict(out=pred, cutoff_index=np.array(ks), cutoff_value=np.array(cuts, f32))


def top_masks(pred: np.ndarray, temp_cond_latents: np.ndarray, mask: np.ndarray, lambda_ts: np.ndarray, step_index: int,
              weight_clamp: float, index_list=None, trailing_ones: bool = False) -> dict:
    """scheduling_euler_discrete.py:548-589 (`step_single`; `step_interp` :735-778 is the same with index_list to 23):
    the selection of guided_replace with `torch.clamp(weight, weight_clamp, 1)` (:577), no blend; top_masks is the
    reference's [1, 1 + len(index_list) * b, c, h, w] bool tensor (frames concatenated on dim 0, unsqueezed, a frame
    of ones prepended, :586-589).  trailing_ones: `step_interp` (index_list defaults to 1..F-2, a second frame of
    ones appended, :775-777).  mask holds at least max(index_list) frames (row tau - 1 is frame tau's)."""
    pred = pred.astype(f32)
    b, F, c, h, w = pred.shape
    index_list = list(range(1, F - 1 if trailing_ones else F)) if index_list is None else list(index_list)
    valid = (f32(1) - mask.astype(f32)) > f32(0.5)
    valid = np.concatenate([np.ones_like(valid[:, :1]), valid], axis=1)
    tops, ks, cuts = [], [], []
    with np.errstate(invalid="ignore", over="ignore"):
        for tau in index_list:
            fea = temp_cond_latents[1:2, tau].astype(f32)
            cm = valid.shape[2]
            m_ = (valid[:, tau].astype(f32).sum(axis=1, keepdims=True, dtype=f32) * (f32(1) / f32(cm))).astype(f32)
            m = m_ > f32(0.5)
            nz = int((~m).sum())
            mf = m.astype(f32)
            d = np.abs(((pred[:, tau] * mf).astype(f32) - (fea * mf).astype(f32)).astype(f32))
            flat = np.sort(d.ravel())
            wt = f32(min(max(f32(lambda_ts[step_index, tau]), f32(weight_clamp)), f32(1)))
            k = int(f32(wt * f32(flat.size - nz))) + nz
            cutoff = flat[k - 1]
            tops.append((d <= cutoff) & m)
            ks.append(k)
            cuts.append(cutoff)
    t = np.concatenate(tops, axis=0)[None]
    t = np.concatenate([np.ones_like(t[:, :1]), t] + ([np.ones_like(t[:, :1])] if trailing_ones else []), axis=1)
    return dict(top_masks=t, cutoff_index=np.array(ks), cutoff_value=np.array(cuts, f32))
