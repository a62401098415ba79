"""Golden fixtures of the scheduler's guided latent replacement (scheduler half):
tests/golden/latent_*.npz, latent_top_*.npz, latent_interp_*.npz.

Produced by CALLING THE LIVE, PATCHED SCHEDULER of the reference
(/root/reference/thirdparty/diffusers/src/diffusers/schedulers/scheduling_euler_discrete.py, class
`EulerDiscreteScheduler`): `step_interp_prob_uncertain` (:1343-1513, the soft blend of :1442-1499), `step_single`
(:445-631, selection of :548-589) and `step_interp` (:633-812, selection of :735-778).  The class imports here once two
absent third-party modules are stubbed in `sys.modules` (ref_exec.live_scheduler); nothing of the reference is
retyped.  The scheduler is configured with `prediction_type="sample"`, so `pred_original_sample` IS the `model_output`
tensor passed in (:1433) and the replacement block is the only thing that touches it; the method returns it as
`EulerDiscreteSchedulerOutput.pred_original_sample`.  The methods hard-code `index_list = [1..23]` / `[1..24]`, so
every case has F = 25 frames (the production clip length) and a small latent grid.  `top_masks`, `cutoff_index_e` and
`cutoff_value` are locals of the methods: a `sys.settrace` hook reads them from the live frame.

    python tests/golden/make_golden_latent.py
"""
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
import ref_exec  # noqa: E402

SCHED = "thirdparty/diffusers/src/diffusers/schedulers/scheduling_euler_discrete.py"


class LocalsTap:
    """Records (cutoff_index_e, cutoff_value) per loop iteration and the final `top_masks` of a scheduler method."""

    def __init__(self, func_name):
        self.func_name, self.cut, self.top_masks = func_name, {}, None

    def __enter__(self):
        def local(frame, event, arg):
            loc = frame.f_locals
            if "ii" in loc and "cutoff_value" in loc and "cutoff_index_e" in loc and "top_mask" in loc:
                self.cut[int(loc["ii"])] = (int(loc["cutoff_index_e"]), float(loc["cutoff_value"]))
            if event == "return" and "top_masks" in loc and torch.is_tensor(loc["top_masks"]):
                self.top_masks = loc["top_masks"].clone()
            return local

        def glob(frame, event, arg):
            return local if frame.f_code.co_name == self.func_name else None
        sys.settrace(glob)
        return self

    def __exit__(self, *a):
        sys.settrace(None)

    def cutoffs(self):
        ks = sorted(self.cut)
        return np.array([self.cut[k][0] for k in ks]), np.array([self.cut[k][1] for k in ks], np.float32)


def make_scheduler(n_steps):
    Sched = ref_exec.live_scheduler()
    s = Sched(prediction_type="sample")
    s.set_timesteps(n_steps)
    s.is_scale_input_called = True
    return s


def call(method, pred, tcl, mask, lambda_ts, step, **kw):
    n_steps = lambda_ts.shape[0]
    s = make_scheduler(n_steps)
    model_output = pred.clone()                      # == pred_original_sample inside (prediction_type "sample")
    sample = torch.zeros_like(pred)
    with LocalsTap(method) as tap:
        out = getattr(s, method)(model_output, float(s.timesteps[step]), sample, temp_cond_latents=tcl, mask=mask.clone(),
                                 lambda_ts=lambda_ts, step_i=step, **kw)
    assert out.pred_original_sample is model_output
    return out.pred_original_sample, tap


def main():
    _, tag_blend = ref_exec.source(SCHED, 1442, 1499)
    _, tag_single = ref_exec.source(SCHED, 548, 589)
    _, tag_interp = ref_exec.source(SCHED, 735, 778)
    g = torch.Generator().manual_seed(77)
    F = 25
    cases = [  # b, c, h, w, mask batch, mask channels, hole fraction, step
        (1, 4, 9, 16, 1, 1, 0.3, 0),
        (2, 4, 8, 8, 2, 4, 0.5, 2),
        (1, 4, 9, 16, 1, 4, 0.0, 1),       # nothing masked out
        (1, 2, 6, 10, 1, 1, 1.0, 3),       # everything masked out
        (1, 4, 18, 32, 1, 1, 0.25, 5),
    ]
    for i, (b, c, h, w, bm, cm, frac, step) in enumerate(cases):
        pred = torch.randn((b, F, c, h, w), generator=g)
        tcl = torch.randn((2, F, c, h, w), generator=g)
        tcl[1] = pred[0] + 0.3 * torch.randn((F, c, h, w), generator=g)       # the conditioning latents are close
        # 1 = hole (the pipeline's convention); F - 2 = 23 frames as SVD_2pass_prob_uncertain*.py passes it (:549-550)
        mask23 = (torch.rand((bm, F - 2, cm, h, w), generator=g) < frac).float()
        if cm == 4:
            mask23[:, :, 2:] = mask23[:, :, :1]                                # channels mostly agree, some 2-of-4 ties
        lambda_ts = torch.rand((6, F), generator=g) * 1.2                     # some below 0.4, some above 1
        if i == 0:
            pred[0, 2, 1, 3, 4] = float("nan")
            pred[0, 3, 0, 0, 0] = float("inf")
            tcl[1, 1, 0, 2, 2] = pred[0, 1, 0, 2, 2]                          # an exact zero difference
        out, tap = call("step_interp_prob_uncertain", pred, tcl, mask23, lambda_ts, step)
        ks, cv = tap.cutoffs()
        path = os.path.join(HERE, f"latent_{i}.npz")
        np.savez_compressed(path, versions=ref_exec.versions("live EulerDiscreteScheduler.step_interp_prob_uncertain", tag_blend),
                            pred=pred.numpy(), temp_cond_latents=tcl.numpy(), mask=mask23.numpy(), lambda_ts=lambda_ts.numpy(),
                            step_index=np.int64(step), out=out.numpy(), aux_cutoff_index=ks, aux_cutoff_value=cv)
        print(os.path.basename(path), os.path.getsize(path) // 1024, "KiB", "replaced fraction",
              float((out[:, 1:-1] != pred[:, 1:-1]).float().mean()))
        if i in (0, 1, 4):
            wc = [0.5, 0.2, 0.8][(0, 1, 4).index(i)]
            # step_single indexes mask frames up to 24 (:551): the mask has F - 1 frames there
            mask24 = torch.cat([mask23, (torch.rand((bm, 1, cm, h, w), generator=g) < frac).float()], 1)
            out_s, tap = call("step_single", pred, tcl, mask24, lambda_ts, step, weight_clamp=wc)
            assert torch.equal(torch.nan_to_num(out_s), torch.nan_to_num(pred))   # selection only: pred untouched
            ks, cv = tap.cutoffs()
            path = os.path.join(HERE, f"latent_top_{i}.npz")
            np.savez_compressed(path, versions=ref_exec.versions("live EulerDiscreteScheduler.step_single", tag_single),
                                pred=pred.numpy(), temp_cond_latents=tcl.numpy(), mask=mask24.numpy(), lambda_ts=lambda_ts.numpy(),
                                step_index=np.int64(step), weight_clamp=np.float64(wc), out_top_masks=tap.top_masks.numpy(),
                                aux_cutoff_index=ks, aux_cutoff_value=cv)
            print(os.path.basename(path), os.path.getsize(path) // 1024, "KiB", "selected fraction",
                  float(tap.top_masks.float().mean()))
            # step_interp: frames 1..23, clamp 0.4 hard-coded (:766), a trailing frame of ones (:777)
            out_i, tap = call("step_interp", pred, tcl, mask23, lambda_ts, step)
            ks, cv = tap.cutoffs()
            path = os.path.join(HERE, f"latent_interp_{i}.npz")
            np.savez_compressed(path, versions=ref_exec.versions("live EulerDiscreteScheduler.step_interp", tag_interp),
                                pred=pred.numpy(), temp_cond_latents=tcl.numpy(), mask=mask23.numpy(), lambda_ts=lambda_ts.numpy(),
                                step_index=np.int64(step), weight_clamp=np.float64(0.4), out_top_masks=tap.top_masks.numpy(),
                                aux_cutoff_index=ks, aux_cutoff_value=cv)
            print(os.path.basename(path), os.path.getsize(path) // 1024, "KiB", "selected fraction",
                  float(tap.top_masks.float().mean()))


if __name__ == "__main__":
    sys.exit(main())
