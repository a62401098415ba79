"""Generate the golden fixtures in tests/golden/*.npz by RUNNING THE LIVE REFERENCE.

Run in the build container only (needs /root/reference, which does not exist on the GPU box):

    python tests/golden/make_golden.py

Each fixture stores the seeded inputs and the outputs of the unmodified reference functions
(`solver_utils/forward_warp.py`, `solver_utils/consistency.py` of DecaYale/SYN3R @ 402f4e9), plus the
numpy / torch / BLAS versions they were produced with (NumPy's small-matmul rounding order is
host-BLAS specific).  `inverse_warp` hard-codes `.to('cuda')`
(/root/reference/solver_utils/forward_warp.py:225-256); a harness-side shim maps 'cuda'->'cpu'
without touching the reference tree.  Nothing from the reference is copied: only its outputs.
"""
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
REF = os.environ.get("SYN3R_REFERENCE", "/root/reference")
sys.path.insert(0, REF)

from syn3r_b200 import synth  # noqa: E402

# ---- device shim so the reference's hard-coded .to('cuda') runs on a CPU-only box -------------
_orig_to = torch.Tensor.to


def _to_shim(self, *a, **k):
    a = tuple("cpu" if (isinstance(x, str) and x.startswith("cuda")) else x for x in a)
    if isinstance(k.get("device"), str) and k["device"].startswith("cuda"):
        k["device"] = "cpu"
    return _orig_to(self, *a, **k)


torch.Tensor.to = _to_shim

import solver_utils.forward_warp as ref_fw  # noqa: E402
import solver_utils.consistency as ref_cc  # noqa: E402


def versions():
    try:
        blas = np.show_config(mode="dicts")["Build Dependencies"]["blas"]
        blas = f"{blas.get('name')} {blas.get('version')}"
    except Exception:
        blas = "unknown"
    return np.array([f"numpy {np.__version__}", f"torch {torch.__version__}", f"blas {blas}",
                     "reference DecaYale/SYN3R@402f4e9"])


def save(name, **arrs):
    path = os.path.join(HERE, name + ".npz")
    np.savez_compressed(path, versions=versions(), **arrs)
    print(f"{name}: {os.path.getsize(path) / 1024:.1f} KiB")


def run_forward(name, frame, mask1, depth, T1, T2, K, K2=None):
    warped, mask2, flow = ref_fw.forward_warp(frame, mask1, depth, T1, T2, K, K2)
    extra = {}
    if mask1 is not None:
        extra["mask1"] = mask1
    if K2 is not None:
        extra["K2"] = K2
    save(name, frame=frame, depth=depth, T1=T1, T2=T2, K=K, warped=warped, mask2=mask2, flow12=flow, **extra)


def look_at(cam, target, up):  # demo-style f64 view matrix (own restatement of the usual look-at)
    f = (target - cam) / np.linalg.norm(target - cam)
    r = np.cross(up, f)
    r /= np.linalg.norm(r)
    u = np.cross(f, r)
    Rm = np.eye(4)
    Rm[:3, 0], Rm[:3, 1], Rm[:3, 2] = r, u, f
    Tm = np.eye(4)
    Tm[:3, 3] = -cam
    return Rm.T @ Tm


def forward_cases():
    # generic small cases: f32 image in [0,255], f32 depth, f32 poses, f64 K (as called from
    # model/diffusionGS.py:1571)
    tr = synth.make_trajectory(3, 4, 48, 64, seed=11)
    for i, (s, t) in enumerate([(0, 1), (1, 3), (2, 0)]):
        run_forward(f"fw_generic_{i}", tr.rgb[s], None, tr.depth[s], tr.pose_src[s], tr.pose_tgt[t], tr.K)
    # mid-size with the occluder producing real depth competition + holes
    tr = synth.make_trajectory(2, 3, 120, 160, seed=12, pose_step=2.0)
    run_forward("fw_mid", tr.rgb[0], None, tr.depth[0], tr.pose_src[0], tr.pose_tgt[2], tr.K)
    # uint8 image + f64 poses + demo intrinsics (model/SVD_1pass.py:941-968 regime)
    rng = np.random.default_rng(13)
    H, W = 36, 64
    img = rng.integers(0, 256, (H, W, 3), dtype=np.uint8)
    d = np.clip(1e4 / (rng.random((H, W)) * 4000 + 1500), 1e-4, 500).astype(np.float32)
    Kd = np.array([[260 / 16, 0, 512 / 16], [0, 260 / 16, 288 / 16], [0, 0, 1]], dtype=np.float64)
    ang = np.deg2rad(4.0)
    P0 = look_at(np.array([0.0, 0, -3.0]), np.zeros(3), np.array([0.0, 1, 0]))
    P1 = look_at(np.array([-3 * np.sin(ang), 0, -3 * np.cos(ang)]), np.zeros(3), np.array([0.0, 1, 0]))
    run_forward("fw_u8_f64pose", img, None, d, P0, P1, Kd)
    # source mask + distinct K2
    tr = synth.make_trajectory(2, 2, 40, 56, seed=14)
    m1 = rng.random((40, 56)) > 0.3
    K2 = tr.K.copy()
    K2[0, 0] *= 1.1
    K2[1, 1] *= 0.95
    K2[0, 2] += 1.3
    K2[0, 1] = 0.01
    run_forward("fw_mask_K2", tr.rgb[1], m1, tr.depth[1], tr.pose_src[1], tr.pose_tgt[0], tr.K, K2)
    # ---- degenerate set (correctness only) ----
    tr = synth.make_trajectory(1, 1, 32, 48, seed=15)
    I4 = np.eye(4, dtype=np.float32)
    run_forward("fw_identity", tr.rgb[0], None, tr.depth[0], tr.pose_src[0], tr.pose_src[0].copy(), tr.K)
    # pure y-rotation, integer principal point: a whole row lands on exact integers (N3)
    Kint = np.array([[40.0, 0, 24.0], [0, 40.0, 16.0], [0, 0, 1]])
    Ty = np.eye(4)
    Ty[:3, :3] = synth.euler_deg_to_R(0, 3.0, 0)
    run_forward("fw_integer_row", tr.rgb[0], None, tr.depth[0], I4, Ty.astype(np.float32), Kint)
    # pure integer translation in pixels at constant depth: every pixel lands on an exact integer
    dconst = np.full((32, 48), 2.0, dtype=np.float32)
    Tt = np.eye(4, dtype=np.float32)
    Tt[0, 3] = 0.25   # 40 * 0.25 / 2 = 5 px exactly
    Tt[1, 3] = -0.1   # -2 px (0.1 is not exact in f32 -> near-integer, not exact)
    run_forward("fw_integer_shift", tr.rgb[0], None, dconst, I4, Tt, Kint)
    # z<=0 points, zero depth, huge translation (mostly out of frame)
    dneg = tr.depth[0].copy()
    dneg[::3, ::5] = -1.5
    dneg[1::4, 2::7] = 0.0
    Tb = np.eye(4, dtype=np.float32)
    Tb[:3, :3] = synth.euler_deg_to_R(2, -4, 1).astype(np.float32)
    Tb[:3, 3] = [0.3, -0.2, -1.0]
    run_forward("fw_behind", tr.rgb[0], None, dneg, I4, Tb, tr.K)
    Tfar = np.eye(4, dtype=np.float32)
    Tfar[0, 3] = 500.0
    run_forward("fw_all_out", tr.rgb[0], None, tr.depth[0], I4, Tfar, tr.K)
    # all z' <= 0 -> Lmax = 0 -> 0/0 -> all-False mask (Appendix A.1 step 12)
    Tflip = np.eye(4, dtype=np.float32)
    Tflip[:3, :3] = synth.euler_deg_to_R(0, 180.0, 0).astype(np.float32)
    run_forward("fw_all_behind", tr.rgb[0], None, tr.depth[0], I4, Tflip, tr.K)
    # NaN depth poisons the global max (np.max propagates NaN)
    dnan = tr.depth[0].copy()
    dnan[5, 7] = np.nan
    run_forward("fw_nan_depth", tr.rgb[0], None, dnan, I4, Tb, tr.K)


def helper_cases():
    rng = np.random.default_rng(21)
    H, W = 24, 40
    d = (rng.random((H, W)) * 3 + 0.5).astype(np.float32)
    K = rng.random((3, 3)) + np.eye(3) * 50
    K2 = rng.random((3, 3)) + np.eye(3) * 40
    T1 = rng.random((4, 4)).astype(np.float32) + np.eye(4, dtype=np.float32) * 2
    T2 = rng.random((4, 4)).astype(np.float32) + np.eye(4, dtype=np.float32) * 2
    q, X = ref_fw.compute_transformed_points(d, T1, T2, K, K2)
    save("ctp_dense", depth=d, T1=T1, T2=T2, K=K, K2=K2, q=q, X=X)
    T1d, T2d = T1.astype(np.float64), T2.astype(np.float64)
    q, X = ref_fw.compute_transformed_points(d.astype(np.float64), T1d, T2d, K, None)
    save("ctp_f64", depth=d.astype(np.float64), T1=T1d, T2=T2d, K=K, q=q, X=X)
    # bilinear_splatting called directly: C=2 channels, masks, arbitrary flow, is_image False
    frame = rng.standard_normal((H, W, 2))
    flow = rng.standard_normal((H, W, 2)) * 3
    flow[3, 4] = [0.0, 0.0]
    flow[7, 9] = [2.0, -1.0]
    z = rng.random((H, W)) * 10 - 1
    m1 = rng.random((H, W)) > 0.2
    fm = rng.random((H, W)) > 0.2
    out, mask = ref_fw.bilinear_splatting(frame, m1, z, flow, fm, is_image=False)
    save("bs_generic", frame=frame, mask1=m1, depth=z, flow12=flow, flow_mask=fm, out=out, mask=mask)
    save("grid", grid=ref_fw.create_grid(5, 7))


def backward_cases():
    tr = synth.make_trajectory(2, 3, 60, 80, seed=31, pose_step=1.5)
    rng = np.random.default_rng(32)
    Kf = tr.K.astype(np.float32)
    for i, (s, t, bw, use_bg) in enumerate([(0, 1, 20, False), (1, 2, 10, True), (0, 0, 20, True)]):
        img = torch.from_numpy(np.ascontiguousarray(tr.rgb[s].transpose(2, 0, 1)))
        depth = torch.from_numpy(tr.depth[s])[None]
        dps = torch.from_numpy(tr.depth_tgt[t])[None]
        if i == 2:  # exercise zero / negative source depth and zero pseudo depth
            depth = depth.clone()
            depth[0, ::4, ::6] = 0.0
            depth[0, 1::5, 3::7] = -0.5
            dps = dps.clone()
            dps[0, 2::9, ::8] = 0.0
        bg = torch.from_numpy(rng.random((60, 80)) > 0.5)[None] if use_bg else None
        p1, p2 = torch.from_numpy(tr.pose_src[s]), torch.from_numpy(tr.pose_tgt[t])
        out = ref_fw.inverse_warp(img, depth, dps, p1, p2, torch.from_numpy(Kf), bg_mask=bg, bandwidth=bw)
        arrs = {"out_" + k: (v.numpy() if v is not None else np.zeros(0)) for k, v in out.items()}
        if bg is not None:
            arrs["bg_mask"] = bg.numpy()
        save(f"iw_generic_{i}", img=img.numpy(), depth=depth.numpy(), depth_pseudo=dps.numpy(),
             pose1=p1.numpy(), pose2=p2.numpy(), K=Kf, bandwidth=np.float64(bw), **arrs)
    # consistency_check_with_depth alone, with distinct K1/K2 (skewed) and the pieces
    s, t = 1, 0
    K1 = torch.from_numpy(Kf.copy())
    K2 = torch.from_numpy(Kf.copy())
    K2[0, 0] *= 1.05
    K2[0, 1] = 0.02
    d1, d2 = torch.from_numpy(tr.depth_tgt[t]), torch.from_numpy(tr.depth[s])
    p1, p2 = torch.from_numpy(tr.pose_tgt[t]), torch.from_numpy(tr.pose_src[s])
    err = ref_cc.consistency_check_with_depth(d1, p1, K1, d2, p2, K2)
    pts = ref_cc.get_points_from_depth(d1, K1)
    pts2 = ref_cc.transform_points(pts, p1, p2)
    save("cc_generic", depth1=d1.numpy(), pose1=p1.numpy(), K1=K1.numpy(), depth2=d2.numpy(), pose2=p2.numpy(),
         K2=K2.numpy(), err=err.numpy(), pts=pts.numpy(), pts2=pts2.numpy())


if __name__ == "__main__":
    forward_cases()
    helper_cases()
    backward_cases()
