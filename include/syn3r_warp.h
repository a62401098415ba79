/* syn3r_warp.h — C ABI of libsyn3r_warp.so (hand-written sm_100a CUDA, no torch types).
 *
 * Drop-in boundary for ONE hot path of DecaYale/SYN3R @ 402f4e9: the geometric conditioning step
 * (solver_utils/forward_warp.py + solver_utils/consistency.py).  The reference has no FFI for this
 * path — its module-level Python functions ARE the interface — so each entry point
 * below names the reference function(s) it replaces (paths relative to the reference root); the
 * Python mirror of those functions (package syn3r_b200.solver_utils) binds them with ctypes, and
 * INTEGRATION.md shows the stub a maintainer would drop into the reference tree.
 *
 * Conventions
 *   - Every pointer is a DEVICE pointer unless the comment says "host".  Arrays are dense, C order.
 *   - `stream` is a cudaStream_t passed as void*.  All work is enqueued on it; no entry point
 *     allocates, synchronises, or keeps mutable global state (apart from a thread-local error
 *     string).  Outputs and the workspace are caller-owned.
 *   - Return value: 0 = SYN3R_OK, otherwise one of the SYN3R_E* codes; syn3r_last_error() gives
 *     a human-readable reason.  No C++ exception crosses the boundary.
 *   - Poses are world-to-camera 4x4 [R t; 0 1] (forward_warp.py:151-152, consistency.py:30-31);
 *     K is a pixel-unit pinhole 3x3.  Small matrices (inverses, pose products) are prepared by
 *     the caller with the SAME library calls the reference makes (np.linalg.inv / np.matmul in
 *     the caller's dtype, torch.inverse / @ in fp32) so their rounding is the reference's.
 *   - A "pair" is one (source view, target pose) warp = one call of the reference's per-frame
 *     function.  A "slot" is one output frame: PAIR mode -> slot == pair (reference semantics);
 *     FUSED mode -> several pairs (sources) accumulate into one slot (extension).
 */
#ifndef SYN3R_WARP_H
#define SYN3R_WARP_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define SYN3R_ABI_VERSION 17

enum {
    SYN3R_OK = 0,
    SYN3R_EARG = 1,       /* bad argument (null pointer, non-positive size, bad flag combination) */
    SYN3R_EWORKSPACE = 2, /* workspace too small or misaligned (needs 256-byte alignment)        */
    SYN3R_ECUDA = 3,      /* a CUDA runtime call / kernel launch failed                           */
    SYN3R_EDEVICE = 4     /* current device is not compute capability 10.x                        */
};

/* flags of the forward path */
enum {
    SYN3R_FWD_GATE_0P6 = 1 << 0, /* mask &= (sum of weights <= 0.6): model/SVD_1pass.py:853-855         */
    SYN3R_FWD_WINNER = 1 << 1,   /* emit the deterministic winner index (extension, 64-bit atomicMin)   */
    SYN3R_FWD_IS_IMAGE = 1 << 2, /* range check + clip[0,255] + round-half-even (forward_warp.py:119-128)*/
    SYN3R_FWD_EXACT_WEIGHT = 1 << 3, /* soft-depth weight with IEEE log / exp / div like the reference.  Default:
                                      a short fp64 table+polynomial path, relative error < 3e-7 (masks and
                                      winners unchanged, colours move by < 3e-7 of full scale)            */
    SYN3R_FWD_LEGACY_GLOBAL = 1 << 4, /* first-generation path: global fp32 accumulators + REDG, separate
                                      resolve pass; kept for A/B measurements.  Default: per-tile
                                      shared-memory accumulators resolved in place                         */
    SYN3R_FWD_FP32_TILE = 1 << 5     /* second-generation path: per-tile fp32 accumulators (CAS atomics,
                                      forward_tile.cuh); kept for A/B measurements and used automatically for
                                      tiles the default integer block-floating-point kernel (forward_fx.cuh)
                                      hands back.  EXACT_WEIGHT implies it                                 */
    /* measurement aid: run only the named phase(s) of syn3r_forward_warp_batched so a caller can
       bracket each kernel with its own CUDA events; no bit set = all three phases */
    , SYN3R_FWD_PHASE_ZMAX = 1 << 8    /* workspace init, block statistics, per-pair max z' + cell boxes */
    , SYN3R_FWD_PHASE_SPLAT = 1 << 9   /* project + splat                             */
    , SYN3R_FWD_PHASE_RESOLVE = 1 << 10 /* normalise + masks + winner (legacy path only; fused otherwise) */
};

int syn3r_version(void);
const char *syn3r_last_error(void);

/* Device self-test: the shared-reciprocal double division used by the forward projection (two quotients
 * over one denominator) against div.rn.f64 on n pseudo-random operand triples; adds the number of
 * differing quotients to *mismatches (device u64, caller-zeroed).  mode 0: moderate exponents,
 * mode 1: arbitrary bit patterns (denormals, inf, NaN included). */
int syn3r_selftest_ddiv2(unsigned long long n, unsigned long long seed, int mode, unsigned long long *mismatches,
                         void *stream);

/* Device self-test: the constant-divisor division of the lean consistency kernel (x / c through the correctly
 * rounded reciprocal and two Markstein corrections; the divisors are (W-1)/2 and (H-1)/2 of consistency.py:71-73)
 * against div.rn.f32 for the `count` float bit patterns x = first, first+1, ...; adds the number of differing
 * quotients to mismatch_dev[0] and leaves the bit pattern of one such x in mismatch_dev[1] (device u64[2],
 * caller-zeroed).  finite_only != 0 skips x = +-inf / NaN and quotients below 2^-30 (where the kernel's next steps
 * give the same result either way, inverse_kernels.cuh).
 * *rc_out (host, nullable) receives the reciprocal the library uses, 0 if it keeps the IEEE division for this c. */
int syn3r_selftest_div_const(float c, unsigned long long first, unsigned long long count, int finite_only,
                             unsigned long long *mismatch_dev, float *rc_out, void *stream);

/* Number of SMs / max opt-in shared memory of the current device (host query, cached). */
int syn3r_device_info(int *sm_count, int *smem_optin_bytes, int *cc_major, int *cc_minor);

/* ---------------------------------------------------------------------------------------------
 * Forward path: compute_transformed_points (forward_warp.py:7-37) + forward_warp (:141-182) +
 * bilinear_splatting (:41-129), batched over pairs.
 *
 *   rgbd      [S,H,W,4] f32   source colour (0..255 scale) + z-depth, one float4 per pixel
 *   depth64   [S,H,W]   f64   nullable; when given it replaces rgbd[...,3] (callers with f64 depth)
 *   src_mask  [S,H,W]   u8    nullable; mask1 of forward_warp.py:56 (0 = unknown pixel)
 *   Kinv_src  [S,9]     f64   np.linalg.inv(intrinsic1) (:25)
 *   pair_par  [P,24]    f64   per pair: M = (T2 @ inv(T1))[:3,:] row-major (12), intrinsic2 (9), 3 pad
 *   pair_src  [P]       i32   source index of each pair
 *   pair_slot [P]       i32   nullable -> slot = pair index (PAIR mode)
 *   out_rgb_f32 [n_slots,H,W,3] f32  nullable: normalised colour before rounding (:116)
 *   out_rgb_u8  [n_slots,H,W,3] u8   nullable: clipped + rounded (:127-128), needs IS_IMAGE
 *   out_mask    [n_slots,H,W]   u8   mask2 (:112), 0/1
 *   out_winner  [n_slots,H,W]   i32  nullable, needs WINNER: src*H*W + y*W + x of the contributing
 *                                     source pixel with the smallest z' key (high 32 bits of the fp64 z':
 *                                     sign, exponent, 20 mantissa bits), ties -> smallest
 *                                     index; -1 = hole
 *   out_flow    [P,H,W,2]  f64  nullable: flow12 (:177)
 *   out_status  [n_slots]  i32  nullable: bit0 = colour-range assert of :121-125 failed
 *   workspace   >= syn3r_fwd_workspace_bytes(...) bytes, 256-byte aligned: block statistics, per-(pair, block) cell
 *               boxes and, for frames of up to 2304 source blocks (1024 x 576), per-(pair, target tile) block lists
 *               (768 B per pair and 128 x 48 tile: 5.5 MB for 75 pairs at 576 x 1024); no per-pixel accumulators
 * ------------------------------------------------------------------------------------------- */
size_t syn3r_fwd_workspace_bytes(int H, int W, int S, int n_pairs, int n_slots, int flags);

int syn3r_forward_warp_batched(const float *rgbd, const double *depth64, const uint8_t *src_mask,
                               const double *Kinv_src, const double *pair_par, const int32_t *pair_src,
                               const int32_t *pair_slot, int S, int P, int n_slots, int H, int W, int flags,
                               float *out_rgb_f32, uint8_t *out_rgb_u8, uint8_t *out_mask, int32_t *out_winner,
                               double *out_flow, int32_t *out_status, void *workspace, size_t workspace_bytes,
                               void *stream);

/* compute_transformed_points (forward_warp.py:7-37) on its own: q = K2 (M [d K1inv p; 1])[:3] and
 * X = d K1inv p, both [H,W,3] f64.  depth [H,W] f64, par = one pair_par record, Kinv [9]. */
int syn3r_compute_transformed_points(const double *depth, const double *Kinv, const double *par, int H, int W,
                                     double *out_q, double *out_X, void *stream);

/* bilinear_splatting (forward_warp.py:41-129) on its own, from caller-supplied flow and depth:
 *   frame [H,W,C] f32 (C <= 3), mask1 / flow_mask [H,W] u8 nullable, z [H,W] f64 (the `depth1`
 *   argument, i.e. z'), flow [H,W,2] f64.  out [H,W,C] f32 (nullable), out_u8 [H,W,C] (nullable,
 *   IS_IMAGE), out_mask [H,W] u8, out_status [1] i32 nullable.  Workspace as for P = n_slots = 1. */
int syn3r_bilinear_splat(const float *frame, const uint8_t *mask1, const double *z, const double *flow,
                         const uint8_t *flow_mask, int H, int W, int C, int flags, float *out, uint8_t *out_u8,
                         uint8_t *out_mask, int32_t *out_status, void *workspace, size_t workspace_bytes,
                         void *stream);

/* ---------------------------------------------------------------------------------------------
 * Backward path: inverse_warp (forward_warp.py:187-279) incl. its consistency_check_with_depth
 * call (:259 -> consistency.py:44-91), batched over pairs.  All fp32.
 *
 *   src_img   [S,3,H,W] f32 (nullable when no colour output is requested)
 *   src_depth [S,H,W] f32 with src_depth_stride = 1, or the depth lane of the forward path's
 *             [S,H,W,4] RGB-D buffer (pointer = rgbd + 3, src_depth_stride = 4)
 *   src_bg    [S,H,W] u8 nullable (bg_mask, :232)
 *   tgt_depth [T,H,W] f32  (depth_pseudo)
 *   K, Kinv   [9] f32 HOST (Kinv = torch.inverse(K), consistency.py:21); passed by value to the kernels
 *   pair_P    [P,2,16] f32 per pair: pose1 @ inverse(pose2) (:217) then pose2 @ inverse(pose1)
 *   pair_src / pair_tgt [P] i32
 *   outputs, each nullable, one plane set per pair: warped_img [P,3,H,W], warped_depth [P,H,W],
 *   warped_masked_img [P,3,H,W], soft_mask_reproj [P,H,W], reproj_err [P,H,W] f32; mask_warp,
 *   mask_depth, mask, mask_inv, mask_depth_strict, warped_bg, mask_reproj [P,H,W] u8 (0/1).
 *   The depth-normalisation masks need a per-pair min/max pre-pass (:236-240); it only runs when
 *   one of mask_depth / mask / mask_inv / mask_depth_strict / warped_masked_img is requested.
 * ------------------------------------------------------------------------------------------- */
typedef struct syn3r_inv_outputs {
    float *warped_img, *warped_depth, *warped_masked_img, *soft_mask_reproj, *reproj_err;
    uint8_t *mask_warp, *mask_depth, *mask, *mask_inv, *mask_depth_strict, *warped_bg, *mask_reproj;
} syn3r_inv_outputs;

size_t syn3r_inv_workspace_bytes(int n_pairs);

int syn3r_inverse_warp_batched(const float *src_img, const float *src_depth, int src_depth_stride,
                               const uint8_t *src_bg,
                               const float *tgt_depth, const float *K, const float *Kinv, const float *pair_P,
                               const int32_t *pair_src, const int32_t *pair_tgt, int P, int H, int W,
                               float bandwidth, float thr_depth, float thr_depth_strict,
                               const syn3r_inv_outputs *out /* host struct of device pointers */,
                               void *workspace, size_t workspace_bytes, void *stream);

/* consistency_check_with_depth (consistency.py:44-91), batched: frame 1 = depth1[pair_1[p]],
 * frame 2 = depth2[pair_2[p]]; K1, K1inv, K2 [9] f32 HOST, shared; pair_T [P,2,16] f32 = pose2 @ inv(pose1)
 * then pose1 @ inv(pose2) (consistency.py:39).  err [P,H,W] f32 (pixels). */
int syn3r_consistency_batched(const float *depth1, const float *depth2, const float *K1, const float *K1inv,
                              const float *K2, const float *pair_T, const int32_t *pair_1, const int32_t *pair_2,
                              int P, int H, int W, float *err, void *stream);

/* get_points_from_depth (consistency.py:6-23): pts [H,W,3] = (Kinv [x,y,1]) * depth.  Kinv [9] HOST. */
int syn3r_points_from_depth(const float *depth, const float *Kinv, int H, int W, float *pts, void *stream);
/* transform_points (consistency.py:25-42): T [16] HOST = pose2 @ inverse(pose1); out = (T [p;1])[:3] / w. */
int syn3r_transform_points(const float *pts, const float *T, int H, int W, float *out, void *stream);

/* np.packbits(mask != 0) for a device byte mask of n bytes (MSB first; a trailing partial byte is zero-padded):
 * bits [(n + 7) / 8] u8.  The masks of forward_warp (forward_warp.py:112, mask2) and inverse_warp (:260-267,
 * mask_reproj) are 0/1 bytes; packed they cross PCIe / NVLink at one bit per pixel and np.unpackbits(bits)[:n]
 * restores them on the host. */
int syn3r_pack_mask_bits(const uint8_t *mask, size_t n, uint8_t *bits, void *stream);

/* Caller-side post-processing of P warped frames (model/diffusionGS.py:1572-1597 forward branch, :1452-1491 backward
 * branch), all device pointers, one launch:
 *   hole = 1 - valid -> cv2.dilate 5x5 (outside ignored) -> out_mask_png [P,H,W] 0/255 (the reference's three
 *   channels are identical); frame with the dilated holes blanked: out_cond_u8 [P,H,W,3] = uint8(frame * (1 - m))
 *   and / or out_cond_f32 = that / 255 (float32); out_cond_ori = frame / 255 (:1467); out_latent [P,H/8,W/8] 0/1 =
 *   (8x8 block mean of the dilated mask >= 0.2); out_soft_pooled [P,H/8,W/8] = 8x8 block mean of 1 - soft in
 *   NumPy's float32 summation order (:1484-1487).
 * valid [P,H,W] (mask2 of forward_warp or mask_reproj of inverse_warp; 1 = covered).  The frame is either img_u8
 * [P,H,W,3] (forward_warp's uint8 frame) or img_chw [P,3,H,W] float32 in [0,256) (inverse_warp's warped_img); the
 * other is NULL.  soft [P,H,W] or NULL.  Any output may be NULL.  H and W must be multiples of 8. */
int syn3r_condition_post_batched(const uint8_t *valid, const uint8_t *img_u8, const float *img_chw, const float *soft,
                                 int P, int H, int W, uint8_t *out_mask_png, uint8_t *out_cond_u8, float *out_cond_f32,
                                 float *out_cond_ori, uint8_t *out_latent, float *out_soft_pooled, void *stream);

/* Uncertainty fusion of N frames (model/diffusionGS.py:821-854 and the clip of :876):
 *   conf = exp(-(|cond_ori - gs| / 0.5)^3) * (cond_ori.sum(-1) > 0);  unc = 1 - conf * (1 - soft_ori)   [float64]
 *   out_masks [N,H/8,W/8] f32 = 8x8 block mean of unc;  out_cond [N,H,W,3] f32 = clip(unc > 0.5 ? gs : cond_ori, 0, 1);
 *   out_unc [N,H,W] f64 = geo_inten_uncertainty.  cond_ori, gs [N,H,W,3] f32; soft_ori [N,H,W] f32 = 1 - soft_mask_reproj.
 * Any output may be NULL (not all).  H and W must be multiples of 8. */
int syn3r_fuse_uncertainty_batched(const float *cond_ori, const float *gs, const float *soft_ori, int N, int H, int W,
                                   float *out_cond, float *out_masks, double *out_unc, void *stream);

/* Nearby-frame consistency (model/diffusionGS.py:1347-1355, the tail of consistency_check_from_nearby_images_bw;
 * SURVEY 8f rank 2): for target frame n, whose neighbour warps are pairs [pair_first[n], pair_first[n+1]) of one
 * syn3r_inverse_warp_batched call (bandwidth 10, pairs ordered by target),
 *   out_unc[n]     = 1 - mean_k soft[k]                                               [N,H,W] f32
 *   out_int_unc[n] = 1 - exp(-(|mean_k warped[k] - images[n]| * inv_sigma)^3)          [N,H,W] f32, norm over channels
 * soft [P,H,W] = soft_mask_reproj, warped [P,3,H,W] = warped_img, images [N,3,H,W] f32 (0..255), pair_first [N+1] i32
 * DEVICE, every frame with at least one pair (the reference raises on an empty torch.stack).  inv_sigma = 1 / 0.1 as
 * float32 (= 10.0f).  Either output may be NULL. */
int syn3r_nearby_uncertainty_batched(const float *soft, const float *warped, const float *images, const int *pair_first,
                                     int N, int H, int W, float inv_sigma, float *out_unc, float *out_int_unc,
                                     void *stream);

/* Forward-backward flow consistency (thirdparty/gmflow/gmflow/geometry.py:75-96 forward_backward_consistency_check,
 * with flow_warp :65-72 and bilinear_sample :41-62), both directions in one launch:
 *   out_fwd_occ = |fwd + grid_sample(bwd, x + fwd)| > alpha (|fwd| + |bwd|) + beta   as float32 0 / 1, [B,H,W]
 *   out_bwd_occ = |bwd + grid_sample(fwd, x + bwd)| > the same threshold
 * fwd_flow, bwd_flow [B,2,H,W] f32 (channel 0 = x displacement), device.  out_diff_* [B,H,W] f32 (extension: the
 * two norms before thresholding).  Any output may be NULL (not all).  H, W >= 2. */
int syn3r_flow_fb_consistency_batched(const float *fwd_flow, const float *bwd_flow, int B, int H, int W, float alpha,
                                      float beta, float *out_fwd_occ, float *out_bwd_occ, float *out_diff_fwd,
                                      float *out_diff_bwd, void *stream);

/* Guided latent replacement of the patched Euler scheduler (thirdparty/diffusers/src/diffusers/schedulers/
 * scheduling_euler_discrete.py:1442-1499, step_interp_prob_uncertain), all frames in one launch, in
 * place on pred [b,F,c,h,w] f32 (pred_original_sample):
 *   for tau in [tau0, tau0 + n_tau) (the reference's index_list 1..23; frames 0 and F - 1 are skipped, see below):
 *     m_ = channel mean of ((1 - mask) > 0.5), m = m_ > 0.5, d = |pred[:,tau] m - cond[tau] m|,
 *     k = int(clamp(lambda_row[tau], 0.4, 1) * (b c h w - count(~m))) + count(~m), cutoff = k-th smallest d
 *     (the reference's torch.sort + index; a radix select here), w = x / (1 + x) with x = 1 / (1 - m_ + 1e-6),
 *     zeroed below 0.51 and where d > cutoff; pred[:,tau] = (1 - w) pred[:,tau] + w cond[tau];
 *   pred[:,0] = cond[0], pred[:,F-1] = cond[F-1] (:1496-1499).
 * cond_latents [F,c,h,w] = temp_cond_latents[1]; mask [mask_batch, mask_frames, mask_channels, h, w] f32, 1 = hole, as
 * the pipeline passes it (the scheduler prepends the frame of ones itself, so frame tau reads mask row tau - 1;
 * mask_frames is F - 2 = 23 in the SVD pipelines' step_interp* calls, F - 1 = 24 for step_single); lambda_row [F] = lambda_ts[step_index].  All
 * device pointers.  out_cutoff_index / out_cutoff_value [n_tau] optional (cutoff_index_e, cutoff_value). */
int syn3r_latent_guided_replace(float *pred, const float *cond_latents, const float *mask, const float *lambda_row, int b,
                                int F, int c, int h, int w, int mask_batch, int mask_frames, int mask_channels, int tau0,
                                int n_tau, int *out_cutoff_index, float *out_cutoff_value, void *stream);

/* The selection half of the same block as step_single (:548-586) and step_interp (:735-774) use it: identical mask
 * binarisation, sort and cutoff, with the caller's weight_clamp as the lower clamp of lambda (:577, :766) and the
 * boolean top_mask = (d <= cutoff) & m kept instead of a blend (:583, :772).  pred is only read; tau may reach F - 1
 * (their index_list runs to 24).  out_top_masks [n_tau, b, c, h, w] u8 0 / 1 (the reference concatenates the frames
 * and prepends a frame of ones, :586-589). */
int syn3r_latent_top_masks(const float *pred, const float *cond_latents, const float *mask, const float *lambda_row, int b,
                           int F, int c, int h, int w, int mask_batch, int mask_frames, int mask_channels, int tau0, int n_tau,
                           float weight_clamp, uint8_t *out_top_masks, int *out_cutoff_index, float *out_cutoff_value,
                           void *stream);

/* Score of a pose candidate (model/diffusionGS.py:759-761): out_mean[p] = mean over the frame of
 * 1 - soft[p] (float32 element arithmetic as in the reference, deterministic fixed-order reduction).  soft [P,H,W] is
 * the soft_mask_reproj output of syn3r_inverse_warp_batched for the (nearest reference, candidate) pairs.
 * workspace: syn3r_mean_uncertainty_workspace_bytes(P) bytes, caller-owned. */
size_t syn3r_mean_uncertainty_workspace_bytes(int P);
int syn3r_mean_uncertainty_batched(const float *soft, int P, int H, int W, float *out_mean, void *workspace,
                                   size_t workspace_bytes, void *stream);

/* A/B switch of the lean consistency kernel (the syn3r_inverse_warp_batched call that asks only for mask_reproj /
 * soft_mask_reproj / mask_warp): 1 (default) = k_reproj_masks2, two rows per thread in packed fp32 instructions;
 * 0 = k_reproj_masks, one row per thread.  Both produce the same bits (tests/test_gpu_parity.py::
 * test_packed_reproj_matches_scalar); the switch exists for that test and for timing.  Process-wide; returns the
 * previous setting.  No reference counterpart. */
int syn3r_set_reproj_packed(int on);

/* A/B switch of the forward tile kernel in PAIR mode with per-(pair, tile) block lists: 1 = k_fwd_tile_fx_ov, whose
 * resolve step is split so that its second phase overlaps the next work item's first sweep (three barriers per item
 * instead of five); 0 (default) = k_fwd_tile_fx.  Measured equal (1.191 vs 1.192 ms on config B), so the simpler kernel
 * stays the default.  Same bits either way
 * (tests/test_gpu_parity.py::test_overlapped_tile_kernel_matches); for that test and for timing.  Process-wide; returns
 * the previous setting.  No reference counterpart. */
int syn3r_set_fx_overlap(int on);

/* syn3r_mean_uncertainty_batched without a workspace, for callers that cannot allocate device memory (code under CUDA
 * graph capture, tight serving loops, harnesses that forbid cudaMalloc).  No counterpart in the reference.
 * The chunk partials live in a file-scope device array inside the library
 * (SYN3R_MEAN_UNCERTAINTY_STATIC_MAX_P x 64 doubles = 2 MiB on every GPU that loads the library, until exit); the call
 * allocates nothing (no cudaMalloc / cudaFree, no allocating library routine).  P above the maximum returns
 * SYN3R_EWORKSPACE with a message.  Same arguments, arithmetic and result bits as syn3r_mean_uncertainty_batched.
 * All calls share the one array, which keeps the previous call's contents (nothing needs to start at zero: every
 * partial that is read is written first); calls that OVERLAP on the GPU - different streams with no ordering between
 * them, on the same device - race on it.  Calls on one stream, or ordered by events, are safe. */
#define SYN3R_MEAN_UNCERTAINTY_STATIC_MAX_P 4096
int syn3r_mean_uncertainty_static_max_p(void);
int syn3r_mean_uncertainty_static(const float *soft, int P, int H, int W, float *out_mean, void *stream);

#ifdef __cplusplus
}
#endif
#endif /* SYN3R_WARP_H */
