// abi_inverse.cu — extern "C" launchers of the backward / consistency path (include/syn3r_warp.h).
#include <math.h>
#include <string.h>

#include "inverse_packed.cuh"

using namespace syn3r;

namespace {

dim3 stride_grid2(size_t n, int per_block, int ny) {
    size_t need = (n + per_block - 1) / per_block;
    size_t cap = (size_t)sm_count() * 8;
    size_t gx = need;
    if ((size_t)ny * need > cap) {
        gx = (cap + ny - 1) / ny;
        if (gx < 1) gx = 1;
        if (gx > need) gx = need;
    }
    return dim3((unsigned)gx, (unsigned)ny, 1);
}

// RN(1 / c) as a float, or 0 when div_const_rn (inverse_kernels.cuh) must not be used for this divisor: c not a
// positive normal number of moderate size, or its mantissa all ones (the one case Markstein's theorem excludes).
// 1 - c r is exact in double (24 x 24 bits), so the candidate with the smallest residual is the nearest float.
float rounded_reciprocal(float c) {
    if (!(c >= 1.0f / 1048576.0f && c <= 1048576.0f)) return 0.0f;
    uint32_t bits;
    memcpy(&bits, &c, 4);
    if ((bits & 0x7fffffu) == 0x7fffffu) return 0.0f;
    float r = (float)(1.0 / (double)c);
    float best = r;
    double be = fabs(1.0 - (double)c * (double)r);
    const float cand[2] = {nextafterf(r, 0.0f), nextafterf(r, INFINITY)};
    for (int k = 0; k < 2; ++k) {
        const double e = fabs(1.0 - (double)c * (double)cand[k]);
        if (e < be) { be = e; best = cand[k]; }
    }
    return best;
}

CamF32 make_cam(const float *K1, const float *K1inv, const float *K2, int H, int W) {
    CamF32 c;
    memcpy(c.K1, K1, sizeof(c.K1));
    memcpy(c.K1inv, K1inv, sizeof(c.K1inv));
    memcpy(c.K2, K2, sizeof(c.K2));
    c.half_wm1 = (float)((double)(W - 1) / 2.0);
    c.half_hm1 = (float)((double)(H - 1) / 2.0);
    c.half_w = (float)W / 2.0f;
    c.half_h = (float)H / 2.0f;
    c.rc_wm1 = rounded_reciprocal(c.half_wm1);
    c.rc_hm1 = rounded_reciprocal(c.half_hm1);
    return c;
}

// smallest float x with sqrt_rn(x) >= bw, so that  sqrt_rn(e2) < bw  <=>  e2 < x  (host sqrtf is correctly rounded).
// bw NaN -> NaN (nothing passes); bw <= 0 -> 0 (no e2 is below 0); bw = inf -> inf (every finite e2 passes).
float reproj_threshold_sq(float bw) {
    if (bw != bw) return bw;
    if (!(bw > 0.0f)) return 0.0f;
    if (bw > 3.4e38f) return bw;
    volatile float x = bw * bw;                        // a few ulps from the answer (inf when bw^2 overflows)
    if (x > 3.4e38f) x = 3.4028234e38f;
    while (x > 0.0f && sqrtf(x) >= bw) x = nextafterf(x, 0.0f);
    while (sqrtf(x) < bw) { if (x >= 3.4028234e38f) return INFINITY; x = nextafterf(x, INFINITY); }
    return x;
}

}  // namespace

static int g_reproj_packed = 1;

extern "C" {

int syn3r_set_reproj_packed(int on) {
    const int was = g_reproj_packed;
    g_reproj_packed = on ? 1 : 0;
    return was;
}

size_t syn3r_inv_workspace_bytes(int n_pairs) {
    if (n_pairs <= 0) return 0;
    return ((size_t)n_pairs * 16 + 255) / 256 * 256;
}

int syn3r_inverse_warp_batched(const float *src_img, const float *src_depth, int src_depth_stride,
                               const uint8_t *src_bg,
                               const float *tgt_depth, const float *K, const float *Kinv, const float *pair_P,
                               const int32_t *pair_src, const int32_t *pair_tgt, int P, int H, int W,
                               float bandwidth, float thr_depth, float thr_depth_strict,
                               const syn3r_inv_outputs *out, void *workspace, size_t workspace_bytes, void *stream) {
    if (P == 0) return SYN3R_OK;                            // empty batch: a valid no-op
    SYN3R_REQUIRE(src_depth && tgt_depth && K && Kinv && pair_P && pair_src && pair_tgt && out,
                  "null required pointer");
    SYN3R_REQUIRE(src_img || !(out->warped_img || out->warped_masked_img), "colour outputs need src_img");
    SYN3R_REQUIRE(src_depth_stride == 1 || src_depth_stride == 4, "src_depth_stride must be 1 or 4");
    SYN3R_REQUIRE(P > 0 && H > 0 && W > 0, "non-positive size");
    SYN3R_REQUIRE((size_t)H * W < 0x7fffffffull, "H*W must fit int32");
    int rc = check_device();
    if (rc != SYN3R_OK) return rc;
    const bool need_minmax = out->mask_depth || out->mask || out->mask_inv || out->mask_depth_strict ||
                             out->warped_masked_img;
    cudaStream_t st = (cudaStream_t)stream;
    uint32_t *ws = nullptr;
    if (need_minmax) {
        if (!workspace || workspace_bytes < syn3r_inv_workspace_bytes(P) || ((uintptr_t)workspace & 255)) {
            set_error("inverse workspace: need %zu bytes, 256-byte aligned (got %zu)", syn3r_inv_workspace_bytes(P),
                      workspace_bytes);
            return SYN3R_EWORKSPACE;
        }
        ws = (uint32_t *)workspace;
    }
    CamF32 cam = make_cam(K, Kinv, K, H, W);
    InvPairs pr{pair_P, pair_src, pair_tgt};
    const size_t hw = (size_t)H * W;
    dim3 g = stride_grid2(hw, 256, P);
    const bool lean = !need_minmax && !out->warped_img && !out->warped_depth && !out->warped_bg &&
                      (out->mask_reproj || out->soft_mask_reproj || out->reproj_err);
    if (lean) {
        // trajectory hot path: only the consistency-side outputs -> the register-resident specialisation
        SYN3R_REQUIRE(P <= 65535, "lean path: at most 65535 pairs per call");
        dim3 gl((W + SYN3R_RM_COLS - 1) / SYN3R_RM_COLS, (H + SYN3R_RM_ROWS - 1) / SYN3R_RM_ROWS, P);
        const float bw_sq = reproj_threshold_sq(bandwidth), inv_bw = 1.0f / bandwidth;
        if (g_reproj_packed && !out->reproj_err && cam.rc_wm1 != 0.0f && cam.rc_hm1 != 0.0f) {
            gl.y = (H + SYN3R_RM2_ROWS - 1) / SYN3R_RM2_ROWS;
            SYN3R_REQUIRE(gl.y <= 65535, "lean path: H too large");
            // two rows per thread in packed fp32 instructions (inverse_packed.cuh): the same bits, half the issue slots
            const PkConst pkc{-0.0f, 1.0f, -1.0f};
            if (src_depth_stride == 4)
                k_reproj_masks2<4><<<gl, SYN3R_RM_COLS, 0, st>>>(cam, pr, src_depth, tgt_depth, H, W, bw_sq, inv_bw, pkc,
                                                               out->mask_reproj, out->soft_mask_reproj, out->mask_warp);
            else
                k_reproj_masks2<1><<<gl, SYN3R_RM_COLS, 0, st>>>(cam, pr, src_depth, tgt_depth, H, W, bw_sq, inv_bw, pkc,
                                                               out->mask_reproj, out->soft_mask_reproj, out->mask_warp);
            return check_cuda(cudaGetLastError(), "k_reproj_masks2");
        }
#define SYN3R_RM_LAUNCH(DS, ERR)                                                                                     \
    k_reproj_masks<DS, ERR><<<gl, SYN3R_RM_COLS, 0, st>>>(cam, pr, src_depth, tgt_depth, H, W, bandwidth, bw_sq, inv_bw, \
                                                          out->mask_reproj, out->soft_mask_reproj, out->reproj_err,   \
                                                          out->mask_warp)
        if (src_depth_stride == 4) { if (out->reproj_err) SYN3R_RM_LAUNCH(4, true); else SYN3R_RM_LAUNCH(4, false); }
        else { if (out->reproj_err) SYN3R_RM_LAUNCH(1, true); else SYN3R_RM_LAUNCH(1, false); }
#undef SYN3R_RM_LAUNCH
        return check_cuda(cudaGetLastError(), "k_reproj_masks");
    }
    if (need_minmax) {
        k_inv_init_minmax<<<(P + 255) / 256, 256, 0, st>>>(ws, P);
        k_inv_minmax<<<g, 256, 0, st>>>(cam, pr, src_depth, src_depth_stride, tgt_depth, H, W, ws);
    }
    k_inverse_warp<<<g, 256, 0, st>>>(cam, pr, src_img, src_depth, src_depth_stride, src_bg, tgt_depth, H, W, bandwidth, thr_depth,
                                      thr_depth_strict, *out, ws);
    return check_cuda(cudaGetLastError(), "inverse_warp_batched launches");
}

int syn3r_selftest_div_const(float c, unsigned long long first, unsigned long long count, int finite_only,
                              unsigned long long *mismatch_dev, float *rc_out, void *stream) {
    SYN3R_REQUIRE(mismatch_dev, "null required pointer");
    int rc = check_device();
    if (rc != SYN3R_OK) return rc;
    const float r = rounded_reciprocal(c);
    if (rc_out) *rc_out = r;
    if (r == 0.0f) return SYN3R_OK;                         // the kernels use the IEEE division for this divisor
    k_selftest_div_const<<<sm_count() * 8, 256, 0, (cudaStream_t)stream>>>(first, count, c, r, mismatch_dev, finite_only);
    return check_cuda(cudaGetLastError(), "k_selftest_div_const");
}

int syn3r_consistency_batched(const float *depth1, const float *depth2, const float *K1, const float *K1inv,
                              const float *K2, const float *pair_T, const int32_t *pair_1, const int32_t *pair_2,
                              int P, int H, int W, float *err, void *stream) {
    SYN3R_REQUIRE(depth1 && depth2 && K1 && K1inv && K2 && pair_T && pair_1 && pair_2 && err,
                  "null required pointer");
    SYN3R_REQUIRE(P > 0 && H > 0 && W > 0, "non-positive size");
    int rc = check_device();
    if (rc != SYN3R_OK) return rc;
    CamF32 cam = make_cam(K1, K1inv, K2, H, W);
    k_consistency<<<stride_grid2((size_t)H * W, 256, P), 256, 0, (cudaStream_t)stream>>>(cam, pair_T, pair_1, pair_2,
                                                                                         depth1, depth2, H, W, err);
    return check_cuda(cudaGetLastError(), "k_consistency");
}

int syn3r_points_from_depth(const float *depth, const float *Kinv, int H, int W, float *pts, void *stream) {
    SYN3R_REQUIRE(depth && Kinv && pts, "null required pointer");
    SYN3R_REQUIRE(H > 0 && W > 0, "non-positive size");
    int rc = check_device();
    if (rc != SYN3R_OK) return rc;
    CamF32 cam = make_cam(Kinv, Kinv, Kinv, H, W);
    k_points_from_depth<<<stride_grid2((size_t)H * W, 256, 1), 256, 0, (cudaStream_t)stream>>>(depth, cam, H, W, pts);
    return check_cuda(cudaGetLastError(), "k_points_from_depth");
}

int syn3r_transform_points(const float *pts, const float *T, int H, int W, float *out, void *stream) {
    SYN3R_REQUIRE(pts && T && out, "null required pointer");
    SYN3R_REQUIRE(H > 0 && W > 0, "non-positive size");
    int rc = check_device();
    if (rc != SYN3R_OK) return rc;
    Mat4F32 m;
    memcpy(m.m, T, sizeof(m.m));
    k_transform_points<<<stride_grid2((size_t)H * W, 256, 1), 256, 0, (cudaStream_t)stream>>>(pts, m, (size_t)H * W,
                                                                                              out);
    return check_cuda(cudaGetLastError(), "k_transform_points");
}

}  // extern "C"
