// inverse_kernels.cuh — backward (target -> source) nearest warp + round-trip reprojection masks.
// Replaces inverse_warp (solver_utils/forward_warp.py:187-279) and consistency_check_with_depth /
// get_points_from_depth / transform_points (solver_utils/consistency.py:6-91).  All fp32, op order
// of the torch-CPU kernels the oracle was pinned against (oracle/syn3r_oracle.c header); step
// numbers refer to SURVEY Appendix A.2 / A.3.
#pragma once
#include <math_constants.h>

#include "common.cuh"

namespace syn3r {

// float -> int32 like cvtps2dq after rounding: NaN / out of range -> INT32_MIN
__device__ __forceinline__ int to_i32(float v) {
    return (v >= -2147483648.0f && v < 2147483648.0f) ? (int)v : INT_MIN;
}
// grid_sample un-normalisation, align_corners=False: ((g + 1) * size/2) - 0.5 as one FMA
__device__ __forceinline__ float unnorm(float g, float half_size) {
    return __fmaf_rn(__fadd_rn(g, 1.0f), half_size, -0.5f);
}
// `ds` = element stride between consecutive pixels of the depth plane (1 = planar [H,W]; 4 = the
// depth lane of an interleaved [H,W,4] RGB-D buffer, so the forward path's input is reused as is)
__device__ __forceinline__ float tap(const float *__restrict__ img, int H, int W, int ds, int yy, int xx) {
    return (yy > -1 && yy < H && xx > -1 && xx < W) ? __ldg(img + ((size_t)yy * W + xx) * ds) : 0.0f;
}

struct CamF32 {            // shared by every pair of a call
    float K1[9], K1inv[9], K2[9];
    float half_wm1, half_hm1;   // (float)((W-1)/2.0), (float)((H-1)/2.0)  consistency.py:71-73
    float half_w, half_h;       // W/2, H/2 (grid_sample un-normalise)
    float rc_wm1, rc_hm1;       // correctly rounded 1 / half_wm1, 1 / half_hm1 (0 = not usable: IEEE division instead)
};

// x / c for a divisor c that is the same for every pixel, with rc = RN(1 / c) found on the host: two Markstein
// corrections.  q0 = RN(x rc) is within 2 ulps of x / c; r = x - c q exactly (one FMA), so q1 = RN(q0 + r0 rc) is a
// faithful quotient and q2 = RN(q1 + r1 rc) the correctly rounded one (Markstein's theorem: rc correctly rounded,
// mantissa of c not all ones - checked on the host), i.e. the bits of __fdiv_rn(x, c), in 5 instructions instead of
// the 11 of the IEEE sequence (MUFU.RCP, FCHK, 5 FFMA, 3 branch / reconvergence).  The theorem needs x / c to stay
// in the normal range and the remainders not to underflow: an x of +-inf gives NaN here (inf - inf) where the division
// gives inf, and a quotient below ~2^-100 may differ in its last bit; in both cases the caller's next steps (- 1,
// un-normalise, floor, bounds test) produce the same values either way (inf and NaN both end as NaN weights and no
// tap; a quotient under 2^-25 disappears in the `- 1`), which is why this is only used where it is (k_reproj_masks:
// the grid normalisation).
// tests/test_gpu_parity.py::test_const_division_exhaustive compares all 2^32 x for the divisors of the benchmark sizes.
__device__ __forceinline__ float div_const_rn(float x, float c, float rc) {
    const float q0 = __fmul_rn(x, rc);
    const float q1 = __fmaf_rn(__fmaf_rn(-c, q0, x), rc, q0);
    return __fmaf_rn(__fmaf_rn(-c, q1, x), rc, q1);
}

// self-test of div_const_rn against div.rn.f32 over a range of x bit patterns; *mismatch counts differing results
// (two NaNs count as equal)
__global__ void k_selftest_div_const(unsigned long long first, unsigned long long count, float c, float rc,
                                     unsigned long long *__restrict__ mismatch, int finite_only) {
    unsigned long long bad = 0;
    for (unsigned long long i = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x; i < count;
         i += (unsigned long long)gridDim.x * blockDim.x) {
        const float x = __uint_as_float((unsigned)(first + i));
        const float a = div_const_rn(x, c, rc), b = __fdiv_rn(x, c);
        const bool same = (__float_as_uint(a) == __float_as_uint(b)) || (a != a && b != b);
        // |quotient| < 2^-30: the caller's `- 1.0f` gives -1 whatever the last bits of the quotient are
        const bool skip = finite_only && (!(fabsf(x) < 3.0e38f) || fabsf(b) < 9.3132257e-10f);
        if (!same && !skip) { ++bad; mismatch[1] = (unsigned)(first + i); }
    }
    if (bad) atomicAdd(mismatch, bad);
}

// A.3: round-trip reprojection error in pixels for frame-1 pixel (x,y) with depth d1.
// T12 = pose2 @ inv(pose1), T21 = pose1 @ inv(pose2) (consistency.py:39), row-major 4x4.
__device__ __forceinline__ float reproj_err_px(const CamF32 &c, const float *T12, const float *T21,
                                               const float *__restrict__ depth2, int ds, int x, int y, float d1,
                                               int H, int W) {
    const float fx_ = (float)x, fy_ = (float)y;
    // consistency.py:58 — points = (K1^-1 [x,y,1]) * depth
    float X0 = __fmul_rn(dot3_t(c.K1inv + 0, fx_, fy_, 1.0f), d1);
    float X1 = __fmul_rn(dot3_t(c.K1inv + 3, fx_, fy_, 1.0f), d1);
    float X2 = __fmul_rn(dot3_t(c.K1inv + 6, fx_, fy_, 1.0f), d1);
    // :61 — to frame 2, divide by w
    float w = dot4_t(T12 + 12, X0, X1, X2, 1.0f);
    float Y0 = __fdiv_rn(dot4_t(T12 + 0, X0, X1, X2, 1.0f), w);
    float Y1 = __fdiv_rn(dot4_t(T12 + 4, X0, X1, X2, 1.0f), w);
    float Y2 = __fdiv_rn(dot4_t(T12 + 8, X0, X1, X2, 1.0f), w);
    // :67-68
    float p0 = dot3_t(c.K2 + 0, Y0, Y1, Y2), p1 = dot3_t(c.K2 + 3, Y0, Y1, Y2), p2 = dot3_t(c.K2 + 6, Y0, Y1, Y2);
    float u2 = __fdiv_rn(p0, p2), v2 = __fdiv_rn(p1, p2);
    // :71-73 — (w-1)/2 normalisation ...
    float gx = __fsub_rn(__fdiv_rn(u2, c.half_wm1), 1.0f);
    float gy = __fsub_rn(__fdiv_rn(v2, c.half_hm1), 1.0f);
    // :76 — ... sampled with align_corners=False (N6), bilinear, zeros padding
    float ix = unnorm(gx, c.half_w), iy = unnorm(gy, c.half_h);
    float xw = floorf(ix), yn = floorf(iy);
    float we = __fsub_rn(ix, xw), ee = __fsub_rn(1.0f, we), nn = __fsub_rn(iy, yn), ss = __fsub_rn(1.0f, nn);
    float c_nw = __fmul_rn(ss, ee), c_ne = __fmul_rn(ss, we), c_sw = __fmul_rn(nn, ee), c_se = __fmul_rn(nn, we);
    int ixw = to_i32(xw), iyn = to_i32(yn);
    int ixe = (ixw == INT_MIN) ? INT_MIN : ixw + 1, iys = (iyn == INT_MIN) ? INT_MIN : iyn + 1;
    float d12 = __fmul_rn(tap(depth2, H, W, ds, iyn, ixw), c_nw);
    d12 = __fmaf_rn(tap(depth2, H, W, ds, iyn, ixe), c_ne, d12);
    d12 = __fmaf_rn(tap(depth2, H, W, ds, iys, ixw), c_sw, d12);
    d12 = __fmaf_rn(tap(depth2, H, W, ds, iys, ixe), c_se, d12);
    // :79 — re-lift with the sampled depth
    float Z0 = __fmul_rn(__fdiv_rn(Y0, Y2), d12), Z1 = __fmul_rn(__fdiv_rn(Y1, Y2), d12),
          Z2 = __fmul_rn(__fdiv_rn(Y2, Y2), d12);
    // :80 — back to frame 1
    float w1 = dot4_t(T21 + 12, Z0, Z1, Z2, 1.0f);
    float B0 = __fdiv_rn(dot4_t(T21 + 0, Z0, Z1, Z2, 1.0f), w1);
    float B1 = __fdiv_rn(dot4_t(T21 + 4, Z0, Z1, Z2, 1.0f), w1);
    float B2 = __fdiv_rn(dot4_t(T21 + 8, Z0, Z1, Z2, 1.0f), w1);
    // :82-83, :86-89
    float r0 = dot3_t(c.K1 + 0, B0, B1, B2), r1 = dot3_t(c.K1 + 3, B0, B1, B2), r2 = dot3_t(c.K1 + 6, B0, B1, B2);
    float ex = __fsub_rn(__fdiv_rn(r0, r2), fx_), ey = __fsub_rn(__fdiv_rn(r1, r2), fy_);
    return __fsqrt_rn(__fmaf_rn(ey, ey, __fmul_rn(ex, ex)));
}

// A.2 steps 3-7 for target pixel (x,y): source sample index (or -1) and mask_warp.
__device__ __forceinline__ void inv_sample(const CamF32 &c, const float *P, int x, int y, float z, int H, int W,
                                           int &sidx, bool &mask_warp) {
    const float fx = c.K1[0], fy = c.K1[4], cx = c.K1[2], cy = c.K1[5];     // :198-201, skew ignored
    float c0 = __fmul_rn(__fdiv_rn(__fsub_rn((float)x, cx), fx), z);          // :208-213
    float c1 = __fmul_rn(__fdiv_rn(__fsub_rn((float)y, cy), fy), z);
    float c2 = z;
    float Q0 = dot4_t(P + 0, c0, c1, c2, 1.0f), Q1 = dot4_t(P + 4, c0, c1, c2, 1.0f),
          Q2 = dot4_t(P + 8, c0, c1, c2, 1.0f);                                // :217-221
    float xs = __fadd_rn(__fdiv_rn(__fmul_rn(fx, Q0), Q2), cx);               // :222-223
    float ys = __fadd_rn(__fdiv_rn(__fmul_rn(fy, Q1), Q2), cy);
    float gx = __fsub_rn(__fdiv_rn(__fmul_rn(2.0f, xs), (float)W), 1.0f);     // :225
    float gy = __fsub_rn(__fdiv_rn(__fmul_rn(2.0f, ys), (float)H), 1.0f);
    int xi = to_i32(rintf(unnorm(gx, c.half_w))), yi = to_i32(rintf(unnorm(gy, c.half_h)));  // nearest, ties-even
    bool inb = (xi > -1 && xi < W && yi > -1 && yi < H);
    sidx = inb ? yi * W + xi : -1;
    mask_warp = (xs >= 0.0f) && (xs < (float)W) && (ys >= 0.0f) && (ys < (float)H);   // :233
}

struct InvPairs {
    const float *pair_P;       // [P,2,16]
    const int *pair_src, *pair_tgt;
};

// pass A (:236-240): per-pair max and positive-min of the nearest-sampled source depth.
// ws_minmax[p] = {ordered max key, ordered min key, nan flag}
__global__ void __launch_bounds__(256)
k_inv_minmax(CamF32 c, InvPairs pr, const float *__restrict__ src_depth, int ds,
             const float *__restrict__ tgt_depth, int H, int W, uint32_t *__restrict__ ws_minmax) {
    const int p = blockIdx.y, s = pr.pair_src[p], t = pr.pair_tgt[p];
    const size_t hw = (size_t)H * W;
    const float *P = pr.pair_P + (size_t)p * 32;
    const float *sd = src_depth + s * hw * ds, *td = tgt_depth + t * hw;
    float dmax = -CUDART_INF_F, dmin = CUDART_INF_F;
    int has_nan = 0;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < hw; i += (size_t)gridDim.x * blockDim.x) {
        int y = (int)(i / W), x = (int)(i - (size_t)y * W), sidx;
        bool mw;
        inv_sample(c, P, x, y, td[i], H, W, sidx, mw);
        float wd = sidx >= 0 ? __ldg(sd + (size_t)sidx * ds) : 0.0f;
        if (wd != wd) has_nan = 1;
        else if (wd > dmax) dmax = wd;
        float v = (wd > 0.0f) ? wd : 1e4f;
        if (v < dmin) dmin = v;
    }
    uint32_t kmax = ordered_u32(dmax), kmin = ordered_u32(dmin);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        kmax = max(kmax, __shfl_xor_sync(0xffffffffu, kmax, o));
        kmin = min(kmin, __shfl_xor_sync(0xffffffffu, kmin, o));
        has_nan |= __shfl_xor_sync(0xffffffffu, has_nan, o);
    }
    if ((threadIdx.x & 31) == 0) {
        atomicMax(ws_minmax + 4 * p + 0, kmax);
        atomicMin(ws_minmax + 4 * p + 1, kmin);
        if (has_nan) atomicOr(ws_minmax + 4 * p + 2, 1u);
    }
}

__global__ void k_inv_init_minmax(uint32_t *ws, int P) {
    int p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p < P) { ws[4 * p] = 0u; ws[4 * p + 1] = 0xffffffffu; ws[4 * p + 2] = 0u; ws[4 * p + 3] = 0u; }
}

// pass B: everything else of A.2, one thread per (pair, target pixel).
__global__ void __launch_bounds__(256)
k_inverse_warp(CamF32 c, InvPairs pr, const float *__restrict__ src_img, const float *__restrict__ src_depth,
               int ds, const uint8_t *__restrict__ src_bg, const float *__restrict__ tgt_depth, int H, int W, float bandwidth,
               float thr_depth, float thr_strict, syn3r_inv_outputs o, const uint32_t *__restrict__ ws_minmax) {
    const int p = blockIdx.y, s = pr.pair_src[p], t = pr.pair_tgt[p];
    const size_t hw = (size_t)H * W;
    const float *P = pr.pair_P + (size_t)p * 32, *Pinv = P + 16;
    const float *sd = src_depth + s * hw * ds, *td = tgt_depth + t * hw;
    const float *si = src_img ? src_img + 3 * s * hw : nullptr;
    float d_min = 0.0f, range = 0.0f;
    if (ws_minmax) {
        float d_max = ws_minmax[4 * p + 2] ? CUDART_NAN_F : unordered_f32(ws_minmax[4 * p]);
        d_min = unordered_f32(ws_minmax[4 * p + 1]);
        range = __fsub_rn(d_max, d_min);
    }
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < hw; i += (size_t)gridDim.x * blockDim.x) {
        int y = (int)(i / W), x = (int)(i - (size_t)y * W), sidx;
        bool mw;
        const float z = td[i];
        inv_sample(c, P, x, y, z, H, W, sidx, mw);
        const size_t q = (size_t)p * hw + i;
        float wd = sidx >= 0 ? __ldg(sd + (size_t)sidx * ds) : 0.0f;
        bool pos = wd > 0.0f;
        float rgb[3];
#pragma unroll
        for (int ch = 0; ch < 3; ++ch) rgb[ch] = (si && sidx >= 0) ? __ldg(si + ch * hw + sidx) : 0.0f;
        bool m = false;
        if (ws_minmax) {
            float nwd = pos ? __fdiv_rn(__fsub_rn(wd, d_min), range) : 0.0f;     // :242-245
            float ndp = __fdiv_rn(__fsub_rn(z, d_min), range);                   // :248
            float dd = fabsf(__fsub_rn(nwd, ndp));
            bool md = dd < thr_depth, mds = dd < thr_strict;                     // :250-251
            m = mw && md;                                                        // :255
            if (o.mask_depth) o.mask_depth[q] = md;
            if (o.mask_depth_strict) o.mask_depth_strict[q] = mds;
            if (o.mask) o.mask[q] = m;
            if (o.mask_inv) o.mask_inv[q] = !m;
            if (o.warped_masked_img)
#pragma unroll
                for (int ch = 0; ch < 3; ++ch)
                    o.warped_masked_img[((size_t)p * 3 + ch) * hw + i] = __fmul_rn(rgb[ch], m ? 1.0f : 0.0f);  // :270
        }
        if (o.warped_img)
#pragma unroll
            for (int ch = 0; ch < 3; ++ch) o.warped_img[((size_t)p * 3 + ch) * hw + i] = rgb[ch];
        if (o.warped_depth) o.warped_depth[q] = pos ? wd : 0.0f;                 // :244
        if (o.mask_warp) o.mask_warp[q] = mw;
        if (o.warped_bg) o.warped_bg[q] = (src_bg && sidx >= 0) ? (src_bg[s * hw + sidx] != 0) : 0;   // :232
        if (o.mask_reproj || o.soft_mask_reproj || o.reproj_err) {
            // :259 — frame 1 = target (depth_pseudo, pose2), frame 2 = source (depth, pose1)
            float e = reproj_err_px(c, P, Pinv, sd, ds, x, y, z, H, W);
            if (o.reproj_err) o.reproj_err[q] = e;
            if (o.mask_reproj) o.mask_reproj[q] = (e < bandwidth) && mw;         // :260, :267
            if (o.soft_mask_reproj) {
                float r = __fdiv_rn(e, bandwidth);
                o.soft_mask_reproj[q] = expf(-__fmul_rn(__fmul_rn(r, r), r));    // :265
            }
        }
    }
}

// ---------------------------------------------------------------------------------------------
// Lean specialisation for the trajectory hot path: only the outputs the reference's callers consume
// from the consistency side (mask_reproj, soft_mask_reproj; optionally mask_warp / reproj_err;
// model/diffusionGS.py:758,1346-1347,1486-1487).  Same arithmetic, same op order and therefore the
// same bits as k_inverse_warp, but organised for instruction count: one thread owns a pixel COLUMN
// of a (pair, 128 x ROWS) tile, the pair's two 4x4 matrices live in registers, everything that only
// depends on x (or on y) is hoisted, the homogeneous divide is skipped when w == 1.0f exactly
// (x / 1.0f == x for every x, NaN included) and Y2/Y2 is formed without a division.
// ---------------------------------------------------------------------------------------------
#define SYN3R_RM_ROWS 8
#define SYN3R_RM_COLS 128

template <int DS>
__device__ __forceinline__ float tap_ds(const float *__restrict__ img, int H, int W, int yy, int xx) {
    return ((unsigned)yy < (unsigned)H && (unsigned)xx < (unsigned)W) ? __ldg(img + (size_t)(yy * W + xx) * DS) : 0.0f;
}

// ERR = false (reproj_err not requested, the trajectory case): the error itself is never formed.
//   mask_reproj   e < bandwidth with e = sqrt_rn(e2)  <=>  e2 < bw_sq, where bw_sq is the smallest float whose
//                 correctly rounded square root reaches the bandwidth (sqrt_rn is monotone; found on the host by
//                 reproj_threshold_sq()): the mask stays bit-exact without the IEEE square root;
//   soft mask     exp(-(e / bw)^3) through sqrt.approx, a reciprocal multiply and ex2.approx: |error| <= 5e-7
//                 (the bar for this output is 2e-6, it is compared against a host exp() anyway).
template <int DS, bool ERR>
__global__ void __launch_bounds__(SYN3R_RM_COLS)
k_reproj_masks(CamF32 c, InvPairs pr, const float *__restrict__ src_depth, const float *__restrict__ tgt_depth,
               int H, int W, float bandwidth, float bw_sq, float inv_bw, uint8_t *__restrict__ o_mask_reproj,
               float *__restrict__ o_soft, float *__restrict__ o_err, uint8_t *__restrict__ o_mask_warp) {
    const int p = blockIdx.z, s = pr.pair_src[p], t = pr.pair_tgt[p];
    const int x = blockIdx.x * SYN3R_RM_COLS + threadIdx.x;
    const int y0 = blockIdx.y * SYN3R_RM_ROWS;
    const size_t hw = (size_t)H * W;
    __shared__ float s_yn[SYN3R_RM_ROWS];
    const float fx = c.K1[0], fy = c.K1[4], cx = c.K1[2], cy = c.K1[5];      // forward_warp.py:198-201
    if (threadIdx.x < SYN3R_RM_ROWS)
        s_yn[threadIdx.x] = __fdiv_rn(__fsub_rn((float)(y0 + threadIdx.x), cy), fy);   // :209, per row
    __syncthreads();
    if (x >= W) return;
    float A[16], B[16];          // A = pose1 @ inv(pose2)  (:217; also T12 of consistency.py:39), B = its inverse pair
    {
        const float4 *Pm = reinterpret_cast<const float4 *>(pr.pair_P + (size_t)p * 32);
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            float4 a = __ldg(Pm + k), b = __ldg(Pm + 4 + k);
            A[4 * k] = a.x; A[4 * k + 1] = a.y; A[4 * k + 2] = a.z; A[4 * k + 3] = a.w;
            B[4 * k] = b.x; B[4 * k + 1] = b.y; B[4 * k + 2] = b.z; B[4 * k + 3] = b.w;
        }
    }
    const float *sd = src_depth + (size_t)s * hw * DS;
    const float *td = tgt_depth + (size_t)t * hw;
    const float xf = (float)x;
    const float xn = __fdiv_rn(__fsub_rn(xf, cx), fx);                        // :208
    const float kx0 = __fmul_rn(c.K1inv[0], xf), kx1 = __fmul_rn(c.K1inv[3], xf), kx2 = __fmul_rn(c.K1inv[6], xf);
    const float Wf = (float)W, Hf = (float)H;
    const bool fastc = c.rc_wm1 != 0.0f && c.rc_hm1 != 0.0f;
    const int rows = min(SYN3R_RM_ROWS, H - y0);
    float z_next = __ldg(td + (size_t)y0 * W + x);            // one row ahead: the DRAM latency hides behind a row of math
    for (int r = 0; r < rows; ++r) {
        const int y = y0 + r;
        const float yf = (float)y;
        const size_t i = (size_t)y * W + x;
        const float z = z_next;
        if (r + 1 < rows) z_next = __ldg(td + i + W);
        // ---- A.2 steps 3-7: where the target pixel falls in the source view -> mask_warp (:233) -----
        const float c0 = __fmul_rn(xn, z), c1 = __fmul_rn(s_yn[r], z);
        const float Q0 = dot4_t(A + 0, c0, c1, z, 1.0f), Q1 = dot4_t(A + 4, c0, c1, z, 1.0f),
                    Q2 = dot4_t(A + 8, c0, c1, z, 1.0f);
        const float xs = __fadd_rn(__fdiv_rn(__fmul_rn(fx, Q0), Q2), cx);    // :222-223
        const float ys = __fadd_rn(__fdiv_rn(__fmul_rn(fy, Q1), Q2), cy);
        const bool mw = (xs >= 0.0f) && (xs < Wf) && (ys >= 0.0f) && (ys < Hf);
        // ---- A.3: round trip target -> source -> target ------------------------------------------------
        float X0 = __fmul_rn(__fmaf_rn(c.K1inv[2], 1.0f, __fmaf_rn(c.K1inv[1], yf, kx0)), z);   // consistency.py:58
        float X1 = __fmul_rn(__fmaf_rn(c.K1inv[5], 1.0f, __fmaf_rn(c.K1inv[4], yf, kx1)), z);
        float X2 = __fmul_rn(__fmaf_rn(c.K1inv[8], 1.0f, __fmaf_rn(c.K1inv[7], yf, kx2)), z);
        float Y0 = dot4_t(A + 0, X0, X1, X2, 1.0f), Y1 = dot4_t(A + 4, X0, X1, X2, 1.0f),
              Y2 = dot4_t(A + 8, X0, X1, X2, 1.0f);
        const float w = dot4_t(A + 12, X0, X1, X2, 1.0f);                                        // :37-40
        if (w != 1.0f) { Y0 = __fdiv_rn(Y0, w); Y1 = __fdiv_rn(Y1, w); Y2 = __fdiv_rn(Y2, w); }
        const float p0 = dot3_t(c.K2 + 0, Y0, Y1, Y2), p1 = dot3_t(c.K2 + 3, Y0, Y1, Y2),
                    p2 = dot3_t(c.K2 + 6, Y0, Y1, Y2);                                           // :67-68
        const float u2 = __fdiv_rn(p0, p2), v2 = __fdiv_rn(p1, p2);
        const float gx = __fsub_rn(fastc ? div_const_rn(u2, c.half_wm1, c.rc_wm1) : __fdiv_rn(u2, c.half_wm1), 1.0f);   // :71-73
        const float gy = __fsub_rn(fastc ? div_const_rn(v2, c.half_hm1, c.rc_hm1) : __fdiv_rn(v2, c.half_hm1), 1.0f);
        const float ix = unnorm(gx, c.half_w), iy = unnorm(gy, c.half_h);                        // :76 (N6)
        const float xw = floorf(ix), yn_ = floorf(iy);
        const float we = __fsub_rn(ix, xw), ee = __fsub_rn(1.0f, we), nn = __fsub_rn(iy, yn_), ss = __fsub_rn(1.0f, nn);
        const int ixw = to_i32(xw), iyn = to_i32(yn_);
        float d12;
        // all four taps of every lane inside the frame (everything but samples that leave it or touch its border):
        // no bounds tests, one address, the taps at fixed offsets of it
        if (__all_sync(0xffffffffu, (unsigned)ixw < (unsigned)(W - 1) && (unsigned)iyn < (unsigned)(H - 1))) {
            const float *t0 = sd + (size_t)(iyn * W + ixw) * DS, *t1 = t0 + (size_t)W * DS;
            const float nw = __ldg(t0), ne = __ldg(t0 + DS), sw = __ldg(t1), se = __ldg(t1 + DS);
            d12 = __fmul_rn(nw, __fmul_rn(ss, ee));
            d12 = __fmaf_rn(ne, __fmul_rn(ss, we), d12);
            d12 = __fmaf_rn(sw, __fmul_rn(nn, ee), d12);
            d12 = __fmaf_rn(se, __fmul_rn(nn, we), d12);
        } else {
            const int ixe = (ixw == INT_MIN) ? INT_MIN : ixw + 1, iys = (iyn == INT_MIN) ? INT_MIN : iyn + 1;
            d12 = __fmul_rn(tap_ds<DS>(sd, H, W, iyn, ixw), __fmul_rn(ss, ee));
            d12 = __fmaf_rn(tap_ds<DS>(sd, H, W, iyn, ixe), __fmul_rn(ss, we), d12);
            d12 = __fmaf_rn(tap_ds<DS>(sd, H, W, iys, ixw), __fmul_rn(nn, ee), d12);
            d12 = __fmaf_rn(tap_ds<DS>(sd, H, W, iys, ixe), __fmul_rn(nn, we), d12);
        }
        // :79 — (X2 / X2.z) * d12; Y2/Y2 is 1 for finite non-zero Y2 and NaN otherwise
        const float Z0 = __fmul_rn(__fdiv_rn(Y0, Y2), d12), Z1 = __fmul_rn(__fdiv_rn(Y1, Y2), d12);
        const float aY2 = fabsf(Y2);
        const float Z2 = (aY2 > 0.0f && aY2 < CUDART_INF_F) ? d12 : CUDART_NAN_F;
        float B0 = dot4_t(B + 0, Z0, Z1, Z2, 1.0f), B1 = dot4_t(B + 4, Z0, Z1, Z2, 1.0f),
              B2 = dot4_t(B + 8, Z0, Z1, Z2, 1.0f);
        const float w1 = dot4_t(B + 12, Z0, Z1, Z2, 1.0f);                                       // :80
        if (w1 != 1.0f) { B0 = __fdiv_rn(B0, w1); B1 = __fdiv_rn(B1, w1); B2 = __fdiv_rn(B2, w1); }
        const float r0 = dot3_t(c.K1 + 0, B0, B1, B2), r1 = dot3_t(c.K1 + 3, B0, B1, B2),
                    r2 = dot3_t(c.K1 + 6, B0, B1, B2);                                           // :82-83
        const float ex = __fsub_rn(__fdiv_rn(r0, r2), xf), ey = __fsub_rn(__fdiv_rn(r1, r2), yf);
        const float e2 = __fmaf_rn(ey, ey, __fmul_rn(ex, ex));                                   // :86-89
        const size_t q = (size_t)p * hw + i;
        if (o_mask_warp) o_mask_warp[q] = mw;
        if (ERR) {
            const float e = __fsqrt_rn(e2);
            o_err[q] = e;
            if (o_mask_reproj) o_mask_reproj[q] = (e < bandwidth) && mw;                         // forward_warp.py:260,267
            if (o_soft) {
                const float rr = __fdiv_rn(e, bandwidth);
                o_soft[q] = expf(-__fmul_rn(__fmul_rn(rr, rr), rr));                             // :265
            }
        } else {
            if (o_mask_reproj) o_mask_reproj[q] = (e2 < bw_sq) && mw;
            if (o_soft) {
                float sr, ex2;
                asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(sr) : "f"(e2));
                const float rr = __fmul_rn(sr, inv_bw);
                asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(ex2) : "f"(__fmul_rn(__fmul_rn(__fmul_rn(rr, rr), rr), -1.44269504f)));
                o_soft[q] = ex2;
            }
        }
    }
}

// consistency_check_with_depth (consistency.py:44-91) batched.
__global__ void __launch_bounds__(256)
k_consistency(CamF32 c, const float *__restrict__ pair_T, const int *__restrict__ pair_1,
              const int *__restrict__ pair_2, const float *__restrict__ depth1, const float *__restrict__ depth2,
              int H, int W, float *__restrict__ err) {
    const int p = blockIdx.y;
    const size_t hw = (size_t)H * W;
    const float *T12 = pair_T + (size_t)p * 32, *T21 = T12 + 16;
    const float *d1 = depth1 + pair_1[p] * hw, *d2 = depth2 + pair_2[p] * hw;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < hw; i += (size_t)gridDim.x * blockDim.x) {
        int y = (int)(i / W), x = (int)(i - (size_t)y * W);
        err[(size_t)p * hw + i] = reproj_err_px(c, T12, T21, d2, 1, x, y, d1[i], H, W);
    }
}

__global__ void __launch_bounds__(256)
k_points_from_depth(const float *__restrict__ depth, CamF32 c, int H, int W, float *__restrict__ pts) {
    const size_t hw = (size_t)H * W;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < hw; i += (size_t)gridDim.x * blockDim.x) {
        int y = (int)(i / W), x = (int)(i - (size_t)y * W);
#pragma unroll
        for (int b = 0; b < 3; ++b) pts[3 * i + b] = __fmul_rn(dot3_t(c.K1inv + 3 * b, (float)x, (float)y, 1.0f), depth[i]);
    }
}

struct Mat4F32 { float m[16]; };
__global__ void __launch_bounds__(256)
k_transform_points(const float *__restrict__ pts, Mat4F32 T, size_t hw, float *__restrict__ out) {
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < hw; i += (size_t)gridDim.x * blockDim.x) {
        float a = pts[3 * i], b = pts[3 * i + 1], cc = pts[3 * i + 2];
        float w = dot4_t(T.m + 12, a, b, cc, 1.0f);
#pragma unroll
        for (int m = 0; m < 3; ++m) out[3 * i + m] = __fdiv_rn(dot4_t(T.m + 4 * m, a, b, cc, 1.0f), w);
    }
}

}  // namespace syn3r
