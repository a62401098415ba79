// forward_kernels.cuh — forward depth warp: unproject -> reproject -> soft-z bilinear splat -> resolve.
// Replaces compute_transformed_points / forward_warp / bilinear_splatting of the reference
// (solver_utils/forward_warp.py:7-37, :141-182, :41-129); step numbers refer to SURVEY Appendix A.1.
#include <math_constants.h>

#pragma once
#include "common.cuh"

namespace syn3r {

struct FwdGeom {           // per-pixel result of A.1 steps 3-9 for one pair
    double ox, oy;         // trans_pos_offset (unclipped), :63
    double z;              // z' = q[2], :174
    double flow_x, flow_y; // :177
};

// A.1 steps 5-9 for one pair from the unprojected point X (steps 3-4 are per source pixel).
__device__ __forceinline__ FwdGeom project_pair(const double *par, double X0, double X1, double X2, int x, int y) {
    const double *M = par, *K2 = par + 12;
    double t0 = dot4_blas_h(M + 0, X0, X1, X2);
    double t1 = dot4_blas_h(M + 4, X0, X1, X2);
    double t2 = dot4_blas_h(M + 8, X0, X1, X2);
    double q0 = dot3_blas(K2 + 0, t0, t1, t2);
    double q1 = dot3_blas(K2 + 3, t0, t1, t2);
    double q2 = dot3_blas(K2 + 6, t0, t1, t2);
    FwdGeom g;
    g.z = q2;
    double cx = __ddiv_rn(q0, q2), cy = __ddiv_rn(q1, q2);          // :172
    g.flow_x = __dsub_rn(cx, (double)x);                             // :177
    g.flow_y = __dsub_rn(cy, (double)y);
    g.ox = __dadd_rn(__dadd_rn(g.flow_x, (double)x), 1.0);           // :61, :63 (double rounding kept, N4)
    g.oy = __dadd_rn(__dadd_rn(g.flow_y, (double)y), 1.0);
    return g;
}

__device__ __forceinline__ double clip_f64(double v, double lo, double hi) {  // np.clip keeps NaN
    return (v != v) ? v : (v < lo ? lo : (v > hi ? hi : v));
}

// One axis of A.1 steps 9-11: kept corner indices (unpadded, -1 = lands on the cropped ring) and
// the two proximity factors.  floor/ceil come from the UNCLIPPED offset (:64-65); a value that
// does not fit int64 (NaN, inf, |v| >= 2^63) becomes INT64_MIN on x86 and is clipped to 0, i.e.
// the ring.  Whenever the clip of :66-71 changes the offset, both corners are on the ring, so the
// kept corners never see a clipped value.
__device__ __forceinline__ void axis_corners(double o, int n, int &i_fl, int &i_ce, double &a_fl, double &a_ce) {
    double fl = floor(o), ce = ceil(o);
    bool fits = (o >= -9223372036854775808.0) && (o < 9223372036854775808.0);
    i_fl = (fits && fl >= 1.0 && fl <= (double)n) ? (int)fl - 1 : -1;
    i_ce = (fits && ce >= 1.0 && ce <= (double)n) ? (int)ce - 1 : -1;
    a_fl = __dsub_rn(1.0, __dsub_rn(o, fl));   // 1 - (off - floor)
    a_ce = __dsub_rn(1.0, __dsub_rn(ce, o));   // 1 - (ceil - off)
}

// soft depth weight 1/dw, :82-84.  Exact variant: IEEE log / div / exp like the reference.
__device__ __forceinline__ double depth_weight_div(double z, double Lmax) {
    double ld = log(__dadd_rn(1.0, clip_f64(z, 0.0, 5000.0)));
    return exp(__dmul_rn(__ddiv_rn(ld, Lmax), 50.0));
}

struct SplatTarget {
    float4 *acc;               // [H*W] (I_r, I_g, I_b, S) scaled by 2^90
    unsigned long long *key;   // [H*W] winner keys or nullptr
    int H, W;
};

// A.1 steps 9-14 for one source pixel: up to four REDG.F32x4 (+ four 64-bit atomicMin).
template <bool WINNER>
__device__ __forceinline__ void splat_pixel(const SplatTarget &t, double ox, double oy, double z, double Lmax,
                                            float r, float g, float b, double m01, uint32_t src_lin) {
    int xf, xc, yf, yc;
    double ax, bx, ay, by;
    axis_corners(ox, t.W, xf, xc, ax, bx);
    axis_corners(oy, t.H, yf, yc, ay, by);
    if (((xf & xc) < 0) || ((yf & yc) < 0)) return;   // every corner on the ring
    double dw = depth_weight_div(z, Lmax);
    unsigned long long key = 0;
    if (WINNER) key = ((unsigned long long)zkey_of(z) << 32) | src_lin;
    // corner order of the reference: NW (fy,fx), SW (cy,fx), NE (fy,cx), SE (cy,cx), :99-107
    const int cxs[4] = {xf, xf, xc, xc}, cys[4] = {yf, yc, yf, yc};
    const double pw[4] = {__dmul_rn(ay, ax), __dmul_rn(by, ax), __dmul_rn(ay, bx), __dmul_rn(by, bx)};
#pragma unroll
    for (int c = 0; c < 4; ++c) {
        if (cxs[c] < 0 || cys[c] < 0) continue;
        double w = __ddiv_rn(__dmul_rn(pw[c], m01), dw);            // :86-89 (mask1, flow mask in m01)
        float ws = __double2float_rn(__dmul_rn(w, SYN3R_W_SCALE));
        size_t o = (size_t)cys[c] * t.W + cxs[c];
        red_add_f4(t.acc + o, __fmul_rn(r, ws), __fmul_rn(g, ws), __fmul_rn(b, ws), ws);
        if (WINNER && w > 0.0) atomicMin(t.key + o, key);
    }
}

// ---------------------------------------------------------------------------------------------
// K1: per-pair max of z' over ALL source pixels (np.max at :84 is global and propagates NaN).
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
k_fwd_zmax(const float4 *__restrict__ rgbd, const double *__restrict__ depth64, const double *__restrict__ Kinv_src,
           const double *__restrict__ pair_par, const int *__restrict__ pair_src, int H, int W,
           unsigned long long *__restrict__ zmax_key, int *__restrict__ znan) {
    const int p = blockIdx.y, s = pair_src[p];
    const size_t hw = (size_t)H * W;
    const double *Ki = Kinv_src + 9 * s, *par = pair_par + 24 * p;
    double zmax = -CUDART_INF;
    int has_nan = 0;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < hw; i += (size_t)gridDim.x * blockDim.x) {
        int y = (int)(i / W), x = (int)(i - (size_t)y * W);
        double d = depth64 ? depth64[s * hw + i] : (double)rgbd[s * hw + i].w;
        double X0 = __dmul_rn(d, dot3_blas(Ki + 0, x, y, 1.0));
        double X1 = __dmul_rn(d, dot3_blas(Ki + 3, x, y, 1.0));
        double X2 = __dmul_rn(d, dot3_blas(Ki + 6, x, y, 1.0));
        const double *M = par, *K2 = par + 12;
        double t0 = dot4_blas_h(M + 0, X0, X1, X2), t1 = dot4_blas_h(M + 4, X0, X1, X2),
               t2 = dot4_blas_h(M + 8, X0, X1, X2);
        double z = dot3_blas(K2 + 6, t0, t1, t2);
        if (z != z) has_nan = 1;
        else if (z > zmax) zmax = z;
    }
    unsigned long long k = ordered_u64(zmax);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        unsigned long long k2 = __shfl_xor_sync(0xffffffffu, k, o);
        k = k2 > k ? k2 : k;
        has_nan |= __shfl_xor_sync(0xffffffffu, has_nan, o);
    }
    if ((threadIdx.x & 31) == 0) {
        atomicMax(zmax_key + p, k);
        if (has_nan) atomicOr(znan + p, 1);
    }
}

__device__ __forceinline__ double pair_Lmax(const unsigned long long *zmax_key, const int *znan, int p) {
    if (znan[p]) return CUDART_NAN;                       // np.max propagates NaN
    double zm = unordered_f64(zmax_key[p]);
    return log(__dadd_rn(1.0, clip_f64(zm, 0.0, 5000.0))); // max of a monotone map = map of the max
}

// ---------------------------------------------------------------------------------------------
// K2: project + splat, one thread per (pair, source pixel).
// ---------------------------------------------------------------------------------------------
template <bool WINNER>
__global__ void __launch_bounds__(256)
k_fwd_splat(const float4 *__restrict__ rgbd, const double *__restrict__ depth64, const uint8_t *__restrict__ src_mask,
            const double *__restrict__ Kinv_src, const double *__restrict__ pair_par,
            const int *__restrict__ pair_src, const int *__restrict__ pair_slot, int H, int W,
            const unsigned long long *__restrict__ zmax_key, const int *__restrict__ znan,
            float4 *__restrict__ acc, unsigned long long *__restrict__ keys, double *__restrict__ out_flow) {
    const int p = blockIdx.y, s = pair_src[p];
    const int slot = pair_slot ? pair_slot[p] : p;
    const size_t hw = (size_t)H * W;
    __shared__ double sLmax;
    if (threadIdx.x == 0) sLmax = pair_Lmax(zmax_key, znan, p);
    __syncthreads();
    const double Lmax = sLmax;
    const double *Ki = Kinv_src + 9 * s, *par = pair_par + 24 * p;
    SplatTarget tgt{acc + (size_t)slot * hw, WINNER ? keys + (size_t)slot * hw : nullptr, H, W};
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < hw; i += (size_t)gridDim.x * blockDim.x) {
        int y = (int)(i / W), x = (int)(i - (size_t)y * W);
        float4 px = rgbd[s * hw + i];
        double d = depth64 ? depth64[s * hw + i] : (double)px.w;
        double X0 = __dmul_rn(d, dot3_blas(Ki + 0, x, y, 1.0));     // :31-32
        double X1 = __dmul_rn(d, dot3_blas(Ki + 3, x, y, 1.0));
        double X2 = __dmul_rn(d, dot3_blas(Ki + 6, x, y, 1.0));
        FwdGeom g = project_pair(par, X0, X1, X2, x, y);
        if (out_flow) {
            double2 *f = reinterpret_cast<double2 *>(out_flow) + (size_t)p * hw + i;
            *f = make_double2(g.flow_x, g.flow_y);
        }
        double m01 = (src_mask && !src_mask[s * hw + i]) ? 0.0 : 1.0;
        splat_pixel<WINNER>(tgt, g.ox, g.oy, g.z, Lmax, px.x, px.y, px.z, m01, (uint32_t)(s * hw + i));
    }
}

// ---------------------------------------------------------------------------------------------
// K3: resolve one slot: crop is implicit, mask = S > 0 (:112), out = I / S (:116), optional range
// check + clip + round-half-even + uint8 (:119-128), winner index.
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
k_fwd_resolve(const float4 *__restrict__ acc, const unsigned long long *__restrict__ keys, size_t hw, int C,
              int flags, float *__restrict__ out_f32, uint8_t *__restrict__ out_u8, uint8_t *__restrict__ out_mask,
              int *__restrict__ out_winner, int *__restrict__ out_status) {
    const int slot = blockIdx.y;
    const float gate = 0.6f * (float)SYN3R_W_SCALE;
    int bad = 0;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < hw; i += (size_t)gridDim.x * blockDim.x) {
        const size_t o = (size_t)slot * hw + i;
        float4 a = acc[o];
        bool m = a.w > 0.0f;                                       // NaN -> false
        if (flags & SYN3R_FWD_GATE_0P6) m = m && (a.w <= gate);
        float v[3] = {m ? __fdiv_rn(a.x, a.w) : 0.0f, m ? __fdiv_rn(a.y, a.w) : 0.0f, m ? __fdiv_rn(a.z, a.w) : 0.0f};
        out_mask[o] = (uint8_t)m;
        for (int c = 0; c < C; ++c) {
            if (out_f32) out_f32[o * C + c] = v[c];
            if (flags & SYN3R_FWD_IS_IMAGE) {
                if (!(v[c] >= 0.0f) || !(v[c] <= 256.0f)) bad = 1;
                if (out_u8) out_u8[o * C + c] = (uint8_t)__float2int_rn(fminf(fmaxf(v[c], 0.0f), 255.0f));
            }
        }
        if (out_winner) {
            unsigned long long k = keys[o];
            out_winner[o] = (k == ~0ull) ? -1 : (int)(uint32_t)(k & 0xffffffffull);
        }
    }
    if (out_status && __syncthreads_or(bad) && threadIdx.x == 0) atomicOr(out_status + slot, 1);
}


// ---------------------------------------------------------------------------------------------
// Stand-alone pieces of the reference API (not on the batched hot path).
// ---------------------------------------------------------------------------------------------

// compute_transformed_points (:7-37): q [H,W,3], X [H,W,3] f64.
__global__ void __launch_bounds__(256)
k_transformed_points(const double *__restrict__ depth, const double *__restrict__ Kinv,
                     const double *__restrict__ par, int H, int W, double *__restrict__ out_q,
                     double *__restrict__ out_X) {
    const size_t hw = (size_t)H * W;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < hw; i += (size_t)gridDim.x * blockDim.x) {
        int y = (int)(i / W), x = (int)(i - (size_t)y * W);
        double d = depth[i];
        double X0 = __dmul_rn(d, dot3_blas(Kinv + 0, x, y, 1.0));
        double X1 = __dmul_rn(d, dot3_blas(Kinv + 3, x, y, 1.0));
        double X2 = __dmul_rn(d, dot3_blas(Kinv + 6, x, y, 1.0));
        const double *M = par, *K2 = par + 12;
        double t0 = dot4_blas_h(M + 0, X0, X1, X2), t1 = dot4_blas_h(M + 4, X0, X1, X2),
               t2 = dot4_blas_h(M + 8, X0, X1, X2);
        if (out_X) { out_X[3 * i] = X0; out_X[3 * i + 1] = X1; out_X[3 * i + 2] = X2; }
        out_q[3 * i + 0] = dot3_blas(K2 + 0, t0, t1, t2);
        out_q[3 * i + 1] = dot3_blas(K2 + 3, t0, t1, t2);
        out_q[3 * i + 2] = dot3_blas(K2 + 6, t0, t1, t2);
    }
}

// bilinear_splatting (:41-129) pre-pass: global max of log-depth input `z` (the function's depth1).
__global__ void __launch_bounds__(256)
k_splat_zmax(const double *__restrict__ z, size_t hw, unsigned long long *__restrict__ zmax_key,
             int *__restrict__ znan) {
    double zmax = -CUDART_INF;
    int has_nan = 0;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < hw; i += (size_t)gridDim.x * blockDim.x) {
        double v = z[i];
        if (v != v) has_nan = 1;
        else if (v > zmax) zmax = v;
    }
    unsigned long long k = ordered_u64(zmax);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        unsigned long long k2 = __shfl_xor_sync(0xffffffffu, k, o);
        k = k2 > k ? k2 : k;
        has_nan |= __shfl_xor_sync(0xffffffffu, has_nan, o);
    }
    if ((threadIdx.x & 31) == 0) {
        atomicMax(zmax_key, k);
        if (has_nan) atomicOr(znan, 1);
    }
}

// bilinear_splatting (:41-107) from caller-supplied flow12 / depth1: frame [H,W,C] f32, C <= 3.
__global__ void __launch_bounds__(256)
k_splat_from_flow(const float *__restrict__ frame, const uint8_t *__restrict__ mask1, const double *__restrict__ z,
                  const double *__restrict__ flow, const uint8_t *__restrict__ flow_mask, int H, int W, int C,
                  const unsigned long long *__restrict__ zmax_key, const int *__restrict__ znan,
                  float4 *__restrict__ acc) {
    const size_t hw = (size_t)H * W;
    __shared__ double sLmax;
    if (threadIdx.x == 0) sLmax = pair_Lmax(zmax_key, znan, 0);
    __syncthreads();
    const double Lmax = sLmax;
    SplatTarget tgt{acc, nullptr, H, W};
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < hw; i += (size_t)gridDim.x * blockDim.x) {
        int y = (int)(i / W), x = (int)(i - (size_t)y * W);
        double ox = __dadd_rn(__dadd_rn(flow[2 * i], (double)x), 1.0);      // :61, :63
        double oy = __dadd_rn(__dadd_rn(flow[2 * i + 1], (double)y), 1.0);
        double m01 = ((mask1 && !mask1[i]) || (flow_mask && !flow_mask[i])) ? 0.0 : 1.0;
        float c0 = frame[i * C], c1 = C > 1 ? frame[i * C + 1] : 0.0f, c2 = C > 2 ? frame[i * C + 2] : 0.0f;
        splat_pixel<false>(tgt, ox, oy, z[i], Lmax, c0, c1, c2, m01, 0u);
    }
}

}  // namespace syn3r
