// Forward-backward flow consistency (gmflow), SURVEY 8f rank 4.
//
// Reference: /root/reference/thirdparty/gmflow/gmflow/geometry.py:75-96 (forward_backward_consistency_check) with
// flow_warp (:65-72) and bilinear_sample (:41-62) = F.grid_sample(bilinear, zeros, align_corners=True).
//
// One thread per pixel does BOTH directions: it reads its own forward and backward flow once (16 B), gathers the
// four neighbours of x + fwd(x) in the backward flow and of x + bwd(x) in the forward flow (flows are a few pixels,
// so the taps hit lines the neighbouring threads also touch: L1 / L2 traffic, not DRAM) and writes the two
// occlusion flags (8 B).  The reference makes ~40 elementwise passes over [B,2,H,W] tensors for the same result.
// HBM-bound: 24 algorithmic bytes per pixel.
//
// Arithmetic in float32 exactly as the ops run on the CUDA device gmflow runs on: `/ (w - 1)` on a Python scalar is a
// multiply by 1.0f / (w - 1), the unnormalisation is ((g + 1) / 2) * (w - 1), the taps are accumulated nw, ne, sw, se
// with fused multiply-adds (ATen grid_sampler_2d), norms are sqrt(a^2 + b^2) of separately rounded squares.
#pragma once
#include "common.cuh"

namespace syn3r {

struct FlowSample { float u, v; };

// F.grid_sample(feature[b], grid(sx, sy)) of a 2-channel field; fu / fv are the channel planes of this batch item
__device__ __forceinline__ FlowSample flow_sample(const float *__restrict__ fu, const float *__restrict__ fv, float sx, float sy,
                                                  int H, int W, float inv_wm1, float inv_hm1) {
    const float xg = __fsub_rn(__fmul_rn(__fmul_rn(2.0f, sx), inv_wm1), 1.0f);              // geometry.py:50-51
    const float yg = __fsub_rn(__fmul_rn(__fmul_rn(2.0f, sy), inv_hm1), 1.0f);
    const float ix = __fmul_rn(__fmul_rn(__fadd_rn(xg, 1.0f), 0.5f), (float)(W - 1));       // align_corners=True
    const float iy = __fmul_rn(__fmul_rn(__fadd_rn(yg, 1.0f), 0.5f), (float)(H - 1));
    const float x0 = floorf(ix), y0 = floorf(iy), x1 = __fadd_rn(x0, 1.0f), y1 = __fadd_rn(y0, 1.0f);
    const float wx0 = __fsub_rn(x1, ix), wx1 = __fsub_rn(ix, x0), wy0 = __fsub_rn(y1, iy), wy1 = __fsub_rn(iy, y0);
    const float w[4] = {__fmul_rn(wx0, wy0), __fmul_rn(wx1, wy0), __fmul_rn(wx0, wy1), __fmul_rn(wx1, wy1)};
    // within_bounds_2d (0 <= tap <= size - 1 on the integer-valued floats) in integers: the conversion saturates for
    // huge / infinite coordinates (out of bounds either way: above 2^24 the float x0 + 1 equals x0, both outside) and
    // gives 0 for NaN, which is excluded explicitly (a NaN coordinate fails every float comparison: no tap, output 0)
    const int jx = __float2int_rz(x0), jy = __float2int_rz(y0);
    const bool fin = (ix == ix) && (iy == iy);
    if (fin && (unsigned)jx < (unsigned)(W - 1) && (unsigned)jy < (unsigned)(H - 1)) {
        // all four taps inside the frame (everything but a one-pixel border and flows leaving the frame): no bounds
        // tests, no clamps, two row pointers per plane
        const float *pu = fu + (jy * W + jx), *pv = fv + (jy * W + jx);
        const float u00 = __ldg(pu), u01 = __ldg(pu + 1), u10 = __ldg(pu + W), u11 = __ldg(pu + W + 1);
        const float v00 = __ldg(pv), v01 = __ldg(pv + 1), v10 = __ldg(pv + W), v11 = __ldg(pv + W + 1);
        FlowSample o;
        o.u = __fmaf_rn(u11, w[3], __fmaf_rn(u10, w[2], __fmaf_rn(u01, w[1], __fmaf_rn(u00, w[0], 0.0f))));
        o.v = __fmaf_rn(v11, w[3], __fmaf_rn(v10, w[2], __fmaf_rn(v01, w[1], __fmaf_rn(v00, w[0], 0.0f))));
        return o;
    }
    const bool okx[2] = {(unsigned)jx < (unsigned)W, (unsigned)(jx + 1) < (unsigned)W && jx != INT_MAX};
    const bool oky[2] = {fin && (unsigned)jy < (unsigned)H, fin && (unsigned)(jy + 1) < (unsigned)H && jy != INT_MAX};
    // the four taps are loaded unconditionally from clamped addresses (so that all loads of a thread are in flight
    // together) and only used when in bounds
    const int cx0 = min(max(jx, 0), W - 1), cx1 = min(max(jx, -1), W - 2) + 1;      // W, H >= 2 (launcher)
    const int cy0 = min(max(jy, 0), H - 1) * W, cy1 = (min(max(jy, -1), H - 2) + 1) * W;
    const int q[4] = {cy0 + cx0, cy0 + cx1, cy1 + cx0, cy1 + cx1};
    const bool ok[4] = {okx[0] && oky[0], okx[1] && oky[0], okx[0] && oky[1], okx[1] && oky[1]};
    float tu[4], tv[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        tu[k] = __ldg(fu + q[k]);
        tv[k] = __ldg(fv + q[k]);
    }
    FlowSample o = {0.0f, 0.0f};
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        o.u = ok[k] ? __fmaf_rn(tu[k], w[k], o.u) : o.u;
        o.v = ok[k] ? __fmaf_rn(tv[k], w[k], o.v) : o.v;
    }
    return o;
}

__device__ __forceinline__ float flow_norm(float a, float b) {
    return __fsqrt_rn(__fadd_rn(__fmul_rn(a, a), __fmul_rn(b, b)));
}

// A warp owns 32 * SYN3R_FLOW_PX consecutive pixels; lane l takes pixels l, l + 32, l + 64, ...: every load and store
// instruction of the warp covers 32 consecutive pixels (coalesced, taps included when the motion is smooth), and a
// thread has 4 + 16 loads per pixel in flight for SYN3R_FLOW_PX pixels.
#ifndef SYN3R_FLOW_PX
#define SYN3R_FLOW_PX 4
#endif
#ifndef SYN3R_FLOW_MINB
#define SYN3R_FLOW_MINB 4
#endif
__global__ void __launch_bounds__(256, SYN3R_FLOW_MINB)
k_flow_fb_consistency(const float *__restrict__ fwd, const float *__restrict__ bwd, int H, int W, float alpha, float beta,
                      float inv_wm1, float inv_hm1, float *__restrict__ o_fwd_occ, float *__restrict__ o_bwd_occ,
                      float *__restrict__ o_diff_fwd, float *__restrict__ o_diff_bwd) {
    const unsigned hw = (unsigned)H * (unsigned)W;             // < 2^31 (checked by the launcher): 32-bit index math
    const unsigned i0 = (blockIdx.x * 256u + (threadIdx.x & ~31u)) * SYN3R_FLOW_PX + (threadIdx.x & 31u);
    if (i0 >= hw) return;
    const int b = blockIdx.y;
    const float *fu = fwd + (size_t)b * 2 * hw, *fv = fu + hw, *bu = bwd + (size_t)b * 2 * hw, *bv = bu + hw;
    float f0[SYN3R_FLOW_PX], f1[SYN3R_FLOW_PX], b0[SYN3R_FLOW_PX], b1[SYN3R_FLOW_PX];
#pragma unroll
    for (int j = 0; j < SYN3R_FLOW_PX; ++j) {
        const unsigned i = min(i0 + 32u * j, hw - 1u);         // past the end: recompute the last pixel, store nothing
        f0[j] = __ldg(fu + i); f1[j] = __ldg(fv + i); b0[j] = __ldg(bu + i); b1[j] = __ldg(bv + i);
    }
    int y = (int)(i0 / (unsigned)W), x = (int)(i0 - (unsigned)y * (unsigned)W);      // one division per thread
#pragma unroll
    for (int j = 0; j < SYN3R_FLOW_PX; ++j) {
        const unsigned i = min(i0 + 32u * j, hw - 1u);
        if (j > 0 && i0 + 32u * j < hw) {                      // 32 pixels on: wrap into the following rows (W may be < 32)
            x += 32;
            while (x >= W) { x -= W; ++y; }
        }
        const float xf = (float)x, yf = (float)y;
        const float mag = __fadd_rn(flow_norm(f0[j], f1[j]), flow_norm(b0[j], b1[j]));                              // :83
        const FlowSample wb = flow_sample(bu, bv, __fadd_rn(xf, f0[j]), __fadd_rn(yf, f1[j]), H, W, inv_wm1, inv_hm1);   // :85
        const FlowSample wf = flow_sample(fu, fv, __fadd_rn(xf, b0[j]), __fadd_rn(yf, b1[j]), H, W, inv_wm1, inv_hm1);   // :86
        const float df = flow_norm(__fadd_rn(f0[j], wb.u), __fadd_rn(f1[j], wb.v));                                 // :88
        const float db = flow_norm(__fadd_rn(b0[j], wf.u), __fadd_rn(b1[j], wf.v));                                 // :89
        const float thr = __fadd_rn(__fmul_rn(alpha, mag), beta);                                                   // :91
        if (i0 + 32u * j < hw) {
            const size_t q = (size_t)b * hw + i;
            if (o_fwd_occ) __stcs(o_fwd_occ + q, df > thr ? 1.0f : 0.0f);                                           // :93-94
            if (o_bwd_occ) __stcs(o_bwd_occ + q, db > thr ? 1.0f : 0.0f);
            if (o_diff_fwd) __stcs(o_diff_fwd + q, df);
            if (o_diff_bwd) __stcs(o_diff_bwd + q, db);
        }
    }
}

}  // namespace syn3r
