// postprocess.cuh — caller-side post-processing of a warped frame, on the device.
// Replaces the per-frame host block of /root/reference/model/diffusionGS.py:1572-1597 (`warp_images`) and
// :1452-1491 (`warp_images_bw`): hole = 1 - valid -> 5x5 dilation (cv2.dilate, outside ignored) -> mask png,
// frame with the dilated holes blanked (uint8 truncation, / 255), 8x8 block mean of the dilated mask thresholded
// at 0.2 (= 13 of 64 pixels), 8x8 block mean of 1 - soft_mask_reproj in NumPy's own summation order.
//
// HBM-bound byte work: one CTA owns a 256 x 16 pixel tile of one frame (32 x 2 latent cells); the valid mask is
// staged in shared memory with its 2-pixel halo, dilated separably (5-wide OR along x, then along y), and every
// input and output byte is touched once.  Algorithmic bytes per pixel: valid 1 + image 3 (uint8) or 12 (fp32) in;
// mask png 1, blanked frame 3 (uint8) and / or 12 (fp32), soft 4 in; 1/64 + 4/64 for the latent outputs.
#pragma once
#include "common.cuh"

namespace syn3r {

#define SYN3R_PP_TW 256
#define SYN3R_PP_TH 16

struct PostArgs {
    const uint8_t *valid;      // [P,H,W] 1 = covered
    const uint8_t *img_u8;     // [P,H,W,3] or null
    const float *img_chw;      // [P,3,H,W] or null (inverse_warp's warped_img layout, values in [0, 256))
    const float *soft;         // [P,H,W] or null
    int H, W;
    uint8_t *mask_png;         // [P,H,W] 0 / 255 or null
    uint8_t *cond_u8;          // [P,H,W,3] or null
    float *cond_f32;           // [P,H,W,3] or null
    float *cond_ori;           // [P,H,W,3] = image / 255 (backward branch, :1467) or null
    uint8_t *latent;           // [P,H/8,W/8] 0 / 1 or null
    float *soft_pooled;        // [P,H/8,W/8] or null
};

__global__ void __launch_bounds__(SYN3R_PP_TW)
k_condition_post(const PostArgs a) {
    __shared__ uint8_t s_hole[SYN3R_PP_TH + 4][SYN3R_PP_TW + 8];
    __shared__ uint8_t s_hor[SYN3R_PP_TH + 4][SYN3R_PP_TW];
    const int t = threadIdx.x, p = blockIdx.z;
    const int x0 = blockIdx.x * SYN3R_PP_TW, y0 = blockIdx.y * SYN3R_PP_TH;
    const int H = a.H, W = a.W;
    const size_t hw = (size_t)H * W;
    const uint8_t *valid = a.valid + (size_t)p * hw;

    // ---- stage hole = !valid with a 2-pixel halo; outside the image there is no hole (cv2.dilate ignores it) ----
    for (int r = 0; r < SYN3R_PP_TH + 4; ++r) {
        const int gy = y0 - 2 + r;
        for (int c = t; c < SYN3R_PP_TW + 4; c += SYN3R_PP_TW) {
            const int gx = x0 - 2 + c;
            uint8_t h = 0;
            if ((unsigned)gy < (unsigned)H && (unsigned)gx < (unsigned)W) h = valid[(size_t)gy * W + gx] == 0;
            s_hole[r][c] = h;
        }
    }
    __syncthreads();
#pragma unroll
    for (int r = 0; r < SYN3R_PP_TH + 4; ++r)
        s_hor[r][t] = s_hole[r][t] | s_hole[r][t + 1] | s_hole[r][t + 2] | s_hole[r][t + 3] | s_hole[r][t + 4];
    __syncthreads();

    const int gx = x0 + t;
    const bool col_in = gx < W;
    unsigned dil = 0;                                       // bit r: pixel (y0 + r, gx) is inside a dilated hole
#pragma unroll
    for (int r = 0; r < SYN3R_PP_TH; ++r) {
        const unsigned d = s_hor[r][t] | s_hor[r + 1][t] | s_hor[r + 2][t] | s_hor[r + 3][t] | s_hor[r + 4][t];
        dil |= (d ? 1u : 0u) << r;
    }

    // ---- per-pixel outputs ------------------------------------------------------------------------------------------
    for (int r = 0; r < SYN3R_PP_TH; ++r) {
        const int gy = y0 + r;
        if (gy >= H || !col_in) break;
        const size_t i = (size_t)gy * W + gx, q = (size_t)p * hw + i;
        const bool d = (dil >> r) & 1u;
        if (a.mask_png) a.mask_png[q] = d ? 255 : 0;                                          // :1580-1581
        if (a.img_u8 || a.img_chw) {
            unsigned c[3];
            float f[3] = {0.f, 0.f, 0.f};
#pragma unroll
            for (int k = 0; k < 3; ++k) {
                if (a.img_u8) {
                    c[k] = a.img_u8[q * 3 + k];
                } else {
                    f[k] = a.img_chw[((size_t)p * 3 + k) * hw + i];
                    c[k] = (unsigned)__float2int_rz(f[k]) & 0xffu;                          // np.uint8(.): truncation
                }
                if (d) c[k] = 0;                                                              // * (1 - mask_erosion_), :1586
            }
#pragma unroll
            for (int k = 0; k < 3; ++k) {
                if (a.cond_u8) a.cond_u8[q * 3 + k] = (uint8_t)c[k];
                if (a.cond_f32) a.cond_f32[q * 3 + k] = __fdiv_rn((float)c[k], 255.0f);       // :1590
                if (a.cond_ori) a.cond_ori[q * 3 + k] = __fdiv_rn(a.img_u8 ? (float)a.img_u8[q * 3 + k] : f[k], 255.0f);
            }
        }
    }

    // ---- latent-resolution outputs: 8x8 blocks = 8 adjacent lanes x 8 rows -------------------------------------------
#pragma unroll
    for (int half = 0; half < SYN3R_PP_TH / 8; ++half) {
        const int gy0 = y0 + 8 * half;
        const bool live = col_in && gy0 < H;                // H, W are multiples of 8: a block is inside or outside
        unsigned cnt = live ? __popc((dil >> (8 * half)) & 0xffu) : 0u;
        float rs = 0.0f;
        if (a.soft_pooled && live) {
            const float *s = a.soft + (size_t)p * hw + (size_t)gy0 * W + gx;
            rs = __fsub_rn(1.0f, s[0]);                     // r[j] = a[j]; r[j] += a[8 i + j]  (NumPy pairwise_sum, n = 64)
#pragma unroll
            for (int i = 1; i < 8; ++i) rs = __fadd_rn(rs, __fsub_rn(1.0f, s[(size_t)i * W]));
        }
#pragma unroll
        for (int o = 1; o < 8; o <<= 1) {                   // ((r0+r1)+(r2+r3))+((r4+r5)+(r6+r7)): fp add is commutative
            cnt += __shfl_xor_sync(0xffffffffu, cnt, o);
            rs = __fadd_rn(rs, __shfl_xor_sync(0xffffffffu, rs, o));
        }
        if (live && (t & 7) == 0) {
            const size_t lq = ((size_t)p * (H / 8) + gy0 / 8) * (W / 8) + gx / 8;
            if (a.latent) a.latent[lq] = cnt >= 13u;        // count / 64 >= 0.2  (:1594-1596)
            if (a.soft_pooled) a.soft_pooled[lq] = __fmul_rn(rs, 0.015625f);
        }
    }
}

}  // namespace syn3r
