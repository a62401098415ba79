// postprocess.cuh — caller-side post-processing of a warped frame, on the device.
// Replaces the per-frame host block of /root/reference/model/diffusionGS.py:1572-1597 (`warp_images`) and
// :1452-1491 (`warp_images_bw`): hole = 1 - valid -> 5x5 dilation (cv2.dilate, outside ignored) -> mask png,
// frame with the dilated holes blanked (uint8 truncation, / 255), 8x8 block mean of the dilated mask thresholded
// at 0.2 (= 13 of 64 pixels), 8x8 block mean of 1 - soft_mask_reproj in NumPy's own summation order.
//
// HBM-bound byte work: one CTA owns a 256 x 16 pixel tile of one frame (32 x 2 latent cells); the valid mask is
// staged in shared memory with its 2-pixel halo (32-bit loads), dilated separably (5-wide OR along x, then along y)
// into a byte plane, and every input and output byte is touched once with 8- or 16-byte accesses: the uint8 frame
// as a flat run of 768 bytes per tile row (byte b belongs to pixel b / 3), the mask png 8 pixels at a time, the
// soft mask as float2 (a thread owns 2 of a block's 8 columns, its xor-1 / xor-2 lanes the other pairs).
// Algorithmic bytes per pixel: valid 1 + image 3 (uint8) or 12 (fp32) in; mask png 1, blanked frame 3 (uint8)
// and / or 12 (fp32), soft 4 in; 1/64 + 4/64 for the latent outputs.
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

__device__ __forceinline__ float u8_over_255(unsigned v) { return __fdiv_rn((float)v, 255.0f); }      // :1590

__global__ void __launch_bounds__(SYN3R_PP_TW, 5)
k_condition_post(const PostArgs a) {
    // four pixels per 32-bit word, one byte (0 / 1) each
    __shared__ uint32_t s_hole_w[SYN3R_PP_TH + 4][SYN3R_PP_TW / 4 + 4];     // columns x0 - 4 .. x0 + 259
    __shared__ uint32_t s_hor_w[SYN3R_PP_TH + 4][SYN3R_PP_TW / 4];
    __shared__ __align__(16) uint32_t s_dil_w[SYN3R_PP_TH][SYN3R_PP_TW / 4];
    uint8_t (*s_dil)[SYN3R_PP_TW] = reinterpret_cast<uint8_t (*)[SYN3R_PP_TW]>(s_dil_w);
    const int t = threadIdx.x, p = blockIdx.z;
    const int x0 = blockIdx.x * SYN3R_PP_TW, y0 = blockIdx.y * SYN3R_PP_TH;
    const int H = a.H, W = a.W;
    const size_t hw = (size_t)H * W;
    const int tw = min(SYN3R_PP_TW, W - x0), th = min(SYN3R_PP_TH, H - y0);        // multiples of 8
    const uint8_t *valid = a.valid + (size_t)p * hw;

    // ---- first (independent of the mask; its loads overlap the staging below): pooled uncertainty, the 8x8 block mean
    //      of 1 - soft in NumPy's pairwise_sum order (:1484-1487):  r[j] = a[j]; r[j] += a[8 i + j];
    //      ((r0+r1)+(r2+r3))+((r4+r5)+(r6+r7)).  A thread owns two adjacent columns of a block (all 256 threads load),
    //      lanes xor 1 and xor 2 hold the other pairs; fp32 addition is commutative, so every lane forms the same sums.
    if (a.soft_pooled) {
        const int by = t >> 7, c = (t & 127) * 2;
        const bool live = c < tw && 8 * by < th;
        float r0 = 0.f, r1 = 0.f;
        if (live) {
            const float *s = a.soft + (size_t)p * hw + (size_t)(y0 + 8 * by) * W + x0 + c;
            float2 v[8];
#pragma unroll
            for (int i = 0; i < 8; ++i) v[i] = __ldg(reinterpret_cast<const float2 *>(s + (size_t)i * W));
            r0 = __fsub_rn(1.0f, v[0].x); r1 = __fsub_rn(1.0f, v[0].y);
#pragma unroll
            for (int i = 1; i < 8; ++i) { r0 = __fadd_rn(r0, __fsub_rn(1.0f, v[i].x)); r1 = __fadd_rn(r1, __fsub_rn(1.0f, v[i].y)); }
        }
        const float pair = __fadd_rn(r0, r1);
        const float quad = __fadd_rn(pair, __shfl_xor_sync(0xffffffffu, pair, 1));
        const float all8 = __fadd_rn(quad, __shfl_xor_sync(0xffffffffu, quad, 2));
        if (live && !(t & 3))
            a.soft_pooled[((size_t)p * (H / 8) + (y0 / 8 + by)) * (W / 8) + (x0 + c) / 8] = __fmul_rn(all8, 0.015625f);
    }

    // ---- stage hole = !valid with its halo; outside the image there is no hole (cv2.dilate ignores it) ----------
    // thread (tx, ty) = (word column, row mod 4); every loop below has a compile-time trip count: the loads of all
    // its iterations are in flight together
    constexpr int WT = SYN3R_PP_TW / 4;                     // 64 words per tile row
    const int tx = t & (WT - 1), ty = t >> 6;
    {
        uint32_t v[(SYN3R_PP_TH + 4) / 4];
#pragma unroll
        for (int n = 0; n < (SYN3R_PP_TH + 4) / 4; ++n) {
            const int gy = y0 - 2 + ty + 4 * n, gx = x0 + 4 * tx;
            v[n] = 0x01010101u;                             // covered
            if ((unsigned)gy < (unsigned)H && gx < W) v[n] = __ldg(reinterpret_cast<const uint32_t *>(valid + (size_t)gy * W + gx));
        }
        uint32_t vh = 0x01010101u;                          // the two halo words of row t / 2 (threads 0..39)
        const int hr = t >> 1, hgx = (t & 1) ? x0 + SYN3R_PP_TW : x0 - 4, hgy = y0 - 2 + hr;
        if (t < 2 * (SYN3R_PP_TH + 4) && (unsigned)hgy < (unsigned)H && (unsigned)hgx < (unsigned)W)
            vh = __ldg(reinterpret_cast<const uint32_t *>(valid + (size_t)hgy * W + hgx));
#pragma unroll
        for (int n = 0; n < (SYN3R_PP_TH + 4) / 4; ++n) s_hole_w[ty + 4 * n][tx + 1] = __vcmpeq4(v[n], 0u) & 0x01010101u;
        if (t < 2 * (SYN3R_PP_TH + 4)) s_hole_w[hr][(t & 1) ? WT + 1 : 0] = __vcmpeq4(vh, 0u) & 0x01010101u;
    }
    __syncthreads();
    // separable 5x5 OR on the packed words: the neighbours at -2 .. +2 are funnel shifts over the adjacent words
#pragma unroll
    for (int n = 0; n < (SYN3R_PP_TH + 4) / 4; ++n) {
        const int r = ty + 4 * n;
        const uint32_t prev = s_hole_w[r][tx], cur = s_hole_w[r][tx + 1], next = s_hole_w[r][tx + 2];
        s_hor_w[r][tx] = cur | __funnelshift_r(cur, next, 8) | __funnelshift_r(cur, next, 16) |
                         __funnelshift_l(prev, cur, 8) | __funnelshift_l(prev, cur, 16);
    }
    __syncthreads();
#pragma unroll
    for (int n = 0; n < SYN3R_PP_TH / 4; ++n) {
        const int r = ty + 4 * n;
        s_dil_w[r][tx] = s_hor_w[r][tx] | s_hor_w[r + 1][tx] | s_hor_w[r + 2][tx] | s_hor_w[r + 3][tx] | s_hor_w[r + 4][tx];
    }
    __syncthreads();

    // ---- mask png: 8 pixels per access (:1580-1581); unit (c8, r8) = 8 pixels of row r8 + 8 n ---------------------------
    const int c8 = (t & 31) * 8, r8 = t >> 5;
    if (a.mask_png) {
#pragma unroll
        for (int n = 0; n < SYN3R_PP_TH / 8; ++n) {
            const int r = r8 + 8 * n;
            if (r < th && c8 < tw) {
                uint2 d = *reinterpret_cast<const uint2 *>(&s_dil[r][c8]);
                d.x *= 255u; d.y *= 255u;                   // bytes 0 / 1 -> 0 / 255, no carries between bytes
                *reinterpret_cast<uint2 *>(a.mask_png + (size_t)p * hw + (size_t)(y0 + r) * W + x0 + c8) = d;
            }
        }
    }

    // ---- the frame, blanked inside the dilated holes (:1586-1590) --------------------------------------------------------
    if (a.img_u8) {
        // a tile row is a flat run of 3 tw bytes; a unit = 8 pixels = 24 bytes = three 8-byte accesses, and the byte
        // mask of the unit is six byte permutations of the eight dilation bytes (byte b belongs to pixel b / 3)
#pragma unroll
        for (int n = 0; n < SYN3R_PP_TH / 8; ++n) {
            const int r = r8 + 8 * n, c = c8;
            if (r >= th || c >= tw) continue;
            const size_t base = ((size_t)p * hw + (size_t)(y0 + r) * W + x0 + c) * 3;
            const uint2 *src = reinterpret_cast<const uint2 *>(a.img_u8 + base);
            const uint2 i0 = __ldg(src), i1 = __ldg(src + 1), i2 = __ldg(src + 2);
            uint2 d = *reinterpret_cast<const uint2 *>(&s_dil[r][c]);
            d.x *= 255u; d.y *= 255u;                       // 0xff where the pixel is blanked
            const uint32_t in[6] = {i0.x, i0.y, i1.x, i1.y, i2.x, i2.y};
            const uint32_t m[6] = {__byte_perm(d.x, 0u, 0x1000), __byte_perm(d.x, 0u, 0x2211), __byte_perm(d.x, 0u, 0x3332),
                                   __byte_perm(d.y, 0u, 0x1000), __byte_perm(d.y, 0u, 0x2211), __byte_perm(d.y, 0u, 0x3332)};
            uint32_t o[6];
#pragma unroll
            for (int k = 0; k < 6; ++k) o[k] = in[k] & ~m[k];
            if (a.cond_u8) {
                uint2 *dst = reinterpret_cast<uint2 *>(a.cond_u8 + base);
                dst[0] = make_uint2(o[0], o[1]); dst[1] = make_uint2(o[2], o[3]); dst[2] = make_uint2(o[4], o[5]);
            }
            if (a.cond_f32) {
                float4 *dst = reinterpret_cast<float4 *>(a.cond_f32 + base);
#pragma unroll
                for (int k = 0; k < 6; ++k)
                    dst[k] = make_float4(u8_over_255(o[k] & 0xffu), u8_over_255((o[k] >> 8) & 0xffu),
                                         u8_over_255((o[k] >> 16) & 0xffu), u8_over_255(o[k] >> 24));
            }
            if (a.cond_ori) {                               // frame / 255 (:1467)
                float4 *dst = reinterpret_cast<float4 *>(a.cond_ori + base);
#pragma unroll
                for (int k = 0; k < 6; ++k)
                    dst[k] = make_float4(u8_over_255(in[k] & 0xffu), u8_over_255((in[k] >> 8) & 0xffu),
                                         u8_over_255((in[k] >> 16) & 0xffu), u8_over_255(in[k] >> 24));
            }
        }
    } else if (a.img_chw) {
        // float32 planes (inverse_warp's warped_img): 4 pixels per thread and row, np.uint8() truncation (:1468)
        for (int u = t; u < SYN3R_PP_TH * (SYN3R_PP_TW / 4); u += SYN3R_PP_TW) {
            const int r = u / (SYN3R_PP_TW / 4), c = (u % (SYN3R_PP_TW / 4)) * 4;
            if (r >= th || c >= tw) continue;
            const size_t i = (size_t)(y0 + r) * W + x0 + c;
            float f[3][4];
#pragma unroll
            for (int ch = 0; ch < 3; ++ch) {
                const float4 v = __ldg(reinterpret_cast<const float4 *>(a.img_chw + ((size_t)p * 3 + ch) * hw + i));
                f[ch][0] = v.x; f[ch][1] = v.y; f[ch][2] = v.z; f[ch][3] = v.w;
            }
            unsigned b[12];
            float ori[12];
#pragma unroll
            for (int j = 0; j < 4; ++j)
#pragma unroll
                for (int ch = 0; ch < 3; ++ch) {
                    const unsigned v = (unsigned)__float2int_rz(f[ch][j]) & 0xffu;
                    b[3 * j + ch] = s_dil[r][c + j] ? 0u : v;
                    ori[3 * j + ch] = __fdiv_rn(f[ch][j], 255.0f);
                }
            const size_t base = ((size_t)p * hw + i) * 3;
            if (a.cond_u8) {
                uint32_t *o = reinterpret_cast<uint32_t *>(a.cond_u8 + base);
#pragma unroll
                for (int k = 0; k < 3; ++k) o[k] = b[4 * k] | (b[4 * k + 1] << 8) | (b[4 * k + 2] << 16) | (b[4 * k + 3] << 24);
            }
            if (a.cond_f32) {
                float4 *o = reinterpret_cast<float4 *>(a.cond_f32 + base);
#pragma unroll
                for (int k = 0; k < 3; ++k)
                    o[k] = make_float4(u8_over_255(b[4 * k]), u8_over_255(b[4 * k + 1]), u8_over_255(b[4 * k + 2]), u8_over_255(b[4 * k + 3]));
            }
            if (a.cond_ori) {
                float4 *o = reinterpret_cast<float4 *>(a.cond_ori + base);
#pragma unroll
                for (int k = 0; k < 3; ++k) o[k] = make_float4(ori[4 * k], ori[4 * k + 1], ori[4 * k + 2], ori[4 * k + 3]);
            }
        }
    }

    // ---- latent mask: 8x8 block count of the dilated mask, >= 13 of 64 (:1592-1596) -----------------------------------
    if (a.latent && t < (SYN3R_PP_TW / 8) * (SYN3R_PP_TH / 8)) {
        const int bx = t % (SYN3R_PP_TW / 8), by = t / (SYN3R_PP_TW / 8);
        if (8 * bx < tw && 8 * by < th) {
            unsigned cnt = 0;
#pragma unroll
            for (int r = 0; r < 8; ++r) cnt += __popcll(*reinterpret_cast<const unsigned long long *>(&s_dil[8 * by + r][8 * bx]));
            a.latent[((size_t)p * (H / 8) + (y0 / 8 + by)) * (W / 8) + x0 / 8 + bx] = cnt >= 13u;
        }
    }

}

// ---------------------------------------------------------------------------------------------------------------------
// Uncertainty fusion: /root/reference/model/diffusionGS.py:821-854 and the clip of :876.
//   conf = exp(-(|cond - gs| / 0.5)^3) [float32] * (cond.sum(-1) > 0);  unc = 1 - conf * (1 - soft_ori) [float64];
//   masks = 8x8 block mean of unc in NumPy's float64 pairwise order, cast float32;  cond' = clip(unc > 0.5 ? gs : cond).
// Elementwise and HBM-bound: 28 B in + 12 B out per pixel.  One CTA = 256 columns x 8 rows (32 latent cells); a thread
// owns two adjacent columns, its xor-1 / xor-2 lanes the other pairs of the 8x8 block (as in k_condition_post).
// The float32 exp / cube are within ~2 ulp of NumPy's SIMD library calls: outputs agree to 1e-6, the `unc > 0.5`
// choice outside the |unc - 0.5| <= 1e-6 band.
// ---------------------------------------------------------------------------------------------------------------------
__device__ __forceinline__ double fuse_pixel(const float c[3], const float g[3], float soft_ori, float out[3]) {
    const float s = __fadd_rn(__fadd_rn(c[0], c[1]), c[2]);                                          // :833
    const float d0 = __fsub_rn(c[0], g[0]), d1 = __fsub_rn(c[1], g[1]), d2 = __fsub_rn(c[2], g[2]);
    const float nrm = __fsqrt_rn(__fadd_rn(__fadd_rn(__fmul_rn(d0, d0), __fmul_rn(d1, d1)), __fmul_rn(d2, d2)));
    const float t = __fdiv_rn(nrm, 0.5f);
    const float conf32 = expf(-__fmul_rn(__fmul_rn(t, t), t));                                       // :826
    const double conf = (double)conf32 * ((s > 0.0f) ? 1.0 : 0.0);                                   // * (1 - uncertainty_mask)
    const double unc = 1.0 - conf * (double)__fsub_rn(1.0f, soft_ori);                               // :836-838
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        const float v = (unc > 0.5) ? g[k] : c[k];                                                   // :854
        out[k] = (v != v) ? v : fminf(fmaxf(v, 0.0f), 1.0f);                                         // :876, NaN kept
    }
    return unc;
}

__global__ void __launch_bounds__(128)
k_fuse_uncertainty(const float *__restrict__ cond_ori, const float *__restrict__ gs, const float *__restrict__ soft_ori,
                   int H, int W, float *__restrict__ out_cond, float *__restrict__ out_masks, double *__restrict__ out_unc) {
    const int t = threadIdx.x, n = blockIdx.z;
    const int x = blockIdx.x * 256 + 2 * t, y0 = blockIdx.y * 8;
    const size_t hw = (size_t)H * W;
    const bool live = x < W;                                 // W is a multiple of 8: a pair of columns is in or out
    double r0 = 0.0, r1 = 0.0;
    if (live) {
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            const size_t q = (size_t)n * hw + (size_t)(y0 + i) * W + x;
            const float2 *c2 = reinterpret_cast<const float2 *>(cond_ori + q * 3), *g2 = reinterpret_cast<const float2 *>(gs + q * 3);
            const float2 ca = __ldg(c2), cb = __ldg(c2 + 1), cc = __ldg(c2 + 2), ga = __ldg(g2), gb = __ldg(g2 + 1), gc = __ldg(g2 + 2);
            const float2 so = __ldg(reinterpret_cast<const float2 *>(soft_ori + q));
            const float cA[3] = {ca.x, ca.y, cb.x}, cB[3] = {cb.y, cc.x, cc.y}, gA[3] = {ga.x, ga.y, gb.x}, gB[3] = {gb.y, gc.x, gc.y};
            float oA[3], oB[3];
            const double uA = fuse_pixel(cA, gA, so.x, oA), uB = fuse_pixel(cB, gB, so.y, oB);
            if (out_cond) {
                float2 *o2 = reinterpret_cast<float2 *>(out_cond + q * 3);
                o2[0] = make_float2(oA[0], oA[1]); o2[1] = make_float2(oA[2], oB[0]); o2[2] = make_float2(oB[1], oB[2]);
            }
            if (out_unc) *reinterpret_cast<double2 *>(out_unc + q) = make_double2(uA, uB);
            r0 = i ? r0 + uA : uA;                           // r[j] = a[j]; r[j] += a[8 i + j]
            r1 = i ? r1 + uB : uB;
        }
    }
    const double pair = r0 + r1;
    const double quad = pair + __shfl_xor_sync(0xffffffffu, pair, 1);
    const double all8 = quad + __shfl_xor_sync(0xffffffffu, quad, 2);
    if (live && out_masks && !(t & 3)) out_masks[((size_t)n * (H / 8) + y0 / 8) * (W / 8) + x / 8] = (float)(all8 / 64.0);
}

// ---------------------------------------------------------------------------------------------
// k_nearby_uncertainty: the tail of consistency_check_from_nearby_images_bw (model/diffusionGS.py:1347-1355),
// one thread per VEC consecutive pixels of a target frame n whose neighbour warps are pairs [first[n], first[n+1]):
//   conf      = mean_k soft[k]                       (:1347  torch.stack(masks).mean(dim=0): sum in order, * 1/cnt)
//   out_unc   = 1 - conf                             (:1348)
//   w_c       = mean_k warped[k][c]                  (:1350)
//   d         = sqrt((d0^2 + d1^2) + d2^2), d_c = w_c - image[n][c]      (:1351  torch.norm(.., dim=0))
//   out_int   = 1 - exp(-(d * inv_sigma)^3)          (:1351-1353; a tensor divided by a Python scalar is a multiply
//               by 1.0f / 0.1f on the CUDA device the reference runs this on, x ** 3 is x * x * x)
// HBM-bound: (16 cnt + 12) bytes read and 8 written per pixel, every byte once.
// ---------------------------------------------------------------------------------------------
template <int VEC>
struct NbVec { float v[VEC]; };
template <int VEC>
__device__ __forceinline__ NbVec<VEC> nb_load(const float *p) {
    NbVec<VEC> r;
    if (VEC == 4) { const float4 t = __ldcs(reinterpret_cast<const float4 *>(p)); r.v[0] = t.x; r.v[1] = t.y; r.v[2] = t.z; r.v[VEC - 1] = t.w; }
    else r.v[0] = __ldcs(p);
    return r;
}
template <int VEC>
__device__ __forceinline__ void nb_store(float *p, const NbVec<VEC> &r) {
    if (VEC == 4) __stcs(reinterpret_cast<float4 *>(p), make_float4(r.v[0], r.v[1], r.v[2], r.v[VEC - 1]));
    else __stcs(p, r.v[0]);
}

template <int VEC>
__global__ void __launch_bounds__(256)
k_nearby_uncertainty(const float *__restrict__ soft, const float *__restrict__ warped, const float *__restrict__ images,
                     const int *__restrict__ first, size_t hw, float inv_sigma, float *__restrict__ out_unc,
                     float *__restrict__ out_int) {
    const int n = blockIdx.y;
    const size_t i = ((size_t)blockIdx.x * 256 + threadIdx.x) * VEC;
    if (i >= hw) return;
    const int k0 = __ldg(first + n), k1 = __ldg(first + n + 1);
    const float factor = __fdiv_rn(1.0f, (float)(k1 - k0));
    NbVec<VEC> cs, w0, w1, w2;
#pragma unroll
    for (int j = 0; j < VEC; ++j) cs.v[j] = w0.v[j] = w1.v[j] = w2.v[j] = 0.0f;
    for (int k = k0; k < k1; ++k) {
        const NbVec<VEC> a = nb_load<VEC>(soft + (size_t)k * hw + i);
        const float *wp = warped + (size_t)k * 3 * hw + i;
        const NbVec<VEC> b0 = nb_load<VEC>(wp), b1 = nb_load<VEC>(wp + hw), b2 = nb_load<VEC>(wp + 2 * hw);
#pragma unroll
        for (int j = 0; j < VEC; ++j) {
            cs.v[j] = (k == k0) ? a.v[j] : __fadd_rn(cs.v[j], a.v[j]);       // NaN and -0 of the first term survive
            w0.v[j] = (k == k0) ? b0.v[j] : __fadd_rn(w0.v[j], b0.v[j]);
            w1.v[j] = (k == k0) ? b1.v[j] : __fadd_rn(w1.v[j], b1.v[j]);
            w2.v[j] = (k == k0) ? b2.v[j] : __fadd_rn(w2.v[j], b2.v[j]);
        }
    }
    const float *ip = images + (size_t)n * 3 * hw + i;
    const NbVec<VEC> t0 = nb_load<VEC>(ip), t1 = nb_load<VEC>(ip + hw), t2 = nb_load<VEC>(ip + 2 * hw);
    NbVec<VEC> u, v;
#pragma unroll
    for (int j = 0; j < VEC; ++j) {
        u.v[j] = __fsub_rn(1.0f, __fmul_rn(cs.v[j], factor));
        const float d0 = __fsub_rn(__fmul_rn(w0.v[j], factor), t0.v[j]), d1 = __fsub_rn(__fmul_rn(w1.v[j], factor), t1.v[j]),
                    d2 = __fsub_rn(__fmul_rn(w2.v[j], factor), t2.v[j]);
        const float nr = __fsqrt_rn(__fadd_rn(__fadd_rn(__fmul_rn(d0, d0), __fmul_rn(d1, d1)), __fmul_rn(d2, d2)));
        const float t = __fmul_rn(nr, inv_sigma);
        v.v[j] = __fsub_rn(1.0f, expf(-__fmul_rn(__fmul_rn(t, t), t)));
    }
    if (out_unc) nb_store<VEC>(out_unc + (size_t)n * hw + i, u);
    if (out_int) nb_store<VEC>(out_int + (size_t)n * hw + i, v);
}

// ---------------------------------------------------------------------------------------------
// k_pack_mask_bits: np.packbits of a 0/1 byte mask (MSB first, the NumPy default), so that a mask travels over PCIe /
// NVLink at 1 bit per pixel instead of 8.  One thread: 32 mask bytes -> 4 bytes; a warp reads 1 KB and writes one
// 128-byte line.  Bytes are tested `!= 0` like astype(bool).  The launcher handles a tail that is not a multiple of
// 32 bytes (and misaligned pointers) with the byte-wise kernel.
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ unsigned pack8_msb_first(unsigned lo, unsigned hi) {
    // lo = mask bytes 0..3 (byte 0 in the low bits), hi = bytes 4..7 -> bit 7 = byte 0 ... bit 0 = byte 7
    unsigned r = 0;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        r |= ((lo >> (8 * k)) & 0xffu) ? (0x80u >> k) : 0u;
        r |= ((hi >> (8 * k)) & 0xffu) ? (0x08u >> k) : 0u;
    }
    return r;
}
__global__ void __launch_bounds__(256)
k_pack_mask_bits(const uint4 *__restrict__ mask, size_t n32, unsigned *__restrict__ bits) {
    // one thread: 32 mask bytes (two 16-byte loads) -> one 32-bit word; a warp writes one whole 128-byte line
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n32; i += (size_t)gridDim.x * blockDim.x) {
        const uint4 v = ld_stream_u4(mask + 2 * i), w = ld_stream_u4(mask + 2 * i + 1);
        bits[i] = pack8_msb_first(v.x, v.y) | (pack8_msb_first(v.z, v.w) << 8) | (pack8_msb_first(w.x, w.y) << 16) |
                  (pack8_msb_first(w.z, w.w) << 24);
    }
}
__global__ void __launch_bounds__(256)
k_pack_mask_bits_tail(const uint8_t *__restrict__ mask, size_t first, size_t n, uint8_t *__restrict__ bits) {
    // output bytes first / 8 ... ; a trailing partial byte is zero-padded like np.packbits
    const size_t b = first / 8 + (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (b * 8 >= n) return;
    unsigned r = 0;
    for (int k = 0; k < 8; ++k)
        if (b * 8 + k < n && mask[b * 8 + k]) r |= 0x80u >> k;
    bits[b] = (uint8_t)r;
}

}  // namespace syn3r
