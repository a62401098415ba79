// Guided latent replacement of the patched Euler scheduler, SURVEY 8f rank 4 (scheduler half).
//
// Reference: /root/reference/thirdparty/diffusers/src/diffusers/schedulers/scheduling_euler_discrete.py:1442-1499
// (step_interp_prob_uncertain).  step_single (:548-586) and step_interp (:735-774) share the selection (mask, sort,
// cutoff) with a caller-given lower clamp and keep the boolean top_mask instead of blending: `blend = 0`.  Per conditioning frame tau the reference
// runs a torch.sort over the b*c*h*w absolute differences between the predicted clean latent and the warped
// conditioning latent, reads ONE order statistic out of it (the cutoff) and blends; about 25 launches per frame, 23
// frames, 100 steps, two passes.
//
// Here: one CTA per frame, all frames in one launch.  The sort becomes a 4-pass radix SELECT on the float bit
// patterns (non-negative floats order like their bits, NaN above inf like torch.sort's "NaN last"), with per-warp
// histograms: the exponent byte (nearly the same for every key) is counted through __match_any_sync, whose cost
// grows with the number of distinct values in the warp, the mantissa bytes through shared-memory atomics.  The keys of a frame (4 x 72 x 128 floats = 147 KB) are staged in shared memory once, so that the four
// passes run at shared-memory latency (recomputing them from global memory in every pass serialises an L2 round trip
// per loop iteration behind the warp-wide match: 0.19 ms instead of 0.05); larger frames fall back to recomputing.
// Results are bit-identical to the reference: a selection and exact float32 arithmetic.
#pragma once
#include "common.cuh"

namespace syn3r {

struct LatentArgs {
    float *pred;                 // [b, F, c, hw] predicted clean latents, updated in place
    const float *cond;           // [F, c, hw] = temp_cond_latents[1]
    const float *mask;           // [bm, Fm, cm, hw], 1 = hole (frame tau of the reference's mask is row tau - 1 here)
    const float *lambda_row;     // [F] = lambda_ts[step_index]
    int b, F, c, hw, bm, cm, tau0, n_tau;
    int Fm;                      // frames in `mask`: F - 2 as the pipelines pass it to step_interp*, F - 1 for step_single
    float clamp_lo;              // lower clamp of lambda: 0.4 (:1476) or the caller's weight_clamp (:577, :766)
    int blend;                   // 1: the soft blend of :1487-1499 in place; 0: selection only (pred is not written)
    unsigned char *out_top;      // [n_tau, b, c, hw] top_mask = (d <= cutoff) & m (:583, :772, :1483), optional
    int cache_keys;              // 1: the frame's b*c*hw keys fit the dynamic shared memory and are staged there once
    int *out_k;                  // [n_tau] cutoff_index_e (optional)
    float *out_cutoff;           // [n_tau] cutoff_value (optional)
};

#define SYN3R_LAT_THREADS 1024

__global__ void __launch_bounds__(SYN3R_LAT_THREADS)
k_latent_guided_replace(const LatentArgs a) {
    extern __shared__ unsigned s_keys[];                       // [b * c * hw] when a.cache_keys
    __shared__ unsigned s_hist[SYN3R_LAT_THREADS / 32][256];
    __shared__ unsigned s_tot[256];
    __shared__ unsigned s_ctl[4];                              // [0] prefix, [1] rank within the prefix, [2] nz
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int chw = a.c * a.hw;
    const unsigned n = (unsigned)a.b * (unsigned)chw;

    if ((int)blockIdx.x >= a.n_tau) {                          // (launched only when blending) :1496-1499: first and last frame = the conditioning latents
        const int fr = ((int)blockIdx.x == a.n_tau) ? 0 : a.F - 1;
        for (unsigned e = tid; e < n; e += SYN3R_LAT_THREADS) {
            const unsigned bi = e / chw, r = e - bi * chw;
            a.pred[((size_t)bi * a.F + fr) * chw + r] = a.cond[(size_t)fr * chw + r];
        }
        return;
    }
    const int tau = a.tau0 + blockIdx.x;
    if (tau <= 0 || tau >= a.F - (a.blend ? 1 : 0)) return;    // blending: frame F - 1 is overwritten by the copy above

    const float inv_cm = __fdiv_rn(1.0f, (float)a.cm);
    // mean over the mask's channels of ((1 - mask) > 0.5) for pixel pix of mask batch item mb (:1447, :1457)
    auto mean_valid = [&](int mb, int pix) -> float {
        const float *mp = a.mask + (((size_t)mb * a.Fm + (tau - 1)) * a.cm) * a.hw + pix;
        float s = 0.0f;
        for (int ch = 0; ch < a.cm; ++ch) s += (__fsub_rn(1.0f, __ldg(mp + (size_t)ch * a.hw)) > 0.5f) ? 1.0f : 0.0f;
        return __fmul_rn(s, inv_cm);
    };
    // |pred * m - FEA * m| of element e (:1464-1469) and the pixel's mean_valid
    auto diff_of = [&](unsigned e, float &p, float &f, float &mv) -> float {
        const unsigned bi = e / chw, r = e - bi * chw, pix = r % a.hw;
        mv = mean_valid(a.bm == 1 ? 0 : (int)bi, (int)pix);
        const float m = mv > 0.5f ? 1.0f : 0.0f;
        p = a.pred[((size_t)bi * a.F + tau) * chw + r];
        f = __ldg(a.cond + (size_t)tau * chw + r);
        return fabsf(__fsub_rn(__fmul_rn(p, m), __fmul_rn(f, m)));
    };

    // ---- num_zero (:1461) ------------------------------------------------------------------------------
    if (tid < 4) s_ctl[tid] = 0;
    __syncthreads();
    {
        unsigned cnt = 0;
#pragma unroll 4
        for (int e = tid; e < a.bm * a.hw; e += SYN3R_LAT_THREADS) cnt += !(mean_valid(e / a.hw, e % a.hw) > 0.5f);
        cnt = __reduce_add_sync(0xffffffffu, cnt);
        if (lane == 0 && cnt) atomicAdd(&s_ctl[2], cnt);
    }
    __syncthreads();
    if (tid == 0) {                                            // :1476-1479
        const unsigned nz = s_ctl[2];
        const float wt = fminf(fmaxf(__ldg(a.lambda_row + tau), a.clamp_lo), 1.0f);
        unsigned k = (unsigned)(int)__fmul_rn(wt, (float)(n - nz)) + nz;
        if (a.out_k) a.out_k[blockIdx.x] = (int)k;
        if (k == 0) k = n;                                     // sorted[-1]: Python wraps to the largest
        s_ctl[1] = k - 1;                                      // zero-based rank
        s_ctl[0] = 0;
    }
    if (a.cache_keys) {                                        // independent iterations: the loads pipeline
#pragma unroll 4
        for (unsigned e = tid; e < n; e += SYN3R_LAT_THREADS) {
            float p, f, mv;
            s_keys[e] = __float_as_uint(diff_of(e, p, f, mv));
        }
    }
    // ---- radix select: the (rank)-th smallest key, 8 bits per pass -------------------------------------
    unsigned known = 0;                                        // bits of the key fixed so far
    for (int shift = 24; shift >= 0; shift -= 8) {
        for (int k = lane; k < 256; k += 32) s_hist[warp][k] = 0;
        __syncthreads();
        const unsigned prefix = s_ctl[0];
        for (unsigned e0 = 0; e0 < n; e0 += SYN3R_LAT_THREADS) {      // whole warps stay in the loop for match_any
            const unsigned e = e0 + tid;
            unsigned bin = 256u;                               // no contribution: one group for all idle lanes (match.any
                                                               // takes a step per distinct value in the warp)
            if (e < n) {
                unsigned key;
                if (a.cache_keys) key = s_keys[e];
                else { float p, f, mv; key = __float_as_uint(diff_of(e, p, f, mv)); }
                if ((key & known) == prefix) bin = (key >> shift) & 255u;
            }
            if (shift == 24) {                                 // sign + exponent byte: a handful of distinct values per warp
                const unsigned peers = __match_any_sync(0xffffffffu, bin);
                if (bin < 256u && lane == __ffs(peers) - 1) s_hist[warp][bin] += __popc(peers);
            } else if (bin < 256u) {                           // mantissa bytes: spread over the bins, few conflicts
                atomicAdd(&s_hist[warp][bin], 1u);
            }
        }
        __syncthreads();
        if (tid < 256) {
            unsigned t = 0;
#pragma unroll 8
            for (int wv = 0; wv < SYN3R_LAT_THREADS / 32; ++wv) t += s_hist[wv][tid];
            s_tot[tid] = t;
        }
        __syncthreads();
        if (warp == 0) {                                       // which bin holds the rank: 8 bins per lane + a warp scan
            unsigned loc[8], sum = 0;
#pragma unroll
            for (int j = 0; j < 8; ++j) { loc[j] = s_tot[lane * 8 + j]; sum += loc[j]; }
            unsigned incl = sum;
#pragma unroll
            for (int off = 1; off < 32; off <<= 1) {
                const unsigned t = __shfl_up_sync(0xffffffffu, incl, off);
                if (lane >= off) incl += t;
            }
            const unsigned excl = incl - sum, r = s_ctl[1];
            __syncwarp();
            if (r >= excl && r < incl) {                       // exactly one lane: the bins hold every key of the prefix
                unsigned rr = r - excl;
                int j = 0;
                while (j < 7 && rr >= loc[j]) { rr -= loc[j]; ++j; }
                s_ctl[1] = rr;
                s_ctl[0] = prefix | ((unsigned)(lane * 8 + j) << shift);
            }
        }
        known |= 255u << shift;
        __syncthreads();
    }
    const float cutoff = __uint_as_float(s_ctl[0]);            // :1480
    if (tid == 0 && a.out_cutoff) a.out_cutoff[blockIdx.x] = cutoff;
    // ---- top mask (:1483) and blend (:1487-1493) --------------------------------------------------------
#pragma unroll 4
    for (unsigned e = tid; e < n; e += SYN3R_LAT_THREADS) {
        float p, f, mv;
        const float d = diff_of(e, p, f, mv);
        const bool sel = d <= cutoff;
        if (a.out_top) a.out_top[(size_t)blockIdx.x * n + e] = sel && (mv > 0.5f);
        if (a.blend) {
            const float x = __fdiv_rn(1.0f, __fadd_rn(__fsub_rn(1.0f, mv), 1e-6f));
            float w = __fdiv_rn(x, __fadd_rn(1.0f, x));
            w = (w >= 0.51f) ? w : 0.0f;
            w = __fmul_rn(sel ? 1.0f : 0.0f, w);
            const unsigned bi = e / chw, r = e - bi * chw;
            a.pred[((size_t)bi * a.F + tau) * chw + r] = __fadd_rn(__fmul_rn(__fsub_rn(1.0f, w), p), __fmul_rn(w, f));
        }
    }
}

}  // namespace syn3r
