// score.cuh — per-pair mean of the reprojection uncertainty.
// The reference scores a perturbed pose candidate by `(1 - warp_dict['soft_mask_reproj']).mean()` after one
// inverse_warp per candidate (/root/reference/model/diffusionGS.py:759-761, 150 calls per view pair); here the
// candidates of all anchors are one batched k_reproj_masks launch and this reduction.  Deterministic: a fixed
// partition of every frame into SYN3R_SC_BLOCKS chunks, fp32 per-thread sums of 1 - soft (the reference's own
// element arithmetic), a fixed-order tree in fp64, then a fixed-order sum of the chunk partials.
#pragma once
#include "common.cuh"

namespace syn3r {

#define SYN3R_SC_BLOCKS 64
#define SYN3R_SC_THREADS 256

__global__ void __launch_bounds__(SYN3R_SC_THREADS)
k_mean_uncertainty_partial(const float *__restrict__ soft, size_t hw, double *__restrict__ partial) {
    const int p = blockIdx.y, b = blockIdx.x, t = threadIdx.x;
    const size_t n4 = hw / 4, per = (n4 + SYN3R_SC_BLOCKS - 1) / SYN3R_SC_BLOCKS;
    const size_t lo = (size_t)b * per, hi = min(lo + per, n4);
    const float4 *s4 = reinterpret_cast<const float4 *>(soft + (size_t)p * hw);
    float acc = 0.0f;
    for (size_t i = lo + t; i < hi; i += SYN3R_SC_THREADS) {
        const float4 v = __ldg(s4 + i);
        acc = __fadd_rn(acc, __fadd_rn(__fadd_rn(__fsub_rn(1.0f, v.x), __fsub_rn(1.0f, v.y)),
                                       __fadd_rn(__fsub_rn(1.0f, v.z), __fsub_rn(1.0f, v.w))));
    }
    if (b == SYN3R_SC_BLOCKS - 1)                            // the up-to-3 elements behind the last float4
        for (size_t i = n4 * 4 + t; i < hw; i += SYN3R_SC_THREADS) acc = __fadd_rn(acc, __fsub_rn(1.0f, soft[(size_t)p * hw + i]));
    double d = (double)acc;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) d += __shfl_xor_sync(0xffffffffu, d, o);
    __shared__ double s_w[SYN3R_SC_THREADS / 32];
    if ((t & 31) == 0) s_w[t >> 5] = d;
    __syncthreads();
    if (t == 0) {
        double tot = 0.0;
#pragma unroll
        for (int k = 0; k < SYN3R_SC_THREADS / 32; ++k) tot += s_w[k];
        partial[(size_t)p * SYN3R_SC_BLOCKS + b] = tot;
    }
}

__global__ void k_mean_uncertainty_final(const double *__restrict__ partial, int P, size_t hw, float *__restrict__ out) {
    const int p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= P) return;
    double tot = 0.0;
    for (int k = 0; k < SYN3R_SC_BLOCKS; ++k) tot += partial[(size_t)p * SYN3R_SC_BLOCKS + k];
    out[p] = (float)(tot / (double)hw);
}

}  // namespace syn3r
