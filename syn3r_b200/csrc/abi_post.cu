// abi_post.cu — extern "C" launcher of the warped-frame post-processing (include/syn3r_warp.h).
#include "postprocess.cuh"
#include "score.cuh"
#include "flow.cuh"
#include "latent.cuh"

#include <mutex>

using namespace syn3r;

extern "C" {

int syn3r_condition_post_batched(const uint8_t *valid, const uint8_t *img_u8, const float *img_chw, const float *soft,
                                 int P, int H, int W, uint8_t *out_mask_png, uint8_t *out_cond_u8, float *out_cond_f32,
                                 float *out_cond_ori, uint8_t *out_latent, float *out_soft_pooled, void *stream) {
    if (P == 0) return SYN3R_OK;                            // empty batch: a valid no-op
    SYN3R_REQUIRE(valid != nullptr, "the valid mask is required");
    SYN3R_REQUIRE(P >= 0 && H > 0 && W > 0, "bad sizes");
    // the reference reshapes to (H/8, 8, W/8, 8) (diffusionGS.py:1593): sizes that do not divide raise there too
    SYN3R_REQUIRE(H % 8 == 0 && W % 8 == 0, "H and W must be multiples of 8 (latent grid of 8x8 blocks)");
    SYN3R_REQUIRE(!(img_u8 && img_chw), "give the frame either as uint8 HWC or as float32 CHW, not both");
    SYN3R_REQUIRE(!(out_cond_u8 || out_cond_f32 || out_cond_ori) || img_u8 || img_chw, "frame outputs need a frame input");
    SYN3R_REQUIRE(!out_soft_pooled || soft, "soft_pooled needs the soft mask");
    SYN3R_REQUIRE(P <= 65535, "at most 65535 frames per call");
    PostArgs a{valid, img_u8, img_chw, soft, H, W, out_mask_png, out_cond_u8, out_cond_f32, out_cond_ori, out_latent,
               out_soft_pooled};
    dim3 grid((W + SYN3R_PP_TW - 1) / SYN3R_PP_TW, (H + SYN3R_PP_TH - 1) / SYN3R_PP_TH, P);
    k_condition_post<<<grid, SYN3R_PP_TW, 0, (cudaStream_t)stream>>>(a);
    return check_cuda(cudaGetLastError(), "k_condition_post");
}

int syn3r_pack_mask_bits(const uint8_t *mask, size_t n, uint8_t *bits, void *stream) {
    if (n == 0) return SYN3R_OK;
    SYN3R_REQUIRE(mask && bits, "null required pointer");
    int rc = check_device();
    if (rc != SYN3R_OK) return rc;
    cudaStream_t st = (cudaStream_t)stream;
    size_t n32 = (((uintptr_t)mask & 15) == 0 && ((uintptr_t)bits & 3) == 0) ? n / 32 : 0;
    if (n32) {
        size_t blocks = (n32 + 255) / 256, cap = (size_t)sm_count() * 16;
        k_pack_mask_bits<<<(unsigned)(blocks < cap ? blocks : cap), 256, 0, st>>>(reinterpret_cast<const uint4 *>(mask), n32,
                                                                               reinterpret_cast<unsigned *>(bits));
    }
    const size_t first = n32 * 32;
    if (first < n) {
        const size_t out_bytes = (n - first + 7) / 8;
        k_pack_mask_bits_tail<<<(unsigned)((out_bytes + 255) / 256), 256, 0, st>>>(mask, first, n, bits);
    }
    return check_cuda(cudaGetLastError(), "k_pack_mask_bits");
}

int syn3r_fuse_uncertainty_batched(const float *cond_ori, const float *gs, const float *soft_ori, int N, int H, int W,
                                   float *out_cond, float *out_masks, double *out_unc, void *stream) {
    if (N == 0) return SYN3R_OK;
    SYN3R_REQUIRE(cond_ori && gs && soft_ori, "null required pointer");
    SYN3R_REQUIRE(N > 0 && N <= 65535 && H > 0 && W > 0, "bad sizes");
    SYN3R_REQUIRE(H % 8 == 0 && W % 8 == 0, "H and W must be multiples of 8 (latent grid of 8x8 blocks)");
    SYN3R_REQUIRE(out_cond || out_masks || out_unc, "no output requested");
    k_fuse_uncertainty<<<dim3((W + 255) / 256, H / 8, N), 128, 0, (cudaStream_t)stream>>>(cond_ori, gs, soft_ori, H, W, out_cond,
                                                                                       out_masks, out_unc);
    return check_cuda(cudaGetLastError(), "k_fuse_uncertainty");
}

int syn3r_nearby_uncertainty_batched(const float *soft, const float *warped, const float *images, const int *pair_first,
                                     int N, int H, int W, float inv_sigma, float *out_unc, float *out_int_unc,
                                     void *stream) {
    if (N == 0) return SYN3R_OK;
    SYN3R_REQUIRE(soft && warped && images && pair_first, "null required pointer");
    SYN3R_REQUIRE(N > 0 && N <= 65535 && H > 0 && W > 0, "bad sizes");
    SYN3R_REQUIRE(out_unc || out_int_unc, "no output requested");
    const size_t hw = (size_t)H * W;
    const bool vec = hw % 4 == 0 && (((uintptr_t)soft | (uintptr_t)warped | (uintptr_t)images | (uintptr_t)out_unc |
                                      (uintptr_t)out_int_unc) & 15) == 0;
    cudaStream_t st = (cudaStream_t)stream;
    if (vec)
        k_nearby_uncertainty<4><<<dim3((unsigned)((hw / 4 + 255) / 256), N), 256, 0, st>>>(soft, warped, images, pair_first, hw,
                                                                                       inv_sigma, out_unc, out_int_unc);
    else
        k_nearby_uncertainty<1><<<dim3((unsigned)((hw + 255) / 256), N), 256, 0, st>>>(soft, warped, images, pair_first, hw,
                                                                                   inv_sigma, out_unc, out_int_unc);
    return check_cuda(cudaGetLastError(), "k_nearby_uncertainty");
}

int syn3r_flow_fb_consistency_batched(const float *fwd_flow, const float *bwd_flow, int B, int H, int W, float alpha,
                                      float beta, float *out_fwd_occ, float *out_bwd_occ, float *out_diff_fwd,
                                      float *out_diff_bwd, void *stream) {
    if (B == 0) return SYN3R_OK;
    SYN3R_REQUIRE(fwd_flow && bwd_flow, "null required pointer");
    SYN3R_REQUIRE(B > 0 && B <= 65535 && H > 1 && W > 1, "bad sizes (H and W must be at least 2: the reference divides by w - 1)");
    SYN3R_REQUIRE(out_fwd_occ || out_bwd_occ || out_diff_fwd || out_diff_bwd, "no output requested");
    const size_t hw = (size_t)H * W;
    SYN3R_REQUIRE(hw < (1ull << 30), "frame too large (32-bit pixel indices)");
    cudaStream_t st = (cudaStream_t)stream;
    const float iw = 1.0f / (float)(W - 1), ih = 1.0f / (float)(H - 1);
    const size_t per_block = 256 * SYN3R_FLOW_PX;
    k_flow_fb_consistency<<<dim3((unsigned)((hw + per_block - 1) / per_block), B), 256, 0, st>>>(
        fwd_flow, bwd_flow, H, W, alpha, beta, iw, ih, out_fwd_occ, out_bwd_occ, out_diff_fwd, out_diff_bwd);
    return check_cuda(cudaGetLastError(), "k_flow_fb_consistency");
}

static int launch_latent(float *pred, const float *cond_latents, const float *mask, const float *lambda_row, int b, int F,
                         int c, int h, int w, int mask_batch, int mask_frames, int mask_channels, int tau0, int n_tau,
                         float clamp_lo, int blend, unsigned char *out_top, int *out_cutoff_index, float *out_cutoff_value, void *stream) {
    SYN3R_REQUIRE(pred && cond_latents && mask && lambda_row, "null required pointer");
    SYN3R_REQUIRE(b > 0 && F >= 2 && c > 0 && h > 0 && w > 0 && mask_channels > 0, "bad sizes");
    SYN3R_REQUIRE(mask_batch == 1 || mask_batch == b, "the mask's batch must be 1 or the latents' batch (broadcast of :1464)");
    SYN3R_REQUIRE(n_tau >= 0 && tau0 >= 1 && tau0 + n_tau <= F, "frame range outside 1 .. F - 1");
    SYN3R_REQUIRE(mask_frames >= tau0 + n_tau - 1, "mask has fewer frames than the range reads (frame tau reads mask row tau - 1)");
    SYN3R_REQUIRE((size_t)b * c * h * w < (1ull << 24), "b*c*h*w must stay below 2^24 (float32 count of :1479)");
    LatentArgs a;
    a.pred = pred; a.cond = cond_latents; a.mask = mask; a.lambda_row = lambda_row;
    a.b = b; a.F = F; a.c = c; a.hw = h * w; a.bm = mask_batch; a.cm = mask_channels; a.tau0 = tau0; a.n_tau = n_tau;
    a.Fm = mask_frames;
    a.clamp_lo = clamp_lo; a.blend = blend; a.out_top = out_top;
    a.out_k = out_cutoff_index; a.out_cutoff = out_cutoff_value;
    const size_t key_bytes = (size_t)b * c * h * w * 4, key_cap = 190 * 1024;   // + 34 KB of histograms <= 227 KB
    a.cache_keys = key_bytes <= key_cap;
    static thread_local int configured_dev = -1;
    int dev = -1;
    SYN3R_TRY_CUDA(cudaGetDevice(&dev));
    if (configured_dev != dev) {
        SYN3R_TRY_CUDA(cudaFuncSetAttribute(k_latent_guided_replace, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)key_cap));
        configured_dev = dev;
    }
    const int blocks = n_tau + (blend ? 2 : 0);
    if (blocks == 0) return SYN3R_OK;
    k_latent_guided_replace<<<blocks, SYN3R_LAT_THREADS, a.cache_keys ? key_bytes : 0, (cudaStream_t)stream>>>(a);
    return check_cuda(cudaGetLastError(), "k_latent_guided_replace");
}

int syn3r_latent_guided_replace(float *pred, const float *cond_latents, const float *mask, const float *lambda_row, int b,
                                int F, int c, int h, int w, int mask_batch, int mask_frames, int mask_channels, int tau0,
                                int n_tau, int *out_cutoff_index, float *out_cutoff_value, void *stream) {
    return launch_latent(pred, cond_latents, mask, lambda_row, b, F, c, h, w, mask_batch, mask_frames, mask_channels, tau0, n_tau, 0.4f, 1,
                         nullptr, out_cutoff_index, out_cutoff_value, stream);
}

int syn3r_latent_top_masks(const float *pred, const float *cond_latents, const float *mask, const float *lambda_row, int b,
                           int F, int c, int h, int w, int mask_batch, int mask_frames, int mask_channels, int tau0, int n_tau,
                           float weight_clamp, uint8_t *out_top_masks, int *out_cutoff_index, float *out_cutoff_value,
                           void *stream) {
    SYN3R_REQUIRE(out_top_masks, "null required pointer");
    return launch_latent(const_cast<float *>(pred), cond_latents, mask, lambda_row, b, F, c, h, w, mask_batch, mask_frames,
                         mask_channels, tau0, n_tau, weight_clamp, 0, out_top_masks, out_cutoff_index, out_cutoff_value, stream);
}

size_t syn3r_mean_uncertainty_workspace_bytes(int P) { return P > 0 ? (size_t)P * SYN3R_SC_BLOCKS * sizeof(double) : 0; }

int syn3r_mean_uncertainty_batched(const float *soft, int P, int H, int W, float *out_mean, void *workspace,
                                   size_t workspace_bytes, void *stream) {
    if (P == 0) return SYN3R_OK;
    SYN3R_REQUIRE(soft && out_mean && workspace, "null required pointer");
    SYN3R_REQUIRE(P > 0 && P <= 65535 && H > 0 && W > 0, "bad sizes");
    SYN3R_REQUIRE(((size_t)H * W) % 4 == 0 || P == 1, "H*W must be a multiple of 4 for P > 1 (16-byte frame alignment)");
    SYN3R_REQUIRE(((uintptr_t)soft & 15) == 0 && ((uintptr_t)workspace & 7) == 0, "soft must be 16-byte aligned");
    if (workspace_bytes < syn3r_mean_uncertainty_workspace_bytes(P)) {
        set_error("mean-uncertainty workspace: need %zu bytes (got %zu)", syn3r_mean_uncertainty_workspace_bytes(P), workspace_bytes);
        return SYN3R_EWORKSPACE;
    }
    cudaStream_t st = (cudaStream_t)stream;
    const size_t hw = (size_t)H * W;
    k_mean_uncertainty_partial<<<dim3(SYN3R_SC_BLOCKS, P), SYN3R_SC_THREADS, 0, st>>>(soft, hw, (double *)workspace);
    k_mean_uncertainty_final<<<(P + 127) / 128, 128, 0, st>>>((const double *)workspace, P, hw, out_mean);
    return check_cuda(cudaGetLastError(), "k_mean_uncertainty");
}

// ---------------------------------------------------------------------------------------------------------------
// syn3r_mean_uncertainty_static: syn3r_mean_uncertainty_batched for callers that cannot allocate (CUDA graph capture,
// serving loops, harnesses that forbid cudaMalloc).  No counterpart in the reference.
// The chunk partials live in a file-scope device array instead of a caller-owned workspace: 2 MiB per GPU
// that loads the library, held until exit.  Every call shares it and it keeps the last call's contents: nothing has to
// start at zero (the first kernel overwrites every partial the second one reads), but two calls that overlap on the
// GPU (different streams, no ordering between them) race on it.
// ---------------------------------------------------------------------------------------------------------------
static __device__ __align__(256) double g_mean_unc_scratch[(size_t)SYN3R_MEAN_UNCERTAINTY_STATIC_MAX_P * SYN3R_SC_BLOCKS];

// In host code `g_mean_unc_scratch` is a host-side stand-in, not a device address: the address a kernel may use comes
// from cudaGetSymbolAddress, once per device, with that device current.
static int mean_unc_scratch_ptr(double **out) {
    enum { MAX_DEV = 64 };
    static std::mutex mu;
    static double *ptrs[MAX_DEV] = {};
    int dev = -1;
    SYN3R_TRY_CUDA(cudaGetDevice(&dev));
    SYN3R_REQUIRE(dev >= 0 && dev < MAX_DEV, "device ordinal out of range");
    std::lock_guard<std::mutex> lock(mu);
    if (!ptrs[dev]) {
        double *p = nullptr;
        SYN3R_TRY_CUDA(cudaGetSymbolAddress((void **)&p, g_mean_unc_scratch));
        SYN3R_REQUIRE(p != nullptr, "cudaGetSymbolAddress returned a null scratch address");
        ptrs[dev] = p;
    }
    *out = ptrs[dev];
    return SYN3R_OK;
}

int syn3r_mean_uncertainty_static_max_p(void) { return SYN3R_MEAN_UNCERTAINTY_STATIC_MAX_P; }

int syn3r_mean_uncertainty_static(const float *soft, int P, int H, int W, float *out_mean, void *stream) {
    if (P == 0) return SYN3R_OK;
    SYN3R_REQUIRE(soft && out_mean, "null required pointer");
    SYN3R_REQUIRE(P > 0 && H > 0 && W > 0, "bad sizes");
    if (P > SYN3R_MEAN_UNCERTAINTY_STATIC_MAX_P) {
        set_error("syn3r_mean_uncertainty_static: P = %d exceeds the static scratch (at most %d frames per call); split the "
                  "batch or use syn3r_mean_uncertainty_batched with a caller-owned workspace", P,
                  SYN3R_MEAN_UNCERTAINTY_STATIC_MAX_P);
        return SYN3R_EWORKSPACE;
    }
    SYN3R_REQUIRE(((size_t)H * W) % 4 == 0 || P == 1, "H*W must be a multiple of 4 for P > 1 (16-byte frame alignment)");
    SYN3R_REQUIRE(((uintptr_t)soft & 15) == 0, "soft must be 16-byte aligned");
    int rc = check_device();
    if (rc != SYN3R_OK) return rc;
    double *scratch = nullptr;
    rc = mean_unc_scratch_ptr(&scratch);
    if (rc != SYN3R_OK) return rc;
    cudaStream_t st = (cudaStream_t)stream;
    const size_t hw = (size_t)H * W;
    k_mean_uncertainty_partial<<<dim3(SYN3R_SC_BLOCKS, P), SYN3R_SC_THREADS, 0, st>>>(soft, hw, scratch);
    k_mean_uncertainty_final<<<(P + 127) / 128, 128, 0, st>>>(scratch, P, hw, out_mean);
    return check_cuda(cudaGetLastError(), "k_mean_uncertainty (static scratch)");
}

}  // extern "C"
