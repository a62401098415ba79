// abi.cu — the extern "C" surface of libsyn3r_warp.so (include/syn3r_warp.h) and its launchers.
#include <stdarg.h>
#include <string.h>

#include "forward_kernels.cuh"

namespace syn3r {

static thread_local char g_err[512] = "";

void set_error(const char *fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

int check_cuda(cudaError_t e, const char *what) {
    if (e == cudaSuccess) return SYN3R_OK;
    set_error("CUDA error %d (%s) at %s", (int)e, cudaGetErrorString(e), what);
    return SYN3R_ECUDA;
}

struct DevInfo { int dev = -1, sms = 0, smem = 0, major = 0, minor = 0; };
static thread_local DevInfo g_dev;

static int device_info(DevInfo &d) {
    int dev = -1;
    SYN3R_TRY_CUDA(cudaGetDevice(&dev));
    if (g_dev.dev != dev) {
        DevInfo n;
        n.dev = dev;
        SYN3R_TRY_CUDA(cudaDeviceGetAttribute(&n.sms, cudaDevAttrMultiProcessorCount, dev));
        SYN3R_TRY_CUDA(cudaDeviceGetAttribute(&n.smem, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev));
        SYN3R_TRY_CUDA(cudaDeviceGetAttribute(&n.major, cudaDevAttrComputeCapabilityMajor, dev));
        SYN3R_TRY_CUDA(cudaDeviceGetAttribute(&n.minor, cudaDevAttrComputeCapabilityMinor, dev));
        g_dev = n;
    }
    d = g_dev;
    return SYN3R_OK;
}

int check_device() {
    DevInfo d;
    int rc = device_info(d);
    if (rc != SYN3R_OK) return rc;
    if (d.major != 10) {
        set_error("libsyn3r_warp is built for sm_100a only; current device is sm_%d%d", d.major, d.minor);
        return SYN3R_EDEVICE;
    }
    return SYN3R_OK;
}

int sm_count() {
    DevInfo d;
    return device_info(d) == SYN3R_OK ? d.sms : 148;
}

static inline size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

// grid.x for a grid-stride kernel over n items with `per_block` threads, grid.y = ny: enough
// blocks for every SM to hold several, a multiple of the SM count when the work allows.
static dim3 stride_grid(size_t n, int per_block, int ny) {
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

struct FwdWorkspace {
    float4 *acc;
    unsigned long long *keys;
    unsigned long long *zmax;
    int *znan;
    size_t bytes, acc_bytes, keys_bytes, small_bytes;
};

static FwdWorkspace fwd_layout(void *base, int H, int W, int n_pairs, int n_slots, int flags) {
    FwdWorkspace w;
    size_t hw = (size_t)H * W;
    w.acc_bytes = align_up((size_t)n_slots * hw * sizeof(float4), 256);
    w.keys_bytes = (flags & SYN3R_FWD_WINNER) ? align_up((size_t)n_slots * hw * 8, 256) : 0;
    size_t zb = align_up((size_t)n_pairs * 8, 256), nb = align_up((size_t)n_pairs * 4, 256);
    w.small_bytes = zb + nb;
    char *b = (char *)base;
    w.acc = (float4 *)b;
    w.keys = w.keys_bytes ? (unsigned long long *)(b + w.acc_bytes) : nullptr;
    w.zmax = (unsigned long long *)(b + w.acc_bytes + w.keys_bytes);
    w.znan = (int *)(b + w.acc_bytes + w.keys_bytes + zb);
    w.bytes = w.acc_bytes + w.keys_bytes + w.small_bytes;
    return w;
}

}  // namespace syn3r

using namespace syn3r;

extern "C" {

int syn3r_version(void) { return SYN3R_ABI_VERSION; }
const char *syn3r_last_error(void) { return g_err; }

int syn3r_device_info(int *sm, int *smem, int *major, int *minor) {
    DevInfo d;
    int rc = device_info(d);
    if (rc != SYN3R_OK) return rc;
    if (sm) *sm = d.sms;
    if (smem) *smem = d.smem;
    if (major) *major = d.major;
    if (minor) *minor = d.minor;
    return SYN3R_OK;
}

size_t syn3r_fwd_workspace_bytes(int H, int W, int n_pairs, int n_slots, int flags) {
    if (H <= 0 || W <= 0 || n_pairs <= 0 || n_slots <= 0) return 0;
    return fwd_layout(nullptr, H, W, n_pairs, n_slots, flags).bytes;
}

int syn3r_forward_warp_batched(const float *rgbd, const double *depth64, const uint8_t *src_mask,
                               const double *Kinv_src, const double *pair_par, const int32_t *pair_src,
                               const int32_t *pair_slot, int S, int P, int n_slots, int H, int W, int flags,
                               float *out_rgb_f32, uint8_t *out_rgb_u8, uint8_t *out_mask, int32_t *out_winner,
                               double *out_flow, int32_t *out_status, void *workspace, size_t workspace_bytes,
                               void *stream) {
    SYN3R_REQUIRE(rgbd && Kinv_src && pair_par && pair_src && out_mask && workspace, "null required pointer");
    SYN3R_REQUIRE(S > 0 && P > 0 && n_slots > 0 && H > 0 && W > 0, "non-positive size");
    SYN3R_REQUIRE((size_t)S * H * W < 0x7fffffffull, "S*H*W must fit the 31-bit winner index");
    SYN3R_REQUIRE(pair_slot || n_slots == P, "PAIR mode (pair_slot == NULL) needs n_slots == P");
    SYN3R_REQUIRE(!out_winner || (flags & SYN3R_FWD_WINNER), "out_winner needs SYN3R_FWD_WINNER");
    SYN3R_REQUIRE(!out_rgb_u8 || (flags & SYN3R_FWD_IS_IMAGE), "out_rgb_u8 needs SYN3R_FWD_IS_IMAGE");
    int rc = check_device();
    if (rc != SYN3R_OK) return rc;
    FwdWorkspace ws = fwd_layout(workspace, H, W, P, n_slots, flags);
    if (workspace_bytes < ws.bytes || ((uintptr_t)workspace & 255)) {
        set_error("forward workspace: need %zu bytes, 256-byte aligned (got %zu at %p)", ws.bytes, workspace_bytes,
                  workspace);
        return SYN3R_EWORKSPACE;
    }
    cudaStream_t st = (cudaStream_t)stream;
    const size_t hw = (size_t)H * W;
    const int phase_bits = SYN3R_FWD_PHASE_ZMAX | SYN3R_FWD_PHASE_SPLAT | SYN3R_FWD_PHASE_RESOLVE;
    const int phases = (flags & phase_bits) ? (flags & phase_bits) : phase_bits;
    dim3 g = stride_grid(hw, 256, P);
    if (phases & SYN3R_FWD_PHASE_ZMAX) {
        SYN3R_TRY_CUDA(cudaMemsetAsync(ws.acc, 0, ws.acc_bytes, st));
        if (ws.keys) SYN3R_TRY_CUDA(cudaMemsetAsync(ws.keys, 0xff, ws.keys_bytes, st));
        SYN3R_TRY_CUDA(cudaMemsetAsync(ws.zmax, 0, ws.small_bytes, st));
        if (out_status) SYN3R_TRY_CUDA(cudaMemsetAsync(out_status, 0, (size_t)n_slots * 4, st));
        k_fwd_zmax<<<g, 256, 0, st>>>((const float4 *)rgbd, depth64, Kinv_src, pair_par, pair_src, H, W, ws.zmax,
                                      ws.znan);
    }
    if (phases & SYN3R_FWD_PHASE_SPLAT) {
        if (flags & SYN3R_FWD_WINNER)
            k_fwd_splat<true><<<g, 256, 0, st>>>((const float4 *)rgbd, depth64, src_mask, Kinv_src, pair_par,
                                                 pair_src, pair_slot, H, W, ws.zmax, ws.znan, ws.acc, ws.keys,
                                                 out_flow);
        else
            k_fwd_splat<false><<<g, 256, 0, st>>>((const float4 *)rgbd, depth64, src_mask, Kinv_src, pair_par,
                                                  pair_src, pair_slot, H, W, ws.zmax, ws.znan, ws.acc, nullptr,
                                                  out_flow);
    }
    if (phases & SYN3R_FWD_PHASE_RESOLVE) {
        dim3 gr = stride_grid(hw, 256, n_slots);
        k_fwd_resolve<<<gr, 256, 0, st>>>(ws.acc, ws.keys, hw, 3, flags, out_rgb_f32, out_rgb_u8, out_mask,
                                          out_winner, out_status);
    }
    return check_cuda(cudaGetLastError(), "forward_warp_batched launches");
}

int syn3r_compute_transformed_points(const double *depth, const double *Kinv, const double *par, int H, int W,
                                     double *out_q, double *out_X, void *stream) {
    SYN3R_REQUIRE(depth && Kinv && par && out_q, "null required pointer");
    SYN3R_REQUIRE(H > 0 && W > 0, "non-positive size");
    int rc = check_device();
    if (rc != SYN3R_OK) return rc;
    k_transformed_points<<<stride_grid((size_t)H * W, 256, 1), 256, 0, (cudaStream_t)stream>>>(depth, Kinv, par, H,
                                                                                               W, out_q, out_X);
    return check_cuda(cudaGetLastError(), "k_transformed_points");
}

int syn3r_bilinear_splat(const float *frame, const uint8_t *mask1, const double *z, const double *flow,
                         const uint8_t *flow_mask, int H, int W, int C, int flags, float *out, uint8_t *out_u8,
                         uint8_t *out_mask, int32_t *out_status, void *workspace, size_t workspace_bytes,
                         void *stream) {
    SYN3R_REQUIRE(frame && z && flow && out_mask && workspace, "null required pointer");
    SYN3R_REQUIRE(H > 0 && W > 0 && C >= 1 && C <= 3, "need H,W > 0 and 1 <= C <= 3");
    SYN3R_REQUIRE(!(flags & SYN3R_FWD_WINNER), "winner index is not defined for bilinear_splat");
    SYN3R_REQUIRE(!out_u8 || (flags & SYN3R_FWD_IS_IMAGE), "out_u8 needs SYN3R_FWD_IS_IMAGE");
    int rc = check_device();
    if (rc != SYN3R_OK) return rc;
    FwdWorkspace ws = fwd_layout(workspace, H, W, 1, 1, flags);
    if (workspace_bytes < ws.bytes || ((uintptr_t)workspace & 255)) {
        set_error("splat workspace: need %zu bytes, 256-byte aligned (got %zu)", ws.bytes, workspace_bytes);
        return SYN3R_EWORKSPACE;
    }
    cudaStream_t st = (cudaStream_t)stream;
    const size_t hw = (size_t)H * W;
    SYN3R_TRY_CUDA(cudaMemsetAsync(ws.acc, 0, ws.acc_bytes, st));
    SYN3R_TRY_CUDA(cudaMemsetAsync(ws.zmax, 0, ws.small_bytes, st));
    if (out_status) SYN3R_TRY_CUDA(cudaMemsetAsync(out_status, 0, 4, st));
    dim3 g = stride_grid(hw, 256, 1);
    k_splat_zmax<<<g, 256, 0, st>>>(z, hw, ws.zmax, ws.znan);
    k_splat_from_flow<<<g, 256, 0, st>>>(frame, mask1, z, flow, flow_mask, H, W, C, ws.zmax, ws.znan, ws.acc);
    k_fwd_resolve<<<g, 256, 0, st>>>(ws.acc, nullptr, hw, C, flags, out, out_u8, out_mask, nullptr, out_status);
    return check_cuda(cudaGetLastError(), "bilinear_splat launches");
}

}  // extern "C"
