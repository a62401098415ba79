// abi.cu — the extern "C" surface of libsyn3r_warp.so (include/syn3r_warp.h) and its launchers.
#include <stdarg.h>
#include <string.h>

#include "forward_fx_overlap.cuh"

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
    // legacy (global accumulator) layout
    float4 *acc;
    unsigned long long *keys;
    // shared by both layouts
    unsigned long long *zmax;
    int *znan;
    // tile layout
    unsigned *counter;          // [0] work counter, [16] cmax_bits, [32] redo_count, [48] redo work counter
    BlockStats *stats;
    CellBox *blk_box, *sb_box;
    PairFast *pair_fast;
    int *redo_list;
    double *blk_zhi;
    unsigned long long *zlo;
    unsigned *tile_cnt;         // per-(pair, tile) block lists (NULL: frame too large for the one-pass cull)
    int *tile_list;
    size_t bytes, acc_bytes, keys_bytes, small_bytes;
};

static FwdWorkspace fwd_layout(void *base, int H, int W, int S, int n_pairs, int n_slots, int flags) {
    FwdWorkspace w{};
    size_t hw = (size_t)H * W;
    char *b = (char *)base;
    size_t zb = align_up((size_t)n_pairs * 8, 256), nb = align_up((size_t)n_pairs * 4, 256);
    if (flags & SYN3R_FWD_LEGACY_GLOBAL) {
        w.acc_bytes = align_up((size_t)n_slots * hw * sizeof(float4), 256);
        w.keys_bytes = (flags & SYN3R_FWD_WINNER) ? align_up((size_t)n_slots * hw * 8, 256) : 0;
        w.small_bytes = zb + nb;
        w.acc = (float4 *)b;
        w.keys = w.keys_bytes ? (unsigned long long *)(b + w.acc_bytes) : nullptr;
        w.zmax = (unsigned long long *)(b + w.acc_bytes + w.keys_bytes);
        w.znan = (int *)(b + w.acc_bytes + w.keys_bytes + zb);
        w.bytes = w.acc_bytes + w.keys_bytes + w.small_bytes;
        return w;
    }
    TileGeom g = make_geom(H, W);
    const bool lists = SYN3R_FX_TILE_LISTS && g.nsb * SYN3R_SB_BLOCKS <= SYN3R_FX_FLAT;
    const size_t cnt_b = lists ? align_up((size_t)n_pairs * g.ntiles * 4, 256) : 0;
    w.small_bytes = zb + nb + 256 + zb + cnt_b;         // ... + zlo keys + tile-list counts (all zeroed per call)
    size_t stats_b = align_up((size_t)S * g.nsb * SYN3R_SB_BLOCKS * sizeof(BlockStats), 256);
    size_t blk_b = align_up((size_t)n_pairs * g.nsb * SYN3R_SB_BLOCKS * sizeof(CellBox), 256);
    size_t sb_b = align_up((size_t)n_pairs * g.nsb * sizeof(CellBox), 256);
    w.zmax = (unsigned long long *)b;
    w.znan = (int *)(b + zb);
    w.counter = (unsigned *)(b + zb + nb);
    size_t pf_b = align_up((size_t)n_pairs * sizeof(PairFast), 256);
    size_t redo_b = align_up((size_t)n_slots * g.ntiles * sizeof(int), 256);
    w.stats = (BlockStats *)(b + w.small_bytes);
    w.blk_box = (CellBox *)(b + w.small_bytes + stats_b);
    w.sb_box = (CellBox *)(b + w.small_bytes + stats_b + blk_b);
    w.pair_fast = (PairFast *)(b + w.small_bytes + stats_b + blk_b + sb_b);
    w.redo_list = (int *)(b + w.small_bytes + stats_b + blk_b + sb_b + pf_b);
    size_t zhi_b = align_up((size_t)n_pairs * g.nsb * SYN3R_SB_BLOCKS * sizeof(double), 256);
    w.blk_zhi = (double *)(b + w.small_bytes + stats_b + blk_b + sb_b + pf_b + redo_b);
    w.zlo = (unsigned long long *)(b + zb + nb + 256);
    const size_t list_b = lists ? align_up((size_t)n_pairs * g.ntiles * SYN3R_TILE_LIST_CAP * 4, 256) : 0;
    w.tile_cnt = lists ? (unsigned *)(b + zb + nb + 256 + zb) : nullptr;
    w.tile_list = lists ? (int *)(b + w.small_bytes + stats_b + blk_b + sb_b + pf_b + redo_b + zhi_b) : nullptr;
    w.bytes = w.small_bytes + stats_b + blk_b + sb_b + pf_b + redo_b + zhi_b + list_b;
    return w;
}

template <bool WINNER, bool EXACT>
static int launch_tile(const TileArgs &ta, int n_items, cudaStream_t st) {
    static thread_local int configured_dev = -1;
    const size_t smem = tile_smem_bytes(WINNER);
    int dev = -1;
    SYN3R_TRY_CUDA(cudaGetDevice(&dev));
    if (configured_dev != dev) {
        SYN3R_TRY_CUDA(cudaFuncSetAttribute(k_fwd_tile<WINNER, EXACT>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                            (int)smem));
        configured_dev = dev;
    }
    int grid = sm_count();
    if (grid > n_items) grid = n_items;
    k_fwd_tile<WINNER, EXACT><<<grid, SYN3R_TILE_THREADS, smem, st>>>(ta);
    return SYN3R_OK;
}

// fixed-point kernel, then the fp32 kernel over whatever work items it handed back (normally none)
template <bool WINNER, bool FLAT, bool AHEAD>
static int launch_tile_fx_as(const TileArgs &ta, int n_items, cudaStream_t st) {
    static thread_local int configured_dev = -1;
    const size_t smem = fx_smem_bytes(WINNER), smem_redo = tile_smem_bytes(WINNER);
    int dev = -1;
    SYN3R_TRY_CUDA(cudaGetDevice(&dev));
    if (configured_dev != dev) {
        SYN3R_TRY_CUDA(cudaFuncSetAttribute(k_fwd_tile_fx<WINNER, FLAT, AHEAD>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                            (int)smem));
        SYN3R_TRY_CUDA(cudaFuncSetAttribute(k_fwd_tile<WINNER, false>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                            (int)smem_redo));
        configured_dev = dev;
    }
    int grid = sm_count() * (SYN3R_FX_TWO_CTAS ? 2 : 1);
    if (grid > n_items) grid = n_items;
    k_fwd_tile_fx<WINNER, FLAT, AHEAD><<<grid, SYN3R_TILE_THREADS, smem, st>>>(ta);
    TileArgs tr = ta;
    tr.redo_mode = 1;
    k_fwd_tile<WINNER, false><<<grid, SYN3R_TILE_THREADS, smem_redo, st>>>(tr);
    return SYN3R_OK;
}
// frames of up to SYN3R_FX_FLAT source blocks (1024 x 576 and smaller) take the one-pass cull
#if SYN3R_FX_OV_AVAILABLE
// PAIR mode with block lists: the kernel that overlaps an item's resolve step with the next item's sweep 1
template <bool WINNER>
static int launch_tile_fx_ov(const TileArgs &ta, int n_items, cudaStream_t st) {
    static thread_local int configured_dev = -1;
    const size_t smem = fx_ov_smem_bytes(WINNER), smem_redo = tile_smem_bytes(WINNER);
    int dev = -1;
    SYN3R_TRY_CUDA(cudaGetDevice(&dev));
    if (configured_dev != dev) {
        SYN3R_TRY_CUDA(cudaFuncSetAttribute(k_fwd_tile_fx_ov<WINNER>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        SYN3R_TRY_CUDA(cudaFuncSetAttribute(k_fwd_tile<WINNER, false>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                            (int)smem_redo));
        configured_dev = dev;
    }
    int grid = sm_count();
    if (grid > n_items) grid = n_items;
    k_fwd_tile_fx_ov<WINNER><<<grid, SYN3R_TILE_THREADS, smem, st>>>(ta);
    TileArgs tr = ta;
    tr.redo_mode = 1;
    k_fwd_tile<WINNER, false><<<grid, SYN3R_TILE_THREADS, smem_redo, st>>>(tr);
    return SYN3R_OK;
}
#endif

static int g_fx_overlap = 0;       // syn3r_set_fx_overlap (A/B switch; off: measured equal to k_fwd_tile_fx, see forward_fx_overlap.cuh)

template <bool WINNER>
static int launch_tile_fx(const TileArgs &ta, int n_items, cudaStream_t st) {
    if (ta.g.nsb * SYN3R_SB_BLOCKS > SYN3R_FX_FLAT) return launch_tile_fx_as<WINNER, false, false>(ta, n_items, st);
#if SYN3R_FX_OV_AVAILABLE
    if (g_fx_overlap && !ta.pair_slot && ta.tile_cnt) return launch_tile_fx_ov<WINNER>(ta, n_items, st);
#endif
    // PAIR mode with block lists: every work item is claimed and staged one item ahead
    if (SYN3R_FX_PREFETCH && !ta.pair_slot && ta.tile_cnt) return launch_tile_fx_as<WINNER, true, true>(ta, n_items, st);
    return launch_tile_fx_as<WINNER, true, false>(ta, n_items, st);
}

}  // namespace syn3r

namespace syn3r {
// self-test of ddiv2_rn against the library division on pseudo-random operands (mode 0: moderate
// exponents like projected coordinates, mode 1: arbitrary bit patterns incl. denormals, inf, NaN)
__global__ void k_selftest_ddiv2(unsigned long long n, unsigned long long seed, int mode,
                                 unsigned long long *__restrict__ mismatches) {
    unsigned long long bad = 0;
    for (unsigned long long i = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x; i < n;
         i += (unsigned long long)gridDim.x * blockDim.x) {
        unsigned long long v[3];
        unsigned long long x = (i + 1) * 0x9E3779B97F4A7C15ull + seed;
#pragma unroll
        for (int k = 0; k < 3; ++k) {
            x ^= x >> 30; x *= 0xBF58476D1CE4E5B9ull; x ^= x >> 27; x *= 0x94D049BB133111EBull; x ^= x >> 31;
            v[k] = x;
            x += 0x9E3779B97F4A7C15ull;
        }
        double d[3];
#pragma unroll
        for (int k = 0; k < 3; ++k) {
            if (mode == 0) {   // sign | exponent in [1023-40, 1023+40] | random mantissa
                unsigned long long e = 1023 - 40 + (v[k] >> 52) % 81;
                d[k] = __longlong_as_double((long long)((v[k] & 0x800fffffffffffffull) | (e << 52)));
            } else {
                d[k] = __longlong_as_double((long long)v[k]);
            }
        }
        double c0, c1;
        ddiv2_rn(d[0], d[1], d[2], c0, c1);
        const double r0 = __ddiv_rn(d[0], d[2]), r1 = __ddiv_rn(d[1], d[2]);
        const bool same0 = (__double_as_longlong(c0) == __double_as_longlong(r0)) || (c0 != c0 && r0 != r0);
        const bool same1 = (__double_as_longlong(c1) == __double_as_longlong(r1)) || (c1 != c1 && r1 != r1);
        bad += (same0 ? 0 : 1) + (same1 ? 0 : 1);
    }
    if (bad) atomicAdd(mismatches, bad);
}
}  // namespace syn3r

using namespace syn3r;

extern "C" {

int syn3r_selftest_ddiv2(unsigned long long n, unsigned long long seed, int mode, unsigned long long *mismatches,
                         void *stream) {
    SYN3R_REQUIRE(mismatches, "null required pointer");
    int rc = check_device();
    if (rc != SYN3R_OK) return rc;
    k_selftest_ddiv2<<<sm_count() * 8, 256, 0, (cudaStream_t)stream>>>(n, seed, mode, mismatches);
    return check_cuda(cudaGetLastError(), "k_selftest_ddiv2");
}


int syn3r_version(void) { return SYN3R_ABI_VERSION; }

int syn3r_set_fx_overlap(int on) {
    const int was = g_fx_overlap;
    g_fx_overlap = on ? 1 : 0;
    return was;
}
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

size_t syn3r_fwd_workspace_bytes(int H, int W, int S, int n_pairs, int n_slots, int flags) {
    if (H <= 0 || W <= 0 || S <= 0 || n_pairs <= 0 || n_slots <= 0) return 0;
    return fwd_layout(nullptr, H, W, S, n_pairs, n_slots, flags).bytes;
}

int syn3r_forward_warp_batched(const float *rgbd, const double *depth64, const uint8_t *src_mask,
                               const double *Kinv_src, const double *pair_par, const int32_t *pair_src,
                               const int32_t *pair_slot, int S, int P, int n_slots, int H, int W, int flags,
                               float *out_rgb_f32, uint8_t *out_rgb_u8, uint8_t *out_mask, int32_t *out_winner,
                               double *out_flow, int32_t *out_status, void *workspace, size_t workspace_bytes,
                               void *stream) {
    // an empty batch is a valid no-op (the reference's per-frame loop simply does not run): nothing is read or written
    if (P == 0 && (n_slots == 0 || pair_slot == nullptr)) return SYN3R_OK;
    SYN3R_REQUIRE(rgbd && Kinv_src && pair_par && pair_src && out_mask && workspace, "null required pointer");
    SYN3R_REQUIRE(S > 0 && P > 0 && n_slots > 0 && H > 0 && W > 0, "non-positive size");
    SYN3R_REQUIRE((size_t)S * H * W < 0x7fffffffull, "S*H*W must fit the 31-bit winner index");
    SYN3R_REQUIRE(pair_slot || n_slots == P, "PAIR mode (pair_slot == NULL) needs n_slots == P");
    SYN3R_REQUIRE(!out_winner || (flags & SYN3R_FWD_WINNER), "out_winner needs SYN3R_FWD_WINNER");
    SYN3R_REQUIRE(!out_rgb_u8 || (flags & SYN3R_FWD_IS_IMAGE), "out_rgb_u8 needs SYN3R_FWD_IS_IMAGE");
    int rc = check_device();
    if (rc != SYN3R_OK) return rc;
    SYN3R_REQUIRE(H <= 32767 && W <= 32767, "H and W must fit int16 cell boxes");
    FwdWorkspace ws = fwd_layout(workspace, H, W, S, P, n_slots, flags);
    if (workspace_bytes < ws.bytes || ((uintptr_t)workspace & 255)) {
        set_error("forward workspace: need %zu bytes, 256-byte aligned (got %zu at %p)", ws.bytes, workspace_bytes,
                  workspace);
        return SYN3R_EWORKSPACE;
    }
    cudaStream_t st = (cudaStream_t)stream;
    const size_t hw = (size_t)H * W;
    const int phase_bits = SYN3R_FWD_PHASE_ZMAX | SYN3R_FWD_PHASE_SPLAT | SYN3R_FWD_PHASE_RESOLVE;
    const int phases = (flags & phase_bits) ? (flags & phase_bits) : phase_bits;

    if (!(flags & SYN3R_FWD_LEGACY_GLOBAL)) {
        // ---- tile path: accumulators stay in shared memory (forward_tile.cuh) ---------------------------
        const TileGeom g = make_geom(H, W);
        SYN3R_REQUIRE(P <= 65535 && S <= 65535, "tile path: at most 65535 pairs / sources per call");
        if (phases & SYN3R_FWD_PHASE_ZMAX) {
            SYN3R_TRY_CUDA(cudaMemsetAsync(ws.zmax, 0, ws.small_bytes, st));
            k_src_stats<<<dim3(g.nsb, S), 1024, 0, st>>>((const float4 *)rgbd, depth64, g, ws.stats, ws.counter + 16);
            // boxes + bounds of z' per block, then the exact max z' over the blocks that can hold it
            k_fwd_prep<<<dim3(g.nsb, P), 32, 0, st>>>((const float4 *)rgbd, depth64, Kinv_src, pair_par, pair_src, g,
                                                      ws.stats, ws.zmax, ws.znan, ws.blk_box, ws.sb_box, ws.pair_fast,
                                                      1, ws.blk_zhi, ws.zlo, ws.tile_cnt, ws.tile_list);
            k_fwd_prep<<<dim3(g.nsb, P), 256, 0, st>>>((const float4 *)rgbd, depth64, Kinv_src, pair_par, pair_src, g,
                                                       ws.stats, ws.zmax, ws.znan, ws.blk_box, ws.sb_box, ws.pair_fast,
                                                       2, ws.blk_zhi, ws.zlo, nullptr, nullptr);
            k_fwd_finalize<<<(P + 127) / 128, 128, 0, st>>>(ws.zmax, ws.znan, P, ws.pair_fast);
        }
        if (phases & SYN3R_FWD_PHASE_SPLAT) {
            if (!(phases & SYN3R_FWD_PHASE_ZMAX)) {      // counters only; cmax_bits ([16]) is phase-1 state
                SYN3R_TRY_CUDA(cudaMemsetAsync(ws.counter, 0, 64, st));
                SYN3R_TRY_CUDA(cudaMemsetAsync(ws.counter + 32, 0, 128, st));
            }
            if (out_status) SYN3R_TRY_CUDA(cudaMemsetAsync(out_status, 0, (size_t)n_slots * 4, st));
            TileArgs ta{(const float4 *)rgbd, depth64, src_mask, Kinv_src, pair_par, pair_src, pair_slot, P, n_slots,
                        flags, g, ws.zmax, ws.znan, ws.blk_box, ws.sb_box, ws.tile_cnt, ws.tile_list, ws.counter, out_rgb_f32, out_rgb_u8,
                        out_mask, out_winner, out_status, ws.pair_fast, ws.counter + 16, ws.redo_list,
                        ws.counter + 32, ws.counter + 48, 0};
            const int n_items = n_slots * g.ntiles;
            const bool win = (flags & SYN3R_FWD_WINNER) != 0, exact = (flags & SYN3R_FWD_EXACT_WEIGHT) != 0;
            const bool fp32acc = (flags & SYN3R_FWD_FP32_TILE) != 0;
            int rc2;
            if (exact) rc2 = win ? launch_tile<true, true>(ta, n_items, st) : launch_tile<false, true>(ta, n_items, st);
            else if (fp32acc || depth64 || src_mask)     // f64 depth / mask1: per-pair API extras, fp32 tile kernel
                rc2 = win ? launch_tile<true, false>(ta, n_items, st) : launch_tile<false, false>(ta, n_items, st);
            else rc2 = win ? launch_tile_fx<true>(ta, n_items, st) : launch_tile_fx<false>(ta, n_items, st);
            if (rc2 != SYN3R_OK) return rc2;
            if (out_flow) {
                SYN3R_REQUIRE(H <= 65535, "flow output: H must be <= 65535");
                k_fwd_flow<<<dim3((W + 255) / 256, H, P), 256, 0, st>>>((const float4 *)rgbd, depth64, Kinv_src,
                                                                        pair_par, pair_src, H, W, out_flow);
            }
        }
        return check_cuda(cudaGetLastError(), "forward_warp_batched (tile path) launches");
    }

    // ---- legacy path: global fp32 accumulators + REDG (kept for A/B measurements) -----------------------
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
    FwdWorkspace ws = fwd_layout(workspace, H, W, 1, 1, 1, flags | SYN3R_FWD_LEGACY_GLOBAL);
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

#if SYN3R_FX_PHASE_CLOCKS
// A/B build only (tools/phase_clocks.py): per-CTA phase cycle sums of the last k_fwd_tile_fx launch
int syn3r_debug_fx_phase(unsigned long long *host_out, int n_ctas) {
    SYN3R_TRY_CUDA(cudaDeviceSynchronize());
    SYN3R_TRY_CUDA(cudaMemcpyFromSymbol(host_out, g_fx_phase, sizeof(unsigned long long) * 8 * (size_t)n_ctas));
    return SYN3R_OK;
}
#endif
}  // extern "C"
