// forward_fx.cuh — forward depth warp, target-tile-owned, INTEGER shared-memory accumulation.
//
// Why: on sm_100a the only native shared-memory atomics are 32-bit integer ones (ATOMS.ADD / MIN / MAX, one
// warp instruction per cycle per SM, tools/microbench/atomics.cu); atomicAdd(float) and every 64-bit
// atomic compile to LDS + ATOMS.CAST.SPIN loops, which is what bounded k_fwd_tile (forward_tile.cuh) at
// 3.1 ms on the headline workload.  This kernel produces the same outputs with native atomics only.
//
// Per work item (output slot, 128x48 target tile), two sweeps over the culled source blocks:
//   sweep 1   every corner that lands in the tile does  ATOMS.MAX Wmax[cell] <- bits(w)      (w > 0 fp32)
//             and, for the winner output,               ATOMS.MIN zkey[cell] <- ordered(float z')
//             and counts itself,                          ATOMS.ADD N[cell] <- 1
//   sweep 2   the corner's weight is expressed relative to the cell's own power of two,
//             q = w * 2^(21 - exponent(Wmax[cell])) in (0, 2^22), and accumulated as integers:
//             S += rint(q), I_c += rint(q * colour_c / cscale)  (4 ATOMS.ADD); rint() is the 1.5 * 2^23
//             magic-number add, and what is accumulated is the raw bit pattern MAGIC_BITS + rint(.) - the
//             N * MAGIC_BITS is taken off once per cell in the resolve step (mod 2^32) instead of once per add.
//             The winner index is an ATOMS.MIN among the corners whose z-key equals the cell's minimum.
//   resolve   colour = I_c / S * cscale, mask = Wmax > 0, winner; planes are cleared for the next item.
// A block-floating-point accumulator with a per-cell exponent: the largest contribution of a cell keeps 21-22
// significant bits, a cell's rounding error is <= N * 2^-22 of full scale for N contributions (typically
// sqrt(N) * 1e-7), sums are order-independent (bitwise reproducible run to run), and the hole mask does not
// depend on the accumulation at all.  A cell with N >= 512 contributions could pass S = 2^31 (each q < 2^22): such
// a work item is handed to k_fwd_tile (fp32 CAS accumulation) through a redo list.
//
// The projection is evaluated as q = d (A p) + b with the composite A = K2 R K1inv (12 DFMA + one reciprocal
// per pixel instead of the reference's ~60-operation chain).  Its floor() can differ from the bit-faithful
// chain only when the coordinate is within the bound k_fwd_prep derives of an
// integer; such pixels (all of them for an identity pose, ~1e-9 of them for a generic one) take
// exact_pixel(), which IS the bit-faithful chain.  Cell assignment, hence the hole mask and the set of
// contributors, is therefore identical to the reference's; weights differ by ~1e-12 relative.
#pragma once
#include "forward_tile.cuh"

namespace syn3r {

#define SYN3R_FX_BITS 21
#ifndef SYN3R_FX_UNIT_ROWS
#define SYN3R_FX_UNIT_ROWS 4            /* rows of a source block one warp takes at a time (2 per iteration) */
#endif
#define SYN3R_FX_UNITS (SYN3R_BLK_H / SYN3R_FX_UNIT_ROWS)
// End game of a sweep: the last SYN3R_FX_TAIL_UNITS work units of a block list are handed out as single iterations
// (2 rows) instead of 4-row units, so that the warps reach the barrier that ends the sweep within one iteration of each
// other (measured with SYN3R_FX_PHASE_CLOCKS: the tails were 4 % + 10 % of the kernel with 4-row units throughout).
#ifndef SYN3R_FX_TAIL_UNITS
#define SYN3R_FX_TAIL_UNITS 0
#endif
// SYN3R_FX_PREFETCH: PAIR mode with block lists - the NEXT work item is claimed while sweep 2 of the current one runs,
// and its per-pair constants, its source index and its block list are copied global -> shared by cp.async during the
// resolve step (no registers held, listB is free by then), so that an item no longer starts with three dependent L2
// round trips (work counter -> pair constants -> list) in front of its first barrier.
#ifndef SYN3R_FX_PREFETCH
#define SYN3R_FX_PREFETCH 1
#endif
__device__ __forceinline__ void cp_async_16(void *smem_dst, const void *gsrc) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"((unsigned)__cvta_generic_to_shared(smem_dst)), "l"(gsrc) : "memory");
}
__device__ __forceinline__ void cp_async_4(void *smem_dst, const void *gsrc) {
    asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"((unsigned)__cvta_generic_to_shared(smem_dst)), "l"(gsrc) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_all;" ::: "memory"); }
// The resolve step is a burst of shared-memory loads, then arithmetic, then stores, and all 24 warps walk through those
// phases together (LSU idle while they compute, issue slots idle while they load).  SYN3R_FX_RESOLVE_STAGGER_NS > 0:
// half of the warps of every scheduler start that many nanoseconds later, so that one half's loads and stores overlap
// the other half's arithmetic.
#ifndef SYN3R_FX_RESOLVE_STAGGER_NS
#define SYN3R_FX_RESOLVE_STAGGER_NS 0
#endif
// rows of a source block a thread takes per iteration (independent dependency chains per thread).  2 = the default
// (768 threads at 80 registers); 3 / 4 = A/B builds for fewer, wider threads (SYN3R_TILE_THREADS=512 at 128 registers)
#ifndef SYN3R_FX_ROWS
#define SYN3R_FX_ROWS 2
#endif
#define SYN3R_FX_FLAT 2304    /* FLAT kernels: frames of up to this many source blocks (a 1024 x 576 frame), culled in one
                                 pass; listB holds them all */
#if SYN3R_FX_TWO_CTAS
#define SYN3R_FX_LISTB_CAP 1024     /* A/B build only: hits of one 128 x 24 tile (no overflow guard) */
#else
#define SYN3R_FX_LISTB_CAP SYN3R_FX_FLAT
#endif
#define SYN3R_FX_FLAT_PER_THREAD ((SYN3R_FX_FLAT + SYN3R_TILE_THREADS - 1) / SYN3R_TILE_THREADS)
// Floor experiments (timing builds only, results are WRONG for every value but 0):
//   1  sweep 2 skipped            2  sweep 2 with an fp32 projection, no exactness filter, approximate weight (what a
//   3  no shared-memory atomics      replay of cached sweep-1 geometry would cost)     4  exact weight in sweep 1 too
//   5  the floor of the formulation: ONE sweep that loads and projects every culled row and stops there (the vote on
//      the projected cells keeps the arithmetic alive), then the resolve step with all its output stores
#ifndef SYN3R_FX_EXPERIMENT
#define SYN3R_FX_EXPERIMENT 0
#endif
#ifndef SYN3R_FX_BATCH_LDS
#define SYN3R_FX_BATCH_LDS 1
#endif
#ifndef SYN3R_FX_TRASH_SPLIT
#define SYN3R_FX_TRASH_SPLIT 1
#endif
// SYN3R_FX_PHASE_CLOCKS (A/B build, tools/phase_clocks.py): per-CTA cycle sums of the phases of a work item
//   [0] claim  [1] cull  [2] sweep 1 until its first warp runs dry  [3] sweep-1 tail (until the barrier releases)
//   [4] sweep 2 until its first warp runs dry  [5] sweep-2 tail  [6] resolve  [7] items
#ifndef SYN3R_FX_PHASE_CLOCKS
#define SYN3R_FX_PHASE_CLOCKS 0
#endif
#if SYN3R_FX_PHASE_CLOCKS
__device__ unsigned long long g_fx_phase[1024 * 8];
#define FX_CLK(slot) do { if (threadIdx.x == 0) sm.ctl[slot] = (int)clock(); } while (0)
#else
#define FX_CLK(slot) do { } while (0)
#endif
#define SYN3R_FX_MAGIC 12582912.0f          /* 1.5 * 2^23: low mantissa bits of (v + MAGIC) are rint(v) */
#define SYN3R_FX_MAGIC_BITS 0x4B400000

// Shared-memory layout: planes of CELLS 32-bit words, addressed as fx_smem[PLANE * CELLS + cell] so that the
// plane and the corner offset fold into the immediate field of LDS / ATOMS.
//   WMAX  fp32 bit pattern of the largest corner weight (0 = untouched)     S   sum of MAGIC_BITS + rint(q)
//   I0..2 sum of MAGIC_BITS + rint(q * c / cscale)     N  number of corner updates     ZKEY, IDX  winner (WINNER only)
extern __shared__ __align__(16) unsigned fx_smem[];
enum { FX_WMAX = 0, FX_S = 1, FX_I0 = 2, FX_I1 = 3, FX_I2 = 4, FX_N = 5, FX_ZKEY = 6, FX_IDX = 7 };
#define FX_MAX_N 512u
#define FX_CELLS (SYN3R_TILE_W * SYN3R_TILE_H)
#define FX_PLANE (SYN3R_TILE_W * (SYN3R_TILE_H + 1))       /* one extra row: per-lane trash cells */

struct FxAux {                  // the rest of the shared memory, behind the planes
    double2 *logtab;
    int *listA, *listB, *ctl;
    unsigned char *stage;       // staged constants of the next work item: PairFast (128 B), list count, source index
};
static_assert(sizeof(PairFast) == 128, "the staging area and its 8 x 16-byte copies assume a 128-byte PairFast");

struct PixelGeom {              // where a source pixel lands relative to the tile
    int lxf, lyf;               // local cell of the floor corner
    float ax, bx, ay, by;       // proximity factors of the floor / ceil corner per axis (A.1 step 11)
    double z;                   // z' (:174)
};

// The bit-faithful chain for one pixel (same code as splat_block_into_tile): out of line, rarely taken.
__device__ __noinline__ PixelGeom exact_pixel(const double *__restrict__ par_g, const double *__restrict__ Ki_g,
                                              int x, int y, double d, int tx0, int ty0, int tw, int th) {
    double par[21], Ki[9];
#pragma unroll
    for (int k = 0; k < 21; ++k) par[k] = par_g[k];
#pragma unroll
    for (int k = 0; k < 9; ++k) Ki[k] = Ki_g[k];
    const double xd = (double)x, yd = (double)y;
    double X0, X1, X2;
    unproject(Ki, xd, yd, d, X0, X1, X2);
    const double *M = par, *K2 = par + 12;
    const double t0 = dot4_blas_h(M + 0, X0, X1, X2), t1 = dot4_blas_h(M + 4, X0, X1, X2),
                 t2 = dot4_blas_h(M + 8, X0, X1, X2);
    const double q0 = dot3_blas(K2 + 0, t0, t1, t2), q1 = dot3_blas(K2 + 3, t0, t1, t2),
                 z = dot3_blas(K2 + 6, t0, t1, t2);
    double cx, cy;
    ddiv2_rn(q0, q1, z, cx, cy);
    const double ox = __dadd_rn(__dadd_rn(__dsub_rn(cx, xd), xd), 1.0);
    const double oy = __dadd_rn(__dadd_rn(__dsub_rn(cy, yd), yd), 1.0);
    int xf, xc, yf, yc;
    PixelGeom g;
    axis_cells(ox, xf, xc, g.ax, g.bx);
    axis_cells(oy, yf, yc, g.ay, g.by);
    g.lxf = xf - tx0; g.lyf = yf - ty0;
    // an exact-integer landing puts both "corners" on one cell (N3): the ceil corner is not a distinct cell,
    // its factor (1.0) joins the floor corner's and its own becomes 0 (a zero weight changes nothing anywhere)
    if (xc == xf) { g.ax = __fadd_rn(g.ax, g.bx); g.bx = 0.0f; }
    if (yc == yf) { g.ay = __fadd_rn(g.ay, g.by); g.by = 0.0f; }
    g.z = z;
    return g;
}

// Local coordinate SHIFTED BY ONE (v1 = coordinate - tile origin + 1) -> n1 = rint(v1) and the signed remainder
// f in [-0.5, 0.5] (as fp32), without any conversion instruction on the fp64 side: adding 1.5 * 2^52 leaves
// rint(v1) in the low word.  For every v1 whose corners can touch the tile n1 is in [0, extent + 1] and the high
// word is exactly the magic number's; NaN, inf, negative and huge values fail that test.
__device__ __forceinline__ void rint_split(double v1, int &n1, float &f, bool &hi_ok) {
    const double t = __dadd_rn(v1, 6755399441055744.0);
    n1 = __double2loint(t);
    hi_ok = __double2hiint(t) == 0x43380000;
    f = __double2float_rn(__dsub_rn(v1, __dsub_rn(t, 6755399441055744.0)));
}

// 2^90 / dw as in fast_weight_scale (forward_tile.cuh), degree-4 polynomial (|r| < 2^-7: the r^5 term is < 1e-11).
// `clamp_hi`: the pair's max z' exceeds 5000, so the upper clip of :82 can matter (uniform per pair).
__device__ __forceinline__ float fx_weight_scale(const double2 *__restrict__ logtab, double cexp, double z,
                                                 bool clamp_hi) {
    double v = __dadd_rn(1.0, z);
    if (__double2hiint(z) < 0) v = 1.0;                                     // clip(z, 0, .): z < 0 or -0.0
    if (clamp_hi && z > 5000.0) v = 5001.0;
    const int idx = (__double2hiint(v) - 0x3ff00000) >> 14;
    const double2 T = logtab[idx];
    const double r = __fma_rn(v, T.x, -1.0);
    double p = __fma_rn(r, -0.36067376022224085, 0.48089834696298783);      // -1/(4 ln2), 1/(3 ln2)
    p = __fma_rn(r, p, -0.72134752044448170);                               // -1/(2 ln2)
    p = __fma_rn(r, p, 1.4426950408889634);                                 //  1/ln2
    const double L2 = __fma_rn(r, p, T.y);
    const double E = __dmul_rn(L2, cexp);                                   // in [0, 72.2]
    const double t = __dadd_rn(E, 6755399441055744.0);
    const int n = __double2loint(t);
    const double f = __dsub_rn(__dsub_rn(t, 6755399441055744.0), E);        // n - E in [-0.5, 0.5]
    float e2;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(e2) : "f"(__double2float_rn(f)));
    return __fmul_rn(e2, __int_as_float((127 + 90 - n) << 23));
}

// Sweep 1 only needs each cell's power-of-two scale, so its weight may be any UPPER bound of the exact one that is
// tight to well under a factor of two: 2^(90 - cexp log2(1 + clip(z', 0, 5000))) through lg2.approx / ex2.approx,
// times (1 + 2^-10).  Error budget of the exponent: cexp * (2^-23 / ln 2 [v in fp32] + 2^-22 [lg2.approx, v < 2])
// + 72 * (2^-22 [lg2.approx, relative] + 2^-23) <= 2.2e-4 for cexp <= 500, i.e. 1.5e-4 relative on the weight,
// well inside the 2^-10 pad: w_exact <= w_upper, hence q = w_exact * 2^(21 - exponent(Wmax_upper)) < 2^22 as before,
// and the cell's largest q is >= 2^21 / (1 + 2^-9).  The hole mask (Wmax != 0) depends on the proximity factors only.
__device__ __forceinline__ float fx_weight_upper(float cexp_f, float zs) {
    const float v = __fadd_rn(1.0f, fminf(fmaxf(zs, 0.0f), 5000.0f));
    float l2, e2;
    asm("lg2.approx.ftz.f32 %0, %1;" : "=f"(l2) : "f"(v));
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(e2) : "f"(__fmaf_rn(-cexp_f, l2, 90.0014084f)));   // + log2(1 + 2^-10)
    return e2;
}

struct FxPairRegs {             // per-pair values a thread keeps in registers over the pair's blocks
    double A[9], b[3];
    float k0, k1, k2, k3, k4;
    double cexp;
    float cexp_f;               // cexp in fp32 (relative error 2^-24: 72 * 2^-24 on the exponent, in the budget above)
    bool clamp_hi;
    bool upper_ok;              // cexp <= 500: sweep 1 may use fx_weight_upper
};

struct RowProj {                // fast projection of one source pixel, relative to the tile
    double q2;                  // z'
    int nx1, ny1;               // rint(local coordinate + 1)
    float fx, fy;               // signed remainders in [-0.5, 0.5]
    bool cand;                  // the floor or the ceil corner can fall inside the tile (decided exactly later)
};

// tx1d / ty1d = (double)(tile origin - 1)
__device__ __forceinline__ RowProj fx_project(const FxPairRegs &pr, double u0, double u1, double u2, double yd,
                                              double d, double tx1d, double ty1d, int tw, int th) {
    RowProj o;
    const double q0 = __fma_rn(d, __fma_rn(pr.A[1], yd, u0), pr.b[0]);
    const double q1 = __fma_rn(d, __fma_rn(pr.A[4], yd, u1), pr.b[1]);
    o.q2 = __fma_rn(d, __fma_rn(pr.A[7], yd, u2), pr.b[2]);
    double r0;
    asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(r0) : "d"(o.q2));
    double e = __fma_rn(-o.q2, r0, 1.0);
    e = __fma_rn(e, e, e);
    const double rc = __fma_rn(r0, e, r0);                                  // 1/q2, relative error ~2^-58
    const double lx = __dsub_rn(__dmul_rn(q0, rc), tx1d), ly = __dsub_rn(__dmul_rn(q1, rc), ty1d);
    bool okx, oky;
    rint_split(lx, o.nx1, o.fx, okx);
    rint_split(ly, o.ny1, o.fy, oky);
    // n1 in [0, extent + 1]  <=>  nearest integer of the local coordinate in [-1, extent]
    o.cand = okx && oky && (unsigned)o.nx1 <= (unsigned)(tw + 1) && (unsigned)o.ny1 <= (unsigned)(th + 1);
    return o;
}

#if SYN3R_FX_EXPERIMENT == 3
#define FX_ATOM(op, addr, val) do { if ((val) == 0x12345u) *(addr) = (val); } while (0)
#else
#define FX_ATOM(op, addr, val) op(addr, val)
#endif

__device__ __forceinline__ RowProj fx_project_f32(const FxPairRegs &pr, float u0, float u1, float u2, float yf, float d,
                                                  float tx1, float ty1, int tw, int th) {
    RowProj o;
    const float q0 = __fmaf_rn(d, __fmaf_rn((float)pr.A[1], yf, u0), (float)pr.b[0]);
    const float q1 = __fmaf_rn(d, __fmaf_rn((float)pr.A[4], yf, u1), (float)pr.b[1]);
    const float q2 = __fmaf_rn(d, __fmaf_rn((float)pr.A[7], yf, u2), (float)pr.b[2]);
    float rc;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(rc) : "f"(q2));
    const float lx = __fsub_rn(__fmul_rn(q0, rc), tx1), ly = __fsub_rn(__fmul_rn(q1, rc), ty1);
    const float rx = rintf(lx), ry = rintf(ly);
    o.nx1 = (int)rx; o.ny1 = (int)ry;
    o.fx = __fsub_rn(lx, rx); o.fy = __fsub_rn(ly, ry);
    o.q2 = (double)q2;
    o.cand = (unsigned)o.nx1 <= (unsigned)(tw + 1) && (unsigned)o.ny1 <= (unsigned)(th + 1);
    return o;
}

// Corner updates of one pixel.  LEAN: all four corners are distinct cells inside the tile (the warp voted so):
// the corner and the plane fold into immediate offsets of one base address, no validity logic.  Otherwise corners
// that are not part of the tile go, with weight 0, to the lane's own cell of the trash row behind the tile, so
// that the four corners still run without a branch.
template <bool WINNER, int SWEEP, bool LEAN>
__device__ __forceinline__ void fx_corners(const PixelGeom &g, float wsc, const float4 px, unsigned i, float cs_inv,
                                           int lane, int tw, int th) {
    const float wx0 = __fmul_rn(g.ax, wsc), wx1 = __fmul_rn(g.bx, wsc);
    float w4[4] = {__fmul_rn(g.ay, wx0), __fmul_rn(g.by, wx0), __fmul_rn(g.ay, wx1), __fmul_rn(g.by, wx1)};   // NW SW NE SE
    const int o00 = g.lyf * SYN3R_TILE_W + g.lxf;
    int oc[4] = {o00, o00 + SYN3R_TILE_W, o00 + 1, o00 + SYN3R_TILE_W + 1};
    if (!LEAN) {
        // (a merged "corner" of an exact-integer landing has weight 0 and is harmless wherever it points)
        const bool vx0 = (unsigned)g.lxf < (unsigned)tw, vx1 = (unsigned)(g.lxf + 1) < (unsigned)tw;
        const bool vy0 = (unsigned)g.lyf < (unsigned)th, vy1 = (unsigned)(g.lyf + 1) < (unsigned)th;
        const bool ok[4] = {vx0 && vy0, vx0 && vy1, vx1 && vy0, vx1 && vy1};
        const int trash = SYN3R_TILE_H * SYN3R_TILE_W + lane;
#pragma unroll
        for (int c = 0; c < 4; ++c) { w4[c] = ok[c] ? w4[c] : 0.0f; oc[c] = ok[c] ? oc[c] : trash; }
    }
    unsigned key = 0;
    if (WINNER) key = zkey_of(g.z);
    if (SWEEP == 1) {
#pragma unroll
        for (int c = 0; c < 4; ++c) {
            FX_ATOM(atomicMax, fx_smem + FX_WMAX * FX_PLANE + oc[c], __float_as_uint(w4[c]));   // positive floats order as ints
            FX_ATOM(atomicAdd, fx_smem + FX_N * FX_PLANE + oc[c], 1u);                          // one per sweep-2 update below
            if (WINNER && (LEAN || w4[c] > 0.0f)) FX_ATOM(atomicMin, fx_smem + FX_ZKEY * FX_PLANE + oc[c], key);
        }
    } else {
        const float c0 = __fmul_rn(px.x, cs_inv), c1 = __fmul_rn(px.y, cs_inv), c2 = __fmul_rn(px.z, cs_inv);
#if SYN3R_FX_BATCH_LDS
        // the planes sweep 2 READS (per-cell scale, winner key) are final: all eight look-ups of the pixel go out
        // together, ahead of the atomics, so that one shared-memory latency is exposed instead of eight
        unsigned wm4[4], zk4[4];
#pragma unroll
        for (int c = 0; c < 4; ++c) {
            wm4[c] = fx_smem[FX_WMAX * FX_PLANE + oc[c]];
            if (WINNER) zk4[c] = fx_smem[FX_ZKEY * FX_PLANE + oc[c]];
        }
#endif
#pragma unroll
        for (int c = 0; c < 4; ++c) {
            unsigned *q4 = fx_smem + oc[c];
#if SYN3R_FX_BATCH_LDS
            const unsigned wm = wm4[c];
#else
            const unsigned wm = q4[FX_WMAX * FX_PLANE];
#endif
            const float sc = __uint_as_float(((254u + SYN3R_FX_BITS) << 23) - (wm & 0x7f800000u));
            const float q = __fmul_rn(w4[c], sc);                           // in [0, 2^22)
            FX_ATOM(atomicAdd, q4 + FX_S * FX_PLANE, __float_as_uint(__fadd_rn(q, SYN3R_FX_MAGIC)));
            FX_ATOM(atomicAdd, q4 + FX_I0 * FX_PLANE, __float_as_uint(__fmaf_rn(q, c0, SYN3R_FX_MAGIC)));
            FX_ATOM(atomicAdd, q4 + FX_I1 * FX_PLANE, __float_as_uint(__fmaf_rn(q, c1, SYN3R_FX_MAGIC)));
            FX_ATOM(atomicAdd, q4 + FX_I2 * FX_PLANE, __float_as_uint(__fmaf_rn(q, c2, SYN3R_FX_MAGIC)));
            if (WINNER)     // branch-free: a corner that does not hold the cell's key offers the neutral element
#if SYN3R_FX_BATCH_LDS
                FX_ATOM(atomicMin, q4 + FX_IDX * FX_PLANE, ((LEAN || w4[c] > 0.0f) && zk4[c] == key) ? i : ~0u);
#else
                FX_ATOM(atomicMin, q4 + FX_IDX * FX_PLANE, ((LEAN || w4[c] > 0.0f) && q4[FX_ZKEY * FX_PLANE] == key) ? i : ~0u);
#endif
            asm volatile("" ::: "memory");          // one corner at a time: keeps the live range of the 16 addends short
        }
    }
}

// Everything after the projection for the TWO rows a thread owns in one iteration, stage by stage for both rows so
// that their independent dependency chains interleave.  All 32 lanes stay converged up to the votes.
// i[j] = s*H*W + y*W + x (the winner index).
template <bool WINNER, int SWEEP>
__device__ __forceinline__ void fx_rows2(const TileArgs &a, const double2 *__restrict__ logtab, const FxPairRegs &pr,
                                         const RowProj (&rp)[2], const float4 (&px)[2], const bool (&cand)[2], int p,
                                         int s, int x, int y, const unsigned (&i)[2], int tx0, int ty0, int tw, int th,
                                         float cs_inv, int lane) {
    if (!__any_sync(0xffffffffu, cand[0] || cand[1])) return;               // nothing of these two rows near the tile
    if (SYN3R_FX_EXPERIMENT == 5) {
        if (__all_sync(0xffffffffu, rp[0].nx1 == 12345 && rp[1].ny1 == 54321 && rp[0].fx == 7.0f && rp[1].fy == 3.0f))
            fx_smem[lane] = 1u;
        return;
    }
    PixelGeom g[2];
    bool any[2], full[2];
#pragma unroll
    for (int j = 0; j < 2; ++j) {
        const float zf = fabsf(__double2float_rn(rp[j].q2)), df = fabsf(px[j].w);
        const float lhs = __fmul_rn(fminf(fabsf(rp[j].fx), fabsf(rp[j].fy)), zf);
        const float rhs = __fmaf_rn(zf, pr.k2, __fmaf_rn(df, pr.k1, pr.k0));
        bool slow = !(lhs > rhs);
        if (SYN3R_FX_EXPERIMENT == 2 && SWEEP == 2) slow = false;
        else if (WINNER)     // z-key = high word of z': ambiguous when the low word is within 2^14 ulps of a carry, or
                        // when the fast z' is not known to 2^-39 relative
            slow = slow || ((unsigned)__double2loint(rp[j].q2) + 16384u < 32768u) ||
                   !(__fmul_rn(zf, 1.8e-12f) > __fmaf_rn(df, pr.k3, pr.k4));
        const bool negx = rp[j].fx < 0.0f, negy = rp[j].fy < 0.0f;
        g[j].lxf = rp[j].nx1 - (negx ? 2 : 1); g[j].lyf = rp[j].ny1 - (negy ? 2 : 1);
        g[j].ax = negx ? -rp[j].fx : __fsub_rn(1.0f, rp[j].fx); g[j].bx = negx ? __fadd_rn(1.0f, rp[j].fx) : rp[j].fx;
        g[j].ay = negy ? -rp[j].fy : __fsub_rn(1.0f, rp[j].fy); g[j].by = negy ? __fadd_rn(1.0f, rp[j].fy) : rp[j].fy;
        g[j].z = rp[j].q2;
        const bool exact = cand[j] && slow;
        if (exact)                                                          // rare: the bit-faithful chain decides
            g[j] = exact_pixel(a.pair_par + (size_t)p * 24, a.Kinv_src + (size_t)s * 9, x, y + j, (double)px[j].w, tx0,
                               ty0, tw, th);
        // some corner column AND row inside the tile: lxf in [-1, tw-1], lyf in [-1, th-1]; all four corners
        // inside (and, not coming from the exact chain, distinct cells with w > 0): lxf in [0, tw-2], lyf in [0, th-2]
        any[j] = cand[j] && (unsigned)(g[j].lxf + 1) <= (unsigned)tw && (unsigned)(g[j].lyf + 1) <= (unsigned)th;
        full[j] = cand[j] && !exact && (unsigned)g[j].lxf < (unsigned)(tw - 1) && (unsigned)g[j].lyf < (unsigned)(th - 1);
    }
    if (!__any_sync(0xffffffffu, any[0] || any[1])) return;
    float wsc[2];
    if ((SWEEP == 1 && pr.upper_ok && SYN3R_FX_EXPERIMENT != 4) || (SYN3R_FX_EXPERIMENT == 2 && SWEEP == 2)) {
#pragma unroll
        for (int j = 0; j < 2; ++j) wsc[j] = fx_weight_upper(pr.cexp_f, __double2float_rn(g[j].z));
    } else {
#pragma unroll
        for (int j = 0; j < 2; ++j)                                         // 2^90 / dw  (:82-84); z of an idle row: 1
            wsc[j] = fx_weight_scale(logtab, pr.cexp, any[j] ? g[j].z : 1.0, pr.clamp_hi);
    }
    if (__all_sync(0xffffffffu, (full[0] || !any[0]) && (full[1] || !any[1]))) {
#pragma unroll
        for (int j = 0; j < 2; ++j)
            if (any[j]) fx_corners<WINNER, SWEEP, true>(g[j], wsc[j], px[j], i[j], cs_inv, lane, tw, th);
    } else {
#pragma unroll
        for (int j = 0; j < 2; ++j)
            if (any[j]) fx_corners<WINNER, SWEEP, false>(g[j], wsc[j], px[j], i[j], cs_inv, lane, tw, th);
    }
}


#if SYN3R_FX_ROWS != 2
// fx_rows2 / fx_block for R = SYN3R_FX_ROWS rows per thread and iteration (same stages, same arithmetic, arrays of R).
template <bool WINNER, int SWEEP, int R>
__device__ __forceinline__ void fx_rowsN(const TileArgs &a, const double2 *__restrict__ logtab, const FxPairRegs &pr,
                                         const RowProj (&rp)[R], const float4 (&px)[R], const bool (&cand)[R], int p,
                                         int s, int x, int y, const unsigned (&i)[R], int tx0, int ty0, int tw, int th,
                                         float cs_inv, int lane) {
    bool c_any = false;
#pragma unroll
    for (int j = 0; j < R; ++j) c_any = c_any || cand[j];
    if (!__any_sync(0xffffffffu, c_any)) return;
    PixelGeom g[R];
    bool any[R], full[R];
#pragma unroll
    for (int j = 0; j < R; ++j) {
        const float zf = fabsf(__double2float_rn(rp[j].q2)), df = fabsf(px[j].w);
        const float lhs = __fmul_rn(fminf(fabsf(rp[j].fx), fabsf(rp[j].fy)), zf);
        const float rhs = __fmaf_rn(zf, pr.k2, __fmaf_rn(df, pr.k1, pr.k0));
        bool slow = !(lhs > rhs);
        if (WINNER)
            slow = slow || ((unsigned)__double2loint(rp[j].q2) + 16384u < 32768u) ||
                   !(__fmul_rn(zf, 1.8e-12f) > __fmaf_rn(df, pr.k3, pr.k4));
        const bool negx = rp[j].fx < 0.0f, negy = rp[j].fy < 0.0f;
        g[j].lxf = rp[j].nx1 - (negx ? 2 : 1); g[j].lyf = rp[j].ny1 - (negy ? 2 : 1);
        g[j].ax = negx ? -rp[j].fx : __fsub_rn(1.0f, rp[j].fx); g[j].bx = negx ? __fadd_rn(1.0f, rp[j].fx) : rp[j].fx;
        g[j].ay = negy ? -rp[j].fy : __fsub_rn(1.0f, rp[j].fy); g[j].by = negy ? __fadd_rn(1.0f, rp[j].fy) : rp[j].fy;
        g[j].z = rp[j].q2;
        const bool exact = cand[j] && slow;
        if (exact)
            g[j] = exact_pixel(a.pair_par + (size_t)p * 24, a.Kinv_src + (size_t)s * 9, x, y + j, (double)px[j].w, tx0,
                               ty0, tw, th);
        any[j] = cand[j] && (unsigned)(g[j].lxf + 1) <= (unsigned)tw && (unsigned)(g[j].lyf + 1) <= (unsigned)th;
        full[j] = cand[j] && !exact && (unsigned)g[j].lxf < (unsigned)(tw - 1) && (unsigned)g[j].lyf < (unsigned)(th - 1);
    }
    bool a_any = false, lean = true;
#pragma unroll
    for (int j = 0; j < R; ++j) { a_any = a_any || any[j]; lean = lean && (full[j] || !any[j]); }
    if (!__any_sync(0xffffffffu, a_any)) return;
    float wsc[R];
    if (SWEEP == 1 && pr.upper_ok) {
#pragma unroll
        for (int j = 0; j < R; ++j) wsc[j] = fx_weight_upper(pr.cexp_f, __double2float_rn(g[j].z));
    } else {
#pragma unroll
        for (int j = 0; j < R; ++j) wsc[j] = fx_weight_scale(logtab, pr.cexp, any[j] ? g[j].z : 1.0, pr.clamp_hi);
    }
    if (__all_sync(0xffffffffu, lean)) {
#pragma unroll
        for (int j = 0; j < R; ++j)
            if (any[j]) fx_corners<WINNER, SWEEP, true>(g[j], wsc[j], px[j], i[j], cs_inv, lane, tw, th);
    } else {
#pragma unroll
        for (int j = 0; j < R; ++j)
            if (any[j]) fx_corners<WINNER, SWEEP, false>(g[j], wsc[j], px[j], i[j], cs_inv, lane, tw, th);
    }
}

template <bool WINNER, int SWEEP>
__device__ __forceinline__ void fx_block(const TileArgs &a, const double2 *__restrict__ logtab, const FxPairRegs &pr,
                                         int p, int s, int bx0, int by0, int nrows, int tx0, int ty0, int tw, int th, int lane,
                                         float cs_inv) {
    constexpr int R = SYN3R_FX_ROWS;
    const int W = a.g.W, H = a.g.H;
    const int x = bx0 + lane;
    const bool live = x < W;
    const unsigned i0 = (unsigned)s * (unsigned)(H * W) + (unsigned)by0 * W + (live ? x : 0);
    const double xd = (double)x;
    const double u0 = __fma_rn(pr.A[0], xd, pr.A[2]), u1 = __fma_rn(pr.A[3], xd, pr.A[5]),
                 u2 = __fma_rn(pr.A[6], xd, pr.A[8]);
    const double tx1d = (double)(tx0 - 1), ty1d = (double)(ty0 - 1);
    const int rows = min(nrows, H - by0);                      // this unit's rows inside the image
    if (rows <= 0) return;
    const float4 *src = a.rgbd + i0;
    const uint64_t keep = l2_policy_evict_last();
    float4 nxt[R];                                              // a row index past the unit re-reads its last row
#pragma unroll
    for (int j = 0; j < R; ++j) nxt[j] = ld_keep_f4(src + (unsigned)(min(j, rows - 1) * W), keep);
    unsigned ia = i0;
#pragma unroll 1
    for (int r = 0; r < rows; r += R, ia += (unsigned)(R * W)) {
        float4 px[R];
#pragma unroll
        for (int j = 0; j < R; ++j) px[j] = nxt[j];
        if (r + R < rows) {
#pragma unroll
            for (int j = 0; j < R; ++j) nxt[j] = ld_keep_f4(src + (unsigned)(min(r + R + j, rows - 1) * W), keep);
        }
        const int y = by0 + r;
        const double yd = (double)y;
        RowProj rp[R];
        bool cand[R];
        unsigned i[R];
#pragma unroll
        for (int j = 0; j < R; ++j) {
            rp[j] = fx_project(pr, u0, u1, u2, __dadd_rn(yd, (double)j), (double)px[j].w, tx1d, ty1d, tw, th);
            cand[j] = rp[j].cand && live && (r + j < rows);
            i[j] = ia + (unsigned)(j * W);
        }
        fx_rowsN<WINNER, SWEEP, R>(a, logtab, pr, rp, px, cand, p, s, x, y, i, tx0, ty0, tw, th, cs_inv, lane);
        __syncwarp();
    }
}
#else
// Half a source block (32 x 4 px from row by0) of one pair, two rows per iteration.
template <bool WINNER, int SWEEP>
__device__ __forceinline__ void fx_block(const TileArgs &a, const double2 *__restrict__ logtab, const FxPairRegs &pr,
                                         int p, int s, int bx0, int by0, int nrows, int tx0, int ty0, int tw, int th, int lane,
                                         float cs_inv) {
    const int W = a.g.W, H = a.g.H;
    const int x = bx0 + lane;
    const bool live = x < W;
    const unsigned i0 = (unsigned)s * (unsigned)(H * W) + (unsigned)by0 * W + (live ? x : 0);
    const double xd = (double)x;
    const double u0 = __fma_rn(pr.A[0], xd, pr.A[2]), u1 = __fma_rn(pr.A[3], xd, pr.A[5]),
                 u2 = __fma_rn(pr.A[6], xd, pr.A[8]);
    const double tx1d = (double)(tx0 - 1), ty1d = (double)(ty0 - 1);
    const int rows = min(nrows, H - by0);                      // this unit's rows inside the image
    if (rows <= 0) return;
    const float4 *src = a.rgbd + i0;
    const unsigned W2 = 2u * (unsigned)W;
    const uint64_t keep = l2_policy_evict_last();
    // rows r, r+1 of this iteration and (prefetched) of the next; a row index past the block re-reads the last row
    const unsigned rb = (rows > 1) ? (unsigned)W : 0u;         // offset of the odd row of a pair
    float4 pxa = ld_keep_f4(src, keep), pxb = ld_keep_f4(src + rb, keep);
    double yd = (double)by0;
    unsigned ia = i0;
    int y = by0;
#pragma unroll 1
    for (int r = 0; r < rows; r += 2, ia += W2, y += 2, yd = __dadd_rn(yd, 2.0)) {
        const float4 px[2] = {pxa, pxb};
        if (r + 3 < rows) {                                     // full blocks: always, except the last iteration
            src += W2;
            pxa = ld_keep_f4(src, keep);
            pxb = ld_keep_f4(src + W, keep);
        } else if (r + 2 < rows) {
            src += W2;
            pxa = ld_keep_f4(src, keep);
            pxb = pxa;
        }
#if SYN3R_FX_EXPERIMENT == 2
        const RowProj rp[2] = {SWEEP == 2 ? fx_project_f32(pr, (float)u0, (float)u1, (float)u2, (float)y, px[0].w, (float)tx1d, (float)ty1d, tw, th)
                                          : fx_project(pr, u0, u1, u2, yd, (double)px[0].w, tx1d, ty1d, tw, th),
                               SWEEP == 2 ? fx_project_f32(pr, (float)u0, (float)u1, (float)u2, (float)(y + 1), px[1].w, (float)tx1d, (float)ty1d, tw, th)
                                          : fx_project(pr, u0, u1, u2, __dadd_rn(yd, 1.0), (double)px[1].w, tx1d, ty1d, tw, th)};
#else
        const RowProj rp[2] = {fx_project(pr, u0, u1, u2, yd, (double)px[0].w, tx1d, ty1d, tw, th),
                               fx_project(pr, u0, u1, u2, __dadd_rn(yd, 1.0), (double)px[1].w, tx1d, ty1d, tw, th)};
#endif
        const bool cand[2] = {rp[0].cand && live, rp[1].cand && live && (r + 1 < rows)};
        const unsigned i[2] = {ia, ia + (unsigned)W};
        fx_rows2<WINNER, SWEEP>(a, logtab, pr, rp, px, cand, p, s, x, y, i, tx0, ty0, tw, th, cs_inv, lane);
        __syncwarp();
    }
}

#endif  // SYN3R_FX_ROWS

// One sweep over the culled block list (sm.listB, nB blocks) as a software-pipelined STREAM of row pairs: while the
// warp works on row pair n, the two float4 rows of pair n+1 are in flight from L2 and the claim of pair n+2 (a
// shared-memory counter) is in flight too, so neither the L2 latency nor the claim is exposed at a work-unit
// boundary, and the unit of the dynamic scheduler is one iteration (2 x 32 px): short tails at the barrier that ends
// the sweep.  (The former fx_block took 4-row units and exposed one L2 round trip + one claim per unit.)
// Measured (round 2): no gain on config B (1.285 vs 1.293 ms alone; 1.250 vs 1.248 ms next to the other changes) - the
// L2 round trip and the claim at a work-unit boundary are not what the warps wait for - and 7 % SLOWER on config E
// (1459 vs 1351 ms: the two-level cull hands over many short block lists, each paying the pipeline's fill and two
// surplus claims per warp).  Off by default; kept for A/B.
#ifndef SYN3R_FX_STREAM
#define SYN3R_FX_STREAM 0
#endif
template <bool WINNER, int SWEEP>
__device__ __forceinline__ void fx_stream(const TileArgs &a, const FxAux &sm, const FxPairRegs &pr, int p, int s, int nB,
                                          int tx0, int ty0, int tw, int th, int lane, float cs_inv) {
    const int W = a.g.W, H = a.g.H;
    const int total = nB * (SYN3R_BLK_H / 2);
    int *ctr = &sm.ctl[SWEEP == 1 ? 4 : 5];
    const uint64_t keep = l2_policy_evict_last();
    const float4 *frame = a.rgbd + (size_t)s * (size_t)(H * W);
    const unsigned ibase = (unsigned)s * (unsigned)(H * W);
    const double tx1d = (double)(tx0 - 1), ty1d = (double)(ty0 - 1);
    // lane 0 claims for the warp: a plain atom.shared (the compiler would wrap atomicAdd() in its own warp-aggregation
    // sequence of a dozen instructions, pointless with one active lane)
    const unsigned ctr_s = (unsigned)__cvta_generic_to_shared(ctr);
    auto claim = [&]() {
        int v = 0;
        if (lane == 0) asm volatile("atom.shared.add.u32 %0, [%1], 1;" : "=r"(v) : "r"(ctr_s) : "memory");
        return v;
    };
    int kraw = claim();
    int k = __shfl_sync(0xffffffffu, kraw, 0);
    kraw = claim();                                             // the claim after that: consumed one iteration later
    if (k >= total) return;
    int x, y;
    float4 pxa, pxb;
    auto fetch = [&](int kk) {
        const int org = sm.listB[kk >> 2];                      // (by0 << 16) | bx0; SYN3R_BLK_H / 2 = 4 row pairs per block
        x = (org & 0xffff) + lane;
        y = (org >> 16) + ((kk & 3) << 1);
        const float4 *r0 = frame + (size_t)(min(y, H - 1) * W + min(x, W - 1));
        pxa = ld_keep_f4(r0, keep);
        pxb = ld_keep_f4(r0 + ((y + 1 < H) ? W : 0), keep);
    };
    fetch(k);
    for (;;) {
        const int kn = __shfl_sync(0xffffffffu, kraw, 0);
        kraw = claim();
        const float4 px[2] = {pxa, pxb};
        const int cx = x, cy = y;
        if (kn < total) fetch(kn);
        {
            const bool live = cx < W;
            // (double) of a non-negative int without a conversion instruction: 2^52 + v has v in its low word
            const double xd = __dsub_rn(__hiloint2double(0x43300000, cx), 4503599627370496.0);
            const double yd = __dsub_rn(__hiloint2double(0x43300000, cy), 4503599627370496.0);
            const double u0 = __fma_rn(pr.A[0], xd, pr.A[2]), u1 = __fma_rn(pr.A[3], xd, pr.A[5]),
                         u2 = __fma_rn(pr.A[6], xd, pr.A[8]);
            const RowProj rp[2] = {fx_project(pr, u0, u1, u2, yd, (double)px[0].w, tx1d, ty1d, tw, th),
                                   fx_project(pr, u0, u1, u2, __dadd_rn(yd, 1.0), (double)px[1].w, tx1d, ty1d, tw, th)};
            const bool cand[2] = {rp[0].cand && live && cy < H, rp[1].cand && live && cy + 1 < H};
            const unsigned ia = ibase + (unsigned)(cy * W + (live ? cx : 0));
            const unsigned i[2] = {ia, ia + (unsigned)W};
            fx_rows2<WINNER, SWEEP>(a, sm.logtab, pr, rp, px, cand, p, s, cx, cy, i, tx0, ty0, tw, th, cs_inv, lane);
            __syncwarp();
        }
        if (kn >= total) break;
    }
}

__host__ __device__ inline size_t fx_smem_bytes(bool winner) {
    const size_t cells = (size_t)SYN3R_TILE_W * SYN3R_TILE_H;
    return (cells + SYN3R_TILE_W) * 4 * (winner ? 8 : 6) + (SYN3R_FX_TWO_CTAS ? 0 : (size_t)SYN3R_LOGTAB * 16) +
           (SYN3R_LISTA + SYN3R_FX_LISTB_CAP) * 4 + 64 + 192;
}

// culls the pair's source blocks against the tile and runs one sweep over the survivors.
// ctl: [1] nA, [2] nB, [4] / [5] dynamic block counters of sweep 1 / 2, [14] / [15] the same for the end game (single
// iterations), [6] listB is complete and reusable, [9..13] phase clocks (SYN3R_FX_PHASE_CLOCKS builds)
template <bool WINNER, int SWEEP, bool FLAT>
__device__ __forceinline__ void fx_pair(const TileArgs &a, const FxAux &sm, const FxPairRegs &pr, int p, int s, int tile,
                                        int tx0, int ty0, int tx1, int ty1, bool reuse_list, bool fresh, float cs_inv,
                                        int staged_n = -2) {
    // staged_n (AHEAD kernels): >= 0 = listB already holds that many blocks (copied while the previous item resolved):
    // no cull, no barrier, in either sweep; -1 = the staged list overflowed its cap: in-kernel cull; -2 = nothing staged
    const TileGeom &g = a.g;
    const int tid = threadIdx.x, lane = tid & 31;
    const int tw = tx1 - tx0 + 1, th = ty1 - ty0 + 1;
    int *ctl = sm.ctl;
    auto run_list = [&](int nB) {
#if SYN3R_FX_STREAM && SYN3R_FX_EXPERIMENT == 0
        fx_stream<WINNER, SWEEP>(a, sm, pr, p, s, nB, tx0, ty0, tw, th, lane, cs_inv);
#else
        const int units = SYN3R_FX_UNITS * nB;                  // work unit = part of a block: shorter tails
        constexpr int SUB = SYN3R_FX_UNIT_ROWS / SYN3R_FX_ROWS; // iterations of a unit
        const int tail_units = min(units, SYN3R_FX_TAIL_UNITS), main_units = units - tail_units;
        bool tail = (SYN3R_FX_TAIL_UNITS != 0) && main_units == 0;
        for (;;) {
            int k = 0;
            if (lane == 0) k = atomicAdd(&ctl[(SWEEP == 1 ? 4 : 5) + (tail ? 10 : 0)], 1);
            k = __shfl_sync(0xffffffffu, k, 0);
            int u = k, r0 = 0, nr = SYN3R_FX_UNIT_ROWS;
            if (!tail) {
                if (k >= main_units) {
                    if (SYN3R_FX_TAIL_UNITS == 0) break;
                    tail = true;                                // this warp's first end-game claim comes next
                    continue;
                }
            } else {
                if (k >= tail_units * SUB) break;
                u = main_units + k / SUB; r0 = (k % SUB) * SYN3R_FX_ROWS; nr = SYN3R_FX_ROWS;
            }
            const int org = sm.listB[u / SYN3R_FX_UNITS];       // (by0 << 16) | bx0
            fx_block<WINNER, SWEEP>(a, sm.logtab, pr, p, s, org & 0xffff, (org >> 16) + (u % SYN3R_FX_UNITS) * SYN3R_FX_UNIT_ROWS + r0,
                                    nr, tx0, ty0, tw, th, lane, cs_inv);
        }
#if SYN3R_FX_PHASE_CLOCKS
        if (lane == 0) atomicMin(reinterpret_cast<unsigned *>(&ctl[SWEEP == 1 ? 12 : 13]), (unsigned)clock() - (unsigned)ctl[9]);
#endif
#endif
    };
    if (staged_n >= 0) {
        if (SWEEP == 1) FX_CLK(10);
        run_list(staged_n);
        return;
    }
    if (SWEEP == 2 && reuse_list) {       // the list sweep 1 built is complete: no culling, no barrier
        run_list(ctl[2]);
        return;
    }
    const CellBox *bbb = a.blk_box + (size_t)p * g.nsb * SYN3R_SB_BLOCKS;
    if (FLAT) {
        // small frame (<= SYN3R_FX_FLAT blocks, all of which fit listB): every block box is tested in one pass, one
        // L2 round trip and one barrier instead of the two-level cull's two and six
        if (!fresh) {                      // counters left by an earlier pair of this item
            __syncthreads();
            if (tid == 0) { ctl[2] = 0; ctl[4] = 0; ctl[5] = 0; ctl[14] = 0; ctl[15] = 0; }
            __syncthreads();
        }
        bool listed = false;
        if (a.tile_cnt && staged_n == -2) {
            // the list k_fwd_prep built for this (pair, tile): one L2 round trip, no box tests
            const size_t t = (size_t)p * g.ntiles + tile;
            const unsigned cnt = a.tile_cnt[t];
            int e = 0;
            if (tid < SYN3R_TILE_LIST_CAP) e = a.tile_list[t * SYN3R_TILE_LIST_CAP + tid];
            listed = cnt <= SYN3R_TILE_LIST_CAP;
            if (listed) {
                if (tid < (int)cnt) sm.listB[tid] = e;
                if (tid == 0) ctl[2] = (int)cnt;
            }
        }
        if (!listed) {
            const int nblk = g.nsb * SYN3R_SB_BLOCKS;
            CellBox bx[SYN3R_FX_FLAT_PER_THREAD];
#pragma unroll
            for (int k = 0; k < SYN3R_FX_FLAT_PER_THREAD; ++k) {
                const int j = tid + k * SYN3R_TILE_THREADS;
                bx[k] = (j < nblk) ? bbb[j] : CellBox{1, 1, 0, 0};
            }
#pragma unroll
            for (int k = 0; k < SYN3R_FX_FLAT_PER_THREAD; ++k) {
                const int j = tid + k * SYN3R_TILE_THREADS;
                if (box_hits(bx[k], tx0, ty0, tx1, ty1)) {
                    int bx0, by0;
                    block_origin(g, j / SYN3R_SB_BLOCKS, j % SYN3R_SB_BLOCKS, bx0, by0);
                    sm.listB[atomicAdd(&ctl[2], 1)] = (by0 << 16) | bx0;
                }
            }
        }
        if (tid == 0 && SWEEP == 1) ctl[6] = 1;                 // read behind the barrier that ends the sweep
        __syncthreads();
        if (SWEEP == 1) FX_CLK(10);
        run_list(ctl[2]);
        return;
    }
    const CellBox *sbb = a.sb_box + (size_t)p * g.nsb;
    for (int base = 0; base < g.nsb; base += SYN3R_LISTA) {
        __syncthreads();
        if (tid == 0) ctl[1] = 0;
        __syncthreads();
        for (int j = base + tid; j < min(base + SYN3R_LISTA, g.nsb); j += SYN3R_TILE_THREADS)
            if (box_hits(sbb[j], tx0, ty0, tx1, ty1)) sm.listA[atomicAdd(&ctl[1], 1)] = j;
        __syncthreads();
        const int nA = ctl[1];
        for (int a0 = 0; a0 < nA; a0 += SYN3R_LISTB / SYN3R_SB_BLOCKS) {
            __syncthreads();
            if (tid == 0) { ctl[2] = 0; ctl[4] = 0; ctl[5] = 0; ctl[14] = 0; ctl[15] = 0; }
            __syncthreads();
            const int cnt = min(SYN3R_LISTB / SYN3R_SB_BLOCKS, nA - a0) * SYN3R_SB_BLOCKS;
            for (int j = tid; j < cnt; j += SYN3R_TILE_THREADS) {
                const int sb = sm.listA[a0 + j / SYN3R_SB_BLOCKS], bi = j % SYN3R_SB_BLOCKS;
                if (box_hits(bbb[sb * SYN3R_SB_BLOCKS + bi], tx0, ty0, tx1, ty1)) {
                    int bx0, by0;
                    block_origin(g, sb, bi, bx0, by0);
                    sm.listB[atomicAdd(&ctl[2], 1)] = (by0 << 16) | bx0;
                }
            }
            __syncthreads();
            run_list(ctl[2]);
        }
        if (tid == 0 && SWEEP == 1) ctl[6] = (g.nsb <= SYN3R_LISTA && nA <= SYN3R_LISTB / SYN3R_SB_BLOCKS) ? 1 : 0;
    }
}

struct CellOut { float v0, v1, v2; bool m; unsigned range; int win; };

// :109-128 for one cell from its integer accumulators.  `range` = the largest bit pattern of the three colours: a
// value is outside [0, 256] (:121-125; negative, NaN and inf included) exactly when its bit pattern, read as an
// unsigned integer, exceeds that of 256.0f (a resolved colour is never -0.0: it is an int -> float conversion times a
// positive factor), so the range test of a whole row is one maximum per cell and one compare per thread.
// GATE = false (no SYN3R_FWD_GATE_0P6): a cell without contributions has S == I == 0 and an untouched winner
// (~0u == -1), so max(S, 1) gives the zeros and the -1 of :116 without a select.
template <bool GATE>
__device__ __forceinline__ CellOut fx_resolve_cell(unsigned wm, unsigned S, int I0, int I1, int I2, unsigned widx,
                                                   float cscale, float gate) {
    CellOut o;
    o.m = wm != 0u;                                            // some corner with w > 0 landed here (:112)
    if (GATE) {
        const float Sf = __uint2float_rn(S);                   // sum of weights <= 0.6: undo the per-cell scale
        o.m = o.m && (__fmul_rn(Sf, __uint_as_float((wm & 0x7f800000u) - ((unsigned)SYN3R_FX_BITS << 23))) <= gate);
        const float inv = __fdividef(cscale, Sf);
        o.v0 = o.m ? __fmul_rn(__int2float_rn(I0), inv) : 0.0f;   // :116
        o.v1 = o.m ? __fmul_rn(__int2float_rn(I1), inv) : 0.0f;
        o.v2 = o.m ? __fmul_rn(__int2float_rn(I2), inv) : 0.0f;
        o.win = o.m ? (int)widx : -1;
    } else {
        const float inv = __fdividef(cscale, __uint2float_rn(max(S, 1u)));
        o.v0 = __fmul_rn(__int2float_rn(I0), inv);
        o.v1 = __fmul_rn(__int2float_rn(I1), inv);
        o.v2 = __fmul_rn(__int2float_rn(I2), inv);
        o.win = (int)widx;
    }
    o.range = max(max(__float_as_uint(o.v0), __float_as_uint(o.v1)), __float_as_uint(o.v2));
    return o;
}
__device__ __forceinline__ unsigned u8_round(float v) {      // clip + round-half-even (:127-128)
    return (unsigned)__float2int_rn(fminf(fmaxf(v, 0.0f), 255.0f));
}

#if SYN3R_FX_TWO_CTAS
__device__ double2 g_fx_logtab[SYN3R_LOGTAB];
#endif
// AHEAD (launcher: PAIR mode, FLAT, block lists present, SYN3R_FX_PREFETCH): every work item starts with its constants
// and its block list already in shared memory.
template <bool WINNER, bool FLAT, bool AHEAD>
__global__ void __launch_bounds__(SYN3R_TILE_THREADS, SYN3R_FX_TWO_CTAS ? 2 : 1)
k_fwd_tile_fx(const TileArgs a) {
    const int CELLS = FX_PLANE, NPL = WINNER ? 8 : 6;          // plane stride incl. the trash row
    FxAux sm;
    unsigned char *ptr = reinterpret_cast<unsigned char *>(fx_smem + NPL * CELLS);
#if SYN3R_FX_TWO_CTAS
    sm.logtab = g_fx_logtab;                                   // A/B build: the table lives in global memory (L1)
#else
    sm.logtab = reinterpret_cast<double2 *>(ptr); ptr += (size_t)SYN3R_LOGTAB * 16;
#endif
    sm.listA = reinterpret_cast<int *>(ptr); ptr += SYN3R_LISTA * 4;
    sm.listB = reinterpret_cast<int *>(ptr); ptr += SYN3R_FX_LISTB_CAP * 4;
    sm.ctl = reinterpret_cast<int *>(ptr); ptr += 64;
    sm.stage = ptr;
    int *ctl = sm.ctl;

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, nwarps = SYN3R_TILE_THREADS / 32;
    const TileGeom g = a.g;
    const int W = g.W, H = g.H;
    const unsigned hw = (unsigned)(H * W);

    for (int k = tid; k < SYN3R_LOGTAB; k += SYN3R_TILE_THREADS) {
        const double c = __hiloint2double(0x3ff00000 + (k << 14) + (1 << 13), 0);    // bin centre
        sm.logtab[k] = make_double2(__ddiv_rn(1.0, c), log2(c));
    }
    for (int k = tid; k < CELLS; k += SYN3R_TILE_THREADS) {
#pragma unroll
        for (int pl = 0; pl < FX_ZKEY; ++pl) fx_smem[pl * CELLS + k] = 0u;
        if (WINNER) { fx_smem[FX_ZKEY * CELLS + k] = ~0u; fx_smem[FX_IDX * CELLS + k] = ~0u; }
        if (k >= FX_CELLS) fx_smem[FX_WMAX * CELLS + k] = 0x3f800000u;      // trash cells: a finite scale, weights 0
    }
    // colour scale: the power of two >= max |colour| over all sources (k_src_stats); exact to divide by
    const unsigned cb = *a.cmax_bits;
    const bool colour_bad = cb >= 0x7f000000u;                 // inf / NaN / > 2^127: fp32 path for everything
    const unsigned ce = min(max((cb >> 23) + ((cb & 0x7fffffu) ? 1u : 0u), 1u), 253u);
    const float cscale = cb ? __uint_as_float(ce << 23) : 1.0f;
    const float cs_inv = cb ? __uint_as_float((254u - ce) << 23) : 1.0f;
    const int n_items = a.n_slots * g.ntiles;
    const float gate = 0.6f * (float)SYN3R_W_SCALE;
    const bool gated = (a.flags & SYN3R_FWD_GATE_0P6) != 0;
    const uint64_t once = l2_policy_evict_first();

#if SYN3R_FX_PHASE_CLOCKS
    __shared__ unsigned long long ph[8];
    if (tid < 8) ph[tid] = 0;
    unsigned ph_tb = 0;
#endif
    // AHEAD: claim a work item and start the copies of its per-pair constants, source index and block list into shared
    // memory (cp.async: no registers held; listB is free whenever this is called)
    auto stage_item = [&](int it) {
        if (it >= n_items) return;
        const int np = it / g.ntiles;                          // PAIR mode: pair = slot
        const size_t nt = (size_t)np * g.ntiles + (size_t)(it - np * g.ntiles);
        if (tid < 8) cp_async_16(sm.stage + tid * 16, reinterpret_cast<const unsigned char *>(a.pair_fast + np) + tid * 16);
        else if (tid == 8) cp_async_4(sm.stage + 128, a.tile_cnt + nt);
        else if (tid == 9) cp_async_4(sm.stage + 132, a.pair_src + np);
        else if (tid >= 32 && tid < 32 + SYN3R_TILE_LIST_CAP / 4)
            cp_async_16(sm.listB + (tid - 32) * 4, a.tile_list + nt * SYN3R_TILE_LIST_CAP + (size_t)(tid - 32) * 4);
        cp_async_commit();
    };
    if (AHEAD) {
        if (tid == 0) ctl[8] = (int)atomicAdd(a.work_counter, 1u);
        __syncthreads();
        stage_item(ctl[8]);
    }
    for (;;) {
        __syncthreads();
        if (tid == 0) {
#if SYN3R_FX_PHASE_CLOCKS
            const unsigned ph_ta = (unsigned)clock();
#endif
            if (!AHEAD) ctl[0] = (int)atomicAdd(a.work_counter, 1u);
            ctl[2] = 0; ctl[3] = 0; ctl[4] = 0; ctl[5] = 0; ctl[6] = 0; ctl[7] = 0;   // empty block list until culled
            ctl[14] = 0; ctl[15] = 0;
#if SYN3R_FX_PHASE_CLOCKS
            if (AHEAD || ctl[0] >= 0) ph_tb = (unsigned)clock();      // (depends on the atomic's result)
            ph[0] += ph_tb - ph_ta;
            ctl[9] = (int)ph_tb; ctl[10] = (int)ph_tb; ctl[12] = -1; ctl[13] = -1;
#endif
        }
        if (AHEAD) cp_async_wait_all();                        // this thread's share of the staged copies
        __syncthreads();
        const int item = ctl[AHEAD ? 8 : 0];                   // (AHEAD: thread 0 replaces ctl[8] behind the next barrier)
        if (item >= n_items) break;
        if (colour_bad) {
            if (tid == 0) a.redo_list[atomicAdd(a.redo_count, 1u)] = item;
            if (AHEAD) {                                       // (nothing of the next item needs staging on this path)
                __syncthreads();
                if (tid == 0) ctl[8] = (int)atomicAdd(a.work_counter, 1u);
            }
            continue;
        }
        const int slot = item / g.ntiles, tile = item - slot * g.ntiles;
        const int tx0 = (tile % g.ntx) * SYN3R_TILE_W, ty0 = (tile / g.ntx) * SYN3R_TILE_H;
        const int tx1 = min(tx0 + SYN3R_TILE_W, W) - 1, ty1 = min(ty0 + SYN3R_TILE_H, H) - 1;
        bool poisoned = false;
        auto load_pair = [&](int p, FxPairRegs &pr) -> bool {
            const PairFast *pf = AHEAD ? reinterpret_cast<const PairFast *>(sm.stage) : a.pair_fast + p;
            pr.cexp = pf->cexp;
            if (!(pr.cexp < CUDART_INF)) return false;         // max log-depth NaN or 0: every weight is NaN (:84)
#pragma unroll
            for (int k = 0; k < 9; ++k) pr.A[k] = pf->A[k];
#pragma unroll
            for (int k = 0; k < 3; ++k) pr.b[k] = pf->b[k];
            pr.k0 = pf->k0; pr.k1 = pf->k1; pr.k2 = pf->k2; pr.k3 = pf->k3; pr.k4 = pf->k4;
            pr.clamp_hi = pf->pad != 0.0f;
            pr.cexp_f = __double2float_rn(pr.cexp);
            pr.upper_ok = pr.cexp <= 500.0;
            return true;
        };
        // After sweep 1 the counts are final: a cell with FX_MAX_N or more updates could pass S = 2^31 in sweep 2, so
        // the item goes to the redo list (the flag is read behind the barrier that ends sweep 2).
        auto scan_crowded = [&]() {
            bool crowded = false;
            for (int k = tid * 4; k < FX_CELLS; k += SYN3R_TILE_THREADS * 4) {
                const uint4 n = *reinterpret_cast<const uint4 *>(fx_smem + FX_N * CELLS + k);
                crowded = crowded || max(max(n.x, n.y), max(n.z, n.w)) >= FX_MAX_N;
            }
            if (crowded) ctl[3] = 1;
        };
        if (AHEAD || !a.pair_slot) {                           // PAIR mode: the slot's one pair, list built once
            FxPairRegs pr;
            const int p = slot, s = AHEAD ? *reinterpret_cast<const int *>(sm.stage + 132) : a.pair_src[p];
            int staged_n = -2;
            if (AHEAD) {
                const unsigned cnt = *reinterpret_cast<const unsigned *>(sm.stage + 128);
                staged_n = cnt <= SYN3R_TILE_LIST_CAP ? (int)cnt : -1;
            }
            if (load_pair(p, pr)) {                            // (a dead pair leaves the tile empty: mask all false)
                fx_pair<WINNER, 1, FLAT>(a, sm, pr, p, s, tile, tx0, ty0, tx1, ty1, false, true, cs_inv, staged_n);
                __syncthreads();
                FX_CLK(11);
                if (AHEAD && tid == 0) ctl[8] = (int)atomicAdd(a.work_counter, 1u);    // the next item, read after sweep 2
                scan_crowded();
                if (SYN3R_FX_EXPERIMENT != 1 && SYN3R_FX_EXPERIMENT != 5)
                    fx_pair<WINNER, 2, FLAT>(a, sm, pr, p, s, tile, tx0, ty0, tx1, ty1, FLAT || ctl[6] != 0, false, cs_inv, staged_n);
            } else if (AHEAD) {
                __syncthreads();                               // (every thread has read this item's ctl[8])
                if (tid == 0) ctl[8] = (int)atomicAdd(a.work_counter, 1u);
            }
            __syncthreads();
        } else {                                               // FUSED mode: every pair of the slot, per sweep
#pragma unroll 1
            for (int sweep = 1; sweep <= 2; ++sweep) {
                if (sweep == 2) scan_crowded();
#pragma unroll 1
                for (int p = 0; p < a.P; ++p) {
                    if (a.pair_slot[p] != slot) continue;
                    FxPairRegs pr;
                    const int s = a.pair_src[p];
                    if (!load_pair(p, pr)) { poisoned = true; continue; }   // NaN weights poison cells: fp32 path
                    if (sweep == 1) fx_pair<WINNER, 1, FLAT>(a, sm, pr, p, s, tile, tx0, ty0, tx1, ty1, false, false, cs_inv);
                    else fx_pair<WINNER, 2, FLAT>(a, sm, pr, p, s, tile, tx0, ty0, tx1, ty1, false, false, cs_inv);
                }
                __syncthreads();
            }
        }
        if (poisoned) ctl[3] = 1;
        __syncthreads();
#if SYN3R_FX_PHASE_CLOCKS
        const unsigned ph_t3 = (unsigned)clock();
#endif
        const bool redo = ctl[3] != 0;                         // poisoned or crowded: k_fwd_tile recomputes the item
        if (redo && tid == 0) a.redo_list[atomicAdd(a.redo_count, 1u)] = item;
        if (AHEAD) stage_item(ctl[8]);     // the next item's constants and block list travel while this one resolves
                                           // (listB is free from here on)

        // ---- resolve from shared memory (:109-128) and clear the planes for the next item -----------------
        unsigned range = 0u;                                    // largest colour bit pattern this thread resolved
        const int tw = tx1 - tx0 + 1, th = ty1 - ty0 + 1;
        const bool vec = (tw == SYN3R_TILE_W) && ((W & 3) == 0);
        const bool u8_wide = vec && ((W & 15) == 0) && ((hw & 15u) == 0u) && ((reinterpret_cast<uintptr_t>(a.out_u8) & 15) == 0);
#if SYN3R_FX_TRASH_SPLIT
        // the trash row (row TILE_H, cleared only) is shared out over all threads: with it in the row loop one warp of
        // 24 would make a third trip for 49 rows
        for (int k = tid; k < NPL * SYN3R_TILE_W; k += SYN3R_TILE_THREADS) {
            const int pl = k / SYN3R_TILE_W;
            fx_smem[pl * CELLS + FX_CELLS + (k - pl * SYN3R_TILE_W)] = (pl >= FX_ZKEY) ? ~0u : (pl == FX_WMAX ? 0x3f800000u : 0u);
        }
#if SYN3R_FX_RESOLVE_STAGGER_NS
        if ((warp >> 2) & 1) __nanosleep(SYN3R_FX_RESOLVE_STAGGER_NS);
#endif
        for (int ly = warp; ly < SYN3R_TILE_H; ly += nwarps) {
#else
        for (int ly = warp; ly <= SYN3R_TILE_H; ly += nwarps) {          // row TILE_H = the trash row (cleared only)
#endif
            const int c4 = ly * SYN3R_TILE_W + lane * 4;       // this lane's 4 consecutive cells of the row
            uint4 pl[8];
#pragma unroll
            for (int k = 0; k < 8; ++k) {
                if (k >= NPL) continue;
                uint4 *pp = reinterpret_cast<uint4 *>(fx_smem + k * CELLS + c4);
                pl[k] = *pp;
                const unsigned clr = (k >= FX_ZKEY) ? ~0u : ((k == FX_WMAX && ly == SYN3R_TILE_H) ? 0x3f800000u : 0u);
                *pp = make_uint4(clr, clr, clr, clr);
            }
            if (ly >= th || redo) continue;
            const size_t q = (size_t)slot * hw + (size_t)((ty0 + ly) * W + tx0 + lane * 4);   // 64-bit: n_slots * H * W may pass 2^32
            const unsigned wm[4] = {pl[0].x, pl[0].y, pl[0].z, pl[0].w}, S[4] = {pl[1].x, pl[1].y, pl[1].z, pl[1].w};
            const unsigned i0[4] = {pl[2].x, pl[2].y, pl[2].z, pl[2].w}, i1[4] = {pl[3].x, pl[3].y, pl[3].z, pl[3].w};
            const unsigned i2[4] = {pl[4].x, pl[4].y, pl[4].z, pl[4].w}, n[4] = {pl[5].x, pl[5].y, pl[5].z, pl[5].w};
            const unsigned wi[4] = {WINNER ? pl[7].x : 0u, WINNER ? pl[7].y : 0u, WINNER ? pl[7].z : 0u, WINNER ? pl[7].w : 0u};
            CellOut o[4];
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const unsigned off = n[j] * (unsigned)SYN3R_FX_MAGIC_BITS;     // every update added MAGIC_BITS + rint(.)
                if (gated)
                    o[j] = fx_resolve_cell<true>(wm[j], S[j] - off, (int)(i0[j] - off), (int)(i1[j] - off), (int)(i2[j] - off),
                                                 wi[j], cscale, gate);
                else
                    o[j] = fx_resolve_cell<false>(wm[j], S[j] - off, (int)(i0[j] - off), (int)(i1[j] - off), (int)(i2[j] - off),
                                                  wi[j], cscale, gate);
                range = max(range, (vec || lane * 4 + j < tw) ? o[j].range : 0u);
            }
            if (vec) {
                st_once_u32(reinterpret_cast<unsigned *>(a.out_mask + q), (unsigned)o[0].m | ((unsigned)o[1].m << 8) |
                                                                          ((unsigned)o[2].m << 16) | ((unsigned)o[3].m << 24), once);
                if (a.out_f32) {
                    float4 *f = reinterpret_cast<float4 *>(a.out_f32 + q * 3);
                    st_once_f4(f, make_float4(o[0].v0, o[0].v1, o[0].v2, o[1].v0), once);
                    st_once_f4(f + 1, make_float4(o[1].v1, o[1].v2, o[2].v0, o[2].v1), once);
                    st_once_f4(f + 2, make_float4(o[2].v2, o[3].v0, o[3].v1, o[3].v2), once);
                }
                if (a.out_u8) {
                    unsigned b[12], wd[3];
#pragma unroll
                    for (int j = 0; j < 4; ++j) { b[3 * j] = u8_round(o[j].v0); b[3 * j + 1] = u8_round(o[j].v1); b[3 * j + 2] = u8_round(o[j].v2); }
#pragma unroll
                    for (int k = 0; k < 3; ++k) wd[k] = b[4 * k] | (b[4 * k + 1] << 8) | (b[4 * k + 2] << 16) | (b[4 * k + 3] << 24);
                    if (u8_wide) {
                        // the row's 384 bytes leave as 24 x 16-byte stores, whole 128-byte lines (one third of the store
                        // instructions and, when out_u8 is a peer's memory, whole NVLink write packets instead of 4 bytes
                        // every 12): staged through this row of the I0 plane, which was read and cleared above
                        unsigned *stage = fx_smem + FX_I0 * CELLS + ly * SYN3R_TILE_W;
#pragma unroll
                        for (int k = 0; k < 3; ++k) stage[3 * lane + k] = wd[k];
                        __syncwarp();
                        if (lane < 24) {
                            const uint4 v = *reinterpret_cast<const uint4 *>(stage + 4 * lane);
                            *reinterpret_cast<uint4 *>(stage + 4 * lane) = make_uint4(0u, 0u, 0u, 0u);
                            st_once_i4(reinterpret_cast<int4 *>(a.out_u8 + (q - (size_t)(lane * 4)) * 3) + lane,
                                       make_int4((int)v.x, (int)v.y, (int)v.z, (int)v.w), once);
                        }
                        __syncwarp();
                    } else {
                        unsigned *u = reinterpret_cast<unsigned *>(a.out_u8 + q * 3);
#pragma unroll
                        for (int k = 0; k < 3; ++k) st_once_u32(u + k, wd[k], once);
                    }
                }
                if (WINNER && a.out_winner)
                    st_once_i4(reinterpret_cast<int4 *>(a.out_winner + q), make_int4(o[0].win, o[1].win, o[2].win, o[3].win), once);
            } else {
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    if (lane * 4 + j >= tw) continue;
                    const size_t qq = q + j;
                    a.out_mask[qq] = (uint8_t)o[j].m;
                    if (a.out_f32) { a.out_f32[3 * qq] = o[j].v0; a.out_f32[3 * qq + 1] = o[j].v1; a.out_f32[3 * qq + 2] = o[j].v2; }
                    if (a.out_u8) {
                        a.out_u8[3 * qq] = (uint8_t)u8_round(o[j].v0); a.out_u8[3 * qq + 1] = (uint8_t)u8_round(o[j].v1);
                        a.out_u8[3 * qq + 2] = (uint8_t)u8_round(o[j].v2);
                    }
                    if (WINNER && a.out_winner) a.out_winner[qq] = o[j].win;
                }
            }
        }
        if ((a.flags & SYN3R_FWD_IS_IMAGE) && range > 0x43800000u) ctl[7] = 1;    // :121-125
        __syncthreads();
        if (tid == 0 && ctl[7] && a.out_status) atomicOr(a.out_status + slot, 1);
#if SYN3R_FX_PHASE_CLOCKS
        if (tid == 0) {                 // PAIR mode, FLAT cull only (ctl[10..13] are written there)
            const unsigned t4 = (unsigned)clock(), cull = (unsigned)ctl[10] - ph_tb, t2 = (unsigned)ctl[11] - ph_tb;
            const unsigned f1 = min((unsigned)ctl[12], t2), t3 = ph_t3 - ph_tb, f2 = max(min((unsigned)ctl[13], t3), t2);
            ph[1] += cull; ph[2] += f1 - cull; ph[3] += t2 - f1; ph[4] += f2 - t2; ph[5] += t3 - f2; ph[6] += t4 - ph_t3;
            ph[7] += 1;
        }
#endif
    }
#if SYN3R_FX_PHASE_CLOCKS
    if (tid == 0)
        for (int k = 0; k < 8; ++k) g_fx_phase[(blockIdx.x & 1023) * 8 + k] = ph[k];
#endif
}

}  // namespace syn3r
