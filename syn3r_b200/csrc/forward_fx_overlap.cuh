// forward_fx_overlap.cuh — k_fwd_tile_fx_ov: the integer tile kernel of forward_fx.cuh (PAIR mode, frames of up to
// SYN3R_FX_FLAT blocks, per-(pair, tile) block lists present) with the RESOLVE step of a work item overlapped with
// sweep 1 of the next one.
//
// Why: the phase clocks of k_fwd_tile_fx (tools/phase_clocks.py) put 15 % of the kernel in the resolve step, a burst in
// which every warp walks one dependent chain per row (shared-memory loads -> arithmetic -> MUFU -> global stores) with
// nothing else to issue, between two barriers; sweep 1 of the next item only needs three of the eight planes (WMAX, N,
// ZKEY) to be clean.  So the resolve step is split:
//   phase 1   (all warps, between two barriers) every warp copies the WMAX and N words of its two rows into registers
//             and clears WMAX, N, ZKEY and the trash row: a dozen instructions per thread;
//   phase 2   (per warp, no barrier) the warp reads and clears S, I0..I2, IDX of its rows, finishes the arithmetic and
//             stores the outputs - and goes straight on to claim sweep-1 units of the NEXT item.  Sweep 1 touches only
//             the three planes phase 1 cleaned, so warps that are still in phase 2 and warps that already splat run
//             side by side, and the barrier that ends sweep 1 also orders every phase 2 before sweep 2.
// Three barriers per item instead of five.  The next item is claimed right after the sweep-1 barrier; warp 0 issues the
// cp.async copies of its constants and block list into the OTHER of two small staging buffers when it has run out of
// sweep-2 units, i.e. inside its own wait for the slowest warp.  Same arithmetic, same outputs (bitwise) as k_fwd_tile_fx.
//
// MEASURED (B200, round 2): 1.191 ms against 1.192 ms for k_fwd_tile_fx on config B (C: 0.692 / 0.694, D: 0.753 / 0.753),
// step 1.613 / 1.615 ms.  Every warp has the same two rows of phase 2, so they all leave it together and the overlap with
// sweep 1 is a few hundred cycles at the edges; the resolve step's cost is the serial chain of a row (shared-memory
// loads, arithmetic, stores) and not its barriers.  Kept behind syn3r_set_fx_overlap(1) (off by default), with its
// parity test, as the record of the experiment.
#pragma once
#include "forward_fx.cuh"

namespace syn3r {

#ifndef SYN3R_FX_OVERLAP
#define SYN3R_FX_OVERLAP 1
#endif
#define SYN3R_FX_OV_STAGE_BYTES 192     /* PairFast (128) + list count (4) + source index (4), padded */
// a warp owns exactly two tile rows in the split resolve step (A/B builds with other thread counts keep k_fwd_tile_fx)
#define SYN3R_FX_OV_AVAILABLE (SYN3R_FX_OVERLAP && SYN3R_FX_PREFETCH && !SYN3R_FX_TWO_CTAS && !SYN3R_FX_PHASE_CLOCKS && \
                               SYN3R_FX_EXPERIMENT == 0 && SYN3R_TILE_H == 2 * (SYN3R_TILE_THREADS / 32))

__host__ __device__ inline size_t fx_ov_smem_bytes(bool winner) {
    return fx_smem_bytes(winner) + SYN3R_FX_OV_STAGE_BYTES + 2 * SYN3R_TILE_LIST_CAP * 4;
}

struct FxOvOut {                // what phase 2 needs of TileArgs (passed by value to the out-of-line function)
    float *out_f32; uint8_t *out_u8, *out_mask; int *out_winner, *out_status;
    int flags, W, H, ntiles, ntx;
};

// outputs of one tile row from its accumulators (the resolve arithmetic and stores of k_fwd_tile_fx, same statements)
template <bool WINNER>
__device__ __forceinline__ unsigned fx_ov_row_out(const FxOvOut &a, const uint4 (&pl)[8], int ly, int lane, size_t q, int tw,
                                                  bool vec, bool u8_wide, bool gated, float cscale, float gate, uint64_t once,
                                                  int CELLS) {
    const unsigned wm[4] = {pl[0].x, pl[0].y, pl[0].z, pl[0].w}, S[4] = {pl[1].x, pl[1].y, pl[1].z, pl[1].w};
    const unsigned i0[4] = {pl[2].x, pl[2].y, pl[2].z, pl[2].w}, i1[4] = {pl[3].x, pl[3].y, pl[3].z, pl[3].w};
    const unsigned i2[4] = {pl[4].x, pl[4].y, pl[4].z, pl[4].w}, n[4] = {pl[5].x, pl[5].y, pl[5].z, pl[5].w};
    const unsigned wi[4] = {WINNER ? pl[7].x : 0u, WINNER ? pl[7].y : 0u, WINNER ? pl[7].z : 0u, WINNER ? pl[7].w : 0u};
    CellOut o[4];
    unsigned range = 0u;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        const unsigned off = n[j] * (unsigned)SYN3R_FX_MAGIC_BITS;     // every update added MAGIC_BITS + rint(.)
        if (gated)
            o[j] = fx_resolve_cell<true>(wm[j], S[j] - off, (int)(i0[j] - off), (int)(i1[j] - off), (int)(i2[j] - off), wi[j],
                                         cscale, gate);
        else
            o[j] = fx_resolve_cell<false>(wm[j], S[j] - off, (int)(i0[j] - off), (int)(i1[j] - off), (int)(i2[j] - off), wi[j],
                                          cscale, gate);
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
                // whole 128-byte lines, staged through this row of the I0 plane (read and cleared by the caller; a
                // sweep-2 plane, which nobody touches before the barrier that ends the next sweep 1)
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
    return range;
}

#if SYN3R_FX_OV_AVAILABLE
// phase 2 of the split resolve step for this warp's rows (warp, warp + NW) of work item `item`
template <bool WINNER>
__device__ __forceinline__ void fx_ov_finish(const FxOvOut a, int item, bool redo, uint4 wm0, uint4 n0, uint4 wm1, uint4 n1,
                                          float cscale) {
    constexpr int CELLS = FX_PLANE, NPL = WINNER ? 8 : 6, NW = SYN3R_TILE_THREADS / 32;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int W = a.W, H = a.H;
    const unsigned hw = (unsigned)(H * W);
    const float gate = 0.6f * (float)SYN3R_W_SCALE;
    const bool gated = (a.flags & SYN3R_FWD_GATE_0P6) != 0;
    const uint64_t once = l2_policy_evict_first();
    const bool u8_ok = ((W & 15) == 0) && ((hw & 15u) == 0u) && ((reinterpret_cast<uintptr_t>(a.out_u8) & 15) == 0);
    const int slot = item / a.ntiles, tile = item - slot * a.ntiles;
    const int tx0 = (tile % a.ntx) * SYN3R_TILE_W, ty0 = (tile / a.ntx) * SYN3R_TILE_H;
    const int tw = min(tx0 + SYN3R_TILE_W, W) - tx0, th = min(ty0 + SYN3R_TILE_H, H) - ty0;
    const bool vec = (tw == SYN3R_TILE_W) && ((W & 3) == 0);
    unsigned range = 0u;
#pragma unroll 1
    for (int h = 0; h < 2; ++h) {
        const int ly = warp + h * NW;
        const int c4 = ly * SYN3R_TILE_W + lane * 4;
        uint4 pl[8];
        pl[0] = h ? wm1 : wm0; pl[5] = h ? n1 : n0;
#pragma unroll
        for (int k = 1; k < 8; ++k) {
            if (k == 5 || k == FX_ZKEY || k >= NPL) continue;
            uint4 *pp = reinterpret_cast<uint4 *>(fx_smem + k * CELLS + c4);
            pl[k] = *pp;
            const unsigned clr = (k == FX_IDX) ? ~0u : 0u;
            *pp = make_uint4(clr, clr, clr, clr);
        }
        if (ly >= th || redo) continue;
        const size_t q = (size_t)slot * hw + (size_t)((ty0 + ly) * W + tx0 + lane * 4);
        range = max(range, fx_ov_row_out<WINNER>(a, pl, ly, lane, q, tw, vec, vec && u8_ok, gated, cscale, gate, once, CELLS));
    }
    if ((a.flags & SYN3R_FWD_IS_IMAGE) && range > 0x43800000u && a.out_status) atomicOr(a.out_status + slot, 1);   // :121-125
}

template <bool WINNER>
__global__ void __launch_bounds__(SYN3R_TILE_THREADS, 1)
k_fwd_tile_fx_ov(const TileArgs a) {
    constexpr int CELLS = FX_PLANE, NPL = WINNER ? 8 : 6, NW = SYN3R_TILE_THREADS / 32;
    static_assert(SYN3R_TILE_H == 2 * NW, "a warp owns exactly two tile rows in the split resolve step");
    FxAux sm;
    unsigned char *ptr = reinterpret_cast<unsigned char *>(fx_smem + NPL * CELLS);
    sm.logtab = reinterpret_cast<double2 *>(ptr); ptr += (size_t)SYN3R_LOGTAB * 16;
    sm.listA = reinterpret_cast<int *>(ptr); ptr += SYN3R_LISTA * 4;
    sm.listB = reinterpret_cast<int *>(ptr); ptr += SYN3R_FX_LISTB_CAP * 4;        // in-kernel cull of an overflowing list
    sm.ctl = reinterpret_cast<int *>(ptr); ptr += 64;
    unsigned char *const stage_base = ptr;                                         // constants of this / the next item
    ptr += 2 * SYN3R_FX_OV_STAGE_BYTES;
    int *const slist_base = reinterpret_cast<int *>(ptr);                          // their block lists
    auto stage_of = [&](int b) { return stage_base + b * SYN3R_FX_OV_STAGE_BYTES; };
    auto slist_of = [&](int b) { return slist_base + b * SYN3R_TILE_LIST_CAP; };
    int *ctl = sm.ctl;

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const TileGeom g = a.g;
    const int W = g.W, H = g.H;
    const int n_items = a.n_slots * g.ntiles;

    const unsigned cb = *a.cmax_bits;
    if (cb >= 0x7f000000u) {               // colours inf / NaN / > 2^127: the fp32 kernel takes every item
        if (tid == 0)
            for (;;) {
                const int it = (int)atomicAdd(a.work_counter, 1u);
                if (it >= n_items) break;
                a.redo_list[atomicAdd(a.redo_count, 1u)] = it;
            }
        return;
    }
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
    const unsigned ce = min(max((cb >> 23) + ((cb & 0x7fffffu) ? 1u : 0u), 1u), 253u);
    const float cscale = cb ? __uint_as_float(ce << 23) : 1.0f;
    const float cs_inv = cb ? __uint_as_float((254u - ce) << 23) : 1.0f;

    // warp 0: claim a work item and start the copies of its constants, source index and block list into buffer `buf`
    auto claim_and_stage = [&](int buf, bool claim = true) {
        int it = 0;
        if (lane == 0) {
            if (claim) { it = (int)atomicAdd(a.work_counter, 1u); ctl[8] = it; }
            else it = ctl[8];
        }
        it = __shfl_sync(0xffffffffu, it, 0);
        if (it < n_items) {
            const int np = it / g.ntiles;                      // PAIR mode: pair = slot
            const size_t nt = (size_t)np * g.ntiles + (size_t)(it - np * g.ntiles);
            if (lane < 8) cp_async_16(stage_of(buf) + lane * 16, reinterpret_cast<const unsigned char *>(a.pair_fast + np) + lane * 16);
            else if (lane == 8) cp_async_4(stage_of(buf) + 128, a.tile_cnt + nt);
            else if (lane == 9) cp_async_4(stage_of(buf) + 132, a.pair_src + np);
            for (int k = lane; k < SYN3R_TILE_LIST_CAP / 4; k += 32)
                cp_async_16(slist_of(buf) + k * 4, a.tile_list + nt * SYN3R_TILE_LIST_CAP + (size_t)k * 4);
        }
        cp_async_commit();
    };
    if (tid == 0) { ctl[2] = 0; ctl[3] = 0; ctl[4] = 0; ctl[5] = 0; ctl[6] = 0; ctl[9] = 0; ctl[10] = -1; ctl[11] = 0; ctl[14] = 0; ctl[15] = 0; }
    if (warp == 0) { claim_and_stage(0); cp_async_wait_all(); }
    uint4 keep_wm[2], keep_n[2];           // WMAX and N of this warp's two rows of the previous item (phase 1 -> phase 2)
    keep_wm[0] = keep_wm[1] = keep_n[0] = keep_n[1] = make_uint4(0u, 0u, 0u, 0u);

    for (;;) {
        __syncthreads();                   // the sweep-1 planes are clean, the counters reset, the staged copies visible
        // Everything warp-uniform about this item and the previous one is read HERE, right behind the barrier: the
        // compiler keeps a value in uniform registers only where it can see that the warp is converged, and with these
        // loads behind phase 2 the pair constants went to vector registers and 250 B of spills instead.
        // ctl[8] = the item, [9] = its staging buffer, [10] / [11] = the previous item (-1: none) and its redo flag
        const int item = ctl[8], buf = ctl[9], prev_item = ctl[10];
        const bool prev_redo = ctl[11] != 0;
        const int slot = item / g.ntiles, tile = item - slot * g.ntiles;
        const int tx0 = (tile % g.ntx) * SYN3R_TILE_W, ty0 = (tile / g.ntx) * SYN3R_TILE_H;
        const int tx1 = min(tx0 + SYN3R_TILE_W, W) - 1, ty1 = min(ty0 + SYN3R_TILE_H, H) - 1;
        FxPairRegs pr;
        const PairFast *pf = reinterpret_cast<const PairFast *>(stage_of(buf));
        const int p = slot, s = *reinterpret_cast<const int *>(stage_of(buf) + 132);
        const unsigned cnt = *reinterpret_cast<const unsigned *>(stage_of(buf) + 128);
        const int staged_n = cnt <= SYN3R_TILE_LIST_CAP ? (int)cnt : -1;       // -1: in-kernel cull into sm.listB
        FxAux smp = sm;
        if (staged_n >= 0) smp.listB = slist_of(buf);
        pr.cexp = pf->cexp;
#pragma unroll
        for (int k = 0; k < 9; ++k) pr.A[k] = pf->A[k];
#pragma unroll
        for (int k = 0; k < 3; ++k) pr.b[k] = pf->b[k];
        pr.k0 = pf->k0; pr.k1 = pf->k1; pr.k2 = pf->k2; pr.k3 = pf->k3; pr.k4 = pf->k4;
        pr.clamp_hi = pf->pad != 0.0f;
        pr.cexp_f = __double2float_rn(pr.cexp);
        pr.upper_ok = pr.cexp <= 500.0;
        if (tid == 0) { ctl[3] = 0; ctl[12] = item; }     // (ctl[3]: every thread read it before the barrier above; next written
                                                         // behind the next one; [12]: this item, for thread 0 after sweep 2)
        // phase 2 of the previous item, per warp, no barrier behind it: the warp goes straight on to sweep 1
        if (prev_item >= 0)
            fx_ov_finish<WINNER>(FxOvOut{a.out_f32, a.out_u8, a.out_mask, a.out_winner, a.out_status, a.flags, W, H, g.ntiles, g.ntx},
                                 prev_item, prev_redo, keep_wm[0], keep_n[0], keep_wm[1], keep_n[1], cscale);
        if (item >= n_items) break;
        if (pr.cexp < CUDART_INF) {        // else: max log-depth NaN or 0, every weight is NaN (:84): the tile stays empty
            fx_pair<WINNER, 1, true>(a, smp, pr, p, s, tile, tx0, ty0, tx1, ty1, false, true, cs_inv, staged_n);
            __syncthreads();               // sweep 1 done, every warp's phase 2 of the previous item done
            if (tid == 0) ctl[8] = (int)atomicAdd(a.work_counter, 1u);      // the next item
            // a cell with FX_MAX_N or more updates could pass S = 2^31 in sweep 2: the item goes to the redo list
            bool crowded = false;
            for (int k = tid * 4; k < FX_CELLS; k += SYN3R_TILE_THREADS * 4) {
                const uint4 n = *reinterpret_cast<const uint4 *>(fx_smem + FX_N * CELLS + k);
                crowded = crowded || max(max(n.x, n.y), max(n.z, n.w)) >= FX_MAX_N;
            }
            if (crowded) ctl[3] = 1;
            fx_pair<WINNER, 2, true>(a, smp, pr, p, s, tile, tx0, ty0, tx1, ty1, true, false, cs_inv, staged_n);
        } else {
            __syncthreads();
            if (tid == 0) ctl[8] = (int)atomicAdd(a.work_counter, 1u);
        }
        // warp 0, done with its own sweep-2 units, copies the next item's constants and block list into the other staging
        // buffer while the slower warps finish (issued between the sweeps instead, the copies' address arithmetic cost
        // sweep 2 its registers: 128 B of spills)
        if (warp == 0) { claim_and_stage(ctl[9] ^ 1, false); cp_async_wait_all(); }
        __syncthreads();                   // sweep 2 done; ctl[3], ctl[8] and the staged copies are visible
        const bool redo = ctl[3] != 0;
        // phase 1: keep WMAX and N of this warp's two rows, clean the planes sweep 1 of the next item writes
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            const int c4 = (warp + h * NW) * SYN3R_TILE_W + lane * 4;
            uint4 *pw = reinterpret_cast<uint4 *>(fx_smem + FX_WMAX * CELLS + c4), *pn = reinterpret_cast<uint4 *>(fx_smem + FX_N * CELLS + c4);
            keep_wm[h] = *pw; keep_n[h] = *pn;
            *pw = make_uint4(0u, 0u, 0u, 0u); *pn = make_uint4(0u, 0u, 0u, 0u);
            if (WINNER) *reinterpret_cast<uint4 *>(fx_smem + FX_ZKEY * CELLS + c4) = make_uint4(~0u, ~0u, ~0u, ~0u);
        }
        for (int k = tid; k < NPL * SYN3R_TILE_W; k += SYN3R_TILE_THREADS) {        // the trash row of every plane
            const int pl = k / SYN3R_TILE_W;
            fx_smem[pl * CELLS + FX_CELLS + (k - pl * SYN3R_TILE_W)] = (pl >= FX_ZKEY) ? ~0u : (pl == FX_WMAX ? 0x3f800000u : 0u);
        }
        if (tid == 0) {                    // (item and buffer re-read from shared memory: nothing of them lives across sweep 2)
            const int done = ctl[12];
            if (redo) a.redo_list[atomicAdd(a.redo_count, 1u)] = done;
            ctl[2] = 0; ctl[4] = 0; ctl[5] = 0; ctl[6] = 0; ctl[14] = 0; ctl[15] = 0;
            ctl[9] ^= 1; ctl[10] = done; ctl[11] = redo ? 1 : 0;
        }
    }
}
#endif  // SYN3R_FX_OV_AVAILABLE

}  // namespace syn3r
