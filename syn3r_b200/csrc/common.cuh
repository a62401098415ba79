// common.cuh — shared device helpers and host-side error plumbing for libsyn3r_warp.so.
// Compiled for sm_100a only, with -fmad=false: every FMA on the parity-critical paths is written
// explicitly so the rounding sequence is the reference's (oracle/syn3r_oracle.c header).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#include "../../include/syn3r_warp.h"

namespace syn3r {

void set_error(const char *fmt, ...);
int check_cuda(cudaError_t e, const char *what);
int check_device();
int sm_count();

#define SYN3R_TRY_CUDA(expr)                                   \
    do {                                                       \
        int _rc = ::syn3r::check_cuda((expr), #expr);          \
        if (_rc != SYN3R_OK) return _rc;                       \
    } while (0)

#define SYN3R_REQUIRE(cond, msg)                               \
    do {                                                       \
        if (!(cond)) {                                         \
            ::syn3r::set_error("%s: %s", __func__, msg);       \
            return SYN3R_EARG;                                 \
        }                                                      \
    } while (0)

// weights are accumulated in fp32 after an exact power-of-two scaling: the smallest positive
// reference weight is ~1e-26 (prox) / e^50 ≈ 2e-48, below fp32's range; x 2^90 it is ~2e-21 and
// the largest possible sum (1e9 contributions of 2^90) stays below FLT_MAX.  The scale cancels
// in I/S, and S > 0 keeps the reference's meaning exactly (REDG.F32 flushes denormals).
#define SYN3R_W_SCALE 1237940039285380274899124224.0 /* 2^90 */

// ---- fp64 BLAS-order dots (oracle: dot3_blas / dot4_blas) -------------------------
__device__ __forceinline__ double dot3_blas(const double *a, double x0, double x1, double x2) {
    double t = __dmul_rn(a[1], x1);
    t = __fma_rn(a[0], x0, t);
    t = __fma_rn(a[2], x2, t);
    return t;
}
// 4-term row dot with x3 == 1.0: (a0*x0 + a2*x2) + (a1*x1 + a3), no FMA
__device__ __forceinline__ double dot4_blas_h(const double *a, double x0, double x1, double x2) {
    return __dadd_rn(__dadd_rn(__dmul_rn(a[0], x0), __dmul_rn(a[2], x2)), __dadd_rn(__dmul_rn(a[1], x1), a[3]));
}

// ---- fp32 torch-CPU-order dots (oracle: dot3_t / dot4_t) -------------------------------------
__device__ __forceinline__ float dot3_t(const float *a, float x0, float x1, float x2) {
    float t = __fmul_rn(a[0], x0);
    t = __fmaf_rn(a[1], x1, t);
    t = __fmaf_rn(a[2], x2, t);
    return t;
}
__device__ __forceinline__ float dot4_t(const float *a, float x0, float x1, float x2, float x3) {
    float t = __fmul_rn(a[0], x0);
    t = __fmaf_rn(a[1], x1, t);
    t = __fmaf_rn(a[2], x2, t);
    t = __fmaf_rn(a[3], x3, t);
    return t;
}

// order-preserving maps of IEEE bit patterns to unsigned integers (for atomicMin / atomicMax)
__device__ __forceinline__ uint32_t ordered_u32(float f) {
    uint32_t b = __float_as_uint(f);
    return (b & 0x80000000u) ? ~b : (b | 0x80000000u);
}
__device__ __forceinline__ float unordered_f32(uint32_t k) {
    return __uint_as_float((k & 0x80000000u) ? (k & 0x7fffffffu) : ~k);
}
__device__ __forceinline__ unsigned long long ordered_u64(double d) {
    unsigned long long b = (unsigned long long)__double_as_longlong(d);
    return (b >> 63) ? ~b : (b | 0x8000000000000000ull);
}
__device__ __forceinline__ double unordered_f64(unsigned long long k) {
    return __longlong_as_double((long long)((k >> 63) ? (k & 0x7fffffffffffffffull) : ~k));
}

// winner depth key (extension output, include/syn3r_warp.h): the high word of the fp64 z' (sign, exponent,
// 20 mantissa bits), order-preserving.  Truncation instead of a float rounding so that a cheap integer test on
// the low word tells when two evaluations of z' could disagree on the key.
__device__ __forceinline__ unsigned zkey_of(double z) { return ordered_u32(__int_as_float(__double2hiint(z))); }

// streaming (read-once / write-once) accesses: keep L1 for the data that is re-used
__device__ __forceinline__ float4 ld_stream_f4(const float4 *p) {
    float4 r;
    asm volatile("ld.global.nc.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];"
                 : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w) : "l"(p));
    return r;
}
__device__ __forceinline__ uint4 ld_stream_u4(const uint4 *p) {
    uint4 r;
    asm volatile("ld.global.cs.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
    return r;
}
__device__ __forceinline__ void st_stream_f4(float4 *p, float4 v) {
    asm volatile("st.global.L1::no_allocate.v4.f32 [%0], {%1,%2,%3,%4};"
                 :: "l"(p), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}
__device__ __forceinline__ void st_stream_u32(uint32_t *p, uint32_t v) {
    asm volatile("st.global.L1::no_allocate.u32 [%0], %1;" :: "l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ void st_stream_f32(float *p, float v) {
    asm volatile("st.global.L1::no_allocate.f32 [%0], %1;" :: "l"(p), "f"(v) : "memory");
}
// L2 residency hints: the source RGB-D (tens of MB, re-read by every target pose) should stay in the 126 MB L2
// while the outputs (about a GB per trajectory, written once) stream through it.
__device__ __forceinline__ uint64_t l2_policy_evict_last() {
    uint64_t p;
    asm volatile("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(p));
    return p;
}
__device__ __forceinline__ uint64_t l2_policy_evict_first() {
    uint64_t p;
    asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(p));
    return p;
}
__device__ __forceinline__ float4 ld_keep_f4(const float4 *p, uint64_t pol) {
    float4 r;
    asm volatile("ld.global.nc.L1::no_allocate.L2::cache_hint.v4.f32 {%0,%1,%2,%3}, [%4], %5;"
                 : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w) : "l"(p), "l"(pol));
    return r;
}
__device__ __forceinline__ void st_once_f4(float4 *p, float4 v, uint64_t pol) {
    asm volatile("st.global.L1::no_allocate.L2::cache_hint.v4.f32 [%0], {%1,%2,%3,%4}, %5;"
                 :: "l"(p), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w), "l"(pol) : "memory");
}
__device__ __forceinline__ void st_once_u32(uint32_t *p, uint32_t v, uint64_t pol) {
    asm volatile("st.global.L1::no_allocate.L2::cache_hint.u32 [%0], %1, %2;" :: "l"(p), "r"(v), "l"(pol) : "memory");
}
__device__ __forceinline__ void st_once_i4(int4 *p, int4 v, uint64_t pol) {
    asm volatile("st.global.L1::no_allocate.L2::cache_hint.v4.s32 [%0], {%1,%2,%3,%4}, %5;"
                 :: "l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w), "l"(pol) : "memory");
}
// vector fp32 reduction into L2 (REDG.E.ADD.F32x4), one 16-byte transaction per corner
__device__ __forceinline__ void red_add_f4(float4 *p, float a, float b, float c, float d) {
    asm volatile("red.global.add.v4.f32 [%0], {%1,%2,%3,%4};"
                 :: "l"(p), "f"(a), "f"(b), "f"(c), "f"(d) : "memory");
}

}  // namespace syn3r
