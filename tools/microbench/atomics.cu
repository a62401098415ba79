// Microbenchmark: cost of the scatter primitives the forward splat can be built from (B200, sm_100a).
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o atomics atomics.cu ; run: ./atomics
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e), __FILE__, __LINE__); exit(1);} } while (0)

__device__ __forceinline__ void red4(float4 *p, float a, float b, float c, float d) {
    asm volatile("red.global.add.v4.f32 [%0], {%1,%2,%3,%4};" :: "l"(p), "f"(a), "f"(b), "f"(c), "f"(d) : "memory");
}
__device__ __forceinline__ uint32_t hash32(uint32_t x) { x ^= x >> 16; x *= 0x7feb352dU; x ^= x >> 15; x *= 0x846ca68bU; x ^= x >> 16; return x; }

// mode 0: coalesced (lane i -> cell i of a per-warp row), 1: coalesced + 1-cell overlap between neighbours (two ops),
// 2: random cells in a 64 MB window, 3: random cells in a 1 GB window
template <int OP>   // 0 red.v4.f32, 1 4x red.f32, 2 atomicMin u64, 3 st.v4 (plain), 4 red.v4 + min64
__global__ void k_glob(float4 *acc, unsigned long long *keys, size_t ncell, int mode, int iters) {
    size_t gtid = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    size_t nthr = (size_t)gridDim.x * blockDim.x;
    for (int it = 0; it < iters; ++it) {
        size_t c;
        if (mode == 0) c = (gtid + (size_t)it * nthr) % ncell;
        else if (mode == 2) c = hash32((uint32_t)(gtid + (size_t)it * nthr)) % (4u << 20);
        else c = (hash32((uint32_t)(gtid + (size_t)it * nthr)) * 16ull + (gtid & 15)) % ncell;
        float v = (float)it;
        if (OP == 0 || OP == 4) red4(acc + c, v, v, v, 1.0f);
        if (OP == 1) { float *f = (float *)(acc + c); atomicAdd(f, v); atomicAdd(f + 1, v); atomicAdd(f + 2, v); atomicAdd(f + 3, 1.0f); }
        if (OP == 2 || OP == 4) atomicMin(keys + c, (unsigned long long)(gtid + it));
        if (OP == 3) acc[c] = make_float4(v, v, v, 1.0f);
    }
}

// shared-memory atomics: each warp hits 32 distinct banks (conflict-free) in a 8K-cell window
template <int OP>  // 0 atomicAdd f32 x4 (4 ATOMS), 1 atomicMin u64, 2 atomicAdd f32 x1, 3 plain RMW x4 (no atomics)
__global__ void k_smem(float *out, int iters) {
    extern __shared__ float sm[];
    const int ncell = 8192;   // floats
    for (int i = threadIdx.x; i < ncell; i += blockDim.x) sm[i] = 0.f;
    __syncthreads();
    unsigned long long *k64 = (unsigned long long *)sm;
    for (int it = 0; it < iters; ++it) {
        int c = (threadIdx.x + it * 37) & 2047;  // cell of 4 floats
        float v = (float)it;
        if (OP == 0) { atomicAdd(sm + c, v); atomicAdd(sm + 2048 + c, v); atomicAdd(sm + 4096 + c, v); atomicAdd(sm + 6144 + c, 1.f); }
        if (OP == 1) atomicMin(k64 + c, (unsigned long long)(threadIdx.x + it));
        if (OP == 2) atomicAdd(sm + c, v);
        if (OP == 3) { sm[c] += v; sm[2048 + c] += v; sm[4096 + c] += v; sm[6144 + c] += 1.f; }
    }
    __syncthreads();
    if (threadIdx.x == 0) out[blockIdx.x] = sm[5] + sm[4097];
}

// native 32-bit integer shared atomics (ATOMS.ADD / ATOMS.MAX, no CAS loop)
template <int OP>  // 0 4x ATOMS.ADD u32, 1 ATOMS.MAX + ATOMS.ADD, 2 4x ATOMS.ADD with 2-way bank conflicts, 3 LDS + 4x ATOMS.ADD
__global__ void k_smem_int(unsigned *out, int iters) {
    extern __shared__ unsigned smi[];
    const int ncell = 8192;
    for (int i = threadIdx.x; i < ncell; i += blockDim.x) smi[i] = 0u;
    __syncthreads();
    unsigned acc = 0;
    for (int it = 0; it < iters; ++it) {
        int c = (threadIdx.x + it * 37) & 2047;
        if (OP == 2) c = ((threadIdx.x * 2) + it * 37) & 2047;
        unsigned v = (unsigned)it;
        if (OP == 3) acc += smi[(c + 5) & 2047];
        if (OP == 0 || OP == 2 || OP == 3) { atomicAdd(smi + c, v); atomicAdd(smi + 2048 + c, v); atomicAdd(smi + 4096 + c, v); atomicAdd(smi + 6144 + c, 1u); }
        if (OP == 1) { atomicMax(smi + c, v); atomicAdd(smi + 2048 + c, 1u); }
    }
    __syncthreads();
    if (threadIdx.x == 0) out[blockIdx.x] = smi[5] + smi[4097] + acc;
}

// red.v4.f32 into a CTA-private, L2-resident 128 KB window (8192 cells x 16 B), rows of 32 consecutive cells per warp
template <int OP>  // 0 red.v4.f32, 1 red.v4.f32 x4 corners (2x2 neighbourhood)
__global__ void k_glob_window(float4 *acc, int iters) {
    float4 *win = acc + (size_t)blockIdx.x * 8192;
    for (int it = 0; it < iters; ++it) {
        int c = (threadIdx.x + it * 37 * 32) & 8191;
        float v = (float)it;
        red4(win + c, v, v, v, 1.0f);
        if (OP == 1) { red4(win + ((c + 1) & 8191), v, v, v, 1.0f); red4(win + ((c + 128) & 8191), v, v, v, 1.0f); red4(win + ((c + 129) & 8191), v, v, v, 1.0f); }
    }
}

// fp64 / conversion pipe throughput
template <int OP>  // 0 DFMA chain x8 independent, 1 F2F.F32.F64 + F2F.F64.F32, 2 FRND floor f64, 3 ddiv, 4 MUFU.RCP64H
__global__ void k_alu(double *out, int iters, double seed) {
    double a[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) a[i] = seed + i + threadIdx.x * 1e-3;
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            if (OP == 0) a[i] = __fma_rn(a[i], 1.0000001, 1e-9);
            if (OP == 1) a[i] = (double)(__double2float_rn(a[i]) + 1.0f);
            if (OP == 2) a[i] = floor(a[i] * 1.5);
            if (OP == 3) a[i] = __ddiv_rn(seed, a[i] + 2.0);
            if (OP == 4) { double r; asm volatile("rcp.approx.ftz.f64 %0, %1;" : "=d"(r) : "d"(a[i])); a[i] = r + 1.5; }
        }
    }
    double s = 0;
#pragma unroll
    for (int i = 0; i < 8; ++i) s += a[i];
    if (s == 123.456) out[0] = s;
}

template <typename F>
float time_ms(F f, int reps = 5) {
    cudaEvent_t e0, e1;
    CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    f();
    CK(cudaDeviceSynchronize());
    float best = 1e30f;
    for (int r = 0; r < reps; ++r) {
        CK(cudaEventRecord(e0));
        f();
        CK(cudaEventRecord(e1));
        CK(cudaEventSynchronize(e1));
        float ms; CK(cudaEventElapsedTime(&ms, e0, e1));
        if (ms < best) best = ms;
    }
    return best;
}

int main() {
    cudaDeviceProp pr; CK(cudaGetDeviceProperties(&pr, 0));
    int sms = pr.multiProcessorCount;
    double ghz = 1.9;
    printf("device %s, %d SMs\n", pr.name, sms);
    size_t ncell = 64u << 20;   // 64M cells x 16 B = 1 GB
    float4 *acc; unsigned long long *keys;
    CK(cudaMalloc(&acc, ncell * 16)); CK(cudaMalloc(&keys, ncell * 8));
    CK(cudaMemset(acc, 0, ncell * 16)); CK(cudaMemset(keys, 0xff, ncell * 8));
    const int iters = 64, blocks = sms * 8, thr = 256;
    double lanes = (double)blocks * thr * iters;
    const char *opn[] = {"red.v4.f32", "4x red.f32", "min.u64", "st.v4 plain", "red.v4+min64"};
    const char *mdn[] = {"coalesced (1 GB sweep)", "-", "random in 64 MB (L2)", "random 16-cell runs in 1 GB"};
    for (int mode : {0, 2, 3}) {
        for (int op = 0; op < 5; ++op) {
            float ms = 0;
            auto run = [&]() {
                switch (op) {
                    case 0: k_glob<0><<<blocks, thr>>>(acc, keys, ncell, mode, iters); break;
                    case 1: k_glob<1><<<blocks, thr>>>(acc, keys, ncell, mode, iters); break;
                    case 2: k_glob<2><<<blocks, thr>>>(acc, keys, ncell, mode, iters); break;
                    case 3: k_glob<3><<<blocks, thr>>>(acc, keys, ncell, mode, iters); break;
                    case 4: k_glob<4><<<blocks, thr>>>(acc, keys, ncell, mode, iters); break;
                }
            };
            ms = time_ms(run);
            printf("global %-14s %-28s %8.3f ms  %7.2f Glane-op/s  %6.3f SM-cyc/lane\n", opn[op], mdn[mode], ms,
                   lanes / ms / 1e6, ms * 1e-3 * ghz * 1e9 * sms / lanes);
        }
    }
    float *out; CK(cudaMalloc(&out, 4 * blocks));
    const char *son[] = {"4x atomicAdd.f32", "atomicMin.u64", "1x atomicAdd.f32", "4x plain RMW"};
    for (int op = 0; op < 4; ++op) {
        const int it2 = 2048;
        auto run = [&]() {
            switch (op) {
                case 0: k_smem<0><<<sms * 4, 256, 32768>>>(out, it2); break;
                case 1: k_smem<1><<<sms * 4, 256, 32768>>>(out, it2); break;
                case 2: k_smem<2><<<sms * 4, 256, 32768>>>(out, it2); break;
                case 3: k_smem<3><<<sms * 4, 256, 32768>>>(out, it2); break;
            }
        };
        float ms = time_ms(run);
        double l2 = (double)sms * 4 * 256 * it2;
        printf("shared %-18s %8.3f ms  %7.2f Glane-op/s  %6.3f SM-cyc/lane\n", son[op], ms, l2 / ms / 1e6,
               ms * 1e-3 * ghz * 1e9 * sms / l2);
    }
    unsigned *uout; CK(cudaMalloc(&uout, 4 * blocks));
    const char *sin[] = {"4x ATOMS.ADD.u32", "ATOMS.MAX+ATOMS.ADD", "4x ATOMS.ADD 2-way conflict", "LDS + 4x ATOMS.ADD"};
    for (int op = 0; op < 4; ++op) {
        const int it2 = 2048;
        auto run = [&]() {
            switch (op) {
                case 0: k_smem_int<0><<<sms * 4, 256, 32768>>>(uout, it2); break;
                case 1: k_smem_int<1><<<sms * 4, 256, 32768>>>(uout, it2); break;
                case 2: k_smem_int<2><<<sms * 4, 256, 32768>>>(uout, it2); break;
                case 3: k_smem_int<3><<<sms * 4, 256, 32768>>>(uout, it2); break;
            }
        };
        float ms = time_ms(run);
        double l2 = (double)sms * 4 * 256 * it2;
        printf("shared %-28s %8.3f ms  %7.2f Glane-iter/s  %6.3f SM-cyc/lane-iter\n", sin[op], ms, l2 / ms / 1e6,
               ms * 1e-3 * ghz * 1e9 * sms / l2);
    }
    for (int op = 0; op < 2; ++op) {
        const int it2 = 512;
        for (int nb : {sms, sms * 2}) {
            auto run = [&]() {
                if (op == 0) k_glob_window<0><<<nb, 512>>>(acc, it2); else k_glob_window<1><<<nb, 512>>>(acc, it2);
            };
            float ms = time_ms(run);
            double l2 = (double)nb * 512 * it2 * (op ? 4 : 1);
            printf("global red.v4.f32 CTA-private 128 KB window (%s, %d CTAs x 512) %8.3f ms  %7.2f Glane-op/s  %6.3f SM-cyc/lane\n",
                   op ? "2x2 corners" : "1 cell", nb, ms, l2 / ms / 1e6, ms * 1e-3 * ghz * 1e9 * sms / l2);
        }
    }
    double *dout; CK(cudaMalloc(&dout, 8));
    const char *aon[] = {"DFMA", "F2F f64<->f32 pair", "floor.f64 (+DMUL)", "ddiv_rn (+DADD)", "rcp.approx.f64 (+DADD)"};
    for (int op = 0; op < 5; ++op) {
        const int it3 = 4096;
        auto run = [&]() {
            switch (op) {
                case 0: k_alu<0><<<sms * 8, 256>>>(dout, it3, 1.0); break;
                case 1: k_alu<1><<<sms * 8, 256>>>(dout, it3, 1.0); break;
                case 2: k_alu<2><<<sms * 8, 256>>>(dout, it3, 1.0); break;
                case 3: k_alu<3><<<sms * 8, 256>>>(dout, it3, 1.0); break;
                case 4: k_alu<4><<<sms * 8, 256>>>(dout, it3, 1.0); break;
            }
        };
        float ms = time_ms(run);
        double l3 = (double)sms * 8 * 256 * it3 * 8;
        printf("alu    %-24s %8.3f ms  %8.2f Gop/s  %6.2f lanes/clk/SM\n", aon[op], ms, l3 / ms / 1e6,
               l3 / (ms * 1e-3 * ghz * 1e9 * sms));
    }
    return 0;
}
