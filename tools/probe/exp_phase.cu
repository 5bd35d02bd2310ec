// Probe: cost of the exponential phase of the attention forward (one query row of 128 scores per thread -> 64 packed bf16
// pairs + the row sum) as the compiler schedules it, for one and for two warps per SM sub-partition, plus the issue cost of
// the packed / scalar FMA-pipe forms it is built from.  No tensor cores, no memory traffic: registers only.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -I avion_b200/csrc -o tools/probe/exp_phase tools/probe/exp_phase.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
#include "common.cuh"
void avn_set_error(const char*, ...) {}

using namespace avn;
constexpr int TILE = 128;

// ---- variants of row_exp ------------------------------------------------------------------------------------------------
// NUM/DEN: share of the score pairs whose exponential runs on the FMA pipe; SCALAR_FMA: x = s*c - m as two FFMA (uniform
// register operand) instead of one FFMA2; SUMMODE 0: FADD2 into two chains, 1: FADD2 into four chains, 2: no row sum (lower
// bound: what a tensor-core row sum would leave)
template <int NUM, int DEN, bool SCALAR_FMA, int SUMMODE, int PACK = 0>
__device__ __forceinline__ float row_exp_v(const uint32_t* sv, uint32_t* pk, float scale_log2, float m_ref) {
    const float2 sc2 = make_float2(scale_log2, scale_log2), nm2 = make_float2(-m_ref, -m_ref);
    float2 r[4];
#pragma unroll
    for (int q = 0; q < 4; ++q) r[q] = make_float2(0.f, 0.f);
#pragma unroll
    for (int i = 0; i < TILE / 2; ++i) {
        float2 x;
        if (SCALAR_FMA) {
            x.x = fmaf(__uint_as_float(sv[2 * i]), scale_log2, -m_ref);
            x.y = fmaf(__uint_as_float(sv[2 * i + 1]), scale_log2, -m_ref);
        } else {
            x = __ffma2_rn(make_float2(__uint_as_float(sv[2 * i]), __uint_as_float(sv[2 * i + 1])), sc2, nm2);
        }
        float2 e;
        if (NUM > 0 && (i % DEN) < NUM) e = exp2_poly2(x);
        else e = make_float2(ex2(x.x), ex2(x.y));
        if (SUMMODE == 0) r[i & 1] = __fadd2_rn(r[i & 1], e);
        if (SUMMODE == 1) r[i & 3] = __fadd2_rn(r[i & 3], e);
        // PACK 1: truncate (one PRMT on the ALU pipe) instead of F2FP (which turns out to share the MUFU's pipe)
        pk[i] = PACK == 1 ? __byte_perm(__float_as_uint(e.x), __float_as_uint(e.y), 0x7632) : pack_bf16(e.x, e.y);
    }
    const float2 t = __fadd2_rn(__fadd2_rn(r[0], r[1]), __fadd2_rn(r[2], r[3]));
    return t.x + t.y;
}

template <int NUM, int DEN, bool SCALAR_FMA, int SUMMODE, int PACK = 0>
__global__ void __launch_bounds__(256, 1) phase_kernel(float* out, long long* cyc, float scale_log2, float seed, int iters) {
    // the S row arrives through real (volatile) shared-memory loads and P leaves through real stores, standing in for the
    // tcgen05.ld / tcgen05.st of the kernel: ptxas cannot hoist or drop anything.  Layout [vector][thread]: conflict-free.
    extern __shared__ uint4 sm4[];
    const int nt = blockDim.x;
    for (int v = 0; v < TILE / 4; ++v) {
        const float b = -seed * (float)((threadIdx.x * 7 + v * 13) % 97);
        sm4[v * nt + threadIdx.x] = make_uint4(__float_as_uint(b), __float_as_uint(b - 1.f), __float_as_uint(b - 2.f), __float_as_uint(b - 3.f));
    }
    float l = 0.f, m_ref = seed;
    __syncthreads();
    const uint32_t base = smem_u32(sm4 + threadIdx.x);
    const long long t0 = clock64();
#pragma unroll 1
    for (int it = 0; it < iters; ++it) {
        uint32_t sv[TILE];
#pragma unroll
        for (int v = 0; v < TILE / 4; ++v)
            asm volatile("ld.volatile.shared.v4.b32 {%0, %1, %2, %3}, [%4];" : "=r"(sv[4 * v]), "=r"(sv[4 * v + 1]), "=r"(sv[4 * v + 2]), "=r"(sv[4 * v + 3])
                         : "r"(base + v * nt * 16));
        uint32_t pk[TILE / 2];
        l += row_exp_v<NUM, DEN, SCALAR_FMA, SUMMODE, PACK>(sv, pk, scale_log2, m_ref);
#pragma unroll
        for (int v = 0; v < TILE / 8; ++v)
            asm volatile("st.volatile.shared.v4.b32 [%4], {%0, %1, %2, %3};" ::"r"(pk[4 * v]), "r"(pk[4 * v + 1]), "r"(pk[4 * v + 2]), "r"(pk[4 * v + 3]),
                         "r"(base + (TILE / 4 + v) * nt * 16));
    }
    const long long t1 = clock64();
    out[blockIdx.x * blockDim.x + threadIdx.x] = l;
    if ((threadIdx.x & 31) == 0) cyc[threadIdx.x >> 5] = t1 - t0;
}

template <int NUM, int DEN, bool SCALAR_FMA, int SUMMODE, int PACK = 0>
void run_phase(const char* name, float* out, long long* cyc) {
    const int iters = 200;
    auto kern = phase_kernel<NUM, DEN, SCALAR_FMA, SUMMODE, PACK>;
    cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 256 * 48 * 16);
    for (int nw : {4, 8}) {
        for (int rep = 0; rep < 2; ++rep) kern<<<1, nw * 32, nw * 32 * 48 * 16>>>(out, cyc, 0.18f, 1.0f, iters);
        long long c[8];
        cudaMemcpy(c, cyc, sizeof(long long) * nw, cudaMemcpyDeviceToHost);
        long long mx = 0;
        for (int i = 0; i < nw; ++i) mx = c[i] > mx ? c[i] : mx;
        printf("%-58s %d warp(s) per sub-partition: %7.1f cycles per row of 128 per warp%s\n", name, nw / 4, (double)mx / iters,
               nw == 8 ? "  (two rows in flight)" : "");
    }
}

// ---- issue cost of single instruction forms (8 independent chains) -----------------------------------------------------
template <int OP>
__global__ void __launch_bounds__(256, 1) rate_kernel(float* out, long long* cyc, float u, float seed) {
    float2 a[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) a[i] = make_float2(seed * (threadIdx.x + i + 1) * 1e-3f, seed * (threadIdx.x + i + 2) * 1e-3f);
    const float2 u2 = make_float2(u, u), c2 = make_float2(0.25f, 0.25f);
    uint32_t acc = 0;
    __syncthreads();
    const long long t0 = clock64();
#pragma unroll 1
    for (int it = 0; it < 256; ++it) {
#pragma unroll
        for (int rep = 0; rep < 8; ++rep)
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                if (OP == 0) a[i] = __ffma2_rn(a[i], a[(i + 1) & 7], a[(i + 2) & 7]);      // three register pairs
                if (OP == 1) a[i] = __ffma2_rn(a[i], u2, a[i]);                            // uniform multiplier
                if (OP == 2) a[i] = __fadd2_rn(a[i], a[(i + 1) & 7]);
                if (OP == 3) a[i] = __fadd2_rn(a[i], c2);                                  // immediate / constant addend
                if (OP == 4) { a[i].x = fmaf(a[i].x, u, a[i].x); }                         // scalar FFMA, uniform multiplier
                if (OP == 5) { a[i].x = fmaf(a[i].x, a[(i + 1) & 7].x, a[(i + 2) & 7].x); }// scalar FFMA, three registers
                if (OP == 6) { a[i].x = a[i].x + a[(i + 1) & 7].x; }                       // scalar FADD
                if (OP == 7) { a[i].x = ex2(a[i].x); }                                     // MUFU.EX2
                if (OP == 8) { a[i].x = __uint_as_float(pack_bf16(a[i].x, a[i].y)); }         // F2FP (chain through its own result)
                if (OP == 9) { a[i].x = __int_as_float(__float_as_int(a[i].x) + (__float_as_int(a[i].y) << 23)); }  // exponent insert
                if (OP == 10) { a[i].x = fmaxf(a[i].x, -126.f); }
                if (OP == 11) a[i] = __fmul2_rn(a[i], a[(i + 1) & 7]);
            }
    }
    const long long t1 = clock64();
    float s = 0;
#pragma unroll
    for (int i = 0; i < 8; ++i) s += a[i].x + a[i].y;
    out[blockIdx.x * blockDim.x + threadIdx.x] = s + __uint_as_float(acc);
    if ((threadIdx.x & 31) == 0) cyc[threadIdx.x >> 5] = t1 - t0;
}
template <int OP>
void run_rate(const char* name, float* out, long long* cyc) {
    for (int nw : {4, 8}) {
        for (int rep = 0; rep < 2; ++rep) rate_kernel<OP><<<1, nw * 32>>>(out, cyc, 1.0001f, 1.0f);
        long long c[8];
        cudaMemcpy(c, cyc, sizeof(long long) * nw, cudaMemcpyDeviceToHost);
        long long mx = 0;
        for (int i = 0; i < nw; ++i) mx = c[i] > mx ? c[i] : mx;
        printf("%-58s %d warp(s) per sub-partition: %5.2f cycles per warp instruction per sub-partition\n", name, nw / 4,
               (double)mx / (256.0 * 64 * (nw / 4)));
    }
}

// ---- do the MUFU and the FMA pipe overlap inside ONE warp?  NM MUFU.EX2 + NF FADD2 + NA FMNMX per group, all independent ---
template <int NM, int NF, int NA>
__global__ void __launch_bounds__(256, 1) mix_kernel(float* out, long long* cyc, float seed) {
    float m[8]; float2 f[8]; float x[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) { m[i] = -seed * (threadIdx.x + i) * 1e-3f; f[i] = make_float2(m[i], seed); x[i] = m[i] * 3.f; }
    const float2 c2 = make_float2(1e-9f, 1e-9f);
    __syncthreads();
    const long long t0 = clock64();
#pragma unroll 1
    for (int it = 0; it < 256; ++it) {
#pragma unroll
        for (int rep = 0; rep < 16; ++rep) {
            // interleave by hand the way a good schedule would: one MUFU, then its share of the other instructions
#pragma unroll
            for (int q = 0; q < NM; ++q) {
                m[(rep * NM + q) & 7] = ex2(m[(rep * NM + q) & 7]);
#pragma unroll
                for (int j = 0; j < (NF + NM - 1) / NM; ++j) if (q * ((NF + NM - 1) / NM) + j < NF) f[(rep + q + j) & 7] = __fadd2_rn(f[(rep + q + j) & 7], c2);
#pragma unroll
                for (int j = 0; j < (NA + NM - 1) / NM; ++j) if (q * ((NA + NM - 1) / NM) + j < NA) x[(rep + q + j) & 7] = fmaxf(x[(rep + q + j) & 7], m[(rep + 3) & 7]);
            }
            if (NM == 0) {
#pragma unroll
                for (int j = 0; j < NF; ++j) f[(rep + j) & 7] = __fadd2_rn(f[(rep + j) & 7], c2);
#pragma unroll
                for (int j = 0; j < NA; ++j) x[(rep + j) & 7] = fmaxf(x[(rep + j) & 7], m[(rep + 3) & 7]);
            }
        }
    }
    const long long t1 = clock64();
    float s = 0;
#pragma unroll
    for (int i = 0; i < 8; ++i) s += m[i] + f[i].x + f[i].y + x[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
    if ((threadIdx.x & 31) == 0) cyc[threadIdx.x >> 5] = t1 - t0;
}
template <int NM, int NF, int NA>
void run_mix(float* out, long long* cyc) {
    for (int nw : {4, 8}) {
        for (int rep = 0; rep < 2; ++rep) mix_kernel<NM, NF, NA><<<1, nw * 32>>>(out, cyc, 1.0f);
        long long c[8];
        cudaMemcpy(c, cyc, sizeof(long long) * nw, cudaMemcpyDeviceToHost);
        long long mx = 0;
        for (int i = 0; i < nw; ++i) mx = c[i] > mx ? c[i] : mx;
        printf("group = %d MUFU.EX2 + %d FADD2 + %d FMNMX                          %d warp(s) per sub-partition: %6.2f cycles per group per warp "
               "(pipes alone: XU %d, FMA %d)\n", NM, NF, NA, nw / 4, (double)mx / (256.0 * 16), 8 * NM, 2 * NF);
    }
}

int main() {
    float* out; long long* cyc;
    cudaMalloc(&out, 4096 * 4); cudaMalloc(&cyc, 64);
    run_rate<0>("FFMA2 r,r,r", out, cyc);
    run_rate<1>("FFMA2 r,u,r", out, cyc);
    run_rate<2>("FADD2 r,r", out, cyc);
    run_rate<3>("FADD2 r,const", out, cyc);
    run_rate<11>("FMUL2 r,r", out, cyc);
    run_rate<4>("FFMA r,u,r", out, cyc);
    run_rate<5>("FFMA r,r,r", out, cyc);
    run_rate<6>("FADD r,r", out, cyc);
    run_rate<7>("MUFU.EX2", out, cyc);
    run_rate<8>("F2FP.BF16 pack (+IADD)", out, cyc);
    run_rate<9>("exponent insert (shift + add)", out, cyc);
    run_rate<10>("FMNMX r,imm", out, cyc);
    printf("\n");
    run_mix<1, 0, 0>(out, cyc);
    run_mix<0, 4, 0>(out, cyc);
    run_mix<0, 0, 4>(out, cyc);
    run_mix<1, 1, 0>(out, cyc);
    run_mix<1, 2, 0>(out, cyc);
    run_mix<1, 3, 0>(out, cyc);
    run_mix<1, 4, 0>(out, cyc);
    run_mix<1, 6, 0>(out, cyc);
    run_mix<1, 3, 2>(out, cyc);
    run_mix<1, 4, 2>(out, cyc);
    run_mix<2, 6, 2>(out, cyc);
    run_mix<2, 8, 2>(out, cyc);
    printf("\n");
    run_phase<0, 1, false, 0>("all MUFU, FFMA2 scale, FADD2 sum x2", out, cyc);
    run_phase<0, 1, true, 0>("all MUFU, scalar FFMA scale, FADD2 sum x2", out, cyc);
    run_phase<0, 1, false, 2>("all MUFU, FFMA2 scale, no sum", out, cyc);
    run_phase<1, 3, false, 0>("1/3 poly (shipped), FFMA2 scale, FADD2 sum x2", out, cyc);
    run_phase<1, 3, false, 1>("1/3 poly, FFMA2 scale, FADD2 sum x4", out, cyc);
    run_phase<1, 3, true, 0>("1/3 poly, scalar FFMA scale, FADD2 sum x2", out, cyc);
    run_phase<1, 3, false, 2>("1/3 poly, FFMA2 scale, no sum", out, cyc);
    run_phase<1, 4, false, 0>("1/4 poly, FFMA2 scale, FADD2 sum x2", out, cyc);
    run_phase<1, 6, false, 0>("1/6 poly, FFMA2 scale, FADD2 sum x2", out, cyc);
    run_phase<1, 8, false, 0>("1/8 poly, FFMA2 scale, FADD2 sum x2", out, cyc);
    run_phase<1, 2, false, 0>("1/2 poly, FFMA2 scale, FADD2 sum x2", out, cyc);
    run_phase<1, 1, false, 0>("all poly, FFMA2 scale, FADD2 sum x2", out, cyc);
    printf("\n");
    run_phase<0, 1, false, 0, 1>("all MUFU, PRMT pack", out, cyc);
    run_phase<1, 3, false, 0, 1>("1/3 poly, PRMT pack", out, cyc);
    run_phase<1, 4, false, 0, 1>("1/4 poly, PRMT pack", out, cyc);
    run_phase<1, 2, false, 0, 1>("1/2 poly, PRMT pack", out, cyc);
    run_phase<2, 5, false, 0, 1>("2/5 poly, PRMT pack", out, cyc);
    run_phase<1, 3, false, 2, 1>("1/3 poly, PRMT pack, no sum", out, cyc);
    run_phase<1, 3, true, 0, 1>("1/3 poly, PRMT pack, scalar FFMA scale", out, cyc);
    const cudaError_t e = cudaDeviceSynchronize();
    if (e != cudaSuccess) { printf("CUDA error: %s\n", cudaGetErrorString(e)); return 1; }
    return 0;
}
