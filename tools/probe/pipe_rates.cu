// Probe: issue cost (cycles per warp instruction per SM sub-partition) of the instructions the attention softmax leans on.
// nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/probe/pipe_rates tools/probe/pipe_rates.cu ; run on a B200.
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
#include <cuda_bf16.h>
template <int OP>
__global__ void k(float* out, long long* cyc, float seed) {
    float a[8], b[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) { a[i] = seed * (threadIdx.x + i + 1) * 1e-3f; b[i] = a[i] * 0.5f; }
    uint32_t acc = 0;
    __syncthreads();
    const long long t0 = clock64();
#pragma unroll 1
    for (int it = 0; it < 256; ++it) {
#pragma unroll
      for (int rep = 0; rep < 8; ++rep)
#pragma unroll
        for (int i = 0; i < 8; i += 2) {
            if (OP == 0) {          // MUFU.EX2
                asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+f"(a[i]));
                asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+f"(a[i + 1]));
            } else if (OP == 1) {   // FFMA2
                asm volatile("{.reg .b64 x, y; mov.b64 x, {%0, %1}; fma.rn.f32x2 y, x, x, x; mov.b64 {%0, %1}, y;}" : "+f"(a[i]), "+f"(a[i + 1]));
                asm volatile("{.reg .b64 x, y; mov.b64 x, {%0, %1}; fma.rn.f32x2 y, x, x, x; mov.b64 {%0, %1}, y;}" : "+f"(a[i]), "+f"(a[i + 1]));
            } else if (OP == 2) {   // F2FP pack
                uint32_t r;
                asm volatile("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(a[i]), "f"(a[i + 1]));
                acc ^= r;
                asm volatile("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(a[i + 1]), "f"(a[i]));
                acc += r;
            } else if (OP == 3) {   // FFMA
                asm volatile("fma.rn.f32 %0, %0, %0, %0;" : "+f"(a[i]));
                asm volatile("fma.rn.f32 %0, %0, %0, %0;" : "+f"(a[i + 1]));
            } else if (OP == 4) {   // FMNMX
                asm volatile("max.f32 %0, %0, %1;" : "+f"(a[i]) : "f"(a[i + 1]));
                asm volatile("max.f32 %0, %0, %1;" : "+f"(a[i + 1]) : "f"(a[i]));
            } else if (OP == 5) {   // 3-input max
                asm volatile("max.f32 %0, %0, %1, %2;" : "+f"(a[i]) : "f"(a[i + 1]), "f"(a[(i + 2) & 7]));
                asm volatile("max.f32 %0, %0, %1, %2;" : "+f"(a[i + 1]) : "f"(a[i]), "f"(a[(i + 3) & 7]));
            } else if (OP == 6) {   // mixed: per element pair 1 FFMA2 + 2 MUFU + 1 FFMA2 + 1 FMUL2 + 2 F2FP (attention bwd inner loop)
                float x0 = a[i], x1 = a[i + 1];
                asm volatile("{.reg .b64 x, y; mov.b64 x, {%0, %1}; fma.rn.f32x2 y, x, x, x; mov.b64 {%0, %1}, y;}" : "+f"(x0), "+f"(x1));
                asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+f"(x0));
                asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+f"(x1));
                float y0 = a[(i + 2) & 7], y1 = a[(i + 3) & 7];
                asm volatile("{.reg .b64 x, y; mov.b64 x, {%0, %1}; fma.rn.f32x2 y, x, x, x; mov.b64 {%0, %1}, y;}" : "+f"(y0), "+f"(y1));
                asm volatile("{.reg .b64 x, y, z; mov.b64 x, {%0, %1}; mov.b64 y, {%2, %3}; mul.rn.f32x2 z, x, y; mov.b64 {%0, %1}, z;}" : "+f"(y0), "+f"(y1) : "f"(x0), "f"(x1));
                uint32_t r;
                asm volatile("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(x0), "f"(x1));
                acc ^= r;
                asm volatile("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(y0), "f"(y1));
                acc += r;
                a[i] = x0; a[i + 1] = x1;
            } else if (OP == 7) {   // attention fwd inner loop: per element pair 1 FFMA2 + 2 MUFU + 1 FADD2 + 1 F2FP
                float x0 = a[i], x1 = a[i + 1];
                asm volatile("{.reg .b64 x, y; mov.b64 x, {%0, %1}; fma.rn.f32x2 y, x, x, x; mov.b64 {%0, %1}, y;}" : "+f"(x0), "+f"(x1));
                asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+f"(x0));
                asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+f"(x1));
                asm volatile("{.reg .b64 x, y; mov.b64 x, {%0, %1}; mov.b64 y, {%2, %3}; add.rn.f32x2 x, x, y; mov.b64 {%0, %1}, x;}" : "+f"(a[i]), "+f"(a[i + 1]) : "f"(x0), "f"(x1));
                uint32_t r;
                asm volatile("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(x0), "f"(x1));
                acc ^= r;
            } else if (OP == 8) {   // 2 MUFU + 3 independent FFMA2 (no data dependence between the two pipes)
                asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+f"(a[i]));
                asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+f"(a[i + 1]));
                asm volatile("{.reg .b64 x, y; mov.b64 x, {%0, %1}; fma.rn.f32x2 y, x, x, x; mov.b64 {%0, %1}, y;}" : "+f"(b[i]), "+f"(b[i + 1]));
                asm volatile("{.reg .b64 x, y; mov.b64 x, {%0, %1}; fma.rn.f32x2 y, x, x, x; mov.b64 {%0, %1}, y;}" : "+f"(b[(i + 2) & 7]), "+f"(b[(i + 3) & 7]));
                asm volatile("{.reg .b64 x, y; mov.b64 x, {%0, %1}; fma.rn.f32x2 y, x, x, x; mov.b64 {%0, %1}, y;}" : "+f"(b[(i + 4) & 7]), "+f"(b[(i + 5) & 7]));
            } else if (OP == 9) {   // 2 MUFU + 6 independent FFMA
                asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+f"(a[i]));
                asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+f"(a[i + 1]));
#pragma unroll
                for (int q = 0; q < 6; ++q) asm volatile("fma.rn.f32 %0, %0, %0, %0;" : "+f"(b[(i + q) & 7]));
            } else if (OP == 10) {  // polynomial exp2 of a pair on the FMA pipe: round (FADD2 magic), fraction (FADD2), 3 FFMA2, 2 integer shift-adds
                float x0 = a[i], x1 = a[i + 1], r0, r1, f0, f1;
                asm volatile("{.reg .b64 x, y, m; mov.b64 x, {%2, %3}; mov.b64 m, {%4, %4}; add.rn.f32x2 y, x, m; mov.b64 {%0, %1}, y;}" : "=f"(r0), "=f"(r1) : "f"(x0), "f"(x1), "f"(12582912.f));
                asm volatile("{.reg .b64 x, y, m; mov.b64 x, {%2, %3}; mov.b64 m, {%4, %4}; add.rn.f32x2 y, x, m; mov.b64 {%0, %1}, y;}" : "=f"(f0), "=f"(f1) : "f"(r0), "f"(r1), "f"(-12582912.f));
                asm volatile("{.reg .b64 x, y, m; mov.b64 x, {%0, %1}; mov.b64 y, {%2, %3}; sub.rn.f32x2 x, y, x; mov.b64 {%0, %1}, x;}" : "+f"(f0), "+f"(f1) : "f"(x0), "f"(x1));
                float p0 = 0.0555f, p1 = 0.0555f;
                asm volatile("{.reg .b64 x, y, c; mov.b64 x, {%0, %1}; mov.b64 y, {%2, %3}; mov.b64 c, {%4, %4}; fma.rn.f32x2 x, x, y, c; mov.b64 {%0, %1}, x;}" : "+f"(p0), "+f"(p1) : "f"(f0), "f"(f1), "f"(0.2402f));
                asm volatile("{.reg .b64 x, y, c; mov.b64 x, {%0, %1}; mov.b64 y, {%2, %3}; mov.b64 c, {%4, %4}; fma.rn.f32x2 x, x, y, c; mov.b64 {%0, %1}, x;}" : "+f"(p0), "+f"(p1) : "f"(f0), "f"(f1), "f"(0.6931f));
                asm volatile("{.reg .b64 x, y, c; mov.b64 x, {%0, %1}; mov.b64 y, {%2, %3}; mov.b64 c, {%4, %4}; fma.rn.f32x2 x, x, y, c; mov.b64 {%0, %1}, x;}" : "+f"(p0), "+f"(p1) : "f"(f0), "f"(f1), "f"(1.0f));
                a[i] = __int_as_float(__float_as_int(p0) + (__float_as_int(r0) << 23));
                a[i + 1] = __int_as_float(__float_as_int(p1) + (__float_as_int(r1) << 23));
            }
        }
    }
    const long long t1 = clock64();
    float s = 0;
#pragma unroll
    for (int i = 0; i < 8; ++i) s += a[i] + b[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s + __uint_as_float(acc);
    if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
}
template <int OP>
void run(const char* name, int per_iter) {
    float* out; long long* cyc;
    cudaMalloc(&out, 1024 * 4); cudaMalloc(&cyc, 8);
    for (int nw : {4, 8, 16}) {
        k<OP><<<1, nw * 32>>>(out, cyc, 1.0f);
        k<OP><<<1, nw * 32>>>(out, cyc, 1.0f);
        long long c; cudaMemcpy(&c, cyc, 8, cudaMemcpyDeviceToHost);
        const double n_inst = 256.0 * 8 * per_iter * (nw / 4);      // warp instructions per sub-partition
        printf("%-28s %2d warps/SM (%d per sub-partition): %8lld cycles, %.2f cycles per warp-instruction per sub-partition\n", name, nw,
               nw / 4, c, c / n_inst);
    }
    cudaFree(out); cudaFree(cyc);
}
int main() {
    run<0>("MUFU.EX2", 8);
    run<1>("FFMA2 (fma.rn.f32x2)", 8);
    run<2>("F2FP (cvt.rn.bf16x2.f32)", 8);
    run<3>("FFMA", 8);
    run<4>("FMNMX", 8);
    run<5>("FMNMX3 (max.f32 a,b,c)", 8);
    run<6>("bwd mix (per 2 pairs: 28 inst)", 4 * 7);
    run<7>("fwd mix (FFMA2 2MUFU FADD2 F2FP)", 4 * 5);
    run<8>("2 MUFU + 3 indep FFMA2", 4 * 5);
    run<9>("2 MUFU + 6 indep FFMA", 4 * 8);
    run<10>("poly exp2 pair (~9 inst)", 4 * 9);
    return cudaDeviceSynchronize() != cudaSuccess;
}
