// V3: LayerNorm forward / backward over bf16 rows with fp32 statistics (HBM-bound: 4 B/element fwd, 6-8 B/element bwd).
// Replaces the ATen layer_norm behind ln_pre / ln_1 / ln_2 / ln_post / ln_final
// (avion/models/transformer.py:35-41,110,124,213,245,264).  One warp per row, 16-byte loads, row kept in registers.
// The backward also folds in the residual-stream add (x + f(x), transformer.py:148-149) and the column sum that is the
// bias gradient of the Linear feeding this residual stream.
#include "common.cuh"
#include "../../include/avion_b200.h"

namespace {
using namespace avn;

template <int VPL>   // 16-byte vectors (8 bf16) per lane; D <= VPL*256
__global__ void __launch_bounds__(256)
ln_fwd_kernel(const __nv_bfloat16* __restrict__ x, const float* __restrict__ gamma, const float* __restrict__ beta,
              __nv_bfloat16* __restrict__ y, float* __restrict__ mean_out, float* __restrict__ rstd_out, long long rows, int D, float eps) {
    const int lane = threadIdx.x & 31;
    const long long row = (long long)blockIdx.x * 8 + (threadIdx.x >> 5);
    if (row >= rows) return;
    const int nvec = D >> 3;
    const uint4* xr = (const uint4*)(x + row * D);
    float v[VPL][8];
    float s = 0.f;
#pragma unroll
    for (int j = 0; j < VPL; ++j) {
        const int i = lane + 32 * j;
        if (i < nvec) {
            uint4 u = __ldg(xr + i);
            const uint32_t w[4] = {u.x, u.y, u.z, u.w};
#pragma unroll
            for (int e = 0; e < 4; ++e) { float2 f = unpack_bf16(w[e]); v[j][2 * e] = f.x; v[j][2 * e + 1] = f.y; s += f.x + f.y; }
        } else {
#pragma unroll
            for (int e = 0; e < 8; ++e) v[j][e] = 0.f;
        }
    }
    const float mean = warp_sum(s) / D;
    float q = 0.f;
#pragma unroll
    for (int j = 0; j < VPL; ++j)
        if (lane + 32 * j < nvec) {
#pragma unroll
            for (int e = 0; e < 8; ++e) { float d = v[j][e] - mean; q += d * d; }
        }
    const float rstd = rsqrtf(warp_sum(q) / D + eps);
    if (lane == 0) {
        if (mean_out) mean_out[row] = mean;
        if (rstd_out) rstd_out[row] = rstd;
    }
    uint4* yr = (uint4*)(y + row * D);
#pragma unroll
    for (int j = 0; j < VPL; ++j) {
        const int i = lane + 32 * j;
        if (i < nvec) {
            const float4 g0 = __ldg((const float4*)(gamma + i * 8)), g1 = __ldg((const float4*)(gamma + i * 8 + 4));
            const float4 b0 = __ldg((const float4*)(beta + i * 8)), b1 = __ldg((const float4*)(beta + i * 8 + 4));
            const float g[8] = {g0.x, g0.y, g0.z, g0.w, g1.x, g1.y, g1.z, g1.w};
            const float bb[8] = {b0.x, b0.y, b0.z, b0.w, b1.x, b1.y, b1.z, b1.w};
            float o[8];
#pragma unroll
            for (int e = 0; e < 8; ++e) o[e] = fmaf((v[j][e] - mean) * rstd, g[e], bb[e]);
            yr[i] = make_uint4(pack_bf16(o[0], o[1]), pack_bf16(o[2], o[3]), pack_bf16(o[4], o[5]), pack_bf16(o[6], o[7]));
        }
    }
}

template <int VPL>
__global__ void __launch_bounds__(256, 2)
ln_bwd_kernel(const __nv_bfloat16* __restrict__ dy, const __nv_bfloat16* __restrict__ x, const float* __restrict__ gamma,
              const float* __restrict__ mean_in, const float* __restrict__ rstd_in, const __nv_bfloat16* __restrict__ dres,
              __nv_bfloat16* __restrict__ dx, float* __restrict__ dgamma, float* __restrict__ dbeta, float* __restrict__ dcolsum,
              long long rows, int D) {
    extern __shared__ float red[];       // [8 warps][3][VPL*256]
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int nvec = D >> 3;
    // per-lane column accumulators (dgamma, dbeta, column sum of dx); x / dy stay packed (bf16x2) to keep the register
    // count low enough for two CTAs per SM - this kernel lives on memory-level parallelism
    float ag[VPL][8], ab[VPL][8], ac[VPL][8];
#pragma unroll
    for (int j = 0; j < VPL; ++j)
#pragma unroll
        for (int e = 0; e < 8; ++e) ag[j][e] = ab[j][e] = ac[j][e] = 0.f;
    for (long long row = (long long)blockIdx.x * 8 + warp; row < rows; row += (long long)gridDim.x * 8) {
        const float mean = __ldg(mean_in + row), rstd = __ldg(rstd_in + row);
        const uint4* xr = (const uint4*)(x + row * D);
        const uint4* dyr = (const uint4*)(dy + row * D);
        const uint4* drr = dres ? (const uint4*)(dres + row * D) : nullptr;
        uint4 ux[VPL], ud[VPL], ur[VPL];
#pragma unroll
        for (int j = 0; j < VPL; ++j) {
            const int i = lane + 32 * j;
            ux[j] = ud[j] = ur[j] = make_uint4(0, 0, 0, 0);
            if (i < nvec) {
                ux[j] = __ldg(xr + i);
                ud[j] = __ldg(dyr + i);
                if (drr) ur[j] = __ldg(drr + i);
            }
        }
        // all the per-element arithmetic is packed f32x2 (FFMA2 / FADD2 / FMUL2): this kernel is as much issue-bound as
        // memory-bound (~800 scalar instructions per row per warp before)
        const float2 rs2 = make_float2(rstd, rstd), nmr2 = make_float2(-mean * rstd, -mean * rstd);
        float2 s1v = make_float2(0.f, 0.f), s2v = make_float2(0.f, 0.f);
#pragma unroll
        for (int j = 0; j < VPL; ++j) {
            const int i = lane + 32 * j;
            if (i < nvec) {
                const uint32_t wx[4] = {ux[j].x, ux[j].y, ux[j].z, ux[j].w}, wd[4] = {ud[j].x, ud[j].y, ud[j].z, ud[j].w};
                const float4 g0 = __ldg((const float4*)(gamma + i * 8)), g1 = __ldg((const float4*)(gamma + i * 8 + 4));
                const float2 g[4] = {make_float2(g0.x, g0.y), make_float2(g0.z, g0.w), make_float2(g1.x, g1.y), make_float2(g1.z, g1.w)};
#pragma unroll
                for (int e = 0; e < 4; ++e) {
                    const float2 fd = unpack_bf16(wd[e]);
                    const float2 h = __ffma2_rn(unpack_bf16(wx[e]), rs2, nmr2);       // (x - mean) * rstd
                    float2 a = make_float2(ag[j][2 * e], ag[j][2 * e + 1]), bsum = make_float2(ab[j][2 * e], ab[j][2 * e + 1]);
                    a = __ffma2_rn(fd, h, a);
                    bsum = __fadd2_rn(bsum, fd);
                    ag[j][2 * e] = a.x; ag[j][2 * e + 1] = a.y; ab[j][2 * e] = bsum.x; ab[j][2 * e + 1] = bsum.y;
                    const float2 t = __fmul2_rn(fd, g[e]);
                    s1v = __fadd2_rn(s1v, t);
                    s2v = __ffma2_rn(t, h, s2v);
                }
            }
        }
        const float c1 = warp_sum(s1v.x + s1v.y) / D, c2 = warp_sum(s2v.x + s2v.y) / D;
        const float2 nc1 = make_float2(-c1, -c1), nc2 = make_float2(-c2, -c2);
        uint4* dxr = (uint4*)(dx + row * D);
#pragma unroll
        for (int j = 0; j < VPL; ++j) {
            const int i = lane + 32 * j;
            if (i < nvec) {
                const uint32_t wx[4] = {ux[j].x, ux[j].y, ux[j].z, ux[j].w}, wd[4] = {ud[j].x, ud[j].y, ud[j].z, ud[j].w};
                const uint32_t wr[4] = {ur[j].x, ur[j].y, ur[j].z, ur[j].w};
                const float4 g0 = __ldg((const float4*)(gamma + i * 8)), g1 = __ldg((const float4*)(gamma + i * 8 + 4));
                const float2 g[4] = {make_float2(g0.x, g0.y), make_float2(g0.z, g0.w), make_float2(g1.x, g1.y), make_float2(g1.z, g1.w)};
                uint32_t po[4];
#pragma unroll
                for (int e = 0; e < 4; ++e) {
                    const float2 h = __ffma2_rn(unpack_bf16(wx[e]), rs2, nmr2);
                    float2 u = __ffma2_rn(unpack_bf16(wd[e]), g[e], nc1);            // dy*g - c1
                    u = __ffma2_rn(h, nc2, u);                                          //       - h*c2
                    const float2 o = __ffma2_rn(u, rs2, unpack_bf16(wr[e]));           // * rstd + residual gradient
                    po[e] = pack_bf16(o.x, o.y);
                    if (dcolsum) {   // column sum of the values actually stored (bf16-rounded), as the next consumer sees them
                        const float2 c = __fadd2_rn(make_float2(ac[j][2 * e], ac[j][2 * e + 1]), unpack_bf16(po[e]));
                        ac[j][2 * e] = c.x; ac[j][2 * e + 1] = c.y;
                    }
                }
                dxr[i] = make_uint4(po[0], po[1], po[2], po[3]);
            }
        }
    }
    // block reduction of the three column accumulators, then one atomic per column per block
    const int W = VPL * 256;
#pragma unroll
    for (int j = 0; j < VPL; ++j)
#pragma unroll
        for (int e = 0; e < 8; ++e) {
            const int c = (lane + 32 * j) * 8 + e;
            red[(warp * 3 + 0) * W + c] = ag[j][e];
            red[(warp * 3 + 1) * W + c] = ab[j][e];
            red[(warp * 3 + 2) * W + c] = ac[j][e];
        }
    __syncthreads();
    for (int c = threadIdx.x; c < D; c += blockDim.x) {
        float a = 0.f, b = 0.f, cc = 0.f;
#pragma unroll
        for (int w = 0; w < 8; ++w) { a += red[(w * 3 + 0) * W + c]; b += red[(w * 3 + 1) * W + c]; cc += red[(w * 3 + 2) * W + c]; }
        if (dgamma) atomicAdd(dgamma + c, a);
        if (dbeta) atomicAdd(dbeta + c, b);
        if (dcolsum) atomicAdd(dcolsum + c, cc);
    }
}

template <int VPL>
int launch_bwd(const void* dy, const void* x, const float* gamma, const float* mean, const float* rstd, const void* dres, void* dx,
               float* dgamma, float* dbeta, float* dcolsum, int64_t rows, int D, cudaStream_t stream) {
    const size_t smem = (size_t)8 * 3 * VPL * 256 * sizeof(float);
    auto kern = ln_bwd_kernel<VPL>;
    AVN_CHECK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    long long blocks = (rows + 7) / 8;
    const long long cap = (long long)avn_num_sms() * 2;      // two resident CTAs per SM, each warp loops over rows
    if (blocks > cap) blocks = cap;
    kern<<<(int)blocks, 256, smem, stream>>>((const __nv_bfloat16*)dy, (const __nv_bfloat16*)x, gamma, mean, rstd,
                                             (const __nv_bfloat16*)dres, (__nv_bfloat16*)dx, dgamma, dbeta, dcolsum, rows, D);
    AVN_CHECK_CUDA(cudaGetLastError());
    return AVN_OK;
}
}  // namespace

extern "C" int avn_layernorm_fwd(const void* x, const float* gamma, const float* beta, void* y, float* mean, float* rstd,
                                 int64_t rows, int D, float eps, avn_stream_t stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    AVN_REQUIRE(x && gamma && beta && y, "layernorm_fwd: null pointer");
    AVN_REQUIRE(rows > 0 && D > 0 && D % 8 == 0 && D <= 1024, "layernorm_fwd: D=%d must be a multiple of 8 and <= 1024", D);
    const int blocks = (int)((rows + 7) / 8);
    const __nv_bfloat16* xp = (const __nv_bfloat16*)x;
    __nv_bfloat16* yp = (__nv_bfloat16*)y;
    if (D <= 256) ln_fwd_kernel<1><<<blocks, 256, 0, stream>>>(xp, gamma, beta, yp, mean, rstd, rows, D, eps);
    else if (D <= 512) ln_fwd_kernel<2><<<blocks, 256, 0, stream>>>(xp, gamma, beta, yp, mean, rstd, rows, D, eps);
    else if (D <= 768) ln_fwd_kernel<3><<<blocks, 256, 0, stream>>>(xp, gamma, beta, yp, mean, rstd, rows, D, eps);
    else ln_fwd_kernel<4><<<blocks, 256, 0, stream>>>(xp, gamma, beta, yp, mean, rstd, rows, D, eps);
    AVN_CHECK_CUDA(cudaGetLastError());
    return AVN_OK;
}

extern "C" int avn_layernorm_bwd(const void* dy, const void* x, const float* gamma, const float* mean, const float* rstd,
                                 const void* dres, void* dx, float* dgamma, float* dbeta, float* dcolsum, int64_t rows, int D,
                                 avn_stream_t stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    AVN_REQUIRE(dy && x && gamma && mean && rstd && dx, "layernorm_bwd: null pointer");
    AVN_REQUIRE(rows > 0 && D > 0 && D % 8 == 0 && D <= 1024, "layernorm_bwd: D=%d must be a multiple of 8 and <= 1024", D);
    if (D <= 256) return launch_bwd<1>(dy, x, gamma, mean, rstd, dres, dx, dgamma, dbeta, dcolsum, rows, D, stream);
    if (D <= 512) return launch_bwd<2>(dy, x, gamma, mean, rstd, dres, dx, dgamma, dbeta, dcolsum, rows, D, stream);
    if (D <= 768) return launch_bwd<3>(dy, x, gamma, mean, rstd, dres, dx, dgamma, dbeta, dcolsum, rows, D, stream);
    return launch_bwd<4>(dy, x, gamma, mean, rstd, dres, dx, dgamma, dbeta, dcolsum, rows, D, stream);
}
