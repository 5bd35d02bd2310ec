// V7 + C1: pooled-token head = gather one row per sample, LayerNorm, @ proj[D][E], L2-normalise.
// Replaces x[:,0] -> ln_post -> @ image_projection (avion/models/transformer.py:368-376), the EOT gather ->
// @ text_projection (:472-476; ln_final is row-wise so applying it to the gathered row only is identical) and
// F.normalize (avion/models/model_clip.py:95-98).  n rows only (one CTA per sample): latency-, not bandwidth-bound.
#include "common.cuh"
#include "../../include/avion_b200.h"

namespace {
using namespace avn;

__device__ __forceinline__ float block_sum(float v, float* red) {
    v = warp_sum(v);
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    __syncthreads();
    if (lane == 0) red[warp] = v;
    __syncthreads();
    float s = 0.f;
    for (int w = 0; w < (int)(blockDim.x >> 5); ++w) s += red[w];
    return s;
}

__global__ void __launch_bounds__(256)
head_fwd_kernel(const __nv_bfloat16* __restrict__ x, const int64_t* __restrict__ row_idx, const float* __restrict__ gamma,
                const float* __restrict__ beta, const float* __restrict__ proj, float* __restrict__ emb, float* __restrict__ u_out,
                float* __restrict__ mean_out, float* __restrict__ rstd_out, int D, int E, float eps) {
    extern __shared__ float sm[];
    float* xn = sm;            // [D]
    float* red = sm + D;       // [8]
    const int i = blockIdx.x;
    const __nv_bfloat16* xr = x + row_idx[i] * (long long)D;
    float s = 0.f;
    for (int d = threadIdx.x; d < D; d += blockDim.x) { const float v = __bfloat162float(xr[d]); xn[d] = v; s += v; }
    const float mean = block_sum(s, red) / D;
    float q = 0.f;
    for (int d = threadIdx.x; d < D; d += blockDim.x) { const float t = xn[d] - mean; q += t * t; }
    const float rstd = rsqrtf(block_sum(q, red) / D + eps);
    for (int d = threadIdx.x; d < D; d += blockDim.x) xn[d] = fmaf((xn[d] - mean) * rstd, gamma[d], beta[d]);
    __syncthreads();
    float nrm = 0.f;
    float uv[4];
    int cnt = 0;
    for (int e = threadIdx.x; e < E; e += blockDim.x, ++cnt) {
        float a = 0.f;
        for (int d = 0; d < D; ++d) a = fmaf(xn[d], __ldg(proj + (long long)d * E + e), a);
        uv[cnt & 3] = a;
        u_out[(long long)i * E + e] = a;
        nrm += a * a;
    }
    const float norm = fmaxf(sqrtf(block_sum(nrm, red)), 1e-12f);     // F.normalize eps
    cnt = 0;
    for (int e = threadIdx.x; e < E; e += blockDim.x, ++cnt) emb[(long long)i * E + e] = uv[cnt & 3] / norm;
    if (threadIdx.x == 0) { mean_out[i] = mean; rstd_out[i] = rstd; }
}

// per sample: du = (demb - emb*(emb.demb))/norm ; dy[d] = sum_e proj[d][e] du[e] ; LN backward on the row.
__global__ void __launch_bounds__(256)
head_bwd_row_kernel(const float* __restrict__ demb, const __nv_bfloat16* __restrict__ x, const int64_t* __restrict__ row_idx,
                    const float* __restrict__ gamma, const float* __restrict__ proj, const float* __restrict__ u,
                    const float* __restrict__ mean_in, const float* __restrict__ rstd_in, __nv_bfloat16* __restrict__ dx,
                    float* __restrict__ dgamma, float* __restrict__ dbeta, float* __restrict__ ws, int D, int E, int normalized) {
    extern __shared__ float sm[];
    float* du = sm;            // [E]
    float* dyv = sm + E;       // [D]
    float* red = sm + E + D;   // [8]
    const int i = blockIdx.x;
    float* xhat_ws = ws + (long long)i * (D + E);     // normalised (pre-affine) row, for the dproj/dgamma kernels
    float* du_ws = xhat_ws + D;
    float n2 = 0.f, dot = 0.f;
    for (int e = threadIdx.x; e < E; e += blockDim.x) {
        const float uu = u[(long long)i * E + e];
        n2 += uu * uu; dot += uu * demb[(long long)i * E + e];
    }
    n2 = block_sum(n2, red);
    dot = block_sum(dot, red);
    const float norm = fmaxf(sqrtf(n2), 1e-12f);
    for (int e = threadIdx.x; e < E; e += blockDim.x) {
        const float uu = u[(long long)i * E + e];
        const float v = normalized ? (demb[(long long)i * E + e] - uu * dot / (norm * norm)) / norm : demb[(long long)i * E + e];
        du[e] = v; du_ws[e] = v;
    }
    __syncthreads();
    const float mean = mean_in[i], rstd = rstd_in[i];
    const long long r = row_idx[i];
    float s1 = 0.f, s2 = 0.f;
    for (int d = threadIdx.x; d < D; d += blockDim.x) {
        float dy = 0.f;
        const float* pr = proj + (long long)d * E;
        for (int e = 0; e < E; ++e) dy = fmaf(__ldg(pr + e), du[e], dy);
        const float xh = (__bfloat162float(x[r * D + d]) - mean) * rstd;
        xhat_ws[d] = xh;
        atomicAdd(dgamma + d, dy * xh);
        atomicAdd(dbeta + d, dy);
        const float g = dy * gamma[d];
        dyv[d] = g;
        s1 += g; s2 += g * xh;
    }
    s1 = block_sum(s1, red) / D;
    s2 = block_sum(s2, red) / D;
    for (int d = threadIdx.x; d < D; d += blockDim.x)
        dx[r * D + d] = __float2bfloat16_rn(rstd * (dyv[d] - s1 - xhat_ws[d] * s2));
}

// dproj[d][e] += sum_i y[i][d] * du[i][e], y = xhat*gamma + beta (the LN output)
__global__ void __launch_bounds__(256)
head_bwd_proj_kernel(const float* __restrict__ ws, const float* __restrict__ gamma, const float* __restrict__ beta,
                     float* __restrict__ dproj, int n, int D, int E) {
    const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= (long long)D * E) return;
    const int d = (int)(idx / E), e = (int)(idx % E);
    const float g = gamma[d], b = beta[d];
    float a = 0.f;
    for (int i = 0; i < n; ++i) {
        const float* w = ws + (long long)i * (D + E);
        a = fmaf(fmaf(w[d], g, b), w[D + e], a);
    }
    dproj[idx] += a;
}
}  // namespace

extern "C" int avn_head_fwd(const void* x, const int64_t* row_idx, const float* gamma, const float* beta, const float* proj,
                            float* emb, float* u, float* mean, float* rstd, int n, int D, int E, float eps, avn_stream_t stream_) {
    AVN_REQUIRE(x && row_idx && gamma && beta && proj && emb && u && mean && rstd, "head_fwd: null pointer");
    AVN_REQUIRE(n > 0 && D > 0 && E > 0 && E <= 1024, "head_fwd: bad sizes n=%d D=%d E=%d", n, D, E);
    const size_t smem = (size_t)(D + 8) * sizeof(float);
    head_fwd_kernel<<<n, 256, smem, (cudaStream_t)stream_>>>((const __nv_bfloat16*)x, row_idx, gamma, beta, proj, emb, u, mean, rstd, D, E, eps);
    AVN_CHECK_CUDA(cudaGetLastError());
    return AVN_OK;
}

extern "C" int avn_head_bwd(const float* demb, const void* x, const int64_t* row_idx, const float* gamma, const float* beta,
                            const float* proj, const float* u, const float* mean, const float* rstd, void* dx, float* dgamma,
                            float* dbeta, float* dproj, float* ws, int n, int D, int E, int normalized, avn_stream_t stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    AVN_REQUIRE(demb && x && row_idx && gamma && beta && proj && u && mean && rstd && dx && dgamma && dbeta && dproj && ws, "head_bwd: null pointer");
    AVN_REQUIRE(n > 0 && D > 0 && E > 0, "head_bwd: bad sizes");
    const size_t smem = (size_t)(E + D + 8) * sizeof(float);
    head_bwd_row_kernel<<<n, 256, smem, stream>>>(demb, (const __nv_bfloat16*)x, row_idx, gamma, proj, u, mean, rstd,
                                                  (__nv_bfloat16*)dx, dgamma, dbeta, ws, D, E, normalized);
    AVN_CHECK_CUDA(cudaGetLastError());
    const long long tot = (long long)D * E;
    head_bwd_proj_kernel<<<(int)((tot + 255) / 256), 256, 0, stream>>>(ws, gamma, beta, dproj, n, D, E);
    AVN_CHECK_CUDA(cudaGetLastError());
    return AVN_OK;
}
