// O1: multi-tensor AdamW step over fp32 master parameters (HBM-bound: 16 B read + 12 B written per parameter).
// Replaces torch.optim.AdamW as configured by scripts/main_lavila_pretrain.py:296-341 (two parameter groups: weight decay /
// no weight decay) and the logit_scale clamp of :880.  One launch updates every tensor of every group: the host passes a
// chunk table (tensor index, element offset) and per-tensor pointer / decay tables that live on the device.
//   p <- p * (1 - lr*wd);  m <- b1 m + (1-b1) g;  v <- b2 v + (1-b2) g^2;  p <- p - (lr / (1-b1^t)) * m / (sqrt(v)/sqrt(1-b2^t) + eps)
// (the decoupled form of torch.optim.AdamW, same operation order in fp32).
#include "common.cuh"
#include "../../include/avion_b200.h"

namespace {

constexpr int CHUNK = 16384;      // elements per CTA

struct AdamArgs {
    float lr, beta1, beta2, eps, bc1, bc2_sqrt;      // bc1 = 1 - beta1^t, bc2_sqrt = sqrt(1 - beta2^t)
};

__device__ __forceinline__ void adam1(float& p, float g, float& m, float& v, float decay, const AdamArgs& a) {
    p = p * decay;
    m = m + (g - m) * (1.f - a.beta1);                 // torch: exp_avg.lerp_(grad, 1 - beta1)
    v = v * a.beta2 + (1.f - a.beta2) * g * g;         // torch: exp_avg_sq.mul_(beta2).addcmul_(grad, grad, value=1-beta2)
    const float denom = sqrtf(v) / a.bc2_sqrt + a.eps;
    p = p - (a.lr / a.bc1) * (m / denom);
}

__global__ void __launch_bounds__(256)
adamw_kernel(float* const* __restrict__ params, const float* const* __restrict__ grads, float* const* __restrict__ exp_avg,
             float* const* __restrict__ exp_avg_sq, const long long* __restrict__ sizes, const float* __restrict__ weight_decay,
             const int* __restrict__ chunk_tensor, const long long* __restrict__ chunk_offset, AdamArgs a) {
    const int t = chunk_tensor[blockIdx.x];
    const long long off = chunk_offset[blockIdx.x];
    const long long n = min((long long)CHUNK, sizes[t] - off);
    float* p = params[t] + off;
    const float* g = grads[t] + off;
    float* m = exp_avg[t] + off;
    float* v = exp_avg_sq[t] + off;
    const float decay = 1.f - a.lr * weight_decay[t];
    const bool vec = ((((uintptr_t)p | (uintptr_t)g | (uintptr_t)m | (uintptr_t)v) & 15) == 0);
    if (vec) {
        const long long n4 = n >> 2;
        for (long long i = threadIdx.x; i < n4; i += blockDim.x) {
            float4 pp = ((float4*)p)[i], mm = ((float4*)m)[i], vv = ((float4*)v)[i];
            const float4 gg = ((const float4*)g)[i];
            adam1(pp.x, gg.x, mm.x, vv.x, decay, a); adam1(pp.y, gg.y, mm.y, vv.y, decay, a);
            adam1(pp.z, gg.z, mm.z, vv.z, decay, a); adam1(pp.w, gg.w, mm.w, vv.w, decay, a);
            ((float4*)p)[i] = pp; ((float4*)m)[i] = mm; ((float4*)v)[i] = vv;
        }
        for (long long i = (n4 << 2) + threadIdx.x; i < n; i += blockDim.x) adam1(p[i], g[i], m[i], v[i], decay, a);
    } else {
        for (long long i = threadIdx.x; i < n; i += blockDim.x) adam1(p[i], g[i], m[i], v[i], decay, a);
    }
}

__global__ void clamp_scalar_kernel(float* x, float lo, float hi) { *x = fminf(fmaxf(*x, lo), hi); }
}  // namespace

extern "C" int avn_adamw_chunk_elems(void) { return CHUNK; }

extern "C" int avn_adamw_step(void* const* params, const void* const* grads, void* const* exp_avg, void* const* exp_avg_sq,
                              const int64_t* sizes, const float* weight_decay, const int32_t* chunk_tensor,
                              const int64_t* chunk_offset, int n_chunks, float lr, float beta1, float beta2, float eps, int step,
                              avn_stream_t stream_) {
    AVN_REQUIRE(params && grads && exp_avg && exp_avg_sq && sizes && weight_decay && chunk_tensor && chunk_offset,
                "adamw_step: null table");
    AVN_REQUIRE(n_chunks > 0 && step > 0, "adamw_step: n_chunks=%d step=%d", n_chunks, step);
    AdamArgs a;
    a.lr = lr; a.beta1 = beta1; a.beta2 = beta2; a.eps = eps;
    a.bc1 = (float)(1.0 - pow((double)beta1, (double)step));
    a.bc2_sqrt = (float)sqrt(1.0 - pow((double)beta2, (double)step));
    adamw_kernel<<<n_chunks, 256, 0, (cudaStream_t)stream_>>>((float* const*)params, (const float* const*)grads, (float* const*)exp_avg,
                                                              (float* const*)exp_avg_sq, (const long long*)sizes, weight_decay,
                                                              chunk_tensor, (const long long*)chunk_offset, a);
    AVN_CHECK_CUDA(cudaGetLastError());
    return AVN_OK;
}

extern "C" int avn_clamp_scalar_f32(float* x, float lo, float hi, avn_stream_t stream_) {
    AVN_REQUIRE(x != nullptr, "clamp_scalar: null pointer");
    clamp_scalar_kernel<<<1, 1, 0, (cudaStream_t)stream_>>>(x, lo, hi);
    AVN_CHECK_CUDA(cudaGetLastError());
    return AVN_OK;
}
