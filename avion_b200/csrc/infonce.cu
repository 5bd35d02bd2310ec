// L2: sharded InfoNCE (ClipLoss, avion/losses/losses.py:103-149).  Each rank owns n local rows of both logit matrices
// (image->text and text->image) against all N gathered columns: the reference's local_loss branch (:116-118,132-133),
// whose parameter gradients equal the default full-matrix mode once the scalar is all-reduced.
// fp32 throughout (the embeddings are only [N][E] = 2 MB at N=2048): logits are recomputed, never stored; the
// backward keeps only d(logit) [2][n][N] as workspace.  Latency-bound (a few tens of microseconds), not HBM-bound.
#include "common.cuh"
#include "../../include/avion_b200.h"

#include <mutex>

namespace {
using namespace avn;
constexpr int ROWS = 8;

// Scratch of avn_infonce_bwd_static: d(logit) [2][n][N] fp32 for callers that may not allocate (CUDA-graph capture, serving
// loops).  4 MiB, held by every GPU that loads the library until the process exits.  In HOST code `g_static_dlogits` is only a
// stand-in for the symbol: its device address comes from cudaGetSymbolAddress (static_dlogits_ptr below).
constexpr long long STATIC_DLOGITS_FLOATS = 1ll << 20;
static __device__ __align__(256) float g_static_dlogits[STATIC_DLOGITS_FLOATS];

// Block = 8 local rows of one direction; thread t owns columns t, t+256, ...
template <bool BWD>
__global__ void __launch_bounds__(256)
infonce_rows_kernel(const float* __restrict__ img, const float* __restrict__ txt, const float* __restrict__ all_img,
                    const float* __restrict__ all_txt, const float* __restrict__ scale_ptr, int n, int N, int E, int row_offset,
                    float* __restrict__ lse, float* __restrict__ out, float gcoef, float* __restrict__ dlogits,
                    float* __restrict__ dscale) {
    extern __shared__ float sm[];
    const float scale = __ldg(scale_ptr);     // logit_scale.exp() stays on the device: no host read of a parameter per step
    float* rows = sm;                         // [ROWS][E]
    float* red = sm + ROWS * E;               // [8 warps][ROWS][4]
    const int dir = blockIdx.y;
    const int r0 = blockIdx.x * ROWS;
    const float* rsrc = dir == 0 ? img : txt;
    const float* cols = dir == 0 ? all_txt : all_img;
    for (int i = threadIdx.x; i < ROWS * E; i += blockDim.x) {
        const int r = r0 + i / E;
        rows[i] = r < n ? rsrc[(long long)r * E + (i % E)] : 0.f;
    }
    __syncthreads();
    float mx[ROWS], se[ROWS], lab[ROWS], ds = 0.f;
    int am[ROWS];
    float row_lse[ROWS];
#pragma unroll
    for (int r = 0; r < ROWS; ++r) {
        mx[r] = -INFINITY; se[r] = 0.f; lab[r] = 0.f; am[r] = 0x7fffffff;
        row_lse[r] = (BWD && r0 + r < n) ? lse[dir * n + r0 + r] : 0.f;
    }
    for (int j = threadIdx.x; j < N; j += blockDim.x) {
        float dot[ROWS];
#pragma unroll
        for (int r = 0; r < ROWS; ++r) dot[r] = 0.f;
        const float4* c = (const float4*)(cols + (long long)j * E);
        for (int e = 0; e < E / 4; ++e) {
            const float4 cv = __ldg(c + e);
#pragma unroll
            for (int r = 0; r < ROWS; ++r) {
                const float4 rv = *(const float4*)(rows + r * E + 4 * e);
                dot[r] = fmaf(rv.x, cv.x, fmaf(rv.y, cv.y, fmaf(rv.z, cv.z, fmaf(rv.w, cv.w, dot[r]))));
            }
        }
#pragma unroll
        for (int r = 0; r < ROWS; ++r) {
            const float l = scale * dot[r];
            const bool is_label = (j == row_offset + r0 + r);
            if (!BWD) {
                if (l > mx[r]) { se[r] = se[r] * __expf(mx[r] - l) + 1.f; mx[r] = l; am[r] = j; }
                else se[r] += __expf(l - mx[r]);
                if (is_label) lab[r] = l;
            } else if (r0 + r < n) {
                const float g = gcoef * (__expf(l - row_lse[r]) - (is_label ? 1.f : 0.f));   // dLoss/dlogit
                ds = fmaf(g, dot[r], ds);
                dlogits[((long long)dir * n + r0 + r) * N + j] = g * scale;                   // dLoss/d(dot)
            }
        }
    }
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (BWD) {
        ds = warp_sum(ds);
        if (lane == 0) red[warp] = ds;
        __syncthreads();
        if (threadIdx.x == 0) {
            float s = 0.f;
            for (int w = 0; w < 8; ++w) s += red[w];
            atomicAdd(dscale, s);
        }
        return;
    }
    // merge (max, sumexp, argmax, label) across the block
#pragma unroll
    for (int r = 0; r < ROWS; ++r) {
        float m = mx[r], s = se[r], lb = lab[r]; int a = am[r];
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            const float m2 = __shfl_xor_sync(0xffffffffu, m, o), s2 = __shfl_xor_sync(0xffffffffu, s, o);
            const int a2 = __shfl_xor_sync(0xffffffffu, a, o);
            lb += __shfl_xor_sync(0xffffffffu, lb, o);
            const float mn = fmaxf(m, m2);
            const float sa = (m == -INFINITY) ? 0.f : s * __expf(m - mn), sb = (m2 == -INFINITY) ? 0.f : s2 * __expf(m2 - mn);
            if (m2 > m || (m2 == m && a2 < a)) a = a2;
            m = mn; s = sa + sb;
        }
        if (lane == 0) { float* q = red + (warp * ROWS + r) * 4; q[0] = m; q[1] = s; q[2] = lb; q[3] = __int_as_float(a); }
    }
    __syncthreads();
    if (threadIdx.x < ROWS && r0 + threadIdx.x < n) {
        const int r = threadIdx.x;
        float m = -INFINITY, s = 0.f, lb = 0.f; int a = 0x7fffffff;
        for (int w = 0; w < 8; ++w) {
            const float* q = red + (w * ROWS + r) * 4;
            const float m2 = q[0], s2 = q[1]; const int a2 = __float_as_int(q[3]);
            lb += q[2];
            const float mn = fmaxf(m, m2);
            const float sa = (m == -INFINITY) ? 0.f : s * __expf(m - mn), sb = (m2 == -INFINITY) ? 0.f : s2 * __expf(m2 - mn);
            if (m2 > m || (m2 == m && a2 < a)) a = a2;
            m = mn; s = sa + sb;
        }
        const float l = m + logf(s);
        lse[dir * n + r0 + r] = l;
        atomicAdd(out, 0.5f * (l - lb));
        if (dir == 0 && a == row_offset + r0 + r) atomicAdd(out + 1, 1.f);
    }
}

// col-side: d_cols[j][e] += sum_r G[dir][r][j] * rows[r][e]   (dir 0 -> d_all_txt, dir 1 -> d_all_img)
__global__ void __launch_bounds__(256)
infonce_dcols_kernel(const float* __restrict__ img, const float* __restrict__ txt, const float* __restrict__ G, float* __restrict__ d_all,
                     int n, int N, int E) {
    const int dir = blockIdx.y;
    const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= (long long)N * E) return;
    const int j = (int)(idx / E), e = (int)(idx % E);
    const float* rows = dir == 0 ? img : txt;
    const float* g = G + (long long)dir * n * N + j;
    float a = 0.f;
    for (int r = 0; r < n; ++r) a = fmaf(__ldg(g + (long long)r * N), __ldg(rows + (long long)r * E + e), a);
    d_all[(long long)(dir == 0 ? 1 : 0) * N * E + idx] += a;
}

// row-side: d_all[dir][row_offset + r][e] += sum_j G[dir][r][j] * cols[j][e]
__global__ void __launch_bounds__(256)
infonce_drows_kernel(const float* __restrict__ all_img, const float* __restrict__ all_txt, const float* __restrict__ G,
                     float* __restrict__ d_all, int n, int N, int E, int row_offset) {
    const int dir = blockIdx.y;
    const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= (long long)n * E) return;
    const int r = (int)(idx / E), e = (int)(idx % E);
    const float* cols = dir == 0 ? all_txt : all_img;
    const float* g = G + ((long long)dir * n + r) * N;
    float a = 0.f;
    for (int j = 0; j < N; ++j) a = fmaf(__ldg(g + j), __ldg(cols + (long long)j * E + e), a);
    d_all[((long long)dir * N + row_offset + r) * E + e] += a;
}
}  // namespace

extern "C" int avn_infonce_fwd(const float* img, const float* txt, const float* all_img, const float* all_txt, const float* scale,
                               int n, int N, int E, int row_offset, float* lse, float* out, avn_stream_t stream_) {
    AVN_REQUIRE(img && txt && all_img && all_txt && scale && lse && out, "infonce_fwd: null pointer");
    AVN_REQUIRE(n > 0 && N >= n && E > 0 && E % 4 == 0 && row_offset >= 0 && row_offset + n <= N, "infonce_fwd: bad sizes n=%d N=%d E=%d off=%d", n, N, E, row_offset);
    const size_t smem = (size_t)(ROWS * E + 8 * ROWS * 4) * sizeof(float);
    auto kern = infonce_rows_kernel<false>;
    AVN_CHECK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    kern<<<dim3((n + ROWS - 1) / ROWS, 2), 256, smem, (cudaStream_t)stream_>>>(img, txt, all_img, all_txt, scale, n, N, E, row_offset, lse,
                                                                            out, 0.f, nullptr, nullptr);
    AVN_CHECK_CUDA(cudaGetLastError());
    return AVN_OK;
}

extern "C" int avn_infonce_bwd(const float* img, const float* txt, const float* all_img, const float* all_txt, const float* scale,
                               int n, int N, int E, int row_offset, const float* lse, float gscale, int col_side, float* dlogits,
                               float* d_all, float* dscale, avn_stream_t stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    AVN_REQUIRE(img && txt && all_img && all_txt && scale && lse && dlogits && d_all && dscale, "infonce_bwd: null pointer");
    AVN_REQUIRE(n > 0 && N >= n && E > 0 && E % 4 == 0 && row_offset >= 0 && row_offset + n <= N, "infonce_bwd: bad sizes");
    const size_t smem = (size_t)(ROWS * E + 8 * ROWS * 4) * sizeof(float);
    auto kern = infonce_rows_kernel<true>;
    AVN_CHECK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    kern<<<dim3((n + ROWS - 1) / ROWS, 2), 256, smem, stream>>>(img, txt, all_img, all_txt, scale, n, N, E, row_offset,
                                                              const_cast<float*>(lse), nullptr, gscale * 0.5f / n, dlogits, dscale);
    AVN_CHECK_CUDA(cudaGetLastError());
    const long long nc = (long long)N * E, nr = (long long)n * E;
    if (col_side) {      // gathered operands that carry no gradient (losses.py:61-72 with local_loss) get no column-side term
        infonce_dcols_kernel<<<dim3((unsigned)((nc + 255) / 256), 2), 256, 0, stream>>>(img, txt, dlogits, d_all, n, N, E);
        AVN_CHECK_CUDA(cudaGetLastError());
    }
    infonce_drows_kernel<<<dim3((unsigned)((nr + 255) / 256), 2), 256, 0, stream>>>(all_img, all_txt, dlogits, d_all, n, N, E, row_offset);
    AVN_CHECK_CUDA(cudaGetLastError());
    return AVN_OK;
}

// ---- no-allocation variant ---------------------------------------------------------------------------------------------
namespace {
constexpr int MAX_DEVICES = 64;
std::mutex g_static_mu;
float* g_static_ptr[MAX_DEVICES] = {};

// Device address of g_static_dlogits on the CURRENT device, resolved once per device.
int static_dlogits_ptr(float** out) {
    int dev = -1;
    AVN_CHECK_CUDA(cudaGetDevice(&dev));
    AVN_REQUIRE(dev >= 0 && dev < MAX_DEVICES, "infonce_bwd_static: device ordinal %d out of range", dev);
    std::lock_guard<std::mutex> lock(g_static_mu);
    if (!g_static_ptr[dev]) {
        float* p = nullptr;
        AVN_CHECK_CUDA(cudaGetSymbolAddress((void**)&p, g_static_dlogits));
        AVN_REQUIRE(p != nullptr, "infonce_bwd_static: cudaGetSymbolAddress returned a null address");
        g_static_ptr[dev] = p;
    }
    *out = g_static_ptr[dev];
    return AVN_OK;
}
}  // namespace

extern "C" int avn_infonce_static_max_floats(void) { return (int)STATIC_DLOGITS_FLOATS; }

extern "C" int avn_infonce_bwd_static(const float* img, const float* txt, const float* all_img, const float* all_txt,
                                      const float* scale, int n, int N, int E, int row_offset, const float* lse, float gscale,
                                      int col_side, float* d_all, float* dscale, avn_stream_t stream) {
    AVN_REQUIRE(n > 0 && N >= n, "infonce_bwd_static: bad sizes n=%d N=%d", n, N);
    AVN_REQUIRE(2ll * n * N <= STATIC_DLOGITS_FLOATS,
                "infonce_bwd_static: 2*n*N = %lld floats exceeds the static scratch of %lld floats (use avn_infonce_bwd with a "
                "caller-provided workspace)", 2ll * n * N, STATIC_DLOGITS_FLOATS);
    float* dlogits = nullptr;
    const int rc = static_dlogits_ptr(&dlogits);
    if (rc != AVN_OK) return rc;
    // every element [dir][r < n][j < N] of the scratch is written by the rows kernel before the dcols / drows kernels read
    // it (same stream), so the previous call's contents need no clearing
    return avn_infonce_bwd(img, txt, all_img, all_txt, scale, n, N, E, row_offset, lse, gscale, col_side, dlogits, d_all, dscale,
                           stream);
}
