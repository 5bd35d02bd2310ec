// MaxMarginRankingLoss (avion/losses/losses.py:152-227) on all-gathered text / image features, fp32, latency-bound
// (n = global batch <= a few thousand rows of E = 256): the EK-100 MIR fine-tune criterion.
//
//   x = sim_matrix(text, image): x[i][j] = <t_i, v_j> / (max(|t_i|, eps) max(|v_j|, eps))            (:152-160)
//   loss = mean over i != j (fix_norm) of relu(m - x[i][i] + x[i][j]) and relu(m - x[i][i] + x[j][i]) (:197-225)
//
// One CTA per row i evaluates row i and column i of x against the L2-resident normalised features, so neither x nor its
// gradient is materialised.  All reductions run in a fixed order (deterministic).
#include "common.cuh"
#include "../../include/avion_b200.h"

namespace {
using namespace avn;

constexpr int MAX_VPL = 32;      // features per lane: E <= 1024
constexpr int WARPS = 8;

// warp per row: an = a / max(|a|, eps), bn likewise, norms, diag[i] = <an_i, bn_i>
__global__ void __launch_bounds__(WARPS * 32)
mmr_normalize_kernel(const float* __restrict__ a, const float* __restrict__ b, float* __restrict__ an, float* __restrict__ bn,
                     float* __restrict__ na, float* __restrict__ nb, float* __restrict__ diag, int n, int E, float eps) {
    const int i = blockIdx.x * WARPS + (threadIdx.x >> 5), lane = threadIdx.x & 31;
    if (i >= n) return;
    float sa = 0.f, sb = 0.f;
    for (int e = lane; e < E; e += 32) {
        const float x = a[(long long)i * E + e], y = b[(long long)i * E + e];
        sa = fmaf(x, x, sa);
        sb = fmaf(y, y, sb);
    }
    const float ra = fmaxf(sqrtf(warp_sum(sa)), eps), rb = fmaxf(sqrtf(warp_sum(sb)), eps);
    float d = 0.f;
    for (int e = lane; e < E; e += 32) {
        const float x = a[(long long)i * E + e] / ra, y = b[(long long)i * E + e] / rb;
        an[(long long)i * E + e] = x;
        bn[(long long)i * E + e] = y;
        d = fmaf(x, y, d);
    }
    d = warp_sum(d);
    if (lane == 0) { na[i] = ra; nb[i] = rb; diag[i] = d; }
}

// CTA i: partial[i] = sum over j (j != i when fix_norm) of relu(m - d_i + x_ij) + relu(m - d_i + x_ji)
__global__ void __launch_bounds__(WARPS * 32)
mmr_fwd_kernel(const float* __restrict__ an, const float* __restrict__ bn, const float* __restrict__ diag, float* __restrict__ partial,
               int n, int E, float margin, int fix_norm) {
    extern __shared__ float sm[];
    float* ai = sm;           // [E]
    float* bi = sm + E;       // [E]
    __shared__ float red[WARPS];
    const int i = blockIdx.x, warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    for (int e = threadIdx.x; e < E; e += blockDim.x) { ai[e] = an[(long long)i * E + e]; bi[e] = bn[(long long)i * E + e]; }
    __syncthreads();
    const float di = diag[i];
    float acc = 0.f;
    for (int j = warp; j < n; j += WARPS) {
        float xij = 0.f, xji = 0.f;
        for (int e = lane; e < E; e += 32) {
            xij = fmaf(ai[e], __ldg(bn + (long long)j * E + e), xij);
            xji = fmaf(__ldg(an + (long long)j * E + e), bi[e], xji);
        }
        xij = warp_sum(xij);
        xji = warp_sum(xji);
        if (!(fix_norm && j == i)) acc += fmaxf(margin - di + xij, 0.f) + fmaxf(margin - di + xji, 0.f);
    }
    if (lane == 0) red[warp] = acc;
    __syncthreads();
    if (threadIdx.x == 0) {
        float s = 0.f;
        for (int w = 0; w < WARPS; ++w) s += red[w];
        partial[i] = s;
    }
}

__global__ void __launch_bounds__(256) mmr_final_kernel(const float* __restrict__ partial, int n, double inv_count, float* loss) {
    __shared__ double red[256];
    double s = 0.0;
    for (int i = threadIdx.x; i < n; i += 256) s += (double)partial[i];
    red[threadIdx.x] = s;
    __syncthreads();
    for (int w = 128; w > 0; w >>= 1) {
        if (threadIdx.x < w) red[threadIdx.x] += red[threadIdx.x + w];
        __syncthreads();
    }
    if (threadIdx.x == 0) loss[0] = (float)(red[0] * inv_count);
}

// CTA i: gradients of the loss w.r.t. a_i (text) and b_i (image), through the normalisation
__global__ void __launch_bounds__(WARPS * 32)
mmr_bwd_kernel(const float* __restrict__ an, const float* __restrict__ bn, const float* __restrict__ na, const float* __restrict__ nb,
               const float* __restrict__ diag, const float* __restrict__ gloss, float coef, float* __restrict__ da,
               float* __restrict__ db, int n, int E, float margin, int fix_norm) {
    extern __shared__ float sm[];
    float* ai = sm;                    // [E]
    float* bi = sm + E;                // [E]
    float* ra = sm + 2 * E;            // [WARPS][E] partial d(an_i)
    float* rb = ra + WARPS * E;        // [WARPS][E] partial d(bn_i)
    __shared__ float cnt_s[WARPS];
    __shared__ float dots[2];
    const int i = blockIdx.x, warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    for (int e = threadIdx.x; e < E; e += blockDim.x) { ai[e] = an[(long long)i * E + e]; bi[e] = bn[(long long)i * E + e]; }
    __syncthreads();
    const float di = diag[i];
    float acc_a[MAX_VPL], acc_b[MAX_VPL];
#pragma unroll
    for (int k = 0; k < MAX_VPL; ++k) { acc_a[k] = 0.f; acc_b[k] = 0.f; }
    float cnt = 0.f;
    const int vpl = (E + 31) / 32;
    for (int j = warp; j < n; j += WARPS) {
        if (fix_norm && j == i) continue;
        float xij = 0.f, xji = 0.f;
        for (int e = lane; e < E; e += 32) {
            xij = fmaf(ai[e], __ldg(bn + (long long)j * E + e), xij);
            xji = fmaf(__ldg(an + (long long)j * E + e), bi[e], xji);
        }
        xij = warp_sum(xij);
        xji = warp_sum(xji);
        const float dj = __ldg(diag + j);
        const float i1_ij = margin - di + xij > 0.f ? 1.f : 0.f;      // relu(m - x_ii + x_ij)   row i, first half
        const float i2_ij = margin - di + xji > 0.f ? 1.f : 0.f;      // relu(m - x_ii + x_ji)   row i, second half
        const float i1_ji = margin - dj + xji > 0.f ? 1.f : 0.f;      // relu(m - x_jj + x_ji)   row j, first half
        const float i2_ji = margin - dj + xij > 0.f ? 1.f : 0.f;      // relu(m - x_jj + x_ij)   row j, second half
        const float wa = i1_ij + i2_ji, wb = i1_ji + i2_ij;
        cnt += i1_ij + i2_ij;
        if (j == i) continue;          // fix_norm == 0: the diagonal terms are the constant relu(m)
#pragma unroll
        for (int k = 0; k < MAX_VPL; ++k) {
            const int e = lane + 32 * k;
            if (k < vpl && e < E) {
                acc_a[k] = fmaf(wa, __ldg(bn + (long long)j * E + e), acc_a[k]);
                acc_b[k] = fmaf(wb, __ldg(an + (long long)j * E + e), acc_b[k]);
            }
        }
    }
#pragma unroll
    for (int k = 0; k < MAX_VPL; ++k) {
        const int e = lane + 32 * k;
        if (k < vpl && e < E) { ra[warp * E + e] = acc_a[k]; rb[warp * E + e] = acc_b[k]; }
    }
    if (lane == 0) cnt_s[warp] = cnt;
    __syncthreads();
    float total = 0.f;
    for (int w = 0; w < WARPS; ++w) total += cnt_s[w];
    if (!fix_norm) total -= 2.f * (margin > 0.f ? 1.f : 0.f);         // the j == i pass above counted the constant terms
    // d(an_i), d(bn_i) into ra[0], rb[0]; then <an_i, d an_i>, <bn_i, d bn_i>
    float pa = 0.f, pb = 0.f;
    for (int e = threadIdx.x; e < E; e += blockDim.x) {
        float sa = 0.f, sb = 0.f;
        for (int w = 0; w < WARPS; ++w) { sa += ra[w * E + e]; sb += rb[w * E + e]; }
        sa -= total * bi[e];
        sb -= total * ai[e];
        pa = fmaf(ai[e], sa, pa);
        pb = fmaf(bi[e], sb, pb);
        ra[e] = sa;                    // thread e only ever touches column e: no hazard with the loop above
        rb[e] = sb;
    }
    __syncthreads();
    __shared__ float reda[WARPS], redb[WARPS];
    pa = warp_sum(pa);
    pb = warp_sum(pb);
    if (lane == 0) { reda[warp] = pa; redb[warp] = pb; }
    __syncthreads();
    if (threadIdx.x == 0) {
        float s = 0.f, t = 0.f;
        for (int w = 0; w < WARPS; ++w) { s += reda[w]; t += redb[w]; }
        dots[0] = s; dots[1] = t;
    }
    __syncthreads();
    const float k = coef * __ldg(gloss);
    const float inv_na = k / na[i], inv_nb = k / nb[i];
    for (int e = threadIdx.x; e < E; e += blockDim.x) {
        da[(long long)i * E + e] = (ra[e] - ai[e] * dots[0]) * inv_na;
        db[(long long)i * E + e] = (rb[e] - bi[e] * dots[1]) * inv_nb;
    }
}
}  // namespace

extern "C" int avn_mmr_workspace_floats(int n, int E) { return 2 * n * E + 4 * n; }

extern "C" int avn_mmr_fwd(const float* text, const float* image, int n, int E, float margin, int fix_norm, float* workspace,
                           float* loss, avn_stream_t stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    AVN_REQUIRE(text && image && workspace && loss, "mmr_fwd: null pointer");
    AVN_REQUIRE(n > 1 && E > 0 && E <= 32 * MAX_VPL, "mmr_fwd: need n > 1 and E <= %d (n=%d E=%d)", 32 * MAX_VPL, n, E);
    float* an = workspace; float* bn = an + (size_t)n * E; float* na = bn + (size_t)n * E; float* nb = na + n;
    float* diag = nb + n; float* partial = diag + n;
    mmr_normalize_kernel<<<(n + WARPS - 1) / WARPS, WARPS * 32, 0, stream>>>(text, image, an, bn, na, nb, diag, n, E, 1e-8f);
    mmr_fwd_kernel<<<n, WARPS * 32, 2 * E * sizeof(float), stream>>>(an, bn, diag, partial, n, E, margin, fix_norm);
    const double count = fix_norm ? 2.0 * n * (double)(n - 1) : 2.0 * n * (double)n;
    mmr_final_kernel<<<1, 256, 0, stream>>>(partial, n, 1.0 / count, loss);
    AVN_CHECK_CUDA(cudaGetLastError());
    return AVN_OK;
}

extern "C" int avn_mmr_bwd(const float* workspace, const float* gloss, int n, int E, float margin, int fix_norm, float* dtext,
                           float* dimage, avn_stream_t stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    AVN_REQUIRE(workspace && gloss && dtext && dimage, "mmr_bwd: null pointer");
    AVN_REQUIRE(n > 1 && E > 0 && E <= 32 * MAX_VPL, "mmr_bwd: need n > 1 and E <= %d", 32 * MAX_VPL);
    const float* an = workspace; const float* bn = an + (size_t)n * E; const float* na = bn + (size_t)n * E;
    const float* nb = na + n; const float* diag = nb + n;
    const size_t smem = (size_t)(2 + 2 * WARPS) * E * sizeof(float);
    auto kern = mmr_bwd_kernel;
    AVN_CHECK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    const double count = fix_norm ? 2.0 * n * (double)(n - 1) : 2.0 * n * (double)n;
    kern<<<n, WARPS * 32, smem, stream>>>(an, bn, na, nb, diag, gloss, (float)(1.0 / count), dtext, dimage, n, E, margin, fix_norm);
    AVN_CHECK_CUDA(cudaGetLastError());
    return AVN_OK;
}
