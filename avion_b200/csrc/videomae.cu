// VideoMAE pretraining step (BASELINE.json configs[4]): the HBM-bound pieces around the
// transformer blocks.  The blocks themselves run on the same tcgen05 GEMM / attention / LayerNorm kernels as the CLIP towers.
//
//   avn_tube_mask_slots       mask [B,N] -> token order "visible first, then masked" (boolean-index order of
//                             x[~mask] / x[mask], avion/models/model_videomae.py:383,575-576)
//   avn_tubelet_transform_u8  uint8 clips -> (a) normalised bf16 tubelet patches of the VISIBLE tokens only, feature order
//                             (c p0 p1 p2) (model_videomae.py:161; kornia Normalize scripts/main_videomae_pretrain.py:206,336)
//                             (b) per-patch per-channel standardised fp32 regression targets of the MASKED tokens, feature
//                             order (p0 p1 p2 c) (scripts/main_videomae_pretrain.py:322-334)
//   avn_gather_rows_f32_bf16  sinusoid position rows in token order (+ mask_token on masked rows) (model_videomae.py:379,574-577)
//   avn_colsum_batched_bf16   gradient of mask_token
//   avn_mse_fwd / avn_mse_bwd torch.nn.MSELoss (scripts/main_videomae_pretrain.py:129,341)
//
// The reference patch-embeds all 1568 tokens and then drops 90 % of them; patch embedding is per token, so gathering first
// gives identical values for a tenth of the bytes and FLOPs.
#include <algorithm>
#include "common.cuh"
#include "../../include/avion_b200.h"

namespace {
using namespace avn;

// ---------------------------------------------------------------------------------------------------------------------
// one warp per clip: slot[b][n] = rank of token n in "visible tokens ascending, then masked tokens ascending";
// order[b][slot] = n (inverse permutation); counts[b] = number of visible tokens.
__global__ void tube_mask_slots_kernel(const uint8_t* __restrict__ mask, int32_t* __restrict__ slot, int32_t* __restrict__ order,
                                       int32_t* __restrict__ counts, int B, int N) {
    const int b = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (b >= B) return;
    const int lane = threadIdx.x & 31;
    const uint8_t* m = mask + (long long)b * N;
    int nvis = 0;
    for (int n0 = 0; n0 < N; n0 += 32) {
        const int n = n0 + lane;
        const bool vis = n < N && m[n] == 0;
        nvis += __popc(__ballot_sync(0xffffffffu, vis));
    }
    int vis_before = 0;
    for (int n0 = 0; n0 < N; n0 += 32) {
        const int n = n0 + lane;
        const bool in = n < N;
        const bool vis = in && m[n] == 0;
        const unsigned bal = __ballot_sync(0xffffffffu, vis);
        const int vb = vis_before + __popc(bal & ((1u << lane) - 1u));
        if (in) {
            const int s = vis ? vb : nvis + (n - vb);
            slot[(long long)b * N + n] = s;
            order[(long long)b * N + s] = n;
        }
        vis_before += __popc(bal);
    }
    if (lane == 0 && counts != nullptr) counts[b] = nvis;
}

// ---------------------------------------------------------------------------------------------------------------------
// One CTA per (clip, tubelet time step, patch row): stages the 2*p image rows (each crop*3 bytes, 16-byte loads), then for
// each of the gw tokens of that row writes either its patch (visible) or its standardised target (masked).
constexpr int TUBE = 2;

template <int P>      // patch size as a compile-time constant (index arithmetic becomes shifts); 0 = run-time value
__global__ void __launch_bounds__(256)
tubelet_transform_kernel(const uint8_t* __restrict__ frames, const int32_t* __restrict__ slot, __nv_bfloat16* __restrict__ patches,
                         float* __restrict__ targets, int T, int crop, int p_rt, int Nv, float m0, float m1, float m2, float s0,
                         float s1, float s2) {
    const int p = P > 0 ? P : p_rt;
    extern __shared__ uint8_t smem[];
    __nv_bfloat16* lut = (__nv_bfloat16*)smem;                 // [3][256]
    float* stat = (float*)(smem + 3 * 256 * 2);                // [gw][3][2] mean, 1/(std+1e-6)
    int* tok_slot = (int*)(smem + 3 * 256 * 2 + 64 * 6 * 4);   // [gw]
    uint8_t* rows = smem + 3 * 256 * 2 + 64 * 6 * 4 + 64 * 4;  // [TUBE*p][row_bytes]
    const int g = crop / p;
    const int Tt = T / TUBE;
    const int hh = blockIdx.x % g;
    const int tt = (blockIdx.x / g) % Tt;
    const int b = blockIdx.x / (g * Tt);
    const int row_bytes = crop * 3;                            // multiple of 16 (crop % 16 == 0)
    const int N = Tt * g * g;
    const int F = 3 * TUBE * p * p;

    for (int i = threadIdx.x; i < 768; i += blockDim.x) {
        const int c = i >> 8;
        const float mean = c == 0 ? m0 : (c == 1 ? m1 : m2), sd = c == 0 ? s0 : (c == 1 ? s1 : s2);
        lut[i] = __float2bfloat16_rn(__fdiv_rn((float)(i & 255) - mean, sd));
    }
    if (threadIdx.x < g) tok_slot[threadIdx.x] = slot[(long long)b * N + (tt * g + hh) * g + threadIdx.x];
    const int chunks = row_bytes / 16;
    for (int i = threadIdx.x; i < TUBE * p * chunks; i += blockDim.x) {
        const int r = i / chunks, ch = i % chunks;             // r = p0 * p + p1
        const int p0 = r / p, p1 = r % p;
        const long long off = (((long long)(b * T + tt * TUBE + p0) * crop + (hh * p + p1)) * crop) * 3 + (long long)ch * 16;
        *(uint4*)(rows + r * row_bytes + ch * 16) = __ldg((const uint4*)(frames + off));
    }
    __syncthreads();

    // Statistics of the masked tokens: one warp per token.  A token is TUBE*p rows of p*3 contiguous staged bytes; lanes walk
    // its 32-bit words (channel of byte t in a row = t % 3, so a word at word index k starts at channel k % 3).
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nwarps = blockDim.x >> 5;
    const int per_tok = TUBE * p * p;                          // 512 positions per channel
    const int wpr = p * 3 / 4;                                 // words per patch row (12)
    const int words = TUBE * p * wpr;                          // words per token (384)
    if (targets != nullptr) {
        for (int ww = warp; ww < g; ww += nwarps) {
            if (tok_slot[ww] < Nv) continue;
            const uint8_t* base = rows + ww * p * 3;
            float s[3] = {0.f, 0.f, 0.f};
            for (int idx = lane; idx < words; idx += 32) {
                const int r = idx / wpr, k = idx - r * wpr, ph = k % 3;
                const uint32_t w = *(const uint32_t*)(base + r * row_bytes + 4 * k);
                const float t0 = (float)((w & 0xffu) + (w >> 24)), t1 = (float)((w >> 8) & 0xffu), t2 = (float)((w >> 16) & 0xffu);
                if (ph == 0) { s[0] += t0; s[1] += t1; s[2] += t2; }
                else if (ph == 1) { s[1] += t0; s[2] += t1; s[0] += t2; }
                else { s[2] += t0; s[0] += t1; s[1] += t2; }
            }
            float mean[3];
#pragma unroll
            for (int c = 0; c < 3; ++c) mean[c] = warp_sum(s[c]) / (float)per_tok;     // exact: integer sums < 2^24
            float q[3] = {0.f, 0.f, 0.f};
            for (int idx = lane; idx < words; idx += 32) {
                const int r = idx / wpr, k = idx - r * wpr, ph = k % 3;
                const uint32_t w = *(const uint32_t*)(base + r * row_bytes + 4 * k);
                const float ma = ph == 0 ? mean[0] : (ph == 1 ? mean[1] : mean[2]);   // channel of bytes 0 and 3
                const float mb = ph == 0 ? mean[1] : (ph == 1 ? mean[2] : mean[0]);   // byte 1
                const float mc = ph == 0 ? mean[2] : (ph == 1 ? mean[0] : mean[1]);   // byte 2
                const float d0 = (float)(w & 0xffu) - ma, d3 = (float)(w >> 24) - ma;
                const float d1 = (float)((w >> 8) & 0xffu) - mb, d2 = (float)((w >> 16) & 0xffu) - mc;
                const float qa = fmaf(d0, d0, d3 * d3), qb = d1 * d1, qc = d2 * d2;
                if (ph == 0) { q[0] += qa; q[1] += qb; q[2] += qc; }
                else if (ph == 1) { q[1] += qa; q[2] += qb; q[0] += qc; }
                else { q[2] += qa; q[0] += qb; q[1] += qc; }
            }
#pragma unroll
            for (int c = 0; c < 3; ++c) {
                const float ss = warp_sum(q[c]);
                if (lane == 0) {
                    stat[(ww * 3 + c) * 2 + 0] = mean[c];
                    stat[(ww * 3 + c) * 2 + 1] = 1.0f / (sqrtf(ss / (float)(per_tok - 1)) + 1e-6f);
                }
            }
        }
    }
    __syncthreads();

    // visible tokens -> patches: feature f = ((c*TUBE + p0)*p + p1)*p + p2, 8 consecutive p2 per 16-byte store
    if (patches != nullptr) {
        const int p8 = p / 8, vec = F / 8;
        for (int i = threadIdx.x; i < g * vec; i += blockDim.x) {
            const int ww = i / vec, o = i - ww * vec;
            const int s = tok_slot[ww];
            if (s >= Nv) continue;
            const int c = o / (TUBE * p * p8), r = (o / p8) % (TUBE * p), sw0 = (o % p8) * 8;
            const uint8_t* src = rows + r * row_bytes + (ww * p + sw0) * 3 + c;
            const __nv_bfloat16* l = lut + c * 256;
            __nv_bfloat16 v[8];
#pragma unroll
            for (int j = 0; j < 8; ++j) v[j] = l[src[j * 3]];
            *(uint4*)(patches + ((long long)b * Nv + s) * F + (long long)o * 8) = *(uint4*)v;
        }
    }
    // masked tokens -> targets: feature f = (p0*p*p + p1*p + p2)*3 + c = the staged bytes of the token in memory order, so one
    // 32-bit word of the staged rows becomes one float4 of the output row
    if (targets != nullptr) {
        for (int i = threadIdx.x; i < g * words; i += blockDim.x) {
            const int ww = i / words, o = i - ww * words;
            const int s = tok_slot[ww];
            if (s < Nv) continue;
            const float* st = stat + ww * 6;
            const int r = o / wpr, k = o - r * wpr, ph = k % 3;
            const uint32_t w = *(const uint32_t*)(rows + ww * p * 3 + r * row_bytes + 4 * k);
            const int ca = ph, cb = ph == 2 ? 0 : ph + 1, cc = ph == 0 ? 2 : ph - 1;      // channels of bytes (0,3), 1, 2
            const float ma = st[2 * ca], ia = st[2 * ca + 1], mb = st[2 * cb], ib = st[2 * cb + 1], mc = st[2 * cc], ic = st[2 * cc + 1];
            float4 v;
            v.x = ((float)(w & 0xffu) - ma) * ia;
            v.y = ((float)((w >> 8) & 0xffu) - mb) * ib;
            v.z = ((float)((w >> 16) & 0xffu) - mc) * ic;
            v.w = ((float)(w >> 24) - ma) * ia;
            *(float4*)(targets + ((long long)b * (N - Nv) + (s - Nv)) * F + 4 * o) = v;
        }
    }
}

// ---------------------------------------------------------------------------------------------------------------------
// out[b][j][:] = bf16(table[idx[b*idx_stride + j]][:] + (j >= add_from ? add_row[:] : 0)); D % 4 == 0
__global__ void gather_rows_kernel(const float* __restrict__ table, const int32_t* __restrict__ idx, int64_t idx_stride,
                                   const float* __restrict__ add_row, int add_from, __nv_bfloat16* __restrict__ out, int n, int D,
                                   long long total_vec) {
    const int dv = D / 4;
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total_vec; i += (long long)gridDim.x * blockDim.x) {
        const int d4 = (int)(i % dv);
        const long long row = i / dv;
        const int j = (int)(row % n);
        const long long b = row / n;
        const int src = idx[b * idx_stride + j];
        float4 v = __ldg((const float4*)(table + (long long)src * D) + d4);
        if (add_row != nullptr && j >= add_from) {
            const float4 a = __ldg((const float4*)add_row + d4);
            v.x += a.x; v.y += a.y; v.z += a.z; v.w += a.w;
        }
        uint2 o;
        o.x = pack_bf16(v.x, v.y);
        o.y = pack_bf16(v.z, v.w);
        *(uint2*)(out + row * D + d4 * 4) = o;
    }
}

// out[c] += sum_b sum_r x[b*batch_stride + r*cols + c]; two columns per thread, blockDim.x = cols/2
__global__ void colsum_batched_kernel(const __nv_bfloat16* __restrict__ x, int64_t batch_stride, int rows, int cols,
                                      int chunks, float* __restrict__ out) {
    const int b = blockIdx.x / chunks, ch = blockIdx.x % chunks;
    const int per = (rows + chunks - 1) / chunks;
    const int r0 = ch * per, r1 = min(rows, r0 + per);
    const int c2 = threadIdx.x;
    if (c2 * 2 >= cols) return;
    const uint32_t* base = (const uint32_t*)(x + (long long)b * batch_stride) + c2;
    float a0 = 0.f, a1 = 0.f;
    const int stride = cols / 2;
#pragma unroll 4
    for (int r = r0; r < r1; ++r) {
        const float2 v = unpack_bf16(__ldg(base + (long long)r * stride));
        a0 += v.x; a1 += v.y;
    }
    atomicAdd(out + 2 * c2, a0);
    atomicAdd(out + 2 * c2 + 1, a1);
}

// ---------------------------------------------------------------------------------------------------------------------
// MSE: deterministic two-stage sum of (pred - target)^2 / n
__global__ void __launch_bounds__(256) mse_partial_kernel(const __nv_bfloat16* __restrict__ pred, const float* __restrict__ target,
                                                          long long n4, float* __restrict__ partial) {
    float acc = 0.f;
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n4; i += (long long)gridDim.x * blockDim.x) {
        const uint2 pv = __ldg((const uint2*)pred + i);
        const float4 t = __ldg((const float4*)target + i);
        const float2 p01 = unpack_bf16(pv.x), p23 = unpack_bf16(pv.y);
        const float d0 = p01.x - t.x, d1 = p01.y - t.y, d2 = p23.x - t.z, d3 = p23.y - t.w;
        acc += (d0 * d0 + d1 * d1) + (d2 * d2 + d3 * d3);
    }
    __shared__ float red[8];
    acc = warp_sum(acc);
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = acc;
    __syncthreads();
    if (threadIdx.x == 0) {
        float s = 0.f;
        for (int w = 0; w < 8; ++w) s += red[w];
        partial[blockIdx.x] = s;
    }
}

__global__ void __launch_bounds__(256) mse_final_kernel(const float* __restrict__ partial, int count, double inv_n,
                                                        float* __restrict__ loss) {
    __shared__ double red[256];
    double s = 0.0;
    for (int i = threadIdx.x; i < count; i += 256) s += (double)partial[i];
    red[threadIdx.x] = s;
    __syncthreads();
    for (int w = 128; w > 0; w >>= 1) {
        if (threadIdx.x < w) red[threadIdx.x] += red[threadIdx.x + w];
        __syncthreads();
    }
    if (threadIdx.x == 0) loss[0] = (float)(red[0] * inv_n);
}

// dpred = (2/n) * gloss * (pred - target)
__global__ void __launch_bounds__(256) mse_bwd_kernel(const __nv_bfloat16* __restrict__ pred, const float* __restrict__ target,
                                                      const float* __restrict__ gloss, float two_over_n, long long n4,
                                                      __nv_bfloat16* __restrict__ dpred) {
    const float k = two_over_n * __ldg(gloss);
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n4; i += (long long)gridDim.x * blockDim.x) {
        const uint2 pv = __ldg((const uint2*)pred + i);
        const float4 t = __ldg((const float4*)target + i);
        const float2 p01 = unpack_bf16(pv.x), p23 = unpack_bf16(pv.y);
        uint2 o;
        o.x = pack_bf16((p01.x - t.x) * k, (p01.y - t.y) * k);
        o.y = pack_bf16((p23.x - t.z) * k, (p23.y - t.w) * k);
        *((uint2*)dpred + i) = o;
    }
}

constexpr int MSE_GRID = 148 * 8;

}  // namespace

extern "C" int avn_tube_mask_slots(const uint8_t* mask, int32_t* slot, int32_t* order, int32_t* counts, int B, int N,
                                   avn_stream_t stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    AVN_REQUIRE(mask && slot && order, "tube_mask_slots: null pointer");
    AVN_REQUIRE(B > 0 && N > 0, "tube_mask_slots: bad sizes");
    const int warps = 4;
    tube_mask_slots_kernel<<<(B + warps - 1) / warps, warps * 32, 0, stream>>>(mask, slot, order, counts, B, N);
    AVN_CHECK_CUDA(cudaGetLastError());
    return AVN_OK;
}

extern "C" int avn_tubelet_transform_u8(const uint8_t* frames, const int32_t* slot, void* patches, float* targets, int B, int T,
                                        int crop, int patch, int n_visible, const float* mean3, const float* std3,
                                        avn_stream_t stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    AVN_REQUIRE(frames && slot && mean3 && std3 && (patches || targets), "tubelet_transform: null pointer");
    AVN_REQUIRE(B > 0 && T > 0 && T % TUBE == 0, "tubelet_transform: the clip length must be a multiple of the tubelet size 2");
    AVN_REQUIRE(patch > 0 && patch % 8 == 0 && patch <= 32 && crop % patch == 0 && crop % 16 == 0 && crop / patch <= 64,
                "tubelet_transform: unsupported crop %d / patch %d", crop, patch);
    AVN_REQUIRE(((uintptr_t)frames % 16) == 0 && ((uintptr_t)patches % 16) == 0 && ((uintptr_t)targets % 16) == 0,
                "tubelet_transform: buffers must be 16-byte aligned");
    const int g = crop / patch;
    AVN_REQUIRE(n_visible >= 0 && n_visible <= (T / TUBE) * g * g, "tubelet_transform: bad visible count");
    const size_t smem = 3 * 256 * 2 + 64 * 6 * 4 + 64 * 4 + (size_t)TUBE * patch * crop * 3;
    AVN_REQUIRE(smem <= 227 * 1024, "tubelet_transform: crop too wide for shared memory");
    auto kern = patch == 16 ? tubelet_transform_kernel<16> : tubelet_transform_kernel<0>;
    AVN_CHECK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    kern<<<B * (T / TUBE) * g, 256, smem, stream>>>(frames, slot, (__nv_bfloat16*)patches, targets, T, crop, patch, n_visible,
                                                    mean3[0], mean3[1], mean3[2], std3[0], std3[1], std3[2]);
    AVN_CHECK_CUDA(cudaGetLastError());
    return AVN_OK;
}

extern "C" int avn_gather_rows_f32_bf16(const float* table, const int32_t* idx, int64_t idx_stride, const float* add_row,
                                        int add_from, void* out, int B, int n, int D, avn_stream_t stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    AVN_REQUIRE(table && idx && out, "gather_rows: null pointer");
    AVN_REQUIRE(B > 0 && n > 0 && D > 0 && D % 4 == 0, "gather_rows: bad sizes (D must be a multiple of 4)");
    const long long total = (long long)B * n * (D / 4);
    const int grid = (int)std::min<long long>((total + 255) / 256, 148LL * 16);
    gather_rows_kernel<<<grid, 256, 0, stream>>>(table, idx, idx_stride, add_row, add_from, (__nv_bfloat16*)out, n, D, total);
    AVN_CHECK_CUDA(cudaGetLastError());
    return AVN_OK;
}

extern "C" int avn_colsum_batched_bf16(const void* x, int B, int64_t batch_stride, int rows, int cols, float* out,
                                       avn_stream_t stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    AVN_REQUIRE(x && out, "colsum_batched: null pointer");
    AVN_REQUIRE(B > 0 && rows > 0 && cols > 0 && cols % 2 == 0 && cols <= 2048, "colsum_batched: cols must be even and <= 2048");
    AVN_REQUIRE(batch_stride % 2 == 0 && ((uintptr_t)x % 4) == 0, "colsum_batched: misaligned input");
    const int chunks = std::max(1, std::min(rows / 32, (148 * 8 + B - 1) / B));
    colsum_batched_kernel<<<B * chunks, (cols / 2 + 31) / 32 * 32, 0, stream>>>((const __nv_bfloat16*)x, batch_stride, rows, cols,
                                                                              chunks, out);
    AVN_CHECK_CUDA(cudaGetLastError());
    return AVN_OK;
}

extern "C" int avn_mse_workspace_floats(void) { return MSE_GRID; }

extern "C" int avn_mse_fwd(const void* pred, const float* target, int64_t n, float* workspace, float* loss, avn_stream_t stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    AVN_REQUIRE(pred && target && workspace && loss, "mse_fwd: null pointer");
    AVN_REQUIRE(n > 0 && n % 4 == 0, "mse_fwd: element count must be a positive multiple of 4");
    AVN_REQUIRE(((uintptr_t)pred % 8) == 0 && ((uintptr_t)target % 16) == 0, "mse_fwd: misaligned input");
    mse_partial_kernel<<<MSE_GRID, 256, 0, stream>>>((const __nv_bfloat16*)pred, target, n / 4, workspace);
    mse_final_kernel<<<1, 256, 0, stream>>>(workspace, MSE_GRID, 1.0 / (double)n, loss);
    AVN_CHECK_CUDA(cudaGetLastError());
    return AVN_OK;
}

extern "C" int avn_mse_bwd(const void* pred, const float* target, const float* gloss, int64_t n, void* dpred,
                           avn_stream_t stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    AVN_REQUIRE(pred && target && gloss && dpred, "mse_bwd: null pointer");
    AVN_REQUIRE(n > 0 && n % 4 == 0, "mse_bwd: element count must be a positive multiple of 4");
    AVN_REQUIRE(((uintptr_t)pred % 8) == 0 && ((uintptr_t)target % 16) == 0 && ((uintptr_t)dpred % 8) == 0, "mse_bwd: misaligned");
    mse_bwd_kernel<<<MSE_GRID, 256, 0, stream>>>((const __nv_bfloat16*)pred, target, gloss, (float)(2.0 / (double)n), n / 4,
                                                 (__nv_bfloat16*)dpred);
    AVN_CHECK_CUDA(cudaGetLastError());
    return AVN_OK;
}
