// Small HBM-bound helpers around the towers: weight casts, bias-gradient column sums, the CLS / positional / temporal
// table of the video tower (avion/models/transformer.py:321-329) and the token embedding of the text tower (:459-461).
#include "common.cuh"
#include "../../include/avion_b200.h"

namespace {
using namespace avn;

__global__ void cast_kernel(const float* __restrict__ src, __nv_bfloat16* __restrict__ dst, long long n) {
    const long long i = ((long long)blockIdx.x * blockDim.x + threadIdx.x) * 8;
    if (i + 8 <= n) {
        const float4 a = __ldg((const float4*)(src + i)), b = __ldg((const float4*)(src + i + 4));
        *(uint4*)(dst + i) = make_uint4(pack_bf16(a.x, a.y), pack_bf16(a.z, a.w), pack_bf16(b.x, b.y), pack_bf16(b.z, b.w));
    } else {
        for (long long k = i; k < n; ++k) dst[k] = __float2bfloat16_rn(src[k]);
    }
}

__global__ void transpose_cast_kernel(const float* __restrict__ src, __nv_bfloat16* __restrict__ dst, int rows, int cols) {
    __shared__ float tile[32][33];
    const int c0 = blockIdx.x * 32, r0 = blockIdx.y * 32;
    for (int i = threadIdx.y; i < 32; i += blockDim.y) {
        const int r = r0 + i, c = c0 + threadIdx.x;
        tile[i][threadIdx.x] = (r < rows && c < cols) ? src[(long long)r * cols + c] : 0.f;
    }
    __syncthreads();
    for (int i = threadIdx.y; i < 32; i += blockDim.y) {
        const int c = c0 + i, r = r0 + threadIdx.x;
        if (c < cols && r < rows) dst[(long long)c * rows + r] = __float2bfloat16_rn(tile[threadIdx.x][i]);
    }
}

// thread (cx, ry): 8 columns, rows ry + 8*(blockIdx.y + k*gridDim.y); four row loads in flight per thread
__global__ void __launch_bounds__(256)
colsum_kernel(const __nv_bfloat16* __restrict__ x, long long ld, float* __restrict__ out, long long rows, int cols) {
    __shared__ float red[8][32][8];
    const int cx = threadIdx.x & 31, ry = threadIdx.x >> 5;
    const int c = (blockIdx.x * 32 + cx) * 8;
    float acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    if (c < cols) {
        const long long step = (long long)gridDim.y * 8;
        const __nv_bfloat16* xp = x + c;
        long long r = (long long)blockIdx.y * 8 + ry;
        for (; r + 3 * step < rows; r += 4 * step) {
            uint4 u[4];
#pragma unroll
            for (int k = 0; k < 4; ++k) u[k] = __ldg((const uint4*)(xp + (r + k * step) * ld));
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                const uint32_t w[4] = {u[k].x, u[k].y, u[k].z, u[k].w};
#pragma unroll
                for (int e = 0; e < 4; ++e) { const float2 f = unpack_bf16(w[e]); acc[2 * e] += f.x; acc[2 * e + 1] += f.y; }
            }
        }
        for (; r < rows; r += step) {
            const uint4 u = __ldg((const uint4*)(xp + r * ld));
            const uint32_t w[4] = {u.x, u.y, u.z, u.w};
#pragma unroll
            for (int e = 0; e < 4; ++e) { const float2 f = unpack_bf16(w[e]); acc[2 * e] += f.x; acc[2 * e + 1] += f.y; }
        }
    }
#pragma unroll
    for (int e = 0; e < 8; ++e) red[ry][cx][e] = acc[e];
    __syncthreads();
    const int cc = threadIdx.x >> 3, e = threadIdx.x & 7;   // 32 column-vectors x 8 elements
    float s = 0.f;
#pragma unroll
    for (int w = 0; w < 8; ++w) s += red[w][cc][e];
    const int col = (blockIdx.x * 32 + cc) * 8 + e;
    if (col < cols) atomicAdd(out + col, s);
}

// table[0] = cls + pos[0]; table[1+p] = pos[1 + p%G] + temporal[p/G]
__global__ void pos_table_fwd_kernel(const float* __restrict__ cls, const float* __restrict__ pos, const float* __restrict__ temporal,
                                     __nv_bfloat16* __restrict__ table, int T, int G, int D) {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    const long long total = (long long)(T * G + 1) * D;
    if (i >= total) return;
    const int s = (int)(i / D), d = (int)(i % D);
    float v;
    if (s == 0) v = cls[d] + pos[d];
    else {
        const int pidx = s - 1;
        v = pos[(long long)(1 + pidx % G) * D + d] + (temporal ? temporal[(long long)(pidx / G) * D + d] : 0.f);
    }
    table[i] = __float2bfloat16_rn(v);
}

__global__ void broadcast_row_kernel(__nv_bfloat16* __restrict__ x, const __nv_bfloat16* __restrict__ row, int B, long long batch_stride, int D) {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= (long long)B * D) return;
    x[(i / D) * batch_stride + (i % D)] = row[i % D];
}

// dtable[s][d] = sum_b dx[b][s][d]   (thread per 8 columns of one s)
__global__ void __launch_bounds__(256)
batch_sum_kernel(const __nv_bfloat16* __restrict__ dx, float* __restrict__ dtable, int B, int S, int D) {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    const int nvec = D >> 3;
    if (i >= (long long)S * nvec) return;
    float acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    for (int b = 0; b < B; ++b) {
        const uint4 u = __ldg((const uint4*)(dx + ((long long)b * S * D) + i * 8));
        const uint32_t w[4] = {u.x, u.y, u.z, u.w};
#pragma unroll
        for (int e = 0; e < 4; ++e) { const float2 f = unpack_bf16(w[e]); acc[2 * e] += f.x; acc[2 * e + 1] += f.y; }
    }
#pragma unroll
    for (int e = 0; e < 8; ++e) dtable[i * 8 + e] = acc[e];
}

// one thread per (row of pos/temporal/cls, d)
__global__ void pos_table_bwd_kernel(const float* __restrict__ dtable, float* __restrict__ dcls, float* __restrict__ dpos,
                                     float* __restrict__ dtemporal, int T, int G, int D) {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    const long long n_pos = (long long)(G + 1) * D, n_tmp = (long long)T * D;
    if (i < n_pos) {
        const int g = (int)(i / D), d = (int)(i % D);
        if (g == 0) {
            const float v = dtable[d];
            dpos[d] += v;
            dcls[d] += v;
        } else {
            float s = 0.f;
            for (int t = 0; t < T; ++t) s += dtable[(long long)(1 + t * G + (g - 1)) * D + d];
            dpos[i] += s;
        }
    } else if (i < n_pos + n_tmp && dtemporal != nullptr) {
        const long long k = i - n_pos;
        const int t = (int)(k / D), d = (int)(k % D);
        float s = 0.f;
        for (int g = 0; g < G; ++g) s += dtable[(long long)(1 + t * G + g) * D + d];
        dtemporal[k] += s;
    }
}

__global__ void token_embed_fwd_kernel(const int64_t* __restrict__ ids, const float* __restrict__ tok, const float* __restrict__ pos,
                                       __nv_bfloat16* __restrict__ x, long long BL, int L, int D) {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;   // one thread per 4 elements
    const int nv = D >> 2;
    if (i >= BL * nv) return;
    const long long r = i / nv; const int d = (int)(i % nv) * 4;
    const float4 a = __ldg((const float4*)(tok + ids[r] * D + d));
    const float4 b = __ldg((const float4*)(pos + (long long)(r % L) * D + d));
    *(uint2*)(x + r * D + d) = make_uint2(pack_bf16(a.x + b.x, a.y + b.y), pack_bf16(a.z + b.z, a.w + b.w));
}

__global__ void token_embed_bwd_kernel(const int64_t* __restrict__ ids, const __nv_bfloat16* __restrict__ dx, float* __restrict__ dtok,
                                       long long BL, int D) {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;   // one thread per 4 elements
    const int nv = D >> 2;
    if (i >= BL * nv) return;
    const long long r = i / nv; const int d = (int)(i % nv) * 4;
    const uint2 u = __ldg((const uint2*)(dx + r * D + d));
    if ((u.x | u.y) == 0u) return;      // positions after EOT carry exactly zero gradient under the causal mask
    const float2 f0 = unpack_bf16(u.x), f1 = unpack_bf16(u.y);
    float* o = dtok + ids[r] * D + d;
    atomicAdd(o, f0.x); atomicAdd(o + 1, f0.y); atomicAdd(o + 2, f1.x); atomicAdd(o + 3, f1.y);
}

// dpos[l][d] += sum_b dx[b][l][d]
__global__ void pos_embed_bwd_kernel(const __nv_bfloat16* __restrict__ dx, float* __restrict__ dpos, int B, int L, int D) {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= (long long)L * D) return;
    float s = 0.f;
    for (int b = 0; b < B; ++b) s += __bfloat162float(dx[(long long)b * L * D + i]);
    dpos[i] += s;
}
}  // namespace

extern "C" int avn_cast_f32_bf16(const float* src, void* dst, int64_t n, avn_stream_t stream_) {
    AVN_REQUIRE(src && dst && n > 0, "cast: bad arguments");
    AVN_REQUIRE(((uintptr_t)src % 16) == 0 && ((uintptr_t)dst % 16) == 0, "cast: pointers must be 16-byte aligned");
    const long long threads = (n + 7) / 8;
    cast_kernel<<<(int)((threads + 255) / 256), 256, 0, (cudaStream_t)stream_>>>(src, (__nv_bfloat16*)dst, n);
    AVN_CHECK_CUDA(cudaGetLastError());
    return AVN_OK;
}

extern "C" int avn_transpose_cast_f32_bf16(const float* src, void* dst, int rows, int cols, avn_stream_t stream_) {
    AVN_REQUIRE(src && dst && rows > 0 && cols > 0, "transpose_cast: bad arguments");
    dim3 grid((cols + 31) / 32, (rows + 31) / 32), block(32, 8);
    transpose_cast_kernel<<<grid, block, 0, (cudaStream_t)stream_>>>(src, (__nv_bfloat16*)dst, rows, cols);
    AVN_CHECK_CUDA(cudaGetLastError());
    return AVN_OK;
}

extern "C" int avn_colsum_bf16(const void* x, int64_t ld, float* colsum, int64_t rows, int cols, avn_stream_t stream_) {
    AVN_REQUIRE(x && colsum && rows > 0 && cols > 0, "colsum: bad arguments");
    AVN_REQUIRE(cols % 8 == 0 && ld % 8 == 0 && ((uintptr_t)x % 16) == 0, "colsum: cols/ld must be multiples of 8 and x 16-byte aligned");
    const int gx = (cols + 255) / 256;
    long long gy = (8LL * avn_num_sms() + gx - 1) / gx;      // eight resident CTAs per SM
    const long long maxy = (rows + 7) / 8;
    if (gy > maxy) gy = maxy;
    colsum_kernel<<<dim3(gx, (unsigned)gy), 256, 0, (cudaStream_t)stream_>>>((const __nv_bfloat16*)x, ld, colsum, rows, cols);
    AVN_CHECK_CUDA(cudaGetLastError());
    return AVN_OK;
}

extern "C" int avn_pos_table_fwd(const float* cls, const float* pos, const float* temporal, void* table, int T, int G, int D,
                                 avn_stream_t stream_) {
    AVN_REQUIRE(cls && pos && table && T > 0 && G > 0 && D > 0, "pos_table_fwd: bad arguments");
    AVN_REQUIRE(T == 1 || temporal != nullptr, "pos_table_fwd: temporal embedding required for T > 1");
    const long long total = (long long)(T * G + 1) * D;
    pos_table_fwd_kernel<<<(int)((total + 255) / 256), 256, 0, (cudaStream_t)stream_>>>(cls, pos, temporal, (__nv_bfloat16*)table, T, G, D);
    AVN_CHECK_CUDA(cudaGetLastError());
    return AVN_OK;
}

extern "C" int avn_broadcast_row(void* x, const void* row, int B, int64_t batch_stride, int D, avn_stream_t stream_) {
    AVN_REQUIRE(x && row && B > 0 && D > 0, "broadcast_row: bad arguments");
    const long long total = (long long)B * D;
    broadcast_row_kernel<<<(int)((total + 255) / 256), 256, 0, (cudaStream_t)stream_>>>((__nv_bfloat16*)x, (const __nv_bfloat16*)row, B, batch_stride, D);
    AVN_CHECK_CUDA(cudaGetLastError());
    return AVN_OK;
}

extern "C" int avn_pos_table_bwd(const void* dx, float* dtable_ws, float* dcls, float* dpos, float* dtemporal, int B, int T, int G, int D,
                                 avn_stream_t stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    AVN_REQUIRE(dx && dtable_ws && dcls && dpos && B > 0 && T > 0 && G > 0 && D % 8 == 0, "pos_table_bwd: bad arguments");
    const int S = T * G + 1;
    const long long nv = (long long)S * (D / 8);
    batch_sum_kernel<<<(int)((nv + 255) / 256), 256, 0, stream>>>((const __nv_bfloat16*)dx, dtable_ws, B, S, D);
    AVN_CHECK_CUDA(cudaGetLastError());
    const long long n2 = (long long)(G + 1 + T) * D;
    pos_table_bwd_kernel<<<(int)((n2 + 255) / 256), 256, 0, stream>>>(dtable_ws, dcls, dpos, T > 1 ? dtemporal : nullptr, T, G, D);
    AVN_CHECK_CUDA(cudaGetLastError());
    return AVN_OK;
}

extern "C" int avn_token_embed_fwd(const int64_t* ids, const float* tok_emb, const float* pos, void* x, int B, int L, int D,
                                   avn_stream_t stream_) {
    AVN_REQUIRE(ids && tok_emb && pos && x && B > 0 && L > 0 && D % 4 == 0, "token_embed_fwd: bad arguments");
    const long long n = (long long)B * L * (D / 4);
    token_embed_fwd_kernel<<<(int)((n + 255) / 256), 256, 0, (cudaStream_t)stream_>>>(ids, tok_emb, pos, (__nv_bfloat16*)x, (long long)B * L, L, D);
    AVN_CHECK_CUDA(cudaGetLastError());
    return AVN_OK;
}

extern "C" int avn_token_embed_bwd(const int64_t* ids, const void* dx, float* dtok_emb, float* dpos, int B, int L, int D,
                                   avn_stream_t stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    AVN_REQUIRE(ids && dx && dtok_emb && dpos && B > 0 && L > 0 && D % 4 == 0, "token_embed_bwd: bad arguments");
    const long long n = (long long)B * L * (D / 4);
    token_embed_bwd_kernel<<<(int)((n + 255) / 256), 256, 0, stream>>>(ids, (const __nv_bfloat16*)dx, dtok_emb, (long long)B * L, D);
    AVN_CHECK_CUDA(cudaGetLastError());
    const long long n2 = (long long)L * D;
    pos_embed_bwd_kernel<<<(int)((n2 + 255) / 256), 256, 0, stream>>>((const __nv_bfloat16*)dx, dpos, B, L, D);
    AVN_CHECK_CUDA(cudaGetLastError());
    return AVN_OK;
}
