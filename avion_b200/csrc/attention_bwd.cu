// V4 backward: flash-style attention backward on tcgen05, head dim 64.
// Replaces the backward of flash_attn_qkvpacked_func (avion/models/transformer.py:148) / nn.MultiheadAttention (:140-146).
//
// One CTA per (clip, head, 128-row K/V tile j); it loops over the query tiles i and issues five MMAs per (i, j):
//   S  = Q_i K_j^T          dP = dO_i V_j^T                          (recompute; TMEM, 128 columns each)
//   dV += P^T dO_i          dK += dS^T Q_i                           (accumulate in TMEM across i, 64 columns each)
//   dQ_i(partial) = dS K_j                                           (TMEM -> fp32 smem -> TMA reduce-add into dq_acc)
// with P = exp(scale*S - LSE), dS = scale * P o (dP - delta) computed by 8 warps (thread = (query row, column half)).
// P and dS are written once to 128B-swizzled smem as [q][kv] bf16 and consumed both as a K-major A operand (dQ) and as
// an MN-major A operand (P^T, dS^T) - the same bytes, two descriptors.
#include "common.cuh"
#include "../../include/avion_b200.h"

namespace {
using namespace avn;

constexpr int TILE = 128, HD = 64;
constexpr int NTHREADS = 384;     // warps 0-7 math, 8 MMA issuer, 9 TMA producer, 10-11 idle (complete the warpgroup)
constexpr int T16 = TILE * HD * 2;                // 16 KB: one [128][64] bf16 tile
constexpr int OFF_K = 0, OFF_V = T16;
constexpr int OFF_Q = 2 * T16;                    // 2 stages
constexpr int OFF_DO = 4 * T16;                   // 2 stages
constexpr int OFF_P = 6 * T16;                    // 2 buffers x 32 KB ([q][kv] bf16, two 64-column sub-tiles each)
constexpr int OFF_DS = 10 * T16;                  // 2 buffers x 32 KB
constexpr int OFF_BAR = 14 * T16;
constexpr int SMEM_TOTAL = OFF_BAR + 256 + 1024;
constexpr uint32_t TM_S = 0, TM_DP = 128, TM_DV = 256, TM_DK = 320, TM_DQ = 384;
// The P buffer of tile t doubles as the fp32 staging tile of dQ_t once the MMAs of tile t have completed.

struct Bars {
    uint64_t kv_full, qdo_full[2], qdo_empty[2], sdp_full, sdp_free, pds_ready, dq_full, dq_free;
    uint32_t tmem_ptr;
};


// One 32-column chunk of a row: P = exp2(S*scale_log2 - lse_l2), dS = scale * P o (dP - delta) -> packed bf16 pairs.
template <bool MASK>
__device__ __forceinline__ void pds_chunk(const uint32_t* sv, const uint32_t* dv, uint32_t* pk, uint32_t* dsk, float scale_log2,
                                          float lse_l2, float dlt, float scale, int col0, int limit) {
    const float2 sc2 = make_float2(scale_log2, scale_log2), nl2 = make_float2(-lse_l2, -lse_l2);
    const float2 nd2 = make_float2(-dlt * scale, -dlt * scale), s2 = make_float2(scale, scale);
#pragma unroll
    for (int i = 0; i < 32; i += 2) {
        const float2 x = __ffma2_rn(make_float2(__uint_as_float(sv[i]), __uint_as_float(sv[i + 1])), sc2, nl2);
        float2 p = make_float2(ex2(x.x), ex2(x.y));
        if (MASK) {
            if (col0 + i >= limit) p.x = 0.f;
            if (col0 + i + 1 >= limit) p.y = 0.f;
        }
        const float2 t = __ffma2_rn(make_float2(__uint_as_float(dv[i]), __uint_as_float(dv[i + 1])), s2, nd2);   // scale*(dP - delta)
        const float2 d = __fmul2_rn(p, t);
        pk[i >> 1] = pack_bf16(p.x, p.y);
        dsk[i >> 1] = pack_bf16(d.x, d.y);
    }
}

__global__ void __launch_bounds__(NTHREADS, 1)
attn_bwd_kernel(const __grid_constant__ CUtensorMap tmQKV, const __grid_constant__ CUtensorMap tmDO,
                const __grid_constant__ CUtensorMap tmDQKV, const __grid_constant__ CUtensorMap tmDQ,
                const float* __restrict__ lse, const float* __restrict__ delta, int S, int H, float scale, int causal) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
    Bars* bars = (Bars*)(smem + OFF_BAR);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int jt = blockIdx.x, h = blockIdx.y, b = blockIdx.z;
    const int q_tiles = (S + TILE - 1) / TILE;
    const int i0 = causal ? jt : 0;
    const int n_i = q_tiles - i0;
    const int kv0 = jt * TILE;

    if (threadIdx.x == 0) {
        tma_prefetch_desc(&tmQKV); tma_prefetch_desc(&tmDO); tma_prefetch_desc(&tmDQKV); tma_prefetch_desc(&tmDQ);
        mbar_init(&bars->kv_full, 1);
        for (int i = 0; i < 2; ++i) { mbar_init(&bars->qdo_full[i], 1); mbar_init(&bars->qdo_empty[i], 1); }
        mbar_init(&bars->sdp_full, 1); mbar_init(&bars->sdp_free, 256);
        mbar_init(&bars->pds_ready, 256); mbar_init(&bars->dq_full, 1); mbar_init(&bars->dq_free, 256);
        fence_barrier_init();
    }
    if (warp == 8) tmem_alloc(&bars->tmem_ptr, 512);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = bars->tmem_ptr;

    // the 8 math warps take the registers of the control warpgroup (setmaxnreg is per warpgroup: warps 8-11 all shrink)
    if (warp >= 10) {
        reg_dec<40>();       // idle warps: only complete the third warpgroup
    } else if (warp == 9) {
        // ===================== TMA producer =====================
        // (whole warp loops and waits; one elected lane issues, so addresses and coordinates stay in uniform registers)
        reg_dec<40>();
        if (elect_one()) {
            mbar_expect_tx(&bars->kv_full, 2 * T16);
            tma_load_3d(smem + OFF_K, &tmQKV, &bars->kv_full, (H + h) * HD, kv0, b);
            tma_load_3d(smem + OFF_V, &tmQKV, &bars->kv_full, (2 * H + h) * HD, kv0, b);
        }
        __syncwarp();
        for (int t = 0; t < n_i; ++t) {
            const int st = t & 1;
            mbar_wait(&bars->qdo_empty[st], ((t >> 1) & 1) ^ 1);
            if (elect_one()) {
                mbar_expect_tx(&bars->qdo_full[st], 2 * T16);
                tma_load_3d(smem + OFF_Q + st * T16, &tmQKV, &bars->qdo_full[st], h * HD, (i0 + t) * TILE, b);
                tma_load_3d(smem + OFF_DO + st * T16, &tmDO, &bars->qdo_full[st], h * HD, (i0 + t) * TILE, b);
            }
            __syncwarp();
        }
    } else if (warp == 8) {
        // ===================== MMA issuer =====================
        // waits are executed by the whole warp, the MMAs by one elected lane (descriptors stay in uniform registers)
        reg_dec<40>();
        if (n_i > 0) {
            constexpr uint32_t idesc_s = umma_idesc_bf16(TILE, TILE, false, false);
            constexpr uint32_t idesc_dkv = umma_idesc_bf16(TILE, HD, true, true);
            constexpr uint32_t idesc_dq = umma_idesc_bf16(TILE, HD, false, true);
            const uint32_t ka = smem_u32(smem + OFF_K), va = smem_u32(smem + OFF_V);
            auto issue_sdp = [&](int t) {
                const int st = t & 1;
                mbar_wait(&bars->qdo_full[st], (t >> 1) & 1);
                if (t > 0) mbar_wait(&bars->sdp_free, (t - 1) & 1);
                tc_fence_after();
                if (elect_one()) {
                    const uint32_t qa = smem_u32(smem + OFF_Q + st * T16), doa = smem_u32(smem + OFF_DO + st * T16);
#pragma unroll
                    for (int k = 0; k < HD / 16; ++k)
                        umma_ss(tmem_base + TM_S, umma_smem_desc(qa + k * 32, 16, 1024), umma_smem_desc(ka + k * 32, 16, 1024), idesc_s, k > 0);
#pragma unroll
                    for (int k = 0; k < HD / 16; ++k)
                        umma_ss(tmem_base + TM_DP, umma_smem_desc(doa + k * 32, 16, 1024), umma_smem_desc(va + k * 32, 16, 1024), idesc_s, k > 0);
                    umma_commit(&bars->sdp_full);
                }
                __syncwarp();
            };
            mbar_wait(&bars->kv_full, 0);
            issue_sdp(0);
            for (int t = 0; t < n_i; ++t) {
                const int st = t & 1;
                if (t + 1 < n_i) issue_sdp(t + 1);
                mbar_wait(&bars->pds_ready, t & 1);
                tc_fence_after();
                const uint32_t qa = smem_u32(smem + OFF_Q + st * T16), doa = smem_u32(smem + OFF_DO + st * T16);
                const uint32_t pa = smem_u32(smem + OFF_P + st * 2 * T16), dsa = smem_u32(smem + OFF_DS + st * 2 * T16);
                if (elect_one()) {
#pragma unroll
                    for (int k = 0; k < TILE / 16; ++k)      // dV += P^T dO   (A = P as MN-major, B = dO as MN-major)
                        umma_ss(tmem_base + TM_DV, umma_smem_desc(pa + k * 2048, 16384, 1024), umma_smem_desc(doa + k * 2048, 8192, 1024),
                                idesc_dkv, (t > 0 || k > 0) ? 1u : 0u);
#pragma unroll
                    for (int k = 0; k < TILE / 16; ++k)      // dK += dS^T Q
                        umma_ss(tmem_base + TM_DK, umma_smem_desc(dsa + k * 2048, 16384, 1024), umma_smem_desc(qa + k * 2048, 8192, 1024),
                                idesc_dkv, (t > 0 || k > 0) ? 1u : 0u);
                }
                __syncwarp();
                if (t > 0) { mbar_wait(&bars->dq_free, (t - 1) & 1); tc_fence_after(); }   // dQ_{t-1} has left TMEM
                if (elect_one()) {
#pragma unroll
                    for (int k = 0; k < TILE / 16; ++k)      // dQ_i = dS K_j  (A = dS K-major, B = K_j as MN-major)
                        umma_ss(tmem_base + TM_DQ, umma_smem_desc(dsa + (k >> 2) * 16384 + (k & 3) * 32, 16, 1024),
                                umma_smem_desc(ka + k * 2048, 8192, 1024), idesc_dq, k > 0 ? 1u : 0u);
                    umma_commit(&bars->dq_full);
                    umma_commit(&bars->qdo_empty[st]);
                }
                __syncwarp();
            }
        }
    } else {
        // ===================== compute warps: thread = (row, column half) =====================
        reg_inc<232>();
        if (n_i > 0) {
        const int half = warp >> 2;
        const int row = (warp & 3) * 32 + lane;
        const uint32_t lane_addr = (uint32_t)((warp & 3) * 32) << 16;
        const float scale_log2 = scale * 1.4426950408889634f;
        const bool leader = (threadIdx.x & 127) == 0;
        const long long stat_base = ((long long)b * H + h) * S;

        // partial dQ of query tile t: TMEM -> fp32 staging (the P buffer of tile t, free now) -> TMA reduce-add
        auto drain_dq = [&](int t) {
            mbar_wait(&bars->dq_full, t & 1);      // all MMAs of tile t are complete
            tc_fence_after();
            uint32_t v[32];
            tmem_ld32(tmem_base + lane_addr + TM_DQ + half * 32, v);
            tmem_ld_wait();
            tc_fence_before();
            mbar_arrive(&bars->dq_free);
            uint8_t* stg = smem + OFF_P + (t & 1) * 2 * T16 + half * T16;
#pragma unroll
            for (int i = 0; i < 8; ++i)
                st_shared_v4(stg + row * 128 + ((i ^ (row & 7)) << 4), make_uint4(v[4 * i], v[4 * i + 1], v[4 * i + 2], v[4 * i + 3]));
            fence_proxy_async_smem();
            named_bar_sync(1 + half, 128);
            if (leader) { tma_reduce_add_3d(&tmDQ, stg, h * HD + half * 32, (i0 + t) * TILE, b); tma_commit_group(); }
        };

        float lse_nx = 0.f, dlt_nx = 0.f;
        { const int qr = i0 * TILE + row; if (qr < S) { lse_nx = __ldg(lse + stat_base + qr); dlt_nx = __ldg(delta + stat_base + qr); } }
        for (int t = 0; t < n_i; ++t) {
            const int q0 = (i0 + t) * TILE;
            const int qrow = q0 + row;
            const bool row_ok = qrow < S;
            const float lse_l2 = lse_nx * 1.4426950408889634f, dlt = dlt_nx;
            if (t + 1 < n_i) {                      // prefetch the next tile's row statistics
                const int qr = qrow + TILE;
                lse_nx = 0.f; dlt_nx = 0.f;
                if (qr < S) { lse_nx = __ldg(lse + stat_base + qr); dlt_nx = __ldg(delta + stat_base + qr); }
            }
            // valid columns of this thread's row: [0, limit) in tile coordinates
            const int limit = row_ok ? (causal ? min(S - kv0, qrow - kv0 + 1) : (S - kv0)) : 0;
            const bool need_mask = (kv0 + TILE > S) || (q0 + TILE > S) || (causal && kv0 + TILE - 1 > q0);   // warp-uniform
            mbar_wait(&bars->sdp_full, t & 1);
            tc_fence_after();
            uint32_t pk[32], dsk[32];     // packed bf16: 64 P values and 64 dS values of this thread
#pragma unroll
            for (int c = 0; c < 2; ++c) {
                uint32_t sv[32], dv[32];
                tmem_ld32(tmem_base + lane_addr + TM_S + half * 64 + c * 32, sv);
                tmem_ld32(tmem_base + lane_addr + TM_DP + half * 64 + c * 32, dv);
                tmem_ld_wait();
                if (c == 1) { tc_fence_before(); mbar_arrive(&bars->sdp_free); }
                if (need_mask) pds_chunk<true>(sv, dv, pk + c * 16, dsk + c * 16, scale_log2, lse_l2, dlt, scale, half * 64 + c * 32, limit);
                else pds_chunk<false>(sv, dv, pk + c * 16, dsk + c * 16, scale_log2, lse_l2, dlt, scale, half * 64 + c * 32, limit);
            }
            // buffer t&1 last held P/dS of tile t-2 (its MMAs completed before drain_dq(t-2) returned) and, in its P half,
            // the staging tile of dQ_{t-2}: the TMA reduce must have finished reading it
            if (t >= 2) {
                if (leader) tma_wait_group_read0();
                named_bar_sync(1 + half, 128);
            }
            uint8_t* prow = smem + OFF_P + (t & 1) * 2 * T16 + half * T16 + row * 128;
            uint8_t* drow = smem + OFF_DS + (t & 1) * 2 * T16 + half * T16 + row * 128;
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                const int off = (i ^ (row & 7)) << 4;
                st_shared_v4(prow + off, make_uint4(pk[4 * i], pk[4 * i + 1], pk[4 * i + 2], pk[4 * i + 3]));
                st_shared_v4(drow + off, make_uint4(dsk[4 * i], dsk[4 * i + 1], dsk[4 * i + 2], dsk[4 * i + 3]));
            }
            fence_proxy_async_smem();
            mbar_arrive(&bars->pds_ready);
            if (t > 0) drain_dq(t - 1);            // overlaps the MMAs of tile t
        }
        drain_dq(n_i - 1);      // also guarantees that dV and dK are complete
        // dV, dK: TMEM -> bf16 -> swizzled smem (the dS buffers are free now and never touched by the dQ reduces) -> TMA store
        {
            uint32_t v[32], k[32];
            tmem_ld32(tmem_base + lane_addr + TM_DV + half * 32, v);
            tmem_ld32(tmem_base + lane_addr + TM_DK + half * 32, k);
            tmem_ld_wait();
            tc_fence_before();
            uint8_t* vrow = smem + OFF_DS + row * 128;
            uint8_t* krow = smem + OFF_DS + 2 * T16 + row * 128;
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                const int off = ((half * 4 + i) ^ (row & 7)) << 4;
                st_shared_v4(vrow + off, make_uint4(pack_bf16(__uint_as_float(v[8 * i]), __uint_as_float(v[8 * i + 1])),
                                                   pack_bf16(__uint_as_float(v[8 * i + 2]), __uint_as_float(v[8 * i + 3])),
                                                   pack_bf16(__uint_as_float(v[8 * i + 4]), __uint_as_float(v[8 * i + 5])),
                                                   pack_bf16(__uint_as_float(v[8 * i + 6]), __uint_as_float(v[8 * i + 7]))));
                st_shared_v4(krow + off, make_uint4(pack_bf16(__uint_as_float(k[8 * i]), __uint_as_float(k[8 * i + 1])),
                                                   pack_bf16(__uint_as_float(k[8 * i + 2]), __uint_as_float(k[8 * i + 3])),
                                                   pack_bf16(__uint_as_float(k[8 * i + 4]), __uint_as_float(k[8 * i + 5])),
                                                   pack_bf16(__uint_as_float(k[8 * i + 6]), __uint_as_float(k[8 * i + 7]))));
            }
            fence_proxy_async_smem();
            named_bar_sync(3, 256);
            if (threadIdx.x == 0) {
                tma_store_3d(&tmDQKV, smem + OFF_DS + 2 * T16, (H + h) * HD, kv0, b);
                tma_store_3d(&tmDQKV, smem + OFF_DS, (2 * H + h) * HD, kv0, b);
                tma_commit_group();
            }
            if (leader) tma_wait_group0();
        }
        }   // n_i > 0
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 8) tmem_dealloc(tmem_base, 512);
}

// delta[b][h][s] = sum_d dO[b][s][h][d] * O[b][s][h][d]
__global__ void __launch_bounds__(128)
attn_delta_kernel(const __nv_bfloat16* __restrict__ o, const __nv_bfloat16* __restrict__ dout, float* __restrict__ delta, int S, int H) {
    const long long rowi = blockIdx.x;          // b*S + s
    const int b = (int)(rowi / S), s = (int)(rowi % S);
    const uint4* op = (const uint4*)(o + rowi * H * HD);
    const uint4* dp = (const uint4*)(dout + rowi * H * HD);
    for (int v = threadIdx.x; v < ((H * 8 + 31) & ~31); v += blockDim.x) {   // warp-uniform trip count
        float acc = 0.f;
        if (v < H * 8) {
            const uint4 a = __ldg(op + v), d = __ldg(dp + v);
            const uint32_t aw[4] = {a.x, a.y, a.z, a.w}, dw[4] = {d.x, d.y, d.z, d.w};
#pragma unroll
            for (int e = 0; e < 4; ++e) { const float2 x = unpack_bf16(aw[e]), y = unpack_bf16(dw[e]); acc = fmaf(x.x, y.x, fmaf(x.y, y.y, acc)); }
        }
        acc += __shfl_xor_sync(0xffffffffu, acc, 1);
        acc += __shfl_xor_sync(0xffffffffu, acc, 2);
        acc += __shfl_xor_sync(0xffffffffu, acc, 4);
        if ((v & 7) == 0 && v < H * 8) delta[((long long)b * H + (v >> 3)) * S + s] = acc;
    }
}

// dq_acc fp32 [B*S][H*64] -> dqkv[..., 0, :, :] bf16 (row stride 3*H*64)
__global__ void dq_convert_kernel(const float* __restrict__ dq, __nv_bfloat16* __restrict__ dqkv, long long rows, int HW) {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;   // one thread per 8 elements
    const int nv = HW >> 3;
    if (i >= rows * nv) return;
    const long long r = i / nv; const int c = (int)(i % nv) * 8;
    const float4 a = __ldg((const float4*)(dq + r * HW + c)), bq = __ldg((const float4*)(dq + r * HW + c + 4));
    *(uint4*)(dqkv + r * 3 * HW + c) = make_uint4(pack_bf16(a.x, a.y), pack_bf16(a.z, a.w), pack_bf16(bq.x, bq.y), pack_bf16(bq.z, bq.w));
}
}  // namespace

extern "C" int avn_attn_bwd(const void* qkv, const void* out, const void* dout, const float* lse, void* dqkv, float* delta,
                            float* dq_acc, int B, int S, int H, float scale, int causal, avn_stream_t stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    AVN_REQUIRE(qkv && out && dout && lse && dqkv && delta && dq_acc, "attn_bwd: null pointer");
    AVN_REQUIRE(B > 0 && S > 0 && H > 0, "attn_bwd: bad sizes");
    CUtensorMap tmQKV, tmDO, tmDQKV, tmDQ;
    int rc;
    const uint64_t W3 = (uint64_t)3 * H * HD, W1 = (uint64_t)H * HD;
    {
        uint64_t dims[3] = {W3, (uint64_t)S, (uint64_t)B};
        uint64_t str[2] = {W3 * 2, (uint64_t)S * W3 * 2};
        uint32_t box[3] = {HD, TILE, 1};
        if ((rc = avn_make_tmap(&tmQKV, qkv, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, 3, dims, str, box, true))) return rc;
        if ((rc = avn_make_tmap(&tmDQKV, dqkv, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, 3, dims, str, box, true))) return rc;
    }
    {
        uint64_t dims[3] = {W1, (uint64_t)S, (uint64_t)B};
        uint64_t str[2] = {W1 * 2, (uint64_t)S * W1 * 2};
        uint32_t box[3] = {HD, TILE, 1};
        if ((rc = avn_make_tmap(&tmDO, dout, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, 3, dims, str, box, true))) return rc;
    }
    {
        uint64_t dims[3] = {W1, (uint64_t)S, (uint64_t)B};
        uint64_t str[2] = {W1 * 4, (uint64_t)S * W1 * 4};
        uint32_t box[3] = {32, TILE, 1};
        if ((rc = avn_make_tmap(&tmDQ, dq_acc, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 4, 3, dims, str, box, true))) return rc;
    }
    const long long rows = (long long)B * S;
    AVN_CHECK_CUDA(cudaMemsetAsync(dq_acc, 0, (size_t)rows * W1 * sizeof(float), stream));
    attn_delta_kernel<<<(unsigned)rows, 128, 0, stream>>>((const __nv_bfloat16*)out, (const __nv_bfloat16*)dout, delta, S, H);
    AVN_CHECK_CUDA(cudaGetLastError());
    const int tiles = (S + TILE - 1) / TILE;
    AVN_CHECK_CUDA(cudaFuncSetAttribute(attn_bwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_TOTAL));
    attn_bwd_kernel<<<dim3(tiles, H, B), NTHREADS, SMEM_TOTAL, stream>>>(tmQKV, tmDO, tmDQKV, tmDQ, lse, delta, S, H, scale, causal);
    AVN_CHECK_CUDA(cudaGetLastError());
    const long long nvec = rows * (W1 / 8);
    dq_convert_kernel<<<(unsigned)((nvec + 255) / 256), 256, 0, stream>>>(dq_acc, (__nv_bfloat16*)dqkv, rows, (int)W1);
    AVN_CHECK_CUDA(cudaGetLastError());
    return AVN_OK;
}
