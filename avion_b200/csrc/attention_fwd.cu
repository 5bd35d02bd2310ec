// V4 forward: flash-style attention on tcgen05, head dim 64 (non-causal for the video tower, causal for the text tower).
// Replaces flash_attn_qkvpacked_func (reached from avion/models/transformer.py:148) and the masked
// nn.MultiheadAttention of the text tower (:140-146,447-453).
//
// One CTA per (clip, head, PAIR of 128-row query tiles).  Roles:
//   warps 0-3 / 4-7 : softmax group 0 / 1 (one query row per thread: S row read once from TMEM, online softmax in fp32
//                     registers with lazy rescaling, P written to swizzled smem as bf16, O accumulated in TMEM)
//   warps 8 / 10    : MMA issuer of group 0 / 1 (S_w = Q_w K_j^T and O_w = P_w V_j on tcgen05, accumulators in TMEM)
//   warp 9          : TMA producer (Q tiles once, then a 3-stage K/V ring shared by both groups)
// The two groups run half a tile period apart: while one exponentiates (MUFU bound at head dim 64) the other one loads its
// S tile, finds the row maxima, stores P and waits; each group has its own MMA issuer so that nothing re-aligns them.
#include "common.cuh"
#include "../../include/avion_b200.h"

namespace {
using namespace avn;

constexpr int TILE = 128, HD = 64, KV_STAGES = 4;
constexpr int NTHREADS = 384;     // warps 0-7 math, 8 MMA issuer, 9 TMA producer, 10-11 idle (complete the warpgroup)
constexpr int Q_BYTES = TILE * HD * 2;            // 16 KB
constexpr int KV_BYTES = TILE * HD * 2;           // 16 KB each for K and V
constexpr int OFF_Q = 0;
constexpr int OFF_K = OFF_Q + 2 * Q_BYTES;
constexpr int OFF_V = OFF_K + KV_STAGES * KV_BYTES;
constexpr int OFF_BAR = OFF_V + KV_STAGES * KV_BYTES;
constexpr int SMEM_TOTAL = OFF_BAR + 256 + 1024;
// TMEM columns: S_w (fp32) at w*128, O_w (fp32) at 256 + w*64, P_w (bf16 pairs, the A operand of P V) at 384 + w*64
constexpr uint32_t TM_S = 0, TM_O = 256, TM_P = 384;
#ifndef AVN_FWD_LAG
#define AVN_FWD_LAG 1000
#endif
constexpr long long G1_LAG_CYCLES = AVN_FWD_LAG;      // head start of softmax group 0 over group 1

// Optional phase trace (debug aid, off unless avn_debug_set_attn_fwd_trace is given a buffer): CTA (0,0,0) records clock64 at
// phase boundaries as trace[(role * 32 + iteration) * 16 + slot]; role 0 / 1 = first warp of softmax group 0 / 1.
// Compiled in only with -DAVN_TRACE (the extra predicated stores cost ~4 % in the softmax loop); the default build has no
// device global and no debug entry point.
#ifdef AVN_TRACE
__device__ long long* g_trace_fwd = nullptr;
#define TRF(role, it, slot)                                                                            \
    do {                                                                                               \
        if (trace && lane == 0 && (it) < 32) trace[((role) * 32 + (it)) * 16 + (slot)] = clock64();   \
    } while (0)
#else
#define TRF(role, it, slot) do { } while (0)
#endif

struct Bars {
    uint64_t q_full;
    uint64_t kv_full[KV_STAGES], kv_empty[KV_STAGES];
    uint64_t s_full[2], s_free[2], p_ready[2], o_full[2];
    uint32_t tmem_ptr;
};


template <bool MASK>
__device__ __forceinline__ float row_max(const uint32_t* sv, int limit) {
    float a = -INFINITY, b = -INFINITY, c = -INFINITY, d = -INFINITY;     // four independent chains
#pragma unroll
    for (int i = 0; i < TILE; i += 4) {
        float v0 = __uint_as_float(sv[i]), v1 = __uint_as_float(sv[i + 1]), v2 = __uint_as_float(sv[i + 2]), v3 = __uint_as_float(sv[i + 3]);
        if (MASK) {
            if (i >= limit) v0 = -INFINITY;
            if (i + 1 >= limit) v1 = -INFINITY;
            if (i + 2 >= limit) v2 = -INFINITY;
            if (i + 3 >= limit) v3 = -INFINITY;
        }
        a = fmaxf(a, v0); b = fmaxf(b, v1); c = fmaxf(c, v2); d = fmaxf(d, v3);
    }
    return fmaxf(fmaxf(a, b), fmaxf(c, d));
}

// Of every AVN_FWD_POLY_DEN score pairs, AVN_FWD_POLY_NUM take the FMA-pipe exponential: the MUFU pipe (16 results per clock
// per SM) is what bounds this kernel at head dim 64, and its instructions queue behind TMEM and shared-memory traffic.
#ifndef AVN_FWD_POLY_NUM
#define AVN_FWD_POLY_NUM 1
#endif
#ifndef AVN_FWD_POLY_DEN
#define AVN_FWD_POLY_DEN 3
#endif
// POLY = false: every exponential on the MUFU.  Used for short sequences (one or two K/V tiles: the tiny parity fixtures, the
// text tower), where the kernel is latency-bound and the FMA-pipe exponentials buy nothing.
template <bool POLY>
__device__ __forceinline__ float2 exp2_pair(float2 x, int pair_index) {
    if (POLY && (pair_index % AVN_FWD_POLY_DEN) < AVN_FWD_POLY_NUM) return exp2_poly2(x);
    return make_float2(ex2(x.x), ex2(x.y));
}

// p = exp2(s * scale_log2 - m_ref) for one row of 128 scores -> packed bf16 pairs; returns the row sum
template <bool MASK, bool POLY>
__device__ __forceinline__ float row_exp(const uint32_t* sv, uint32_t* pk, float scale_log2, float m_ref, int limit) {
    const float2 sc2 = make_float2(scale_log2, scale_log2), nm2 = make_float2(-m_ref, -m_ref);
    float2 r0 = make_float2(0.f, 0.f), r1 = make_float2(0.f, 0.f);
#pragma unroll
    for (int i = 0; i < TILE / 2; i += 2) {
        const float2 x0 = __ffma2_rn(make_float2(__uint_as_float(sv[2 * i]), __uint_as_float(sv[2 * i + 1])), sc2, nm2);
        const float2 x1 = __ffma2_rn(make_float2(__uint_as_float(sv[2 * i + 2]), __uint_as_float(sv[2 * i + 3])), sc2, nm2);
        float2 e0 = exp2_pair<POLY>(x0, i), e1 = exp2_pair<POLY>(x1, i + 1);
        if (MASK) {
            if (2 * i >= limit) e0.x = 0.f;
            if (2 * i + 1 >= limit) e0.y = 0.f;
            if (2 * i + 2 >= limit) e1.x = 0.f;
            if (2 * i + 3 >= limit) e1.y = 0.f;
        }
        r0 = __fadd2_rn(r0, e0); r1 = __fadd2_rn(r1, e1);
        pk[i] = pack_bf16(e0.x, e0.y); pk[i + 1] = pack_bf16(e1.x, e1.y);
    }
    return (r0.x + r0.y) + (r1.x + r1.y);
}

template <bool POLY>
__global__ void __launch_bounds__(NTHREADS, 1)
attn_fwd_kernel(const __grid_constant__ CUtensorMap tmQKV, const __grid_constant__ CUtensorMap tmO, float* __restrict__ lse,
                int S, int H, float scale_log2, int causal) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
    Bars* bars = (Bars*)(smem + OFF_BAR);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int b = blockIdx.z, h = blockIdx.y;
    const int q_tiles = (S + TILE - 1) / TILE;
    const int kv_tiles = q_tiles;
    const int qt[2] = {(int)blockIdx.x * 2, (int)blockIdx.x * 2 + 1};
    int nkv[2];
#pragma unroll
    for (int w = 0; w < 2; ++w) nkv[w] = qt[w] >= q_tiles ? 0 : (causal ? min(kv_tiles, qt[w] + 1) : kv_tiles);
    const int nkv_max = max(nkv[0], nkv[1]);
#ifdef AVN_TRACE
    long long* trace = (blockIdx.x == 0 && blockIdx.y == 0 && blockIdx.z == 0 && (warp & 3) == 0 && warp < 8) ? g_trace_fwd : nullptr;
#endif

    if (threadIdx.x == 0) {
        tma_prefetch_desc(&tmQKV);
        tma_prefetch_desc(&tmO);
        mbar_init(&bars->q_full, 1);
        for (int i = 0; i < KV_STAGES; ++i) { mbar_init(&bars->kv_full[i], 1); mbar_init(&bars->kv_empty[i], 2); }
        for (int w = 0; w < 2; ++w) {
            mbar_init(&bars->s_full[w], 1); mbar_init(&bars->s_free[w], 128);
            mbar_init(&bars->p_ready[w], 128); mbar_init(&bars->o_full[w], 1);
        }
        fence_barrier_init();
    }
    if (warp == 8) tmem_alloc(&bars->tmem_ptr, 512);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = bars->tmem_ptr;

    // the 8 math warps take the registers of the control warpgroup (setmaxnreg is per warpgroup: warps 8-11 all shrink)
    if (warp == 11) {
        reg_dec<40>();       // idle warp: only completes the third warpgroup
    } else if (warp == 9) {
        // ===================== TMA producer =====================
        // (the whole warp runs the loop and one elected lane issues, so addresses and coordinates stay in uniform registers)
        reg_dec<40>();
        const int n_q = (qt[1] < q_tiles) ? 2 : 1;
        if (elect_one()) {
            mbar_expect_tx(&bars->q_full, n_q * Q_BYTES);
            for (int w = 0; w < n_q; ++w) tma_load_3d(smem + OFF_Q + w * Q_BYTES, &tmQKV, &bars->q_full, h * HD, qt[w] * TILE, b);
        }
        __syncwarp();
        for (int j = 0; j < nkv_max; ++j) {
            const int st = j % KV_STAGES;
            mbar_wait(&bars->kv_empty[st], ((j / KV_STAGES) & 1) ^ 1);
            if (elect_one()) {
                mbar_expect_tx(&bars->kv_full[st], 2 * KV_BYTES);
                tma_load_3d(smem + OFF_K + st * KV_BYTES, &tmQKV, &bars->kv_full[st], (H + h) * HD, j * TILE, b);
                tma_load_3d(smem + OFF_V + st * KV_BYTES, &tmQKV, &bars->kv_full[st], (2 * H + h) * HD, j * TILE, b);
            }
            __syncwarp();
        }
    } else if (warp == 8 || warp == 10) {
        // ===================== MMA issuers: one warp per softmax group =====================
        // Two independent issuers keep the groups decoupled: with a single in-order issuer a group that runs behind stalls
        // the other one's next S tile, which pulls both into lockstep - and then their MUFU-free phases (S load, row
        // maxima, P store, waits) coincide and the MUFU pipe idles a third of the time.
        reg_dec<40>();
        const int w = (warp == 8) ? 0 : 1;
        if (nkv_max > 0) {
            constexpr uint32_t idesc_s = umma_idesc_bf16(TILE, TILE, false, false);
            constexpr uint32_t idesc_o = umma_idesc_bf16(TILE, HD, false, true);
            // waits are executed by the whole warp, the MMAs by one elected lane (descriptors stay in uniform registers)
            auto issue_s = [&](int j) {
                if (j > 0) { mbar_wait(&bars->s_free[w], (j - 1) & 1); tc_fence_after(); }
                if (elect_one()) {
                    const uint32_t qa = smem_u32(smem + OFF_Q + w * Q_BYTES);
                    const uint32_t ka = smem_u32(smem + OFF_K + (j % KV_STAGES) * KV_BYTES);
#pragma unroll
                    for (int k = 0; k < HD / 16; ++k)
                        umma_ss(tmem_base + TM_S + w * TILE, umma_smem_desc(qa + k * 32, 16, 1024), umma_smem_desc(ka + k * 32, 16, 1024),
                                idesc_s, k > 0 ? 1u : 0u);
                    umma_commit(&bars->s_full[w]);
                }
                __syncwarp();
            };
            mbar_wait(&bars->q_full, 0);
            mbar_wait(&bars->kv_full[0], 0);
            tc_fence_after();
            if (nkv[w] > 0) issue_s(0);
            for (int j = 0; j < nkv_max; ++j) {
                const int st = j % KV_STAGES;
                if (j + 1 < nkv_max) {
                    mbar_wait(&bars->kv_full[(j + 1) % KV_STAGES], ((j + 1) / KV_STAGES) & 1);
                    tc_fence_after();
                    if (j + 1 < nkv[w]) issue_s(j + 1);
                }
                if (j < nkv[w]) {
                    mbar_wait(&bars->p_ready[w], j & 1);
                    tc_fence_after();
                    if (elect_one()) {
                        const uint32_t va = smem_u32(smem + OFF_V + st * KV_BYTES);
#pragma unroll
                        for (int k = 0; k < TILE / 16; ++k)      // O_w += P_w V_j : A = P from TMEM (8 columns per K=16), B = V (MN-major)
                            umma_ts(tmem_base + TM_O + w * HD, tmem_base + TM_P + w * 64 + k * 8, umma_smem_desc(va + k * 2048, 8192, 1024),
                                    idesc_o, (j > 0 || k > 0) ? 1u : 0u);
                        umma_commit(&bars->o_full[w]);
                    }
                    __syncwarp();
                }
                // both issuers release every K/V stage (one that has no work for tile j arrives as soon as the tile is loaded)
                if (elect_one()) umma_commit(&bars->kv_empty[st]);
                __syncwarp();
            }
        }
    } else {
        // ===================== softmax groups =====================
        reg_inc<232>();
        const int w = warp >> 2;
        const int row = (warp & 3) * 32 + lane;
        const int n_my = nkv[w];
        if (n_my > 0) {
            const int q0 = qt[w] * TILE;
            const int qrow = q0 + row;
            const uint32_t t_s = tmem_base + ((uint32_t)((warp & 3) * 32) << 16) + TM_S + w * TILE;
            const uint32_t t_o = tmem_base + ((uint32_t)((warp & 3) * 32) << 16) + TM_O + w * HD;
            const uint32_t t_p = tmem_base + ((uint32_t)((warp & 3) * 32) << 16) + TM_P + w * 64;
            uint8_t* sO = smem + OFF_Q + w * Q_BYTES;      // O staging for the TMA store: this group's Q tile is dead by then
            // O accumulates in TMEM across K/V tiles (the MMA adds P_j V_j).  The softmax reference m_ref is only raised
            // when the running row maximum grows by more than 2^8 (lazy rescaling): then O (in TMEM) and l are rescaled.
            float m_ref = -INFINITY, l = 0.f;
            for (int j = 0; j < n_my; ++j) {
                const int kv0 = j * TILE;
                const bool need_mask = (kv0 + TILE > S) || (causal && kv0 + TILE - 1 > q0);
                const int limit = causal ? min(S - kv0, qrow - kv0 + 1) : (S - kv0);   // columns [0, limit) are valid
                TRF(w, j, 0);
                mbar_wait(&bars->s_full[w], j & 1);
                // Group 1 starts its first tile half a period late, so that its exponentials fall into group 0's MUFU-free
                // phases (S load, row maxima, P store, barrier round trips) and vice versa.  The offset persists: each group's
                // period is its own dependency chain.  (The delay has to sit here, after the first S tile exists - before
                // the loop it disappears in the ~3500-cycle prologue both groups wait through.)
                if (w == 1 && j == 0) { const long long t_start = clock64(); while (clock64() - t_start < G1_LAG_CYCLES) { } }
                tc_fence_after();
                TRF(w, j, 1);
                uint32_t sv[TILE];                 // this thread's row of S: one TMEM read, then S is handed back at once
#pragma unroll
                for (int c = 0; c < 4; ++c) tmem_ld32(t_s + c * 32, sv + c * 32);
                tmem_ld_wait();
                tc_fence_before();
                mbar_arrive(&bars->s_free[w]);      // the MMA warp may now overwrite S with Q K_{j+1}^T
                TRF(w, j, 2);
                const float mx = need_mask ? row_max<true>(sv, limit) : row_max<false>(sv, limit);
                const float mxs = mx * scale_log2;
                TRF(w, j, 3);
                const bool raise = mxs > m_ref + 8.0f;
                const float m_new = raise ? mxs : m_ref;
                const float alpha = raise ? ex2(m_ref - m_new) : 1.0f;
                const bool fix_o = (j > 0) && __any_sync(0xffffffffu, raise);
                m_ref = m_new;
                uint32_t pk[TILE / 2];
                float rs;
                if (need_mask) rs = row_exp<true, POLY>(sv, pk, scale_log2, m_ref, limit);
                else rs = row_exp<false, POLY>(sv, pk, scale_log2, m_ref, limit);
                l = fmaf(l, alpha, rs);
                TRF(w, j, 4);
                // P V of tile j-1 must have completed: the MMA has finished reading this group's P buffer and O is stable
                if (j > 0) {
                    mbar_wait(&bars->o_full[w], (j - 1) & 1);
                    tc_fence_after();
                    if (fix_o) {                     // warp-uniform; rows whose reference did not move use alpha = 1
                        const float2 a2 = make_float2(alpha, alpha);
#pragma unroll
                        for (int c = 0; c < 2; ++c) {
                            uint32_t v[32];
                            tmem_ld32(t_o + c * 32, v);
                            tmem_ld_wait();
#pragma unroll
                            for (int i = 0; i < 16; ++i) {
                                const float2 r = __fmul2_rn(make_float2(__uint_as_float(v[2 * i]), __uint_as_float(v[2 * i + 1])), a2);
                                v[2 * i] = __float_as_uint(r.x); v[2 * i + 1] = __float_as_uint(r.y);
                            }
                            tmem_st32(t_o + c * 32, v);
                        }
                        tmem_st_wait();
                    }
                }
                TRF(w, j, 5);
                // P -> TMEM as packed bf16 pairs (row = lane, 64 columns): the MMA reads it as its A operand
                tmem_st32(t_p, pk);
                tmem_st32(t_p + 32, pk + 32);
                tmem_st_wait();
                tc_fence_before();
                mbar_arrive(&bars->p_ready[w]);
                TRF(w, j, 6);
            }
            // all P V accumulated
            mbar_wait(&bars->o_full[w], (n_my - 1) & 1);
            tc_fence_after();
            const float inv_l = 1.f / l;
            const float m = m_ref;
            float o[HD];
            {
                const float2 il2 = make_float2(inv_l, inv_l);
#pragma unroll
                for (int c = 0; c < 2; ++c) {
                    uint32_t v[32];
                    tmem_ld32(t_o + c * 32, v);
                    tmem_ld_wait();
#pragma unroll
                    for (int i = 0; i < 16; ++i) {
                        const float2 r = __fmul2_rn(make_float2(__uint_as_float(v[2 * i]), __uint_as_float(v[2 * i + 1])), il2);
                        o[c * 32 + 2 * i] = r.x; o[c * 32 + 2 * i + 1] = r.y;
                    }
                }
            }
            if (qrow < S) lse[((long long)b * H + h) * S + qrow] = (m + log2f(l)) * 0.6931471805599453f;
            // O tile -> swizzled smem (reuses this group's Q buffer: every MMA has completed) -> TMA store
            uint8_t* stg = sO + row * 128;
#pragma unroll
            for (int i = 0; i < 8; ++i)
                st_shared_v4(stg + ((i ^ (row & 7)) << 4), make_uint4(pack_bf16(o[8 * i], o[8 * i + 1]), pack_bf16(o[8 * i + 2], o[8 * i + 3]),
                               pack_bf16(o[8 * i + 4], o[8 * i + 5]), pack_bf16(o[8 * i + 6], o[8 * i + 7])));
            fence_proxy_async_smem();
            named_bar_sync(1 + w, 128);
            if ((threadIdx.x & 127) == 0) {
                tma_store_3d(&tmO, sO, h * HD, q0, b);
                tma_commit_group();
                tma_wait_group0();
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 8) tmem_dealloc(tmem_base, 512);
}
}  // namespace

#ifdef AVN_TRACE
extern "C" int avn_debug_set_attn_fwd_trace(long long* buf) {
    AVN_CHECK_CUDA(cudaMemcpyToSymbol(g_trace_fwd, &buf, sizeof(buf)));
    return AVN_OK;
}
#endif

extern "C" int avn_attn_fwd(const void* qkv, void* out, float* lse, int B, int S, int H, float scale, int causal, avn_stream_t stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    AVN_REQUIRE(qkv && out && lse, "attn_fwd: null pointer");
    AVN_REQUIRE(B > 0 && S > 0 && H > 0, "attn_fwd: bad sizes B=%d S=%d H=%d", B, S, H);
    AVN_REQUIRE(((uintptr_t)qkv % 16) == 0 && ((uintptr_t)out % 16) == 0, "attn_fwd: pointers must be 16-byte aligned");
    CUtensorMap tmQKV, tmO;
    int rc;
    {
        uint64_t dims[3] = {(uint64_t)3 * H * HD, (uint64_t)S, (uint64_t)B};
        uint64_t str[2] = {(uint64_t)3 * H * HD * 2, (uint64_t)S * 3 * H * HD * 2};
        uint32_t box[3] = {HD, TILE, 1};
        if ((rc = avn_make_tmap(&tmQKV, qkv, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, 3, dims, str, box, true))) return rc;
    }
    {
        uint64_t dims[3] = {(uint64_t)H * HD, (uint64_t)S, (uint64_t)B};
        uint64_t str[2] = {(uint64_t)H * HD * 2, (uint64_t)S * H * HD * 2};
        uint32_t box[3] = {HD, TILE, 1};
        if ((rc = avn_make_tmap(&tmO, out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, 3, dims, str, box, true))) return rc;
    }
    const int q_tiles = (S + TILE - 1) / TILE;
    // long sequences: a third of the exponentials on the FMA pipe (MUFU-bound kernel); S <= 256: all on the MUFU
    auto kern = S > 2 * TILE ? attn_fwd_kernel<true> : attn_fwd_kernel<false>;
    AVN_CHECK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_TOTAL));
    dim3 grid((q_tiles + 1) / 2, H, B);
    kern<<<grid, NTHREADS, SMEM_TOTAL, stream>>>(tmQKV, tmO, lse, S, H, scale * 1.4426950408889634f, causal);
    AVN_CHECK_CUDA(cudaGetLastError());
    return AVN_OK;
}
