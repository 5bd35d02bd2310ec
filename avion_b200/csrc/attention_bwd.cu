// V4 backward: flash-style attention backward on tcgen05, head dim 64.
// Replaces the backward of flash_attn_qkvpacked_func (avion/models/transformer.py:148) / nn.MultiheadAttention (:140-146).
//
// One work item per (clip, head, 128-row K/V tile j): a loop over the query tiles i; persistent CTAs (one per SM) walk the items.  Everything is computed TRANSPOSED
// (key rows on the TMEM lanes) so that P^T can feed the dV MMA straight from TMEM and dS^T is written to smem once:
//   S^T  = K_j Q_i^T         dP^T = V_j dO_i^T                       (recompute; TMEM, 128 fp32 columns each)
//   P^T  = exp(scale*S^T - LSE_q)           -> TMEM as bf16 (A operand of dV)
//   dS^T = scale * P^T o (dP^T - delta_q)   -> smem [kv][q] bf16 (K-major A of dK, MN-major A of dQ: same bytes)
//   dV  += P^T dO_i          dK += dS^T Q_i                           (accumulate in TMEM across i, 64 columns each)
//   dQ_i(partial) = dS K_j                                           (TMEM -> fp32 smem -> TMA reduce-add into dq_acc)
// Warp roles (16 warps): 0-7 compute (thread = key row x half of the 128 query columns), 8-11 dQ drain,
// 12 MMA issuer, 13 TMA producer, 14-15 idle (complete the warpgroup for setmaxnreg).
#include <stdlib.h>
#include "common.cuh"
#include "../../include/avion_b200.h"

namespace {
using namespace avn;

constexpr int TILE = 128, HD = 64;
constexpr int NTHREADS = 512;
constexpr int T16 = TILE * HD * 2;                // 16 KB: one [128][64] bf16 tile
constexpr int OFF_K = 0, OFF_V = T16;
constexpr int QDO_STAGES = 3;                     // the TMA latency of a Q/dO tile pair spans more than one iteration
constexpr int OFF_Q = 2 * T16;                    // QDO_STAGES stages
constexpr int OFF_DO = 5 * T16;                   // QDO_STAGES stages
constexpr int OFF_DS = 8 * T16;                   // 2 buffers x 32 KB ([kv][q] bf16, two 64-column sub-tiles each)
constexpr int OFF_STG = 12 * T16;                 // 32 KB: fp32 dQ staging ([q][32 floats] x 2 sub-tiles)
constexpr int OFF_STAT = 14 * T16;                // 2 buffers x {-lse*log2e [128], -delta*scale [128]} floats
constexpr int OFF_BAR = OFF_STAT + 2 * 256 * 4;
constexpr int SMEM_TOTAL = OFF_BAR + 256;
static_assert(SMEM_TOTAL <= 232448, "shared memory budget");
constexpr uint32_t TM_S = 0, TM_DP = 128, TM_DV = 256, TM_DK = 320, TM_DQ = 384, TM_P = 448;

struct Bars {
    uint64_t kv_full, kv_free, acc_free, qdo_full[QDO_STAGES], qdo_empty[QDO_STAGES], sdp_full, sdp_free, p_ready, p_free, ds_ready[2],
        ds_free[2], dq_full, dq_free;
    uint32_t tmem_ptr;
};
static_assert(sizeof(Bars) <= 256, "barrier block");

// Of every AVN_BWD_POLY_DEN element pairs, AVN_BWD_POLY_NUM take the FMA-pipe exponential (common.cuh: exp2_poly2).  Off by
// default: unlike the forward kernel this one is not MUFU-bound (1/4 .. 1/2 polynomial measured 2-4 % slower) - it saturates
// the shared-memory data pipe (ncu: tensor operand fetch 45 % + LSU wavefronts 50 %).
#ifndef AVN_BWD_POLY_NUM
#define AVN_BWD_POLY_NUM 0
#endif
#ifndef AVN_BWD_POLY_DEN
#define AVN_BWD_POLY_DEN 3
#endif
__device__ __forceinline__ float2 exp2_pair_bwd(float2 x, int pair_index) {
    if ((pair_index % AVN_BWD_POLY_DEN) < AVN_BWD_POLY_NUM) return exp2_poly2(x);
    return make_float2(ex2(x.x), ex2(x.y));
}

// One 32-column chunk of a key row: P^T = exp2(S^T*scale_log2 + nl[q]), dS^T = P^T o (dP^T*scale + nd[q]) -> packed bf16.
// nl = -LSE*log2e (or -inf for query rows past the end), nd = -delta*scale; both broadcast from smem.
template <bool MASK>
__device__ __forceinline__ void pds_chunk(const uint32_t* sv, const uint32_t* dv, uint32_t* pk, uint32_t* dsk, const float* nl,
                                          const float* nd, float scale_log2, float scale, int first_valid) {
    const float2 sc2 = make_float2(scale_log2, scale_log2), s2 = make_float2(scale, scale);
#pragma unroll
    for (int i = 0; i < 32; i += 4) {
        const float4 l4 = *(const float4*)(nl + i), d4 = *(const float4*)(nd + i);
        const float2 x0 = __ffma2_rn(make_float2(__uint_as_float(sv[i]), __uint_as_float(sv[i + 1])), sc2, make_float2(l4.x, l4.y));
        const float2 x1 = __ffma2_rn(make_float2(__uint_as_float(sv[i + 2]), __uint_as_float(sv[i + 3])), sc2, make_float2(l4.z, l4.w));
        float2 p0 = exp2_pair_bwd(x0, i >> 1), p1 = exp2_pair_bwd(x1, (i >> 1) + 1);
        if (MASK) {               // causal: query column c is visible to this key row only if c >= first_valid
            if (i < first_valid) p0.x = 0.f;
            if (i + 1 < first_valid) p0.y = 0.f;
            if (i + 2 < first_valid) p1.x = 0.f;
            if (i + 3 < first_valid) p1.y = 0.f;
        }
        const float2 t0 = __ffma2_rn(make_float2(__uint_as_float(dv[i]), __uint_as_float(dv[i + 1])), s2, make_float2(d4.x, d4.y));
        const float2 t1 = __ffma2_rn(make_float2(__uint_as_float(dv[i + 2]), __uint_as_float(dv[i + 3])), s2, make_float2(d4.z, d4.w));
        const float2 d0 = __fmul2_rn(p0, t0), d1 = __fmul2_rn(p1, t1);
        pk[i >> 1] = pack_bf16(p0.x, p0.y); pk[(i >> 1) + 1] = pack_bf16(p1.x, p1.y);
        dsk[i >> 1] = pack_bf16(d0.x, d0.y); dsk[(i >> 1) + 1] = pack_bf16(d1.x, d1.y);
    }
}

// Persistent: a CTA walks the work items (clip, head, K/V tile) item = blockIdx.x, += gridDim.x (K/V tile fastest, so the CTAs
// of a wave share (clip, head) pairs and the dQ accumulator stays in L2).  Barriers, the Q/dO ring and the TMEM buffers are
// set up once; every role counts its iterations globally (G), so barrier parities simply continue across items.  Between two
// items the producer prefetches the next item's first Q/dO tiles while the last iterations still run, reloads K/V as soon as
// the last MMAs have read them (kv_free), and that load overlaps the dK/dV drain of the finished item (acc_free tells the MMA
// warp when dV/dK have left TMEM).  One CTA per item (gridDim.x = n_items) is the non-persistent schedule of the same code.
__global__ void __launch_bounds__(NTHREADS, 1)
attn_bwd_kernel(const __grid_constant__ CUtensorMap tmQKV, const __grid_constant__ CUtensorMap tmDO,
                const __grid_constant__ CUtensorMap tmDQKV, const __grid_constant__ CUtensorMap tmDQ,
                const float* __restrict__ lse, const float* __restrict__ delta, int S, int H, int n_items, float scale, int causal) {
    extern __shared__ __align__(1024) uint8_t smem[];
    Bars* bars = (Bars*)(smem + OFF_BAR);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int q_tiles = (S + TILE - 1) / TILE;
    // item -> (K/V tile, head, clip), first query tile and number of query tiles (causal: tiles on and below the diagonal)
    auto decode = [&](int item, int& jt, int& h, int& b, int& i0, int& n_i) {
        jt = item % q_tiles; const int r = item / q_tiles; h = r % H; b = r / H;
        i0 = causal ? jt : 0; n_i = q_tiles - i0;
    };

    if (threadIdx.x == 0) {
        if (smem_u32(smem) & 1023u) { printf("avion_b200: attn_bwd dynamic smem base not 1024-byte aligned\n"); __trap(); }
        tma_prefetch_desc(&tmQKV); tma_prefetch_desc(&tmDO); tma_prefetch_desc(&tmDQKV); tma_prefetch_desc(&tmDQ);
        mbar_init(&bars->kv_full, 1); mbar_init(&bars->kv_free, 1); mbar_init(&bars->acc_free, 256);
        for (int i = 0; i < QDO_STAGES; ++i) { mbar_init(&bars->qdo_full[i], 1); mbar_init(&bars->qdo_empty[i], 1); }
        for (int i = 0; i < 2; ++i) { mbar_init(&bars->ds_ready[i], 256); mbar_init(&bars->ds_free[i], 1); }
        mbar_init(&bars->sdp_full, 1); mbar_init(&bars->sdp_free, 256);
        mbar_init(&bars->p_ready, 256); mbar_init(&bars->p_free, 1);
        mbar_init(&bars->dq_full, 1); mbar_init(&bars->dq_free, 128);
        fence_barrier_init();
    }
    if (warp == 12) tmem_alloc(&bars->tmem_ptr, 512);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = bars->tmem_ptr;

    if (warp >= 14) {
        reg_dec<40>();       // idle warps: only complete the fourth warpgroup
    } else if (warp == 13) {
        // ===================== TMA producer =====================
        // (whole warp loops and waits; one elected lane issues, so addresses and coordinates stay in uniform registers)
        reg_dec<40>();
        int G0 = 0, it = 0;
        for (int item = blockIdx.x; item < n_items; item += gridDim.x, ++it) {
            int jt, h, b, i0, n_i;
            decode(item, jt, h, b, i0, n_i);
            auto load_qdo = [&](int t) {
                const int G = G0 + t, st = G % QDO_STAGES;
                mbar_wait(&bars->qdo_empty[st], ((G / QDO_STAGES) & 1) ^ 1);
                if (elect_one()) {
                    mbar_expect_tx(&bars->qdo_full[st], 2 * T16);
                    tma_load_3d(smem + OFF_Q + st * T16, &tmQKV, &bars->qdo_full[st], h * HD, (i0 + t) * TILE, b);
                    tma_load_3d(smem + OFF_DO + st * T16, &tmDO, &bars->qdo_full[st], h * HD, (i0 + t) * TILE, b);
                }
                __syncwarp();
            };
            // the first Q/dO tiles go into ring slots as the previous item's last iterations release them; K/V can only
            // follow the previous item's last MMA
            const int head = min(n_i, QDO_STAGES);
            for (int t = 0; t < head; ++t) load_qdo(t);
            if (it > 0) mbar_wait(&bars->kv_free, (it - 1) & 1);
            if (elect_one()) {
                mbar_expect_tx(&bars->kv_full, 2 * T16);
                tma_load_3d(smem + OFF_K, &tmQKV, &bars->kv_full, (H + h) * HD, jt * TILE, b);
                tma_load_3d(smem + OFF_V, &tmQKV, &bars->kv_full, (2 * H + h) * HD, jt * TILE, b);
            }
            __syncwarp();
            for (int t = head; t < n_i; ++t) load_qdo(t);
            G0 += n_i;
        }
    } else if (warp == 12) {
        // ===================== MMA issuer =====================
        // waits are executed by the whole warp, the MMAs by one elected lane (descriptors stay in uniform registers)
        reg_dec<40>();
        constexpr uint32_t idesc_s = umma_idesc_bf16(TILE, TILE, false, false);
        constexpr uint32_t idesc_kn = umma_idesc_bf16(TILE, HD, false, true);    // A K-major (or TMEM), B MN-major
        constexpr uint32_t idesc_nn = umma_idesc_bf16(TILE, HD, true, true);     // A MN-major, B MN-major
        const uint32_t ka = smem_u32(smem + OFF_K), va = smem_u32(smem + OFF_V);
        auto issue_sdp = [&](int G) {
            const int st = G % QDO_STAGES;
            mbar_wait(&bars->qdo_full[st], (G / QDO_STAGES) & 1);
            if (G > 0) mbar_wait(&bars->sdp_free, (G - 1) & 1);
            tc_fence_after();
            if (elect_one()) {
                const uint32_t qa = smem_u32(smem + OFF_Q + st * T16), doa = smem_u32(smem + OFF_DO + st * T16);
#pragma unroll
                for (int k = 0; k < HD / 16; ++k)      // S^T = K Q^T
                    umma_ss(tmem_base + TM_S, umma_smem_desc(ka + k * 32, 16, 1024), umma_smem_desc(qa + k * 32, 16, 1024), idesc_s, k > 0);
#pragma unroll
                for (int k = 0; k < HD / 16; ++k)      // dP^T = V dO^T
                    umma_ss(tmem_base + TM_DP, umma_smem_desc(va + k * 32, 16, 1024), umma_smem_desc(doa + k * 32, 16, 1024), idesc_s, k > 0);
                umma_commit(&bars->sdp_full);
            }
            __syncwarp();
        };
        int G0 = 0, it = 0;
        for (int item = blockIdx.x; item < n_items; item += gridDim.x, ++it) {
            int jt, h, b, i0, n_i;
            decode(item, jt, h, b, i0, n_i);
            mbar_wait(&bars->kv_full, it & 1);
            issue_sdp(G0);
            for (int t = 0; t < n_i; ++t) {
                const int G = G0 + t, st = G % QDO_STAGES, db = G & 1;
                if (t + 1 < n_i) issue_sdp(G + 1);
                const uint32_t qa = smem_u32(smem + OFF_Q + st * T16), doa = smem_u32(smem + OFF_DO + st * T16);
                const uint32_t dsa = smem_u32(smem + OFF_DS + db * 2 * T16);
                mbar_wait(&bars->p_ready, G & 1);
                if (t == 0 && it > 0) mbar_wait(&bars->acc_free, (it - 1) & 1);     // the previous item's dV / dK have left TMEM
                tc_fence_after();
                if (elect_one()) {
#pragma unroll
                    for (int k = 0; k < TILE / 16; ++k)      // dV += P^T dO   (A = P^T from TMEM, 8 columns per K=16; B = dO MN-major)
                        umma_ts(tmem_base + TM_DV, tmem_base + TM_P + k * 8, umma_smem_desc(doa + k * 2048, 8192, 1024),
                                idesc_kn, (t > 0 || k > 0) ? 1u : 0u);
                    umma_commit(&bars->p_free);
                }
                __syncwarp();
                mbar_wait(&bars->ds_ready[db], (G >> 1) & 1);
                tc_fence_after();
                if (elect_one()) {
#pragma unroll
                    for (int k = 0; k < TILE / 16; ++k)      // dK += dS^T Q   (A = dS^T K-major, B = Q MN-major)
                        umma_ss(tmem_base + TM_DK, umma_smem_desc(dsa + (k >> 2) * 16384 + (k & 3) * 32, 16, 1024),
                                umma_smem_desc(qa + k * 2048, 8192, 1024), idesc_kn, (t > 0 || k > 0) ? 1u : 0u);
                    umma_commit(&bars->qdo_empty[st]);       // Q_t / dO_t are not read after dK
                }
                __syncwarp();
                if (G > 0) { mbar_wait(&bars->dq_free, (G - 1) & 1); tc_fence_after(); }   // the previous dQ tile has left TMEM
                if (elect_one()) {
#pragma unroll
                    for (int k = 0; k < TILE / 16; ++k)      // dQ_i = dS K_j  (A = dS^T bytes read MN-major, B = K_j MN-major)
                        umma_ss(tmem_base + TM_DQ, umma_smem_desc(dsa + k * 2048, 16384, 1024),
                                umma_smem_desc(ka + k * 2048, 8192, 1024), idesc_nn, k > 0 ? 1u : 0u);
                    umma_commit(&bars->dq_full);
                    umma_commit(&bars->ds_free[db]);
                    if (t == n_i - 1) umma_commit(&bars->kv_free);       // every MMA of the item has read K / V
                }
                __syncwarp();
            }
            G0 += n_i;
        }
    } else if (warp >= 8) {
        // ===================== dQ drain: TMEM -> fp32 staging -> TMA reduce-add =====================
        reg_dec<88>();
        const int row = (warp & 3) * 32 + lane;
        const uint32_t lane_addr = (uint32_t)((warp & 3) * 32) << 16;
        const bool leader = (threadIdx.x == 256);
        int G0 = 0;
        for (int item = blockIdx.x; item < n_items; item += gridDim.x) {
            int jt, h, b, i0, n_i;
            decode(item, jt, h, b, i0, n_i);
            for (int t = 0; t < n_i; ++t) {
                const int G = G0 + t;
                mbar_wait(&bars->dq_full, G & 1);
                tc_fence_after();
                uint32_t v[64];
                tmem_ld32(tmem_base + lane_addr + TM_DQ, v);
                tmem_ld32(tmem_base + lane_addr + TM_DQ + 32, v + 32);
                tmem_ld_wait();
                tc_fence_before();
                mbar_arrive(&bars->dq_free);
                uint8_t* stg = smem + OFF_STG;
                if (G >= 1) {                            // the reduce of the previous tile must have finished reading the staging tile
                    if (leader) tma_wait_group_read0();
                    named_bar_sync(2, 128);
                }
#pragma unroll
                for (int c = 0; c < 2; ++c)
#pragma unroll
                    for (int i = 0; i < 8; ++i)
                        st_shared_v4(stg + c * T16 + row * 128 + ((i ^ (row & 7)) << 4),
                                     make_uint4(v[c * 32 + 4 * i], v[c * 32 + 4 * i + 1], v[c * 32 + 4 * i + 2], v[c * 32 + 4 * i + 3]));
                fence_proxy_async_smem();
                named_bar_sync(2, 128);
                if (leader) {
                    tma_reduce_add_3d(&tmDQ, stg, h * HD, (i0 + t) * TILE, b);
                    tma_reduce_add_3d(&tmDQ, stg + T16, h * HD + 32, (i0 + t) * TILE, b);
                    tma_commit_group();
                }
            }
            G0 += n_i;
        }
        if (leader) tma_wait_group0();
    } else {
        // ===================== compute warps: thread = (key row, half of the query columns) =====================
        reg_inc<192>();
        const int half = warp >> 2;
        const int row = (warp & 3) * 32 + lane;
        const uint32_t lane_addr = (uint32_t)((warp & 3) * 32) << 16;
        const float scale_log2 = scale * 1.4426950408889634f;
        float* stat = (float*)(smem + OFF_STAT);
        // per-query-row statistics of iteration G live in stat[(G&1)*256 ...]: threads 0-127 own -lse*log2e, 128-255 own -delta*scale
        const int sidx = threadIdx.x & 127;
        const bool own_lse = threadIdx.x < 128;
        const float stat_mul = own_lse ? -1.4426950408889634f : -scale;
        int G0 = 0;
        for (int item = blockIdx.x; item < n_items; item += gridDim.x) {
            int jt, h, b, i0, n_i;
            decode(item, jt, h, b, i0, n_i);
            const int kv0 = jt * TILE;
            // the raw global value is kept in a register for a whole tile and only converted when it is stored: the use of
            // a load result is what stalls, so the conversion must not sit right behind the load
            const float* stat_src = (own_lse ? lse : delta) + ((long long)b * H + h) * S;
            auto load_stat = [&](int t) -> float {
                const int qr = (i0 + t) * TILE + sidx;
                return qr < S ? __ldg(stat_src + qr) : (own_lse ? INFINITY : 0.f);      // -inf after scaling: P = 0 past the end
            };
            // (buffer G0&1 was last read in iteration G0-2; every thread has passed the barrier of iteration G0-1 since)
            stat[(G0 & 1) * 256 + threadIdx.x] = load_stat(0) * stat_mul;
            float stat_nx = n_i > 1 ? load_stat(1) : 0.f;

            for (int t = 0; t < n_i; ++t) {
                const int G = G0 + t;
                const int buf = G & 1;
                const int q0 = (i0 + t) * TILE;
                const bool need_mask = causal && (q0 < kv0 + TILE - 1);      // diagonal tile of the causal text tower (warp-uniform)
                mbar_wait(&bars->sdp_full, G & 1);
                tc_fence_after();
                named_bar_sync(1, 256);           // stat[buf] is complete; nobody still reads stat[buf^1] (previous iteration)
                if (t + 1 < n_i) {
                    stat[(buf ^ 1) * 256 + threadIdx.x] = stat_nx * stat_mul;
                    if (t + 2 < n_i) stat_nx = load_stat(t + 2);
                }
                const float* nl = stat + buf * 256 + half * 64;
                const float* nd = nl + 128;
                uint32_t pk[32], dsk[32];     // packed bf16: 64 P^T values and 64 dS^T values of this thread
#pragma unroll
                for (int c = 0; c < 2; ++c) {
                    uint32_t sv[32], dv[32];
                    tmem_ld32(tmem_base + lane_addr + TM_S + half * 64 + c * 32, sv);
                    tmem_ld32(tmem_base + lane_addr + TM_DP + half * 64 + c * 32, dv);
                    tmem_ld_wait();
                    if (c == 1) { tc_fence_before(); mbar_arrive(&bars->sdp_free); }
                    // causal: key kv0+row is visible to query q0+col iff col >= kv0 + row - q0
                    const int first_valid = kv0 + row - q0 - (half * 64 + c * 32);
                    if (need_mask) pds_chunk<true>(sv, dv, pk + c * 16, dsk + c * 16, nl + c * 32, nd + c * 32, scale_log2, scale, first_valid);
                    else pds_chunk<false>(sv, dv, pk + c * 16, dsk + c * 16, nl + c * 32, nd + c * 32, scale_log2, scale, 0);
                }
                // dS^T -> smem [kv][q]: buffer G&1 was last read by the dK/dQ MMAs of iteration G-2.  The stores are issued first
                // and complete while P^T goes to TMEM.
                if (G >= 2) mbar_wait(&bars->ds_free[buf], ((G >> 1) - 1) & 1);
                uint8_t* drow = smem + OFF_DS + buf * 2 * T16 + half * T16 + row * 128;
#pragma unroll
                for (int i = 0; i < 8; ++i)
                    st_shared_v4(drow + ((i ^ (row & 7)) << 4), make_uint4(dsk[4 * i], dsk[4 * i + 1], dsk[4 * i + 2], dsk[4 * i + 3]));
                // P^T -> TMEM (bf16 pairs, lane = key row): the previous dV MMA must have finished reading the previous P^T
                if (G > 0) { mbar_wait(&bars->p_free, (G - 1) & 1); tc_fence_after(); }
                tmem_st32(tmem_base + lane_addr + TM_P + half * 32, pk);
                tmem_st_wait();
                tc_fence_before();
                mbar_arrive(&bars->p_ready);
                fence_proxy_async_smem();
                mbar_arrive(&bars->ds_ready[buf]);
            }
            // every MMA of the item has completed once the last ds_free fires (it is committed after dQ of the last tile)
            const int Gl = G0 + n_i - 1;
            mbar_wait(&bars->ds_free[Gl & 1], (Gl >> 1) & 1);
            tc_fence_after();
            // dV, dK: TMEM -> bf16 -> swizzled smem (the dS buffers are free now) -> TMA store
            {
                uint32_t v[32], k[32];
                tmem_ld32(tmem_base + lane_addr + TM_DV + half * 32, v);
                tmem_ld32(tmem_base + lane_addr + TM_DK + half * 32, k);
                tmem_ld_wait();
                tc_fence_before();
                mbar_arrive(&bars->acc_free);          // the next item may start accumulating dV / dK
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
                named_bar_sync(1, 256);
                if (threadIdx.x == 0) {
                    tma_store_3d(&tmDQKV, smem + OFF_DS + 2 * T16, (H + h) * HD, kv0, b);
                    tma_store_3d(&tmDQKV, smem + OFF_DS, (2 * H + h) * HD, kv0, b);
                    tma_commit_group();
                    // the dS buffers are rewritten in the next item's first iterations, behind that iteration's barrier (which
                    // this thread only reaches after the stores have read them)
                    // (the last item waits for full completion here: a wait placed after the item loop makes ptxas compile this
                    // branch for the kernel-wide 128 registers instead of the 192 the warpgroups have)
                    if (item + (int)gridDim.x >= n_items) tma_wait_group0(); else tma_wait_group_read0();
                }
            }
            G0 += n_i;
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 12) tmem_dealloc(tmem_base, 512);
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
    // one CTA per SM walking the items; AVN_ATTN_BWD_PERSIST=0: one CTA per item (A/B runs)
    static const bool persist = !(getenv("AVN_ATTN_BWD_PERSIST") && atoi(getenv("AVN_ATTN_BWD_PERSIST")) == 0);
    const long long n_items = (long long)tiles * H * B;
    AVN_REQUIRE(n_items < (1ll << 31), "attn_bwd: too many work items");
    const int grid = persist ? (int)(n_items < avn_num_sms() ? n_items : avn_num_sms()) : (int)n_items;
    attn_bwd_kernel<<<grid, NTHREADS, SMEM_TOTAL, stream>>>(tmQKV, tmDO, tmDQKV, tmDQ, lse, delta, S, H, (int)n_items, scale, causal);
    AVN_CHECK_CUDA(cudaGetLastError());
    const long long nvec = rows * (W1 / 8);
    dq_convert_kernel<<<(unsigned)((nvec + 255) / 256), 256, 0, stream>>>(dq_acc, (__nv_bfloat16*)dqkv, rows, (int)W1);
    AVN_CHECK_CUDA(cudaGetLastError());
    return AVN_OK;
}
