// bf16 GEMM for sm_100a: persistent, warp-specialised.
//   warp 0      : TMA producer   (cp.async.bulk.tensor -> 4-stage smem ring, 128B swizzle)
//   warp 1      : MMA issuer     (one lane issues tcgen05.mma, fp32 accumulators in TMEM, double-buffered)
//   warps 2..9  : epilogue       (tcgen05.ld -> bias / GELU(erf) / residual / dGELU -> swizzled smem -> TMA store,
//                                 or fp32 TMA reduce-add for split-K weight gradients)
// Tile 128 x BN x 64 (BN = 128 or 256).  Operand majors:
//   NT: A K-major, B K-major        (forward:  X[M,K] * W[N,K]^T)
//   NN: A K-major, B MN-major       (dgrad:    dY[M,N'] * W[N',K'])
//   TN: A MN-major, B MN-major      (wgrad:    dY[tok,N]^T * X[tok,K], contraction over tokens, split-K)
// TN can also produce the bias gradient (column sums of dY over the tokens) from the A tiles it stages anyway: the epilogue
// warps, idle during the main loop of a weight-gradient tile, read A stages while the MMAs consume them and add them into
// per-thread fp32 sums; a stage returns to the producer when the MMA commit AND epilogue warp 0 have arrived (no second pass
// over dY in HBM, nothing added to the latency of a stage's release).
// Replaces the cuBLAS calls behind nn.Linear / FlashMHA / FlashMlp in avion/models/transformer.py:116-120,135,227,320.
#include <stdlib.h>
#include "common.cuh"
#include "../../include/avion_b200.h"

namespace {
using namespace avn;

constexpr int BM = 128, BK = 64;
constexpr int EPI_WARPS = 8;

constexpr int NTHREADS = 64 + EPI_WARPS * 32;
constexpr int A_STAGE_BYTES = BM * BK * 2;       // 16 KB
constexpr int STG_BYTES = 128 * 128;             // one 128-row x 128-byte staging tile per epilogue group
// which kernels may use a second pair of staging tiles - for the R operand read by the epilogue, or for the second output of
// the dual epilogue (the single-CTA 256-wide kernel has no room for it)
constexpr bool r_stageable(int BN, int EPI, int CG) {
    return (EPI == AVN_EPI_BIAS_RESIDUAL || EPI == AVN_EPI_DGELU || EPI == AVN_EPI_BIAS_GELU_DUAL) && (CG == 2 || BN == 128);
}

struct KParams {
    int M, N, K, batch;
    int m_tiles, n_tiles, splits, kblocks, kb_per_split, total_kb;
    const float* bias;
    const __nv_bfloat16* R;
    long long ldr, r_batch_stride;
    float* colsum;      // TN only: fp32 [M] += column sums of A over the contraction, or nullptr
    int a_box4, b_box4; // MN-major operand staged by ONE 4-D box per stage (64-column chunks as the third dimension)
};

// CG = CTAs per MMA (cta_group): with CG = 2 a CTA pair computes a 256 x BN tile, each CTA staging its own 128 rows of A and
// HALF of the B tile - the tensor core's shared-memory operand reads, which bound the single-CTA kernel at about two
// thirds of the MMA rate, drop from 96 to 64 bytes per clock per SM.
// RSTG: a second 128 x 64 staging tile per epilogue group - the residual / pre-activation tile (EPI_BIAS_RESIDUAL, EPI_DGELU)
// that TMA brings in for the epilogue, or the second output of EPI_BIAS_GELU_DUAL on its way out; the CTA-pair kernels pay
// for it with one ring stage (5 instead of 6).
template <int BN, int CG, bool RSTG>
struct Smem {
    static constexpr int STAGES = (CG == 2) ? (RSTG ? 5 : 6) : 4;
    static constexpr int B_STAGE_BYTES = (BN / CG) * BK * 2;
    static constexpr int OFF_A = 0;
    static constexpr int OFF_B = OFF_A + STAGES * A_STAGE_BYTES;
    static constexpr int OFF_STG = OFF_B + STAGES * B_STAGE_BYTES;
    static constexpr int OFF_RSTG = OFF_STG + 2 * STG_BYTES;
    static constexpr int OFF_BIAS = OFF_RSTG + (RSTG ? 2 * STG_BYTES : 0);
    static constexpr int OFF_BAR = OFF_BIAS + BN * 4;
    static constexpr int TOTAL = OFF_BAR + 256 + 1024;   // + alignment slack
};

template <int BN, int MODE, int EPI, int CG, bool RS>
__global__ void __launch_bounds__(NTHREADS, 1)
gemm_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
            const __grid_constant__ CUtensorMap tmC, const __grid_constant__ CUtensorMap tmC2, const KParams p) {
    constexpr bool A_MN = (MODE == AVN_GEMM_TN);
    constexpr bool B_MN = (MODE != AVN_GEMM_NT);
    using L = Smem<BN, CG, RS>;
    constexpr int STAGES = L::STAGES;
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
    uint8_t* sA = smem + L::OFF_A;
    uint8_t* sB = smem + L::OFF_B;
    uint8_t* sStg = smem + L::OFF_STG;
    float* sBias = (float*)(smem + L::OFF_BIAS);
    uint64_t* full_bar = (uint64_t*)(smem + L::OFF_BAR);
    uint64_t* empty_bar = full_bar + STAGES;
    uint64_t* tfull_bar = empty_bar + STAGES;
    uint64_t* tempty_bar = tfull_bar + 2;
    uint64_t* pfull_bar = tempty_bar + 2;       // column-sum mode, peer CTA of a pair: "stage is full", forwarded by the leader
    uint64_t* rfull_bar = pfull_bar + STAGES;   // staged R tile of epilogue group 0 / 1 has landed
    uint32_t* tmem_ptr = (uint32_t*)(rfull_bar + 2);
    const bool colsum_on = (MODE == AVN_GEMM_TN) && p.colsum != nullptr;

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int rank = (CG == 2) ? (int)cluster_ctarank() : 0;      // position inside the CTA pair; rank 0 issues the MMAs

    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&tmA);
        tma_prefetch_desc(&tmB);
        tma_prefetch_desc(&tmC);
        // column-sum mode: a stage is released by the MMA commit plus the arrive of epilogue warp 0
        for (int i = 0; i < STAGES; ++i) { mbar_init(&full_bar[i], 1); mbar_init(&empty_bar[i], colsum_on ? 2 : 1); mbar_init(&pfull_bar[i], 1); }
        for (int i = 0; i < 2; ++i) { mbar_init(&tfull_bar[i], 1); mbar_init(&tempty_bar[i], EPI_WARPS * CG); mbar_init(&rfull_bar[i], 1); }
        fence_barrier_init();
    }
    if (warp == 1) { if (CG == 2) tmem_alloc_2cta(tmem_ptr, 2 * BN); else tmem_alloc(tmem_ptr, 2 * BN); }
    tc_fence_before();
    if (CG == 2) cluster_sync_all(); else __syncthreads();      // the peer's barriers must exist before anything signals them
    tc_fence_after();
    const uint32_t tmem_base = *tmem_ptr;

    // tiles are enumerated per CTA pair: a unit covers CG * 128 rows, CTA `rank` owns rows (unit * CG + rank) * 128 ..
    const int m_units = (p.m_tiles + CG - 1) / CG;
    const int total_tiles = (MODE == AVN_GEMM_TN ? p.splits : p.batch) * m_units * p.n_tiles;
    const int tile0 = blockIdx.x / CG, tile_step = gridDim.x / CG;

    if (warp == 0) {
        // ===================== TMA producer =====================
        if (lane == 0) {
            int stage = 0; uint32_t phase = 0;
            for (int t = tile0; t < total_tiles; t += tile_step) {
                int n_blk = t % p.n_tiles; int r = t / p.n_tiles;
                int m_blk = (r % m_units) * CG + rank; int outer = r / m_units;       // batch (NT/NN) or split (TN)
                int kb0 = 0, kb1 = p.kblocks;
                if (MODE == AVN_GEMM_TN) { kb0 = outer * p.kb_per_split; kb1 = min(p.total_kb, kb0 + p.kb_per_split); }
                const int m0 = m_blk * BM, n0 = n_blk * BN + rank * (BN / CG);      // this CTA's rows of A and of B
                auto load = [&](void* dst, const CUtensorMap* tm, uint64_t* bar, int c0, int c1, int c2) {
                    if (CG == 2) tma_load_3d_2cta(dst, tm, bar, c0, c1, c2); else tma_load_3d(dst, tm, bar, c0, c1, c2);
                };
                // MN-major operand: [k rows][64 columns] chunks; the 4-D map views the row as (chunk, 64 columns), so that one
                // box brings all chunks of a stage ([chunk][k][128 B] in shared memory, as the per-chunk loads lay it out)
                auto load4 = [&](void* dst, const CUtensorMap* tm, uint64_t* bar, int col0, int krow, int kbatch) {
                    if (CG == 2) tma_load_4d_2cta(dst, tm, bar, 0, krow, col0 >> 6, kbatch); else tma_load_4d(dst, tm, bar, 0, krow, col0 >> 6, kbatch);
                };
                for (int kb = kb0; kb < kb1; ++kb) {
                    mbar_wait(&empty_bar[stage], phase ^ 1);
                    uint8_t* a = sA + stage * A_STAGE_BYTES;
                    uint8_t* b = sB + stage * L::B_STAGE_BYTES;
                    // the leader's barrier counts the bytes of both CTAs (the peer's loads signal it, too)
                    if (rank == 0) mbar_expect_tx(&full_bar[stage], CG * (A_STAGE_BYTES + L::B_STAGE_BYTES));
                    if (MODE == AVN_GEMM_TN) {
                        int kbatch = kb / p.kblocks, krow = (kb % p.kblocks) * BK;
                        if (p.a_box4) load4(a, &tmA, &full_bar[stage], m0, krow, kbatch);
                        else {
#pragma unroll
                            for (int j = 0; j < BM / 64; ++j) load(a + j * 8192, &tmA, &full_bar[stage], m0 + 64 * j, krow, kbatch);
                        }
                        if (p.b_box4) load4(b, &tmB, &full_bar[stage], n0, krow, kbatch);
                        else {
#pragma unroll
                            for (int j = 0; j < BN / CG / 64; ++j) load(b + j * 8192, &tmB, &full_bar[stage], n0 + 64 * j, krow, kbatch);
                        }
                    } else {
                        load(a, &tmA, &full_bar[stage], kb * BK, m0, outer);
                        if (MODE == AVN_GEMM_NT) {
                            load(b, &tmB, &full_bar[stage], kb * BK, n0, 0);
                        } else if (p.b_box4) {
                            load4(b, &tmB, &full_bar[stage], n0, kb * BK, 0);
                        } else {
#pragma unroll
                            for (int j = 0; j < BN / CG / 64; ++j) load(b + j * 8192, &tmB, &full_bar[stage], n0 + 64 * j, kb * BK, 0);
                        }
                    }
                    if (++stage == STAGES) { stage = 0; phase ^= 1; }
                }
            }
        }
    } else if (warp == 1) {
        // ===================== MMA issuer (leader CTA only) =====================
        if (rank == 0) {
        constexpr uint32_t idesc = umma_idesc_bf16(BM * CG, BN, A_MN, B_MN);
        constexpr uint32_t a_lbo = A_MN ? 8192 : 16, b_lbo = B_MN ? 8192 : 16;
        constexpr uint32_t a_adv = A_MN ? 2048 : 32, b_adv = B_MN ? 2048 : 32;    // bytes per UMMA_K = 16
        int stage = 0; uint32_t phase = 0;
        int acc = 0; uint32_t acc_phase = 0;
        for (int t = tile0; t < total_tiles; t += tile_step) {
            int outer = (t / p.n_tiles) / m_units;
            int kb0 = 0, kb1 = p.kblocks;
            if (MODE == AVN_GEMM_TN) { kb0 = outer * p.kb_per_split; kb1 = min(p.total_kb, kb0 + p.kb_per_split); }
            mbar_wait(&tempty_bar[acc], acc_phase ^ 1);
            tc_fence_after();
            const uint32_t d_tmem = tmem_base + acc * BN;
            for (int kb = kb0; kb < kb1; ++kb) {
                mbar_wait(&full_bar[stage], phase);
                tc_fence_after();
                if (elect_one()) {
                    const uint32_t a_addr = smem_u32(sA + stage * A_STAGE_BYTES);
                    const uint32_t b_addr = smem_u32(sB + stage * L::B_STAGE_BYTES);
#pragma unroll
                    for (int k = 0; k < BK / 16; ++k) {
                        uint64_t ad = umma_smem_desc(a_addr + k * a_adv, a_lbo, 1024);
                        uint64_t bd = umma_smem_desc(b_addr + k * b_adv, b_lbo, 1024);
                        if (CG == 2) umma_ss_2cta(d_tmem, ad, bd, idesc, (kb > kb0 || k > 0) ? 1u : 0u);
                        else umma_ss(d_tmem, ad, bd, idesc, (kb > kb0 || k > 0) ? 1u : 0u);
                    }
                    if (CG == 2) {       // both CTAs' producers / epilogues are released
                        umma_commit_2cta(&empty_bar[stage]);
                        if (kb == kb1 - 1) umma_commit_2cta(&tfull_bar[acc]);
                    } else {
                        umma_commit(&empty_bar[stage]);
                        if (kb == kb1 - 1) umma_commit(&tfull_bar[acc]);
                    }
                }
                __syncwarp();
                if (++stage == STAGES) { stage = 0; phase ^= 1; }
            }
            if (++acc == 2) { acc = 0; acc_phase ^= 1; }
        }
        }   // rank == 0
    } else {
        // ===================== epilogue =====================
        const int ew = warp - 2;
        const int g = ew >> 2;                  // column half handled by this group of 4 warps
        const int q = warp & 3;                 // TMEM lane quarter this warp may access
        const int row = q * 32 + lane;          // row inside the tile
        const int gtid = (ew & 3) * 32 + lane;  // thread index inside the group
        const int bar_id = 1 + g;
        uint8_t* stg = sStg + g * STG_BYTES;
        float* bias_s = sBias + g * (BN / 2);
        constexpr int UNITS = BN / 2 / 32;      // 32-column units per group
        int acc = 0; uint32_t acc_phase = 0;
        int cs_stage = 0; uint32_t cs_phase = 0;
        uint32_t r_phase = 0;
        for (int t = tile0; t < total_tiles; t += tile_step) {
            int n_blk = t % p.n_tiles; int r = t / p.n_tiles;
            int m_blk = (r % m_units) * CG + rank; int outer = r / m_units;
            const int mb = (MODE == AVN_GEMM_TN) ? 0 : outer;
            const int m0 = m_blk * BM, n0 = n_blk * BN + g * (BN / 2);
            const int m = m0 + row;
            if (MODE == AVN_GEMM_TN && colsum_on) {
                // ---- bias gradient: column sums of this CTA's 128 columns of A over the contraction ----
                // The n_tiles tiles of a row of tiles stage the same A panel: each sums the k-blocks kb = n_blk (mod n_tiles), so
                // the extra shared-memory reads (they compete with the tensor core's operand fetch: one tile reading every
                // stage ran 17 % longer) are spread over all CTAs.  Epilogue warp 0 walks the ring - it waits for a stage to
                // land and is the one extra arrival a stage's release counts in this mode; on a k-block of this tile all
                // eight warps read the stage (thread = 16-byte chunk of 8 columns x 4 of the 64 k-rows) before it arrives.
                const int et = threadIdx.x - 64, mc = et & 15, kg = et >> 4;
                const uint32_t a_off = (uint32_t)((mc >> 3) * 8192 + kg * 4 * 128);
                float2 cs[4];
#pragma unroll
                for (int i = 0; i < 4; ++i) cs[i] = make_float2(0.f, 0.f);
                const int kb0 = outer * p.kb_per_split, kb1 = min(p.total_kb, kb0 + p.kb_per_split);
                int next_mine = kb0 + n_blk;
                for (int kb = kb0; kb < kb1; ++kb) {
                    const bool mine = (kb == next_mine);
                    if (mine) next_mine += p.n_tiles;
                    if (ew == 0) {
                        if (CG == 2 && rank == 1) {
                            mbar_wait(&pfull_bar[cs_stage], cs_phase);
                        } else {
                            mbar_wait(&full_bar[cs_stage], cs_phase);    // the stage has landed; it is not refilled before we arrive
                            // the peer's full barrier is never signalled (its loads count on this one): tell its warp 0
                            if (CG == 2 && lane == 0) mbar_arrive_rank(&pfull_bar[cs_stage], 1);
                        }
                    }
                    if (mine) {
                        named_bar_sync(3, EPI_WARPS * 32);               // warp 0 has seen the stage
                        const uint32_t a = smem_u32(sA + cs_stage * A_STAGE_BYTES) + a_off;
#pragma unroll
                        for (int rr = 0; rr < 4; ++rr) {
                            const int k7 = (kg * 4 + rr) & 7;
                            uint4 v;
                            asm volatile("ld.shared.v4.b32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w)
                                         : "r"(a + rr * 128 + (((mc & 7) ^ k7) << 4)));
                            cs[0] = __fadd2_rn(cs[0], unpack_bf16(v.x)); cs[1] = __fadd2_rn(cs[1], unpack_bf16(v.y));
                            cs[2] = __fadd2_rn(cs[2], unpack_bf16(v.z)); cs[3] = __fadd2_rn(cs[3], unpack_bf16(v.w));
                        }
                        // the loads above have returned (their values were consumed): warps 1-7 check in, warp 0 collects
                        if (ew == 0) named_bar_sync(4, EPI_WARPS * 32); else named_bar_arrive(4, EPI_WARPS * 32);
                    }
                    if (ew == 0) {
                        __syncwarp();
                        if (lane == 0) mbar_arrive(&empty_bar[cs_stage]);
                    }
                    if (++cs_stage == STAGES) { cs_stage = 0; cs_phase ^= 1; }
                }
                // 16 k-groups -> one sum per column: through the (idle) staging tiles, then one atomic per column
                if (gtid == 0) tma_wait_group_read0();               // a previous tile's reduce-add may still read the staging tiles
                named_bar_sync(3, EPI_WARPS * 32);
                float* red = (float*)sStg;                           // [16 k-groups][128 columns]
                *(float4*)(red + kg * 128 + mc * 8) = make_float4(cs[0].x, cs[0].y, cs[1].x, cs[1].y);
                *(float4*)(red + kg * 128 + mc * 8 + 4) = make_float4(cs[2].x, cs[2].y, cs[3].x, cs[3].y);
                named_bar_sync(3, EPI_WARPS * 32);
                if (et < BM && m0 + et < p.M) {
                    float tot = 0.f;
#pragma unroll
                    for (int i = 0; i < 16; ++i) tot += red[i * 128 + et];
                    atomicAdd(p.colsum + m0 + et, tot);
                }
                named_bar_sync(3, EPI_WARPS * 32);                   // the epilogue below reuses the staging tiles
            }
            if (EPI == AVN_EPI_BIAS || EPI == AVN_EPI_BIAS_GELU || EPI == AVN_EPI_BIAS_GELU_DUAL || EPI == AVN_EPI_BIAS_RESIDUAL) {
                named_bar_sync(bar_id, 128);    // previous tile's readers of bias_s are done
                if (gtid < BN / 2) bias_s[gtid] = (p.bias != nullptr && n0 + gtid < p.N) ? p.bias[n0 + gtid] : 0.f;
                named_bar_sync(bar_id, 128);
            }
            // Residual / gelu' tile (R).  Staged path (RS, chosen at launch when R meets TMA's alignment rules): TMA brings one 128 x 64
            // tile per epilogue group into shared memory - the first one of an output tile is requested while the PREVIOUS
            // tile's epilogue ends, the next one as soon as every thread holds its row of the current one - and a thread reads
            // its row with eight 16-byte shared loads.  (Row-per-thread GLOBAL loads cost one L1 wavefront per row and load:
            // ncu showed the L1 data pipe 75 % busy with them and the tensor pipe starved down to 74 % in the dgrad x gelu'
            // GEMM.)  Fallback: those global loads, one 32-column unit ahead.
            constexpr bool HAS_R = (EPI == AVN_EPI_BIAS_RESIDUAL || EPI == AVN_EPI_DGELU);
            uint8_t* rstg = smem + L::OFF_RSTG + g * STG_BYTES;
            auto request_r = [&](int tile, int s64) {          // one thread per group: stage columns s64*64.. of `tile`
                const int nb = tile % p.n_tiles, rr = tile / p.n_tiles;
                const int mblk = (rr % m_units) * CG + rank, ob = rr / m_units;
                mbar_expect_tx(&rfull_bar[g], STG_BYTES);
                tma_load_3d(rstg, &tmC2, &rfull_bar[g], nb * BN + g * (BN / 2) + s64 * 64, mblk * BM, ob);
            };
            if (RS && t == tile0 && gtid == 0) request_r(t, 0);
            uint4 rnext[HAS_R ? 4 : 1];
            uint4 rrow[RS ? 8 : 1];
            auto fetch_r = [&](int u) {
                if (HAS_R) {
                    const int nc = n0 + u * 32;
                    const __nv_bfloat16* rp = p.R + (long long)mb * p.r_batch_stride + (long long)m * p.ldr + nc;
#pragma unroll
                    for (int i = 0; i < 4; ++i) {
                        rnext[i] = make_uint4(0, 0, 0, 0);
                        if (m < p.M && nc + 8 * i + 8 <= p.N) rnext[i] = __ldg((const uint4*)(rp + 8 * i));
                    }
                }
            };
            if (!RS) fetch_r(0);
            mbar_wait(&tfull_bar[acc], acc_phase);
            tc_fence_after();
            const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + acc * BN + g * (BN / 2);
            uint32_t keep[(EPI == AVN_EPI_BIAS_GELU_DUAL && !RS) ? 32 : 1];
#pragma unroll
            for (int u = 0; u < UNITS; ++u) {
                uint32_t v[32];
                tmem_ld32(taddr + u * 32, v);
                tmem_ld_wait();
                if (u == UNITS - 1) {           // accumulator fully read: hand the TMEM buffer back to the MMA warp
                    tc_fence_before();
                    __syncwarp();
                    if (lane == 0) { if (CG == 2) mbar_arrive_leader(&tempty_bar[acc]); else mbar_arrive(&tempty_bar[acc]); }
                }
                const int nc = n0 + u * 32;     // first global column of this unit
                if (EPI == AVN_EPI_F32_ADD) {
                    if (gtid == 0) tma_wait_group_read0();
                    named_bar_sync(bar_id, 128);
#pragma unroll
                    for (int i = 0; i < 8; ++i) {
                        uint4 w = make_uint4(v[4 * i], v[4 * i + 1], v[4 * i + 2], v[4 * i + 3]);
                        st_shared_v4(stg + row * 128 + ((i ^ (row & 7)) << 4), w);
                    }
                    fence_proxy_async_smem();
                    named_bar_sync(bar_id, 128);
                    if (gtid == 0) { tma_reduce_add_3d(&tmC, stg, nc, m0, mb); tma_commit_group(); }
                } else {
                    float f[32];
#pragma unroll
                    for (int j = 0; j < 32; ++j) f[j] = __uint_as_float(v[j]);
                    if (EPI == AVN_EPI_BIAS || EPI == AVN_EPI_BIAS_GELU || EPI == AVN_EPI_BIAS_GELU_DUAL || EPI == AVN_EPI_BIAS_RESIDUAL) {
#pragma unroll
                        for (int j = 0; j < 32; j += 4) {
                            const float4 b4 = *(const float4*)(bias_s + u * 32 + j);
                            const float2 r0 = __fadd2_rn(make_float2(f[j], f[j + 1]), make_float2(b4.x, b4.y));
                            const float2 r1 = __fadd2_rn(make_float2(f[j + 2], f[j + 3]), make_float2(b4.z, b4.w));
                            f[j] = r0.x; f[j + 1] = r0.y; f[j + 2] = r1.x; f[j + 3] = r1.y;
                        }
                    }
                    if (HAS_R) {
                        uint4 rcur[4];
                        if (RS) {
                            if ((u & 1) == 0) {        // a new 64-column R tile: take this row of it, then let the next one come
                                mbar_wait(&rfull_bar[g], r_phase);
                                r_phase ^= 1;
#pragma unroll
                                for (int c = 0; c < 8; ++c) {
                                    asm volatile("ld.shared.v4.b32 {%0, %1, %2, %3}, [%4];"
                                                 : "=r"(rrow[c].x), "=r"(rrow[c].y), "=r"(rrow[c].z), "=r"(rrow[c].w)
                                                 : "r"(smem_u32(rstg) + row * 128 + ((c ^ (row & 7)) << 4)));
                                }
                                named_bar_sync(bar_id, 128);
                                if (gtid == 0) {
                                    if (u + 2 < UNITS) request_r(t, (u >> 1) + 1);
                                    else if (t + tile_step < total_tiles) request_r(t + tile_step, 0);
                                }
                            }
#pragma unroll
                            for (int i = 0; i < 4; ++i) rcur[i] = rrow[(u & 1) * 4 + i];
                        } else {
#pragma unroll
                            for (int i = 0; i < 4; ++i) rcur[i] = rnext[i];
                            if (u + 1 < UNITS) fetch_r(u + 1);
                        }
#pragma unroll
                        for (int i = 0; i < 4; ++i) {
                            const uint4 rv = rcur[i];
                            const uint32_t rr[4] = {rv.x, rv.y, rv.z, rv.w};
#pragma unroll
                            for (int e = 0; e < 4; ++e) {
                                const float2 x = unpack_bf16(rr[e]);
                                const float2 cur = make_float2(f[8 * i + 2 * e], f[8 * i + 2 * e + 1]);
                                // DGELU: R holds the bf16 pre-activation the forward kept (what autograd saves for GELU's backward)
                                const float2 r = (EPI == AVN_EPI_BIAS_RESIDUAL) ? __fadd2_rn(cur, x) : __fmul2_rn(cur, gelu_erf_grad2(x));
                                f[8 * i + 2 * e] = r.x; f[8 * i + 2 * e + 1] = r.y;
                            }
                        }
                    }
                    if (EPI == AVN_EPI_BIAS_GELU_DUAL) {        // C = gelu(pre), C2 = pre: what the backward of the MLP needs
                        // (gelu' is evaluated by the EPI_DGELU epilogue of the dgrad GEMM, which has the issue slots to spare;
                        // computing both here made this epilogue the bottleneck of the kernel: 3.57 ms vs 2.82 ms for gelu alone)
#pragma unroll
                        uint32_t xp[16];
#pragma unroll
                        for (int j = 0; j < 16; ++j) {
                            xp[j] = pack_bf16(f[2 * j], f[2 * j + 1]);
                            if (!RS) keep[(u & 1) * 16 + j] = xp[j];
                            const float2 g = gelu_erf2(make_float2(f[2 * j], f[2 * j + 1]));
                            f[2 * j] = g.x; f[2 * j + 1] = g.y;
                        }
                        if (RS) {       // second output -> its own staging tile; both tiles leave together below
                            if ((u & 1) == 0) {
                                if (gtid == 0) tma_wait_group_read0();
                                named_bar_sync(bar_id, 128);
                            }
#pragma unroll
                            for (int i = 0; i < 4; ++i)
                                st_shared_v4(rstg + row * 128 + ((((u & 1) * 4 + i) ^ (row & 7)) << 4),
                                             make_uint4(xp[4 * i], xp[4 * i + 1], xp[4 * i + 2], xp[4 * i + 3]));
                        }
                    }
                    if (EPI == AVN_EPI_BIAS_GELU) {
#pragma unroll
                        for (int j = 0; j < 32; j += 2) {
                            const float2 g = gelu_erf2(make_float2(f[j], f[j + 1]));
                            f[j] = g.x; f[j + 1] = g.y;
                        }
                    }
                    if ((u & 1) == 0 && !(EPI == AVN_EPI_BIAS_GELU_DUAL && RS)) {         // starting a new 64-column staging tile
                        if (gtid == 0) tma_wait_group_read0();
                        named_bar_sync(bar_id, 128);
                    }
#pragma unroll
                    for (int i = 0; i < 4; ++i) {
                        uint4 w = make_uint4(pack_bf16(f[8 * i], f[8 * i + 1]), pack_bf16(f[8 * i + 2], f[8 * i + 3]),
                                             pack_bf16(f[8 * i + 4], f[8 * i + 5]), pack_bf16(f[8 * i + 6], f[8 * i + 7]));
                        const int c16 = (u & 1) * 4 + i;
                        st_shared_v4(stg + row * 128 + ((c16 ^ (row & 7)) << 4), w);
                    }
                    if ((u & 1) == 1) {
                        fence_proxy_async_smem();
                        named_bar_sync(bar_id, 128);
                        if (gtid == 0) {
                            if (EPI == AVN_EPI_BIAS_GELU_DUAL && RS) {
                                // two 4.9 GB outputs stream through L2: marked evict-first they stop pushing the A panels the
                                // other n-tiles still need out of it (ncu: DRAM read 3.7 GB vs 1.2 GB algorithmic); 3.65 -> 3.41 ms.
                                // The single-output epilogues measured 0-5 % SLOWER with the hint and keep the default policy.
                                const uint64_t pol = l2_policy_evict_first();
                                tma_store_3d_hint(&tmC, stg, nc - 32, m0, mb, pol);
                                tma_store_3d_hint(&tmC2, rstg, nc - 32, m0, mb, pol);
                            } else {
                                tma_store_3d(&tmC, stg, nc - 32, m0, mb);
                            }
                            tma_commit_group();
                        }
                        if (EPI == AVN_EPI_BIAS_GELU_DUAL && !RS) {
                            if (gtid == 0) tma_wait_group_read0();
                            named_bar_sync(bar_id, 128);
#pragma unroll
                            for (int i = 0; i < 8; ++i) {
                                uint4 w = make_uint4(keep[4 * i], keep[4 * i + 1], keep[4 * i + 2], keep[4 * i + 3]);
                                st_shared_v4(stg + row * 128 + ((i ^ (row & 7)) << 4), w);
                            }
                            fence_proxy_async_smem();
                            named_bar_sync(bar_id, 128);
                            if (gtid == 0) { tma_store_3d(&tmC2, stg, nc - 32, m0, mb); tma_commit_group(); }
                        }
                    }
                }
            }
            if (++acc == 2) { acc = 0; acc_phase ^= 1; }
        }
        if (gtid == 0) tma_wait_group0();
    }
    tc_fence_before();
    if (CG == 2) cluster_sync_all(); else __syncthreads();      // neither CTA may leave while its peer can still signal it
    if (warp == 1) { if (CG == 2) tmem_dealloc_2cta(tmem_base, 2 * BN); else tmem_dealloc(tmem_base, 2 * BN); }
}

template <int BN, int MODE, int EPI, int CG, bool RS>
int launch(const avn_gemm_desc* d, cudaStream_t stream) {
    using L = Smem<BN, CG, RS>;
    CUtensorMap tmA, tmB, tmC, tmC2;
    const bool tn = (MODE == AVN_GEMM_TN);
    int rc;
    // one 4-D box per MN-major operand and stage instead of one 3-D box per 64-column chunk (fewer TMA operations for the one
    // producer thread).  Needs whole chunks (the view would wrap a ragged row into the next one); AVN_GEMM_BOX4=0 disables it.
    static const bool box4_on = !(getenv("AVN_GEMM_BOX4") && atoi(getenv("AVN_GEMM_BOX4")) == 0);
    auto box4_ok = [&](int cols) { return box4_on && cols % 64 == 0; };
    bool a_box4 = false, b_box4 = false;
    // ---- A
    if (!tn) {
        uint64_t dims[3] = {(uint64_t)d->K, (uint64_t)d->M, (uint64_t)d->batch};
        uint64_t str[2] = {(uint64_t)d->lda * 2, (uint64_t)(d->batch > 1 ? d->a_batch_stride : d->lda * (int64_t)d->M) * 2};
        uint32_t box[3] = {64, 128, 1};
        if ((rc = avn_make_tmap(&tmA, d->A, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, 3, dims, str, box, true))) return rc;
    } else {
        const uint64_t bstr = (uint64_t)(d->batch > 1 ? d->a_batch_stride : d->lda * (int64_t)d->K) * 2;
        if (box4_ok(d->M)) {
            uint64_t dims[4] = {64, (uint64_t)d->K, (uint64_t)d->M / 64, (uint64_t)d->batch};
            uint64_t str[3] = {(uint64_t)d->lda * 2, 128, bstr};
            uint32_t box[4] = {64, 64, BM / 64, 1};
            a_box4 = avn_make_tmap(&tmA, d->A, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, 4, dims, str, box, true) == AVN_OK;
        }
        if (!a_box4) {
            uint64_t dims[3] = {(uint64_t)d->M, (uint64_t)d->K, (uint64_t)d->batch};
            uint64_t str[2] = {(uint64_t)d->lda * 2, bstr};
            uint32_t box[3] = {64, 64, 1};
            if ((rc = avn_make_tmap(&tmA, d->A, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, 3, dims, str, box, true))) return rc;
        }
    }
    // ---- B
    if (MODE == AVN_GEMM_NT) {
        uint64_t dims[3] = {(uint64_t)d->K, (uint64_t)d->N, 1};
        uint64_t str[2] = {(uint64_t)d->ldb * 2, (uint64_t)d->ldb * 2 * (uint64_t)d->N};
        uint32_t box[3] = {64, (uint32_t)(BN / CG), 1};
        if ((rc = avn_make_tmap(&tmB, d->B, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, 3, dims, str, box, true))) return rc;
    } else {
        const int bb = tn ? d->batch : 1;
        const uint64_t bstr = (uint64_t)(bb > 1 ? d->b_batch_stride : d->ldb * (int64_t)d->K) * 2;
        if (box4_ok(d->N)) {
            uint64_t dims[4] = {64, (uint64_t)d->K, (uint64_t)d->N / 64, (uint64_t)bb};
            uint64_t str[3] = {(uint64_t)d->ldb * 2, 128, bstr};
            uint32_t box[4] = {64, 64, (uint32_t)(BN / CG / 64), 1};
            b_box4 = avn_make_tmap(&tmB, d->B, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, 4, dims, str, box, true) == AVN_OK;
        }
        if (!b_box4) {
            uint64_t dims[3] = {(uint64_t)d->N, (uint64_t)d->K, (uint64_t)bb};
            uint64_t str[2] = {(uint64_t)d->ldb * 2, bstr};
            uint32_t box[3] = {64, 64, 1};
            if ((rc = avn_make_tmap(&tmB, d->B, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, 3, dims, str, box, true))) return rc;
        }
    }
    // ---- C (and C2)
    const int cb = tn ? 1 : d->batch;
    if (EPI == AVN_EPI_F32_ADD) {
        uint64_t dims[3] = {(uint64_t)d->N, (uint64_t)d->M, 1};
        uint64_t str[2] = {(uint64_t)d->ldc * 4, (uint64_t)d->ldc * 4 * (uint64_t)d->M};
        uint32_t box[3] = {32, 128, 1};
        if ((rc = avn_make_tmap(&tmC, d->C, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 4, 3, dims, str, box, true))) return rc;
        tmC2 = tmC;
    } else {
        uint64_t dims[3] = {(uint64_t)d->N, (uint64_t)d->M, (uint64_t)cb};
        uint64_t str[2] = {(uint64_t)d->ldc * 2, (uint64_t)(cb > 1 ? d->c_batch_stride : d->ldc * (int64_t)d->M) * 2};
        uint32_t box[3] = {64, 128, 1};
        if ((rc = avn_make_tmap(&tmC, d->C, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, 3, dims, str, box, true))) return rc;
        tmC2 = tmC;
        if (EPI == AVN_EPI_BIAS_GELU_DUAL)
            if ((rc = avn_make_tmap(&tmC2, d->C2, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, 3, dims, str, box, true))) return rc;
    }
    KParams p;
    p.M = d->M; p.N = d->N; p.K = d->K; p.batch = d->batch;
    p.m_tiles = (d->M + BM - 1) / BM;
    p.n_tiles = (d->N + BN - 1) / BN;
    p.kblocks = (d->K + BK - 1) / BK;
    p.total_kb = p.kblocks * d->batch;
    int splits = 1;
    if (tn) {
        splits = d->splits;
        if (splits <= 0) {
            // auto: the smallest split count (<= 16) whose tiles fill whole waves of CTAs (or CTA pairs) to >= 95 %, else the best one
            const int units = ((p.m_tiles + CG - 1) / CG) * p.n_tiles, slots = avn_num_sms() / CG;
            double best = 0.0;
            splits = 1;
            for (int sp = 1; sp <= 16; ++sp) {
                const int tiles = units * sp, waves = (tiles + slots - 1) / slots;
                const double eff = (double)tiles / ((double)waves * slots);
                if (eff > best + 1e-9) { best = eff; splits = sp; }
                if (eff >= 0.95) { splits = sp; break; }
            }
        }
    }
    if (splits > p.total_kb) splits = p.total_kb;
    p.kb_per_split = (p.total_kb + splits - 1) / splits;
    p.splits = (p.total_kb + p.kb_per_split - 1) / p.kb_per_split;
    p.bias = d->bias;
    p.R = (const __nv_bfloat16*)d->R; p.ldr = d->ldr; p.r_batch_stride = d->r_batch_stride;
    p.colsum = tn ? d->colsum : nullptr;
    p.a_box4 = a_box4; p.b_box4 = b_box4;
    if (RS && EPI != AVN_EPI_BIAS_GELU_DUAL) {        // the R tile is staged by TMA (r_stage_ok() checked the alignment)
        const int rb = d->batch;
        uint64_t dims[3] = {(uint64_t)d->N, (uint64_t)d->M, (uint64_t)rb};
        uint64_t str[2] = {(uint64_t)d->ldr * 2, (uint64_t)(rb > 1 ? d->r_batch_stride : d->ldr * (int64_t)d->M) * 2};
        uint32_t box[3] = {64, 128, 1};
        if ((rc = avn_make_tmap(&tmC2, d->R, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, 3, dims, str, box, true))) return rc;
    }
    const int total_tiles = (tn ? p.splits : d->batch) * ((p.m_tiles + CG - 1) / CG) * p.n_tiles;
    int grid = avn_num_sms() / CG;          // CTA pairs (or CTAs)
    if (grid > total_tiles) grid = total_tiles;
    grid *= CG;
    auto kern = gemm_kernel<BN, MODE, EPI, CG, RS>;
    AVN_CHECK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, L::TOTAL));
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(grid); cfg.blockDim = dim3(NTHREADS); cfg.dynamicSmemBytes = L::TOTAL; cfg.stream = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = CG; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr; cfg.numAttrs = 1;
    AVN_CHECK_CUDA(cudaLaunchKernelEx(&cfg, kern, tmA, tmB, tmC, tmC2, p));
    return AVN_OK;
}

// R through shared memory?  Only where TMA can address it (16-byte aligned rows; a table shared by all batch entries has
// stride 0) and where it pays: the staged kernels give up one ring stage, and with a long contraction (fc2: K = 3072) the
// row-per-thread loads are rare enough not to matter (2.74 ms vs 2.82 ms staged), so those keep the deeper ring.
static bool r_stage_ok(const avn_gemm_desc* d, int epi) {
    static const bool on = !(getenv("AVN_GEMM_RTMA") && atoi(getenv("AVN_GEMM_RTMA")) == 0);
    if (on && epi == AVN_EPI_BIAS_GELU_DUAL) return true;        // second output staged in its own tile: nothing to align
    if (!on || d->R == nullptr || ((uintptr_t)d->R % 16) != 0 || d->ldr % 8 != 0) return false;
    if (d->batch > 1 && (d->r_batch_stride == 0 || d->r_batch_stride % 8 != 0)) return false;
    return epi == AVN_EPI_DGELU || d->K <= 1536;
}
template <int BN, int MODE, int EPI, int CG>
int launch_rs(const avn_gemm_desc* d, cudaStream_t stream) {
    if constexpr (r_stageable(BN, EPI, CG)) {
        if (r_stage_ok(d, EPI)) return launch<BN, MODE, EPI, CG, true>(d, stream);
    }
    return launch<BN, MODE, EPI, CG, false>(d, stream);
}

template <int MODE, int EPI>
int launch_bn(const avn_gemm_desc* d, cudaStream_t stream) {
    // 256-wide tiles when N fills them and there are enough tiles to occupy the SMs; else 128-wide.
    const long long tiles256 = (long long)((d->M + 127) / 128) * ((d->N + 255) / 256) * (MODE == AVN_GEMM_TN ? 1 : d->batch);
    // big problems: CTA pairs (cta_group::2) on 256 x 256 tiles; AVN_GEMM_CG=1 forces the single-CTA kernel (A/B comparisons)
    static const int force_cg = getenv("AVN_GEMM_CG") ? atoi(getenv("AVN_GEMM_CG")) : 0;
    // thresholds in units of the SM count (AVN_GEMM_T256 / AVN_GEMM_TPAIR override them for A/B runs).  1.5 / 1.5: measured on
    // the VideoMAE encoder (M = 10240, N = 768: 240 tiles) - 256-wide CTA-pair tiles beat 128-wide single-CTA tiles there
    // (dgrad 3.06 -> 2.48 ms, bias+residual 2.15 -> 1.81 ms per step); the former 2.0 / 4.0 kept those on the 128-wide path.
    static const double t256 = getenv("AVN_GEMM_T256") ? atof(getenv("AVN_GEMM_T256")) : 1.5;
    static const double tpair = getenv("AVN_GEMM_TPAIR") ? atof(getenv("AVN_GEMM_TPAIR")) : 1.5;
    if (d->N % 256 == 0 && (tiles256 >= t256 * avn_num_sms() || MODE == AVN_GEMM_TN)) {
        const bool pair_ok = (MODE == AVN_GEMM_TN) ? (d->M % 256 == 0) : (tiles256 >= tpair * avn_num_sms());
        if (force_cg != 1 && pair_ok) return launch_rs<256, MODE, EPI, 2>(d, stream);
        return launch_rs<256, MODE, EPI, 1>(d, stream);
    }
    return launch_rs<128, MODE, EPI, 1>(d, stream);
}
}  // namespace

extern "C" int avn_gemm_bf16(const avn_gemm_desc* d, avn_stream_t stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    AVN_REQUIRE(d != nullptr, "gemm: null descriptor");
    AVN_REQUIRE(d->M > 0 && d->N > 0 && d->K > 0 && d->batch > 0, "gemm: bad sizes M=%d N=%d K=%d batch=%d", d->M, d->N, d->K, d->batch);
    AVN_REQUIRE(d->A && d->B && d->C, "gemm: null operand");
    AVN_REQUIRE(d->lda % 8 == 0 && d->ldb % 8 == 0, "gemm: lda/ldb must be multiples of 8 elements (16 B)");
    AVN_REQUIRE(((uintptr_t)d->A % 16) == 0 && ((uintptr_t)d->B % 16) == 0 && ((uintptr_t)d->C % 16) == 0, "gemm: operands must be 16-byte aligned");
    AVN_REQUIRE(d->N % 8 == 0, "gemm: N must be a multiple of 8");
    const int e = d->epi;
    if (d->mode == AVN_GEMM_NT) {
        AVN_REQUIRE(d->ldc % 8 == 0, "gemm: ldc must be a multiple of 8");
        if (e == AVN_EPI_BIAS_RESIDUAL) AVN_REQUIRE(d->R != nullptr, "gemm: residual epilogue needs R");
        if (e == AVN_EPI_BIAS_GELU_DUAL) AVN_REQUIRE(d->C2 != nullptr, "gemm: dual epilogue needs C2");
        switch (e) {
            case AVN_EPI_NONE: return launch_bn<AVN_GEMM_NT, AVN_EPI_NONE>(d, stream);
            case AVN_EPI_BIAS: return launch_bn<AVN_GEMM_NT, AVN_EPI_BIAS>(d, stream);
            case AVN_EPI_BIAS_GELU: return launch_bn<AVN_GEMM_NT, AVN_EPI_BIAS_GELU>(d, stream);
            case AVN_EPI_BIAS_GELU_DUAL: return launch_bn<AVN_GEMM_NT, AVN_EPI_BIAS_GELU_DUAL>(d, stream);
            case AVN_EPI_BIAS_RESIDUAL: return launch_bn<AVN_GEMM_NT, AVN_EPI_BIAS_RESIDUAL>(d, stream);
        }
    } else if (d->mode == AVN_GEMM_NN) {
        AVN_REQUIRE(d->ldc % 8 == 0, "gemm: ldc must be a multiple of 8");
        if (e == AVN_EPI_DGELU) AVN_REQUIRE(d->R != nullptr, "gemm: dgelu epilogue needs R (the pre-activation)");
        switch (e) {
            case AVN_EPI_NONE: return launch_bn<AVN_GEMM_NN, AVN_EPI_NONE>(d, stream);
            case AVN_EPI_DGELU: return launch_bn<AVN_GEMM_NN, AVN_EPI_DGELU>(d, stream);
        }
    } else if (d->mode == AVN_GEMM_TN) {
        AVN_REQUIRE(d->ldc % 4 == 0, "gemm: fp32 ldc must be a multiple of 4");
        AVN_REQUIRE(d->M % 8 == 0, "gemm: TN needs M %% 8 == 0");
        if (e == AVN_EPI_F32_ADD) return launch_bn<AVN_GEMM_TN, AVN_EPI_F32_ADD>(d, stream);
    }
    avn_set_error("gemm: unsupported mode/epilogue combination mode=%d epi=%d", d->mode, e);
    return AVN_ERR_ARG;
}
