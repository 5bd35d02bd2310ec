// V3: LayerNorm forward / backward over bf16 rows with fp32 statistics (HBM-bound: 4 B/element fwd, 6-8 B/element bwd).
// Replaces the ATen layer_norm behind ln_pre / ln_1 / ln_2 / ln_post / ln_final
// (avion/models/transformer.py:35-41,110,124,213,245,264).
//
// Both kernels are persistent and TMA-staged: tiles of 8 consecutive rows (one contiguous block of 8*D*2 bytes per tensor ->
// ONE cp.async.bulk each) stream through a multi-stage shared-memory ring guarded by full/empty mbarriers; the 8 warps take
// one row each, and warp 0 also issues the load that runs stages-1 tiles ahead (no dedicated producer warp: 256 threads x 2
// CTAs leave 128 registers per thread for the column accumulators).  The ring keeps >= 100 KB of loads in flight per SM
// whatever the warps are doing
// (the previous one-warp-per-row kernels stopped loading while a warp reduced / normalised / stored: 58-62 % of the copy
// bandwidth).  Results leave through 16-byte global stores straight from registers.
// The backward also folds in the residual-stream add (x + f(x), transformer.py:148-149) and the column sum that is the
// bias gradient of the Linear feeding this residual stream; per-lane column accumulators (dgamma, dbeta, column sum) live in
// registers across all rows of the CTA and are reduced once at the end.
#include "common.cuh"
#include "../../include/avion_b200.h"

namespace {
using namespace avn;

constexpr int ROWS_PER_TILE = 8;      // = warps
constexpr int NTHREADS = ROWS_PER_TILE * 32;

__device__ __forceinline__ void bulk_load(void* smem_dst, const void* gsrc, uint32_t bytes, uint64_t* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(smem_u32(smem_dst)), "l"(gsrc), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ uint4 lds128(const void* p) {
    uint4 v;
    asm volatile("ld.shared.v4.b32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(smem_u32(p)));
    return v;
}

struct Ring {
    uint64_t* full;
    uint64_t* empty;
    uint8_t* data;
    int stages, stage_bytes;
};

// Carve the dynamic shared memory: [barriers 128 B][stages * stage_bytes].  Thread 0 initialises the barriers.
__device__ __forceinline__ Ring ring_init(uint8_t* smem, int stages, int stage_bytes) {
    Ring r;
    r.full = (uint64_t*)smem;
    r.empty = r.full + 8;
    r.data = smem + 128;
    r.stages = stages;
    r.stage_bytes = stage_bytes;
    if (threadIdx.x == 0) {
        for (int i = 0; i < stages; ++i) { mbar_init(&r.full[i], 1); mbar_init(&r.empty[i], ROWS_PER_TILE); }
        fence_barrier_init();
    }
    __syncthreads();
    return r;
}

// ---------------------------------------------------------------------------------------------------------------------
template <int VPL>   // 16-byte vectors (8 bf16) per lane; D <= VPL*256
__global__ void __launch_bounds__(NTHREADS, 2)
ln_fwd_kernel(const __nv_bfloat16* __restrict__ x, const float* __restrict__ gamma, const float* __restrict__ beta,
              __nv_bfloat16* __restrict__ y, float* __restrict__ mean_out, float* __restrict__ rstd_out, long long rows, int D, float eps,
              int stages) {
    extern __shared__ __align__(128) uint8_t smem[];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int row_bytes = D * 2;
    Ring ring = ring_init(smem, stages, ROWS_PER_TILE * row_bytes);
    const long long n_tiles = (rows + ROWS_PER_TILE - 1) / ROWS_PER_TILE;
    const int nvec = D >> 3;

    // warp 0: load number j (tile blockIdx.x + j * gridDim.x) goes into slot j % stages once its previous occupant was released
    auto produce = [&](int j) {
        const long long tile = blockIdx.x + (long long)j * gridDim.x;
        if (tile >= n_tiles) return;
        const int s = j % stages;
        mbar_wait_parked(&ring.empty[s], ((j / stages) & 1) ^ 1);
        if (elect_one()) {
            const long long r0 = tile * ROWS_PER_TILE;
            const uint32_t bytes = (uint32_t)(min((long long)ROWS_PER_TILE, rows - r0) * row_bytes);
            mbar_expect_tx(&ring.full[s], bytes);
            bulk_load(ring.data + s * ring.stage_bytes, x + r0 * D, bytes, &ring.full[s]);
        }
        __syncwarp();
    };
    if (warp == 0)
        for (int j = 0; j < stages - 1; ++j) produce(j);
    // ===================== one row per warp =====================
    // gamma / beta of this lane's columns stay in registers for all rows of the CTA (as f32x2 pairs: all the per-element
    // arithmetic below is packed FADD2 / FMUL2 / FFMA2 - the kernel is bound by instruction issue, not by HBM, once the ring
    // hides the load latency: 433 -> ~150 warp instructions per row)
    float2 g[VPL][4], bt[VPL][4];
#pragma unroll
    for (int j = 0; j < VPL; ++j) {
        const int i = lane + 32 * j;
#pragma unroll
        for (int e = 0; e < 4; ++e) { g[j][e] = make_float2(0.f, 0.f); bt[j][e] = make_float2(0.f, 0.f); }
        if (i < nvec) {
            const float4 g0 = __ldg((const float4*)(gamma + i * 8)), g1 = __ldg((const float4*)(gamma + i * 8 + 4));
            const float4 b0 = __ldg((const float4*)(beta + i * 8)), b1 = __ldg((const float4*)(beta + i * 8 + 4));
            g[j][0] = make_float2(g0.x, g0.y); g[j][1] = make_float2(g0.z, g0.w); g[j][2] = make_float2(g1.x, g1.y); g[j][3] = make_float2(g1.z, g1.w);
            bt[j][0] = make_float2(b0.x, b0.y); bt[j][1] = make_float2(b0.z, b0.w); bt[j][2] = make_float2(b1.x, b1.y); bt[j][3] = make_float2(b1.z, b1.w);
        }
    }
    const float inv_d = 1.0f / (float)D;
    int it = 0;
    for (long long tile = blockIdx.x; tile < n_tiles; tile += gridDim.x, ++it) {
        const int s = it % stages;
        const long long row = tile * ROWS_PER_TILE + warp;
        if (warp == 0) produce(it + stages - 1);             // refills the slot every warp released in iteration it-1
        mbar_wait_parked(&ring.full[s], (it / stages) & 1);
        float2 v[VPL][4];
        float2 acc[VPL];                                     // one accumulation chain per vector: independent FADD2 streams
        if (row < rows) {
            const uint8_t* xr = ring.data + s * ring.stage_bytes + warp * row_bytes;
#pragma unroll
            for (int j = 0; j < VPL; ++j) {
                const int i = lane + 32 * j;
                acc[j] = make_float2(0.f, 0.f);
                if (i < nvec) {
                    const uint4 u = lds128(xr + i * 16);
                    v[j][0] = unpack_bf16(u.x); v[j][1] = unpack_bf16(u.y); v[j][2] = unpack_bf16(u.z); v[j][3] = unpack_bf16(u.w);
                    acc[j] = __fadd2_rn(__fadd2_rn(v[j][0], v[j][1]), __fadd2_rn(v[j][2], v[j][3]));
                } else {
#pragma unroll
                    for (int e = 0; e < 4; ++e) v[j][e] = make_float2(0.f, 0.f);
                }
            }
        }
        __syncwarp();
        if (lane == 0) mbar_arrive(&ring.empty[s]);          // the row is in registers: the slot can be refilled
        if (row >= rows) continue;
        float2 tot = acc[0];
#pragma unroll
        for (int j = 1; j < VPL; ++j) tot = __fadd2_rn(tot, acc[j]);
        const float mean = warp_sum(tot.x + tot.y) * inv_d;
        const float2 nmean = make_float2(-mean, -mean);
#pragma unroll
        for (int j = 0; j < VPL; ++j) {                      // v <- v - mean (kept: the output needs it again), acc <- sum of squares
            acc[j] = make_float2(0.f, 0.f);
            if (lane + 32 * j < nvec) {
#pragma unroll
                for (int e = 0; e < 4; ++e) { v[j][e] = __fadd2_rn(v[j][e], nmean); acc[j] = __ffma2_rn(v[j][e], v[j][e], acc[j]); }
            }
        }
        tot = acc[0];
#pragma unroll
        for (int j = 1; j < VPL; ++j) tot = __fadd2_rn(tot, acc[j]);
        const float rstd = rsqrtf(warp_sum(tot.x + tot.y) * inv_d + eps);
        if (lane == 0) {
            if (mean_out) mean_out[row] = mean;
            if (rstd_out) rstd_out[row] = rstd;
        }
        const float2 rs2 = make_float2(rstd, rstd);
        uint4* yr = (uint4*)(y + row * D);
#pragma unroll
        for (int j = 0; j < VPL; ++j) {
            const int i = lane + 32 * j;
            if (i < nvec) {
                uint32_t o[4];
#pragma unroll
                for (int e = 0; e < 4; ++e) {
                    const float2 r = __ffma2_rn(__fmul2_rn(v[j][e], rs2), g[j][e], bt[j][e]);      // ((x - mean) * rstd) * g + b
                    o[e] = pack_bf16(r.x, r.y);
                }
                yr[i] = make_uint4(o[0], o[1], o[2], o[3]);
            }
        }
    }
}

// ---------------------------------------------------------------------------------------------------------------------
template <int VPL>
__global__ void __launch_bounds__(NTHREADS, (VPL <= 3) ? 2 : 1)
ln_bwd_kernel(const __nv_bfloat16* __restrict__ dy, const __nv_bfloat16* __restrict__ x, const float* __restrict__ gamma,
              const float* __restrict__ mean_in, const float* __restrict__ rstd_in, const __nv_bfloat16* __restrict__ dres,
              __nv_bfloat16* __restrict__ dx, float* __restrict__ dgamma, float* __restrict__ dbeta, float* __restrict__ dcolsum,
              long long rows, int D, int stages) {
    extern __shared__ __align__(128) uint8_t smem[];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int row_bytes = D * 2, tens_bytes = ROWS_PER_TILE * row_bytes;
    const int ntens = dres ? 3 : 2;
    Ring ring = ring_init(smem, stages, 3 * tens_bytes);      // a stage = [x tile][dy tile][dres tile]
    const long long n_tiles = (rows + ROWS_PER_TILE - 1) / ROWS_PER_TILE;
    const int nvec = D >> 3;

    auto produce = [&](int j) {                               // warp 0, see ln_fwd_kernel
        const long long tile = blockIdx.x + (long long)j * gridDim.x;
        if (tile >= n_tiles) return;
        const int s = j % stages;
        mbar_wait_parked(&ring.empty[s], ((j / stages) & 1) ^ 1);
        if (elect_one()) {
            const long long r0 = tile * ROWS_PER_TILE;
            const uint32_t bytes = (uint32_t)(min((long long)ROWS_PER_TILE, rows - r0) * row_bytes);
            uint8_t* dst = ring.data + s * ring.stage_bytes;
            mbar_expect_tx(&ring.full[s], ntens * bytes);
            bulk_load(dst, x + r0 * D, bytes, &ring.full[s]);
            bulk_load(dst + tens_bytes, dy + r0 * D, bytes, &ring.full[s]);
            if (dres) bulk_load(dst + 2 * tens_bytes, dres + r0 * D, bytes, &ring.full[s]);
        }
        __syncwarp();
    };
    if (warp == 0)
        for (int j = 0; j < stages - 1; ++j) produce(j);
    {
        // per-lane column accumulators (dgamma, dbeta, column sum of dx) across all rows of this warp
        float ag[VPL][8], ab[VPL][8], ac[VPL][8];
#pragma unroll
        for (int j = 0; j < VPL; ++j)
#pragma unroll
            for (int e = 0; e < 8; ++e) ag[j][e] = ab[j][e] = ac[j][e] = 0.f;
        int it = 0;
        for (long long tile = blockIdx.x; tile < n_tiles; tile += gridDim.x, ++it) {
            const int s = it % stages;
            const long long row = tile * ROWS_PER_TILE + warp;
            float mean = 0.f, rstd = 0.f;
            if (row < rows) { mean = __ldg(mean_in + row); rstd = __ldg(rstd_in + row); }      // issued before the waits
            if (warp == 0) produce(it + stages - 1);
            mbar_wait_parked(&ring.full[s], (it / stages) & 1);
            if (row < rows) {
                const uint8_t* xr = ring.data + s * ring.stage_bytes + warp * row_bytes;
                const uint8_t* dyr = xr + tens_bytes;
                const uint8_t* drr = xr + 2 * tens_bytes;
                // all the per-element arithmetic is packed f32x2 (FFMA2 / FADD2 / FMUL2)
                const float2 rs2 = make_float2(rstd, rstd), nmr2 = make_float2(-mean * rstd, -mean * rstd);
                float2 s1a[VPL], s2a[VPL];                    // one chain per vector (independent FADD2 / FFMA2 streams)
                // x and dy are read from the staged tile in both passes (shared-memory bandwidth is idle here; keeping them in
                // registers next to the 72 column accumulators spilled)
#pragma unroll
                for (int j = 0; j < VPL; ++j) {
                    const int i = lane + 32 * j;
                    float2 s1v = make_float2(0.f, 0.f), s2v = make_float2(0.f, 0.f);
                    if (i < nvec) {
                        const uint4 ux = lds128(xr + i * 16), ud = lds128(dyr + i * 16);
                        const uint32_t wx[4] = {ux.x, ux.y, ux.z, ux.w}, wd[4] = {ud.x, ud.y, ud.z, ud.w};
                        const float4 g0 = __ldg((const float4*)(gamma + i * 8)), g1 = __ldg((const float4*)(gamma + i * 8 + 4));
                        const float2 g[4] = {make_float2(g0.x, g0.y), make_float2(g0.z, g0.w), make_float2(g1.x, g1.y), make_float2(g1.z, g1.w)};
#pragma unroll
                        for (int e = 0; e < 4; ++e) {
                            const float2 fd = unpack_bf16(wd[e]);
                            const float2 h = __ffma2_rn(unpack_bf16(wx[e]), rs2, nmr2);       // (x - mean) * rstd
                            float2 a = make_float2(ag[j][2 * e], ag[j][2 * e + 1]), bsum = make_float2(ab[j][2 * e], ab[j][2 * e + 1]);
                            a = __ffma2_rn(fd, h, a);
                            bsum = __fadd2_rn(bsum, fd);
                            ag[j][2 * e] = a.x; ag[j][2 * e + 1] = a.y; ab[j][2 * e] = bsum.x; ab[j][2 * e + 1] = bsum.y;
                            const float2 t = __fmul2_rn(fd, g[e]);
                            s1v = __fadd2_rn(s1v, t);
                            s2v = __ffma2_rn(t, h, s2v);
                        }
                    }
                    s1a[j] = s1v; s2a[j] = s2v;
                }
                float2 s1v = s1a[0], s2v = s2a[0];
#pragma unroll
                for (int j = 1; j < VPL; ++j) { s1v = __fadd2_rn(s1v, s1a[j]); s2v = __fadd2_rn(s2v, s2a[j]); }
                const float inv_d = 1.0f / (float)D;
                const float c1 = warp_sum(s1v.x + s1v.y) * inv_d, c2 = warp_sum(s2v.x + s2v.y) * inv_d;
                const float2 nc1 = make_float2(-c1, -c1), nc2 = make_float2(-c2, -c2);
                uint4* dxr = (uint4*)(dx + row * D);
#pragma unroll
                for (int j = 0; j < VPL; ++j) {
                    const int i = lane + 32 * j;
                    if (i < nvec) {
                        const uint4 ux = lds128(xr + i * 16), ud = lds128(dyr + i * 16);
                        const uint32_t wx[4] = {ux.x, ux.y, ux.z, ux.w}, wd[4] = {ud.x, ud.y, ud.z, ud.w};
                        uint4 ur = make_uint4(0, 0, 0, 0);
                        if (dres) ur = lds128(drr + i * 16);
                        const uint32_t wr[4] = {ur.x, ur.y, ur.z, ur.w};
                        const float4 g0 = __ldg((const float4*)(gamma + i * 8)), g1 = __ldg((const float4*)(gamma + i * 8 + 4));
                        const float2 g[4] = {make_float2(g0.x, g0.y), make_float2(g0.z, g0.w), make_float2(g1.x, g1.y), make_float2(g1.z, g1.w)};
                        uint32_t po[4];
#pragma unroll
                        for (int e = 0; e < 4; ++e) {
                            const float2 h = __ffma2_rn(unpack_bf16(wx[e]), rs2, nmr2);
                            float2 u = __ffma2_rn(unpack_bf16(wd[e]), g[e], nc1);            // dy*g - c1
                            u = __ffma2_rn(h, nc2, u);                                          //       - h*c2
                            const float2 o = __ffma2_rn(u, rs2, unpack_bf16(wr[e]));           // * rstd + residual gradient
                            po[e] = pack_bf16(o.x, o.y);
                            if (dcolsum) {   // column sum of the values actually stored (bf16-rounded), as the next consumer sees them
                                const float2 c = __fadd2_rn(make_float2(ac[j][2 * e], ac[j][2 * e + 1]), unpack_bf16(po[e]));
                                ac[j][2 * e] = c.x; ac[j][2 * e + 1] = c.y;
                            }
                        }
                        dxr[i] = make_uint4(po[0], po[1], po[2], po[3]);
                    }
                }
            }
            __syncwarp();
            if (lane == 0) mbar_arrive(&ring.empty[s]);      // every shared-memory read of this row has completed
        }
        // this warp's column accumulators -> shared memory (the ring is quiescent after the barrier below)
        __syncwarp();
        __syncthreads();                                     // all warps are out of the ring (every issued load was consumed)
        float* red = (float*)ring.data;                      // [8 warps][3][VPL*256]
        const int W = VPL * 256;
#pragma unroll
        for (int j = 0; j < VPL; ++j)
#pragma unroll
            for (int e = 0; e < 8; ++e) {
                const int c = (lane + 32 * j) * 8 + e;
                red[(warp * 3 + 0) * W + c] = ag[j][e];
                red[(warp * 3 + 1) * W + c] = ab[j][e];
                red[(warp * 3 + 2) * W + c] = ac[j][e];
            }
        __syncthreads();
        // block reduction of the three column accumulators, then one atomic per column per block
        for (int c = threadIdx.x; c < D; c += ROWS_PER_TILE * 32) {
            float a = 0.f, b = 0.f, cc = 0.f;
#pragma unroll
            for (int w = 0; w < ROWS_PER_TILE; ++w) { a += red[(w * 3 + 0) * W + c]; b += red[(w * 3 + 1) * W + c]; cc += red[(w * 3 + 2) * W + c]; }
            if (dgamma) atomicAdd(dgamma + c, a);
            if (dbeta) atomicAdd(dbeta + c, b);
            if (dcolsum) atomicAdd(dcolsum + c, cc);
        }
    }
}

int pick_stages(size_t stage_bytes, size_t min_bytes, int ctas_per_sm, int max_stages) {
    // as many stages as fit beside (ctas_per_sm - 1) sibling CTAs in 227 KB (1 KB per CTA is reserved by the runtime)
    const size_t budget = (size_t)227 * 1024 / ctas_per_sm - 1024 - 128;
    int st = (int)(budget / stage_bytes);
    if (st > max_stages) st = max_stages;
    if (st < 2) st = 2;
    while ((size_t)st * stage_bytes < min_bytes) ++st;
    return st;
}

template <int VPL>
int launch_fwd(const void* x, const float* gamma, const float* beta, void* y, float* mean, float* rstd, int64_t rows, int D, float eps,
               cudaStream_t stream) {
    const size_t stage = (size_t)ROWS_PER_TILE * D * 2;
    const int stages = pick_stages(stage, 0, 2, 8);
    const size_t smem = 128 + stages * stage;
    auto kern = ln_fwd_kernel<VPL>;
    AVN_CHECK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    const long long tiles = (rows + ROWS_PER_TILE - 1) / ROWS_PER_TILE;
    const long long cap = (long long)avn_num_sms() * 2;
    kern<<<(int)(tiles < cap ? tiles : cap), NTHREADS, smem, stream>>>((const __nv_bfloat16*)x, gamma, beta, (__nv_bfloat16*)y, mean, rstd,
                                                                     rows, D, eps, stages);
    AVN_CHECK_CUDA(cudaGetLastError());
    return AVN_OK;
}

template <int VPL>
int launch_bwd(const void* dy, const void* x, const float* gamma, const float* mean, const float* rstd, const void* dres, void* dx,
               float* dgamma, float* dbeta, float* dcolsum, int64_t rows, int D, cudaStream_t stream) {
    const int ctas = VPL <= 3 ? 2 : 1;
    const size_t stage = (size_t)3 * ROWS_PER_TILE * D * 2;
    const size_t red = (size_t)ROWS_PER_TILE * 3 * VPL * 256 * sizeof(float);      // final reduction buffer aliases the ring
    const int stages = pick_stages(stage, red, ctas, 6);
    const size_t smem = 128 + stages * stage;
    AVN_REQUIRE(smem <= 227 * 1024, "layernorm_bwd: D=%d does not fit shared memory", D);
    auto kern = ln_bwd_kernel<VPL>;
    AVN_CHECK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    const long long tiles = (rows + ROWS_PER_TILE - 1) / ROWS_PER_TILE;
    const long long cap = (long long)avn_num_sms() * ctas;
    kern<<<(int)(tiles < cap ? tiles : cap), NTHREADS, smem, stream>>>((const __nv_bfloat16*)dy, (const __nv_bfloat16*)x, gamma, mean, rstd,
                                                                     (const __nv_bfloat16*)dres, (__nv_bfloat16*)dx, dgamma, dbeta, dcolsum,
                                                                     rows, D, stages);
    AVN_CHECK_CUDA(cudaGetLastError());
    return AVN_OK;
}
}  // namespace

extern "C" int avn_layernorm_fwd(const void* x, const float* gamma, const float* beta, void* y, float* mean, float* rstd,
                                 int64_t rows, int D, float eps, avn_stream_t stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    AVN_REQUIRE(x && gamma && beta && y, "layernorm_fwd: null pointer");
    AVN_REQUIRE(rows > 0 && D > 0 && D % 8 == 0 && D <= 1024, "layernorm_fwd: D=%d must be a multiple of 8 and <= 1024", D);
    AVN_REQUIRE(((uintptr_t)x % 16) == 0 && ((uintptr_t)y % 16) == 0, "layernorm_fwd: x and y must be 16-byte aligned");
    if (D <= 256) return launch_fwd<1>(x, gamma, beta, y, mean, rstd, rows, D, eps, stream);
    if (D <= 512) return launch_fwd<2>(x, gamma, beta, y, mean, rstd, rows, D, eps, stream);
    if (D <= 768) return launch_fwd<3>(x, gamma, beta, y, mean, rstd, rows, D, eps, stream);
    return launch_fwd<4>(x, gamma, beta, y, mean, rstd, rows, D, eps, stream);
}

extern "C" int avn_layernorm_bwd(const void* dy, const void* x, const float* gamma, const float* mean, const float* rstd,
                                 const void* dres, void* dx, float* dgamma, float* dbeta, float* dcolsum, int64_t rows, int D,
                                 avn_stream_t stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    AVN_REQUIRE(dy && x && gamma && mean && rstd && dx, "layernorm_bwd: null pointer");
    AVN_REQUIRE(rows > 0 && D > 0 && D % 8 == 0 && D <= 1024, "layernorm_bwd: D=%d must be a multiple of 8 and <= 1024", D);
    AVN_REQUIRE(((uintptr_t)x % 16) == 0 && ((uintptr_t)dy % 16) == 0 && ((uintptr_t)dx % 16) == 0 && ((uintptr_t)dres % 16) == 0,
                "layernorm_bwd: tensors must be 16-byte aligned");
    if (D <= 256) return launch_bwd<1>(dy, x, gamma, mean, rstd, dres, dx, dgamma, dbeta, dcolsum, rows, D, stream);
    if (D <= 512) return launch_bwd<2>(dy, x, gamma, mean, rstd, dres, dx, dgamma, dbeta, dcolsum, rows, D, stream);
    if (D <= 768) return launch_bwd<3>(dy, x, gamma, mean, rstd, dres, dx, dgamma, dbeta, dcolsum, rows, D, stream);
    return launch_bwd<4>(dy, x, gamma, mean, rstd, dres, dx, dgamma, dbeta, dcolsum, rows, D, stream);
}
