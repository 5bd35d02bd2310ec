// T1+T2: fused uint8 -> bf16 crop / (resize) / normalise / patchify.  HBM-bound: 1 B read + 2 B written per output element.
// Replaces frames.astype(float32) (avion/data/clip_dataset.py:104,164), permute + kornia Normalize
// (scripts/main_lavila_pretrain.py:816-817) and the einops patchify (avion/models/transformer.py:319); the resize variant
// also covers the "scale=W:H" stage of decord's fused filter graph "crop=w:h:x:y,[hflip,]scale=W:H"
// (third_party/decord/src/video/ffmpeg/threaded_decoder.cc:92) and the non-fused path's
// torchvision RandomResizedCrop (scripts/main_lavila_pretrain.py:437-441).
//
// One CTA per (clip, frame, patch row).  Data movement is bulk-asynchronous on both sides (TMA, 1-D):
//   global --cp.async.bulk--> smem image rows --(unpack, normalise, bf16)--> smem token tile --cp.async.bulk--> global
// The output of a patch row is gw CONSECUTIVE token rows, so each token row leaves as one bulk store of 3*p*p*2 bytes.
//   clip_transform_kernel        : crop at the output size (+ flip).  A thread unpacks 2 pixels (6 bytes, 32-bit shared loads +
//                                  funnel shift), (x - mean) / std in fp32 with the reference's operation order (bit-exact
//                                  bf16(fp32 reference)), three 4-byte stores into the token tile.
//   clip_resize_transform_kernel : crop box (x, y, w, h) of any size resized to the output size with the separable antialiased
//                                  triangle filter of torch's _upsample_bilinear2d_aa (support = max(scale, 1); the float path
//                                  torchvision.transforms.v2.functional.resized_crop(..., antialias=True) takes), rounded to
//                                  the uint8 grid like the frames the reference receives, then normalised.  Image rows of the
//                                  crop stream through a double-buffered bulk-load ring; each thread owns a few (column,
//                                  channel) pairs, filters them horizontally and adds the row into the <= 3 output rows it
//                                  touches (fp32 accumulators private to the thread, in shared memory).
// Box coordinates are clamped into the source like ffmpeg's crop filter does (decord's sampler draws x in [0, W-w+1], one too
// far: threaded_decoder.cc:66-67), so a box can never make the kernel read a neighbouring row or frame.
#include "common.cuh"
#include "../../include/avion_b200.h"

namespace {
using namespace avn;

__device__ __forceinline__ void bulk_load(void* smem_dst, const void* gsrc, uint32_t bytes, uint64_t* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(smem_u32(smem_dst)), "l"(gsrc), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void bulk_store(void* gdst, const void* smem_src, uint32_t bytes) {
    asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(gdst), "r"(smem_u32(smem_src)), "r"(bytes) : "memory");
}

// byte k (0-3) of w as an exact float: 0x4B000000 | byte is 2^23 + byte
__device__ __forceinline__ float byte_to_float(uint32_t w, int k) {
    return __uint_as_float(__byte_perm(w, 0x4B000000u, 0x7440u | (uint32_t)k)) - 8388608.0f;
}

struct Norm { float m[3], s[3], inv[3]; int exact_mul; };

__device__ __forceinline__ float normalise(float x, const Norm& nm, int c) {
    const float t = x - nm.m[c];                     // the reference's order: (x - mean) / std, fp32
    return nm.exact_mul ? t * nm.inv[c] : __fdiv_rn(t, nm.s[c]);
}

// Aligned 16-byte superset of `bytes` bytes starting at byte offset `start` of the frames buffer: [start - mis, ...)
struct RowSpan { long long off; int mis; uint32_t len; bool bulk_ok; };
__device__ __forceinline__ RowSpan row_span(const uint8_t* frames, long long start, int bytes, long long total_bytes) {
    RowSpan r;
    r.mis = (int)(((long long)(uintptr_t)frames + start) & 15);
    r.off = start - r.mis;
    r.len = (uint32_t)((r.mis + bytes + 15) & ~15);
    r.bulk_ok = r.off >= 0 && r.off + (long long)r.len <= total_bytes;      // else: guarded scalar copy (first / last bytes of the buffer)
    return r;
}
__device__ __forceinline__ void guarded_row_copy(uint8_t* dst, const uint8_t* frames, const RowSpan& r, long long total_bytes) {
    for (uint32_t i = threadIdx.x; i < r.len; i += blockDim.x) {
        const long long o = r.off + i;
        dst[i] = (o >= 0 && o < total_bytes) ? frames[o] : (uint8_t)0;
    }
}

// token tile -> global: one bulk store per token row (issued by the first gw lanes of warp 0), then wait until shared memory
// has been read (the CTA must not exit before that)
__device__ __forceinline__ void store_token_tile(__nv_bfloat16* out, const uint8_t* tile, int gw, int pitch_out, int feat_pad) {
    fence_proxy_async_smem();
    __syncthreads();
    if (threadIdx.x < 32) {
        for (int ww = threadIdx.x; ww < gw; ww += 32) bulk_store(out + (long long)ww * feat_pad, tile + ww * pitch_out, feat_pad * 2);
        tma_commit_group();
        tma_wait_group_read0();
    }
}

// ---------------------------------------------------------------------------------------------------------------------
// PC = patch size known at compile time (16, 14: divisions by it become multiplies) or 0 (any even patch, runtime divisions)
template <int PC>
__global__ void __launch_bounds__(256)
clip_transform_kernel(const uint8_t* __restrict__ frames, __nv_bfloat16* __restrict__ patches, int T, int Hs, int Ws, int crop,
                      int p_rt, const int32_t* __restrict__ boxes, Norm nm, int pitch_in, int pitch_out, long long total_bytes) {
    const int p = PC ? PC : p_rt;
    extern __shared__ __align__(128) uint8_t smem[];
    uint64_t* bar = (uint64_t*)smem;
    uint8_t* rows = smem + 128;                                  // [p][pitch_in]
    uint8_t* tile = rows + p * pitch_in;                         // [gw][pitch_out]
    const int gw = crop / p, gh = crop / p;
    const int hh = blockIdx.x % gh;
    const int t = (blockIdx.x / gh) % T;
    const int b = blockIdx.x / (gh * T);
    const int row_bytes = crop * 3;
    int x0 = 0, y0 = 0, flip = 0;
    if (boxes != nullptr) {
        x0 = min(max(boxes[b * 3 + 0], 0), Ws - crop);           // ffmpeg's crop filter clamps the offset into the frame
        y0 = min(max(boxes[b * 3 + 1], 0), Hs - crop);
        flip = boxes[b * 3 + 2] != 0;
    }
    const long long row0 = (((long long)(b * T + t) * Hs + (y0 + hh * p)) * Ws + x0) * 3;
    const long long row_stride = (long long)Ws * 3;

    if (threadIdx.x == 0) {
        mbar_init(bar, 1);
        fence_barrier_init();
        uint32_t tx = 0;
        for (int sh = 0; sh < p; ++sh) {
            const RowSpan r = row_span(frames, row0 + sh * row_stride, row_bytes, total_bytes);
            if (r.bulk_ok) tx += r.len;
        }
        mbar_expect_tx(bar, tx);
    }
    __syncthreads();
    if (threadIdx.x < 32)
        for (int sh = threadIdx.x; sh < p; sh += 32) {
            const RowSpan r = row_span(frames, row0 + sh * row_stride, row_bytes, total_bytes);
            if (r.bulk_ok) bulk_load(rows + sh * pitch_in, frames + r.off, r.len, bar);
        }
    for (int sh = 0; sh < p; ++sh) {                             // at most the first / last row of the whole buffer
        const RowSpan r = row_span(frames, row0 + sh * row_stride, row_bytes, total_bytes);
        if (!r.bulk_ok) guarded_row_copy(rows + sh * pitch_in, frames, r, total_bytes);
    }
    // zero padding of the token rows (3*p*p rounded up to 8 features: 588 -> 592 for patch 14)
    const int feat = 3 * p * p, feat_pad = (feat + 7) & ~7;
    for (int i = threadIdx.x; i < gw * (feat_pad - feat); i += blockDim.x)
        *(__nv_bfloat16*)(tile + (i / (feat_pad - feat)) * pitch_out + (feat + i % (feat_pad - feat)) * 2) = __float2bfloat16_rn(0.f);
    __syncthreads();
    mbar_wait(bar, 0);

    // A thread keeps ONE pixel pair (2 output pixels, same patch: p is even) and walks down the p rows: everything that depends
    // on the column - source byte offset, token, position inside the patch - is computed once (the first version recomputed it
    // per element with runtime divisions: 1676 warp instructions per warp and 82 % issue utilisation in ncu - instruction-bound).
    const int pairs = crop >> 1, pp = p * p;
    const int mis0 = (int)(((long long)(uintptr_t)frames + row0) & 15);
    const int stride_mis = (int)(row_stride & 15);
    for (int q = threadIdx.x; q < pairs; q += blockDim.x) {
        const int px = 2 * q;                                    // output pixels px, px+1
        const int src = flip ? crop - 2 - px : px;               // source pixels src, src+1 (swapped when flipped)
        const int ww = px / p, sw = px % p;
        uint8_t* dst0 = tile + ww * pitch_out + sw * 2;
        for (int sh = 0; sh < p; ++sh) {
            const int o = ((mis0 + sh * stride_mis) & 15) + 3 * src;
            const uint32_t* wp = (const uint32_t*)(rows + sh * pitch_in + (o & ~3));
            const uint32_t w0 = wp[0], w1 = wp[1], w2 = wp[2];
            const int sft = 8 * (o & 3);
            const uint32_t lo = __funnelshift_r(w0, w1, sft), hi = __funnelshift_r(w1, w2, sft);     // bytes 0-3, 4-7 from o
            float a[3], c[3];                                    // pixel src, pixel src+1
            a[0] = byte_to_float(lo, 0); a[1] = byte_to_float(lo, 1); a[2] = byte_to_float(lo, 2);
            c[0] = byte_to_float(lo, 3); c[1] = byte_to_float(hi, 0); c[2] = byte_to_float(hi, 1);
            uint8_t* dst = dst0 + sh * p * 2;
#pragma unroll
            for (int ch = 0; ch < 3; ++ch) {
                const float v0 = normalise(flip ? c[ch] : a[ch], nm, ch), v1 = normalise(flip ? a[ch] : c[ch], nm, ch);
                *(uint32_t*)(dst + ch * pp * 2) = pack_bf16(v0, v1);
            }
        }
    }
    __nv_bfloat16* out = patches + ((long long)(b * T + t) * gh * gw + (long long)hh * gw) * feat_pad;
    store_token_tile(out, tile, gw, pitch_out, feat_pad);
}

// ---------------------------------------------------------------------------------------------------------------------
// Antialiased triangle filter (torch: aten/src/ATen/native/cpu/UpSampleKernel.cpp, _compute_indices_min_size_weights_aa with
// the bilinear filter): output index i <- taps j in [max(0, int(c - sup + .5)), min(n, int(c + sup + .5))), c = scale (i + .5),
// sup = max(scale, 1), weight max(0, 1 - |(j - c + .5) / max(scale, 1)|), normalised by their sum.
struct Axis {
    float scale, sup, inv;
    __device__ __forceinline__ void init(int in_size, int out_size) {
        scale = (float)in_size / (float)out_size;
        sup = scale >= 1.f ? scale : 1.f;
        inv = scale >= 1.f ? 1.f / scale : 1.f;
    }
    __device__ __forceinline__ float center(int i) const { return scale * ((float)i + 0.5f); }
    __device__ __forceinline__ int lo(float c) const { return max((int)(c - sup + 0.5f), 0); }
    __device__ __forceinline__ int hi(float c, int n) const { return min((int)(c + sup + 0.5f), n); }
    __device__ __forceinline__ float weight(int j, float c) const { return fmaxf(0.f, 1.f - fabsf(((float)j - c + 0.5f) * inv)); }
};

constexpr int RS_MAXK = 6;        // (column, channel) pairs per thread: crop <= 512
constexpr int RS_MAXP = 16;       // patch rows per CTA (ViT-B/16, ViT-L/14)

__global__ void __launch_bounds__(256)
clip_resize_transform_kernel(const uint8_t* __restrict__ frames, __nv_bfloat16* __restrict__ patches, int T, int Hs, int Ws, int crop,
                             int p, const int32_t* __restrict__ boxes, Norm nm, int pitch_in, int pitch_out, long long total_bytes) {
    extern __shared__ __align__(128) uint8_t smem[];
    uint64_t* bars = (uint64_t*)smem;                            // [2]
    uint8_t* rowbuf = smem + 128;                                // [2][pitch_in]
    uint8_t* tile = rowbuf + 2 * pitch_in;                       // [gw][pitch_out]
    float* acc = (float*)(tile + (crop / p) * pitch_out);        // [p][3*crop], entry (sh, k) private to thread k % 256
    const int gw = crop / p, gh = crop / p;
    const int hh = blockIdx.x % gh;
    const int t = (blockIdx.x / gh) % T;
    const int b = blockIdx.x / (gh * T);
    const int bw = min(max(boxes[b * 5 + 2], 1), Ws), bh = min(max(boxes[b * 5 + 3], 1), Hs);
    const int bx = min(max(boxes[b * 5 + 0], 0), Ws - bw), by = min(max(boxes[b * 5 + 1], 0), Hs - bh);
    const int flip = boxes[b * 5 + 4] != 0;
    Axis ax, ay;
    ax.init(bw, crop);
    ay.init(bh, crop);
    const int oy0 = hh * p;
    const int r_lo = ay.lo(ay.center(oy0)), r_hi = ay.hi(ay.center(oy0 + p - 1), bh);       // crop rows this patch row reads
    const int row_bytes = bw * 3, npairs = crop * 3;
    const long long row0 = (((long long)(b * T + t) * Hs + by) * Ws + bx) * 3;
    const long long row_stride = (long long)Ws * 3;

    // per pair: tap window and 1 / (sum of weights)
    float cx[RS_MAXK], inv_tot[RS_MAXK];
    int jlo[RS_MAXK], jhi[RS_MAXK];
#pragma unroll
    for (int k = 0; k < RS_MAXK; ++k) {
        const int pair = threadIdx.x + k * 256;
        const int ox = min(pair / 3, crop - 1);
        cx[k] = ax.center(flip ? crop - 1 - ox : ox);            // hflip sits between crop and scale: mirror the output column
        jlo[k] = ax.lo(cx[k]);
        jhi[k] = ax.hi(cx[k], bw);
        float tot = 0.f;
        for (int j = jlo[k]; j < jhi[k]; ++j) tot += ax.weight(j, cx[k]);
        inv_tot[k] = 1.f / tot;
    }
    for (int i = threadIdx.x; i < p * npairs; i += blockDim.x) acc[i] = 0.f;      // (sh, pair): pair % 256 == threadIdx.x - own entries

    if (threadIdx.x == 0) { mbar_init(&bars[0], 1); mbar_init(&bars[1], 1); fence_barrier_init(); }
    __syncthreads();
    auto issue = [&](int r) {                                    // thread 0: start the bulk load of crop row r into buffer r & 1
        const RowSpan s = row_span(frames, row0 + r * row_stride, row_bytes, total_bytes);
        if (s.bulk_ok) {
            mbar_expect_tx(&bars[(r - r_lo) & 1], s.len);
            bulk_load(rowbuf + ((r - r_lo) & 1) * pitch_in, frames + s.off, s.len, &bars[(r - r_lo) & 1]);
        }
    };
    if (threadIdx.x == 0 && r_lo < r_hi) issue(r_lo);
    float wsum[RS_MAXP];                                         // vertical weight totals (uniform across the CTA)
#pragma unroll
    for (int sh = 0; sh < RS_MAXP; ++sh) wsum[sh] = 0.f;

    for (int r = r_lo; r < r_hi; ++r) {
        const int buf = (r - r_lo) & 1;
        const RowSpan s = row_span(frames, row0 + r * row_stride, row_bytes, total_bytes);
        if (threadIdx.x == 0 && r + 1 < r_hi) issue(r + 1);      // buffer buf^1 was released by the barrier at the end of r-1
        if (s.bulk_ok) mbar_wait(&bars[buf], ((r - r_lo) >> 1) & 1);
        else { guarded_row_copy(rowbuf + buf * pitch_in, frames, s, total_bytes); __syncthreads(); }
        const uint8_t* src = rowbuf + buf * pitch_in + s.mis;
        float hv[RS_MAXK];
#pragma unroll
        for (int k = 0; k < RS_MAXK; ++k) {
            const int pair = threadIdx.x + k * 256;
            hv[k] = 0.f;
            if (pair < npairs) {
                const int c = pair % 3;
                float a = 0.f;
                for (int j = jlo[k]; j < jhi[k]; ++j) a = fmaf(ax.weight(j, cx[k]), (float)src[j * 3 + c], a);
                hv[k] = a * inv_tot[k];
            }
        }
#pragma unroll
        for (int sh = 0; sh < RS_MAXP; ++sh) {
            if (sh < p) {
                const float wy = ay.weight(r, ay.center(oy0 + sh));        // uniform across the CTA: no divergence
                if (wy > 0.f) {
                    wsum[sh] += wy;
#pragma unroll
                    for (int k = 0; k < RS_MAXK; ++k) {
                        const int pair = threadIdx.x + k * 256;
                        if (pair < npairs) acc[sh * npairs + pair] = fmaf(wy, hv[k], acc[sh * npairs + pair]);
                    }
                }
            }
        }
        __syncthreads();                                         // everyone is done with rowbuf[buf]: it may be refilled
    }
    // resized value -> uint8 grid (round half to even, clamp) -> normalise -> token tile
    const int feat = 3 * p * p, feat_pad = (feat + 7) & ~7, pp = p * p;
#pragma unroll
    for (int sh = 0; sh < RS_MAXP; ++sh) {
        if (sh < p) {
            const float inv_w = 1.f / wsum[sh];
#pragma unroll
            for (int k = 0; k < RS_MAXK; ++k) {
                const int pair = threadIdx.x + k * 256;
                if (pair < npairs) {
                    const int ox = pair / 3, c = pair % 3;
                    const float v = fminf(fmaxf(rintf(acc[sh * npairs + pair] * inv_w), 0.f), 255.f);
                    *(__nv_bfloat16*)(tile + (ox / p) * pitch_out + (c * pp + sh * p + ox % p) * 2) = __float2bfloat16_rn(normalise(v, nm, c));
                }
            }
        }
    }
    for (int i = threadIdx.x; i < gw * (feat_pad - feat); i += blockDim.x)
        *(__nv_bfloat16*)(tile + (i / (feat_pad - feat)) * pitch_out + (feat + i % (feat_pad - feat)) * 2) = __float2bfloat16_rn(0.f);
    __nv_bfloat16* out = patches + ((long long)(b * T + t) * gh * gw + (long long)hh * gw) * feat_pad;
    store_token_tile(out, tile, gw, pitch_out, feat_pad);
}

// host: (x - mean) * (1/std) is used only if it rounds to the same bf16 as the reference's (x - mean) / std for all 256 x
Norm make_norm(const float* mean3, const float* std3) {
    Norm nm;
    nm.exact_mul = 1;
    for (int c = 0; c < 3; ++c) {
        nm.m[c] = mean3[c]; nm.s[c] = std3[c]; nm.inv[c] = 1.0f / std3[c];
        for (int x = 0; x < 256; ++x) {
            const float t = (float)x - nm.m[c];
            if (__bfloat162float(__float2bfloat16_rn(t * nm.inv[c])) != __bfloat162float(__float2bfloat16_rn(t / nm.s[c]))) nm.exact_mul = 0;
        }
    }
    return nm;
}

// token-row pitch in shared memory: rows shifted by ~p/2 words against each other so that the 4-byte stores of a warp
// (consecutive pixel pairs, then the next token) spread over all banks
int tile_pitch(int feat_pad, int p) {
    int want = 4;
    while (want < p / 2 && want < 16) want *= 2;
    const int cur = (feat_pad * 2 / 4) % 32;
    return feat_pad * 2 + (((want - cur) % 32 + 32) % 32) * 4;
}
}  // namespace

extern "C" int avn_clip_transform_u8_bf16(const uint8_t* frames, void* patches, int B, int T, int Hs, int Ws, int crop, int patch,
                                          const int32_t* boxes, const float* mean3, const float* std3, avn_stream_t stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    AVN_REQUIRE(frames && patches && mean3 && std3, "clip_transform: null pointer");
    AVN_REQUIRE(B > 0 && T > 0 && crop > 0 && patch > 0, "clip_transform: bad sizes");
    AVN_REQUIRE(crop % patch == 0 && patch <= 64 && patch % 2 == 0, "clip_transform: patch must be even, divide crop and be <= 64");
    AVN_REQUIRE(crop <= Hs && crop <= Ws && crop <= 512, "clip_transform: crop %d exceeds source %dx%d or 512", crop, Hs, Ws);
    AVN_REQUIRE(((uintptr_t)patches % 16) == 0, "clip_transform: output must be 16-byte aligned");
    const int feat_pad = (3 * patch * patch + 7) & ~7;
    const int pitch_in = ((crop * 3 + 15 + 15) & ~15) + 16;       // aligned superset of a row + 3 spare words for the funnel loads
    const int pitch_out = tile_pitch(feat_pad, patch);
    const size_t smem = 128 + (size_t)patch * pitch_in + (size_t)(crop / patch) * pitch_out;
    AVN_REQUIRE(smem <= 200 * 1024, "clip_transform: tile does not fit shared memory");
    const long long total = (long long)B * T * Hs * Ws * 3;
    auto kern = patch == 16 ? clip_transform_kernel<16> : (patch == 14 ? clip_transform_kernel<14> : clip_transform_kernel<0>);
    AVN_CHECK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    const int grid = B * T * (crop / patch);
    const int threads = crop / 2 <= 128 ? 128 : 256;             // one pixel pair per thread
    kern<<<grid, threads, smem, stream>>>(frames, (__nv_bfloat16*)patches, T, Hs, Ws, crop, patch, boxes, make_norm(mean3, std3), pitch_in,
                                          pitch_out, total);
    AVN_CHECK_CUDA(cudaGetLastError());
    return AVN_OK;
}

extern "C" int avn_clip_resize_transform_u8_bf16(const uint8_t* frames, void* patches, int B, int T, int Hs, int Ws, int crop,
                                                 int patch, const int32_t* boxes, const float* mean3, const float* std3,
                                                 avn_stream_t stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    AVN_REQUIRE(frames && patches && boxes && mean3 && std3, "clip_resize_transform: null pointer");
    AVN_REQUIRE(B > 0 && T > 0 && crop > 0 && patch > 0 && Hs > 0 && Ws > 0, "clip_resize_transform: bad sizes");
    AVN_REQUIRE(crop % patch == 0 && patch <= RS_MAXP, "clip_resize_transform: patch must divide crop and be <= 16");
    AVN_REQUIRE(crop * 3 <= RS_MAXK * 256 && Ws <= 8192, "clip_resize_transform: output width <= 512, source width <= 8192");
    AVN_REQUIRE(((uintptr_t)patches % 16) == 0, "clip_resize_transform: output must be 16-byte aligned");
    const int feat_pad = (3 * patch * patch + 7) & ~7;
    const int pitch_in = ((Ws * 3 + 15 + 15) & ~15) + 16;
    const int pitch_out = tile_pitch(feat_pad, patch);
    const size_t smem = 128 + 2 * (size_t)pitch_in + (size_t)(crop / patch) * pitch_out + (size_t)patch * crop * 3 * sizeof(float);
    AVN_REQUIRE(smem <= 200 * 1024, "clip_resize_transform: tile does not fit shared memory");
    const long long total = (long long)B * T * Hs * Ws * 3;
    auto kern = clip_resize_transform_kernel;
    AVN_CHECK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    const int grid = B * T * (crop / patch);
    kern<<<grid, 256, smem, stream>>>(frames, (__nv_bfloat16*)patches, T, Hs, Ws, crop, patch, boxes, make_norm(mean3, std3), pitch_in,
                                      pitch_out, total);
    AVN_CHECK_CUDA(cudaGetLastError());
    return AVN_OK;
}
