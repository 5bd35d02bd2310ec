// T1+T2: fused uint8 -> bf16 crop / normalise / patchify (HBM-bound: 1 B read + 2 B written per element).
// Replaces frames.astype(float32) (avion/data/clip_dataset.py:104,164), permute + kornia Normalize
// (scripts/main_lavila_pretrain.py:816-817) and the einops patchify (avion/models/transformer.py:319).
//
// One CTA per (clip, frame, patch-row): it stages the p image rows of the crop (p x 3*crop bytes) in shared memory
// with 16-byte global loads, then writes the gw consecutive output tokens (one contiguous block of gw token rows, each
// 3*p*p features rounded up to a multiple of 8 with zero padding) with 16-byte stores.  Normalisation goes through a 3x256 bf16 table built with the same fp32 (x-mean)/std division as
// the reference, so every output equals bf16(reference fp32 value) bit for bit.
#include "common.cuh"
#include "../../include/avion_b200.h"

namespace {

constexpr int MAX_ROW_BYTES = 3 * 512 + 32;   // crop width up to 512 px

__global__ void __launch_bounds__(256)
clip_transform_kernel(const uint8_t* __restrict__ frames, __nv_bfloat16* __restrict__ patches, int T, int Hs, int Ws, int crop,
                      int p, const int32_t* __restrict__ boxes, float m0, float m1, float m2, float s0, float s1, float s2,
                      long long total_bytes) {
    extern __shared__ uint8_t smem[];
    __nv_bfloat16* lut = (__nv_bfloat16*)smem;               // [3][256]
    int* mis = (int*)(smem + 3 * 256 * 2);                   // [p] misalignment of each staged row
    uint8_t* rows = smem + 3 * 256 * 2 + 64 * 4;             // [p][pitch]
    const int gw = crop / p, gh = crop / p;
    const int hh = blockIdx.x % gh;
    const int t = (blockIdx.x / gh) % T;
    const int b = blockIdx.x / (gh * T);
    const int row_bytes = crop * 3;
    const int pitch = ((row_bytes + 15) / 16) * 16 + 32;
    int x0 = 0, y0 = 0, flip = 0;
    if (boxes != nullptr) { x0 = boxes[b * 3 + 0]; y0 = boxes[b * 3 + 1]; flip = boxes[b * 3 + 2]; }

    for (int i = threadIdx.x; i < 768; i += blockDim.x) {
        const int c = i >> 8;
        const float mean = c == 0 ? m0 : (c == 1 ? m1 : m2), sd = c == 0 ? s0 : (c == 1 ? s1 : s2);
        lut[i] = __float2bfloat16_rn(__fdiv_rn((float)(i & 255) - mean, sd));
    }
    // stage p rows
    const int chunks_per_row = pitch / 16;
    for (int i = threadIdx.x; i < p * chunks_per_row; i += blockDim.x) {
        const int sh = i / chunks_per_row, ch = i % chunks_per_row;
        const long long start = (((long long)(b * T + t) * Hs + (y0 + hh * p + sh)) * Ws + x0) * 3;
        const long long abs_addr = (long long)(uintptr_t)frames + start;
        const int m = (int)(abs_addr & 15);
        if (ch == 0) mis[sh] = m;
        const long long off = start - m + (long long)ch * 16;   // byte offset of this aligned 16B chunk
        if (off >= start + row_bytes) continue;
        uint4 v;
        if (off >= 0 && off + 16 <= total_bytes) {
            v = __ldg((const uint4*)(frames + off));
        } else {
            uint8_t tmp[16];
#pragma unroll
            for (int k = 0; k < 16; ++k) tmp[k] = (off + k >= 0 && off + k < total_bytes) ? frames[off + k] : 0;
            v = *(uint4*)tmp;
        }
        *(uint4*)(rows + sh * pitch + ch * 16) = v;
    }
    __syncthreads();
    // write gw tokens x (3*p*p) features, 8 features (16 B) per thread-iteration.  A token row is 3*p*p features rounded
    // up to a multiple of 8 (zero padding: ViT-L/14 has 588 -> 592), so every row starts on a 16-byte boundary.
    const int feat = 3 * p * p;
    const int feat_pad = (feat + 7) & ~7;
    const int vec_per_tok = feat_pad / 8;
    __nv_bfloat16* out = patches + ((long long)(b * T + t) * gh * gw + (long long)hh * gw) * feat_pad;
    if ((p & 7) == 0) {
        const int p8 = p / 8;                                    // 8-pixel groups per patch row
        for (int o = threadIdx.x; o < gw * vec_per_tok; o += blockDim.x) {
            const int ww = o / vec_per_tok, f8 = o % vec_per_tok;
            const int c = f8 / (p * p8), sh = (f8 / p8) % p, sw0 = (f8 % p8) * 8;
            const uint8_t* r = rows + sh * pitch + mis[sh];
            const __nv_bfloat16* l = lut + c * 256;
            __nv_bfloat16 v[8];
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                int px = ww * p + sw0 + i;
                if (flip) px = crop - 1 - px;
                v[i] = l[r[px * 3 + c]];
            }
            *(uint4*)(out + (long long)o * 8) = *(uint4*)v;
        }
    } else {
        // patch sizes that are not a multiple of 8 (14): a 16-byte output vector straddles patch rows, so each of its 8
        // features is located on its own (feature f = c*p*p + sh*p + sw)
        const int pp = p * p;
        for (int o = threadIdx.x; o < gw * vec_per_tok; o += blockDim.x) {
            const int ww = o / vec_per_tok, f0 = (o % vec_per_tok) * 8;
            __nv_bfloat16 v[8];
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                const int f = f0 + i;
                if (f < feat) {
                    const int c = f / pp, sh = (f % pp) / p, sw = f % p;
                    int px = ww * p + sw;
                    if (flip) px = crop - 1 - px;
                    v[i] = lut[c * 256 + rows[sh * pitch + mis[sh] + px * 3 + c]];
                } else {
                    v[i] = __float2bfloat16_rn(0.f);
                }
            }
            *(uint4*)(out + (long long)o * 8) = *(uint4*)v;
        }
    }
}
}  // namespace

extern "C" int avn_clip_transform_u8_bf16(const uint8_t* frames, void* patches, int B, int T, int Hs, int Ws, int crop, int patch,
                                          const int32_t* boxes, const float* mean3, const float* std3, avn_stream_t stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    AVN_REQUIRE(frames && patches && mean3 && std3, "clip_transform: null pointer");
    AVN_REQUIRE(B > 0 && T > 0 && crop > 0 && patch > 0, "clip_transform: bad sizes");
    AVN_REQUIRE(crop % patch == 0 && patch <= 64, "clip_transform: patch must divide crop (<= 64)");
    AVN_REQUIRE(crop <= Hs && crop <= Ws && crop <= 512, "clip_transform: crop %d exceeds source %dx%d or 512", crop, Hs, Ws);
    AVN_REQUIRE(((uintptr_t)patches % 16) == 0, "clip_transform: output must be 16-byte aligned");
    const int pitch = ((crop * 3 + 15) / 16) * 16 + 32;
    const size_t smem = 3 * 256 * 2 + 64 * 4 + (size_t)patch * pitch;
    const long long total = (long long)B * T * Hs * Ws * 3;
    auto kern = clip_transform_kernel;
    AVN_CHECK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    const int grid = B * T * (crop / patch);
    kern<<<grid, 256, smem, stream>>>(frames, (__nv_bfloat16*)patches, T, Hs, Ws, crop, patch, boxes, mean3[0], mean3[1], mean3[2],
                                      std3[0], std3[1], std3[2], total);
    AVN_CHECK_CUDA(cudaGetLastError());
    return AVN_OK;
}
