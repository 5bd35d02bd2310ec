/* avion_b200 C-ABI: hand-written sm_100a kernels for AVION's video-CLIP train-step hot path.
 *
 * Every entry point takes raw device pointers, sizes and a cudaStream_t (passed as void*), allocates nothing,
 * keeps no global state, launches asynchronously on the given stream and returns 0 on success; on failure it
 * returns non-zero and avn_last_error() describes why (the Python host raises).  No torch types appear here.
 *
 * Each function names the reference code it replaces (paths relative to the AVION tree).
 */
#ifndef AVION_B200_H
#define AVION_B200_H
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef void* avn_stream_t; /* cudaStream_t */

const char* avn_last_error(void);
int avn_version(void);
/* 0 when the current device is compute capability 10.x (B200); the host refuses to run otherwise. */
int avn_device_check(void);

/* ---- T0: random-resized-crop box sampler (host) ---------------------------------------------------------------
 * Replaces FFMPEGThreadedDecoder::SetCodecContext's use_rrc branch
 * (third_party/decord/src/video/ffmpeg/threaded_decoder.cc:49-93) with an explicit seed (the reference seeds
 * mt19937_64 from random_device; here clip i uses mt19937_64(seed + i)).  out: host int32 [n][5] = (x, y, w, h, hflip). */
int avn_sample_rrc_boxes(uint64_t seed, int n, int src_w, int src_h, double scale_min, double scale_max,
                         double ratio_min, double ratio_max, double hflip_prob, int32_t* out);

/* ---- T1+T2: fused uint8 -> bf16 crop / normalise / patchify ---------------------------------------------------
 * Replaces frames.astype(float32) (avion/data/clip_dataset.py:104,164), permute + kornia Normalize
 * (scripts/main_lavila_pretrain.py:816-817) and the einops patchify (avion/models/transformer.py:319) in one pass.
 * frames: u8 [B][T][Hs][Ws][3]; boxes: device int32 [B][3] = (x, y, hflip) or NULL (x=y=0, no flip); x, y are clamped
 * into [0, Ws-crop] x [0, Hs-crop] (ffmpeg's crop filter does the same with decord's off-by-one upper bound);
 * patches: bf16 [B][T*gh*gw][F], F = 3*patch*patch rounded up to a multiple of 8 (zero padding; 588 -> 592 for ViT-L/14),
 * token = t*gh*gw + hh*gw + ww, feature = c*p*p + sh*p + sw. */
int avn_clip_transform_u8_bf16(const uint8_t* frames, void* patches, int B, int T, int Hs, int Ws, int crop, int patch,
                               const int32_t* boxes, const float* mean3 /* host */, const float* std3 /* host */,
                               avn_stream_t stream);

/* Same output contract, but each clip's crop box (x, y, w, h) is RESIZED to crop x crop: the "scale=W:H" stage of decord's
 * fused filter graph "crop=w:h:x:y,[hflip,]scale=W:H" (threaded_decoder.cc:92) and the non-fused path's
 * torchvision RandomResizedCrop (scripts/main_lavila_pretrain.py:437-441).  boxes: device int32 [B][5] = (x, y, w, h, hflip),
 * the layout avn_sample_rrc_boxes writes; offsets are clamped into the frame like ffmpeg's crop filter does.  Filter:
 * separable antialiased bilinear (torch _upsample_bilinear2d_aa, support max(scale, 1)) on the uint8 values, rounded to the
 * uint8 grid (what the reference's decoder hands over), then (x - mean) / std.  swscale's own kernel has no source in the
 * AVION tree: pixel parity with decord after a resize is unpinned; parity is against torchvision's resized_crop. */
int avn_clip_resize_transform_u8_bf16(const uint8_t* frames, void* patches, int B, int T, int Hs, int Ws, int crop, int patch,
                                      const int32_t* boxes, const float* mean3 /* host */, const float* std3 /* host */,
                                      avn_stream_t stream);

/* ---- V3: LayerNorm (nn.LayerNorm / LayerNorm subclass, avion/models/transformer.py:35-41,213) ------------------
 * x,y bf16 [rows][D]; gamma,beta fp32 [D]; mean,rstd fp32 [rows] (saved for backward; may be NULL in fwd). */
int avn_layernorm_fwd(const void* x, const float* gamma, const float* beta, void* y, float* mean, float* rstd,
                      int64_t rows, int D, float eps, avn_stream_t stream);
/* dx = LN'(dy) (+ dres if dres != NULL: the residual-stream gradient that bypasses the LN); dgamma/dbeta fp32 [D]
 * are ACCUMULATED (+=) with atomics, so zero them first.  dcolsum (optional, fp32 [D]) accumulates the column sum
 * of the produced dx (the bias gradient of the Linear that wrote this residual stream). */
int avn_layernorm_bwd(const void* dy, const void* x, const float* gamma, const float* mean, const float* rstd,
                      const void* dres, void* dx, float* dgamma, float* dbeta, float* dcolsum, int64_t rows, int D,
                      avn_stream_t stream);

/* ---- V1,V4,V5,V6: bf16 GEMM on tcgen05 (TMA -> smem ring -> tcgen05.mma -> TMEM -> epilogue -> TMA store) -------
 * Replaces nn.Linear / FlashMHA.Wqkv / out_proj / FlashMlp fc1,fc2 forward, dgrad and wgrad
 * (avion/models/transformer.py:116-120,135,227,320) incl. their bias / GELU(erf) / residual-add epilogues (:148-149). */
enum { AVN_GEMM_NT = 0, AVN_GEMM_NN = 1, AVN_GEMM_TN = 2 };
enum {
    AVN_EPI_NONE = 0,          /* C = acc                                   (bf16) */
    AVN_EPI_BIAS = 1,          /* C = acc + bias[n]                         (bf16) */
    AVN_EPI_BIAS_GELU = 2,     /* C = gelu(acc + bias[n])                   (bf16) */
    AVN_EPI_BIAS_GELU_DUAL = 3,/* x = acc + bias[n]; C = gelu(x), C2 = x (bf16, bf16): C2 (the pre-activation autograd would
                                  have saved, transformer.py:135) feeds AVN_EPI_DGELU */
    AVN_EPI_BIAS_RESIDUAL = 4, /* C = acc + bias[n] + R[m][n]               (bf16); bias may be NULL */
    AVN_EPI_DGELU = 5,         /* C = acc * gelu'(R[m][n]), R = pre-activation from AVN_EPI_BIAS_GELU_DUAL (bf16); exact-erf GELU */
    AVN_EPI_F32_ADD = 6        /* C += acc                                  (fp32, TMA reduce-add; split-K safe) */
};
typedef struct {
    int mode;  /* NT: C[b][m][n] = sum_k A[b][m][k] * B[n][k]      A:[batch][M][K]  B:[N][K]
                  NN: C[b][m][n] = sum_k A[b][m][k] * B[k][n]      A:[batch][M][K]  B:[K][N]
                  TN: C[m][n]   += sum_b sum_k A[b][k][m] * B[b][k][n]   A:[batch][K][M]  B:[batch][K][N]  (C fp32) */
    int epi;
    int M, N, K, batch;
    const void* A; int64_t lda, a_batch_stride;   /* element strides */
    const void* B; int64_t ldb, b_batch_stride;
    void* C;       int64_t ldc, c_batch_stride;
    void* C2;                                      /* same geometry as C */
    const float* bias;                             /* fp32 [N] */
    const void* R; int64_t ldr, r_batch_stride;    /* bf16, indexed like C */
    int splits;                                    /* TN only: split of the contraction across CTAs; 0 = chosen by the library */
    float* colsum;                                 /* TN only, may be NULL: fp32 [M], colsum[m] += sum_b sum_k A[b][k][m] - the bias
                                                      gradient of the Linear whose weight gradient this call computes (the
                                                      `.sum(0)` autograd runs on dY), taken from the A tiles the kernel stages anyway */
} avn_gemm_desc;
int avn_gemm_bf16(const avn_gemm_desc* d, avn_stream_t stream);
/* sizeof(avn_gemm_desc) as this library was built: a binding checks its mirror of the struct against it when it loads */
int avn_sizeof_gemm_desc(void);

/* ---- V4: flash-style attention on tcgen05 (replaces flash_attn_qkvpacked_func reached from transformer.py:148, and
 * nn.MultiheadAttention with the causal mask for the text tower, transformer.py:140-146,447-453); head dim 64.
 * qkv bf16 [B][S][3][H][64]; out bf16 [B][S][H*64]; lse fp32 [B][H][S] (natural-log LSE of scale*QK^T). */
int avn_attn_fwd(const void* qkv, void* out, float* lse, int B, int S, int H, float scale, int causal,
                 avn_stream_t stream);
/* dqkv bf16 [B][S][3][H][64]; delta fp32 [B][H][S] workspace; dq_acc fp32 [B][S][H][64] workspace (zeroed inside). */
int avn_attn_bwd(const void* qkv, const void* out, const void* dout, const float* lse, void* dqkv, float* delta,
                 float* dq_acc, int B, int S, int H, float scale, int causal, avn_stream_t stream);

/* ---- elementwise / small kernels -------------------------------------------------------------------------------- */
/* fp32 -> bf16 cast (weights, once per step). */
int avn_cast_f32_bf16(const float* src, void* dst, int64_t n, avn_stream_t stream);
/* fp32 [rows][cols] -> bf16 [cols][rows] (pre-transposed weight copy). */
int avn_transpose_cast_f32_bf16(const float* src, void* dst, int rows, int cols, avn_stream_t stream);
/* colsum[n] += sum_m X[m][n], X bf16 [rows][cols] with row stride ld (bias gradients). */
int avn_colsum_bf16(const void* x, int64_t ld, float* colsum, int64_t rows, int cols, avn_stream_t stream);
/* V2 (transformer.py:321-329): table bf16 [T*G+1][D]; row 0 = cls + pos[0], row 1+p = pos[1 + p%G] + temporal[p/G].
 * The patch-embed GEMM adds rows 1.. through its residual epilogue; avn_broadcast_row writes row 0 of every clip. */
int avn_pos_table_fwd(const float* cls, const float* pos, const float* temporal, void* table, int T, int G, int D,
                      avn_stream_t stream);
/* x[b*batch_stride + d] = row[d] for b < B (bf16). */
int avn_broadcast_row(void* x, const void* row, int B, int64_t batch_stride, int D, avn_stream_t stream);
/* dx bf16 [B][T*G+1][D] -> dcls[D], dpos[G+1][D], dtemporal[T][D] (all +=); dtable_ws fp32 [T*G+1][D] workspace. */
int avn_pos_table_bwd(const void* dx, float* dtable_ws, float* dcls, float* dpos, float* dtemporal, int B, int T, int G,
                      int D, avn_stream_t stream);
/* X1: x[b][l][:] = tok_emb[ids[b][l]] + pos[l]  (transformer.py:459-461); ids int64 [B][L]. */
int avn_token_embed_fwd(const int64_t* ids, const float* tok_emb, const float* pos, void* x, int B, int L, int D,
                        avn_stream_t stream);
int avn_token_embed_bwd(const int64_t* ids, const void* dx, float* dtok_emb, float* dpos, int B, int L, int D,
                        avn_stream_t stream);

/* ---- V7 + C1: head = gather row, LayerNorm, @ proj[D][E], L2-normalise (transformer.py:368-376,472-476;
 * model_clip.py:95-98).  x bf16 [rows_total][D]; row_idx int64 [n] (row of x per sample); emb fp32 [n][E];
 * saved: xhat-stats mean,rstd [n], pre-normalise vector u fp32 [n][E]. */
int avn_head_fwd(const void* x, const int64_t* row_idx, const float* gamma, const float* beta, const float* proj,
                 float* emb, float* u, float* mean, float* rstd, int n, int D, int E, float eps, avn_stream_t stream);
/* demb fp32 [n][E] -> dx rows (bf16, written at row_idx with row stride D; other rows are left untouched),
 * dgamma,dbeta [D] (+=), dproj [D][E] (+=). */
int avn_head_bwd(const float* demb, const void* x, const int64_t* row_idx, const float* gamma, const float* beta,
                 const float* proj,
                 const float* u, const float* mean, const float* rstd, void* dx, float* dgamma, float* dbeta,
                 float* dproj, float* ws /* fp32 [n][D+E] workspace */, int n, int D, int E,
                 int normalized /* demb is the gradient of the normalised (1) or pre-normalise (0) output */,
                 avn_stream_t stream);

/* ---- L2: sharded InfoNCE (ClipLoss, avion/losses/losses.py:103-149, local rows x gathered columns) -------------
 * img,txt fp32 [n][E] local rows; all_img, all_txt fp32 [N][E] gathered; labels = row_offset + r.
 * scale: DEVICE pointer to one fp32 (logit_scale.exp(), model_clip.py:98) - read by the kernel, never by the host.
 * out[0] += sum over local rows of (CE_i + CE_t)/2, out[1] += number of correct image->text argmaxes.
 * lse fp32 [2][n] saved for backward. */
int avn_infonce_fwd(const float* img, const float* txt, const float* all_img, const float* all_txt, const float* scale,
                    int n, int N, int E, int row_offset, float* lse, float* out, avn_stream_t stream);
/* gscale = dLoss/d(mean local loss).  Produces d_all [2][N][E] fp32 (+=; image block first, then text; contains,
 * at rows row_offset.., the local row-side gradients and - when col_side != 0 - the column-side gradients of every
 * gathered row) and dscale[1] (+=).  col_side = 0 is the reference's local_loss=True, gather_with_grad=False mode, whose
 * gathered copies carry no gradient (losses.py:61-72).  dlogits workspace fp32 [2][n][N]. */
int avn_infonce_bwd(const float* img, const float* txt, const float* all_img, const float* all_txt, const float* scale,
                    int n, int N, int E, int row_offset, const float* lse, float gscale, int col_side, float* dlogits,
                    float* d_all, float* dscale, avn_stream_t stream);
/* No counterpart in the reference: added for callers that cannot allocate device
 * memory (code under CUDA-graph capture, tight serving loops, harnesses that forbid cudaMalloc).  Same computation and
 * arguments as avn_infonce_bwd minus the dlogits workspace, which lives in a file-scope `static __device__` array of
 * avn_infonce_static_max_floats() = 2^20 floats (4 MiB per GPU that loads the library, held until exit).  Requires
 * 2*n*N <= that maximum (e.g. n = 256 local rows x N = 2048 gathered rows), otherwise returns AVN_ERR_ARG with a message.
 * Allocates nothing (no cudaMalloc / cudaFree / library allocation); the first call on each device resolves the array's
 * address with cudaGetSymbolAddress.  The scratch is fully rewritten by every call (nothing to clear), but it is SHARED: two
 * calls that overlap on one GPU (different streams, or concurrent graph branches) race - serialise them. */
int avn_infonce_static_max_floats(void);
int avn_infonce_bwd_static(const float* img, const float* txt, const float* all_img, const float* all_txt, const float* scale,
                           int n, int N, int E, int row_offset, const float* lse, float gscale, int col_side, float* d_all,
                           float* dscale, avn_stream_t stream);

/* ---- O1: optimizer step (scripts/main_lavila_pretrain.py:296-341 AdamW with weight-decay / no-decay groups; :880 clamp) --
 * One launch updates every tensor: device tables of n tensors (params, grads, exp_avg, exp_avg_sq: fp32 pointers; sizes in
 * elements; weight_decay per tensor) and a device chunk table (tensor index, element offset; avn_adamw_chunk_elems()
 * elements per chunk).  Decoupled AdamW in the operation order of torch.optim.AdamW; `step` counts from 1. */
int avn_adamw_chunk_elems(void);
int avn_adamw_step(void* const* params, const void* const* grads, void* const* exp_avg, void* const* exp_avg_sq,
                   const int64_t* sizes, const float* weight_decay, const int32_t* chunk_tensor, const int64_t* chunk_offset,
                   int n_chunks, float lr, float beta1, float beta2, float eps, int step, avn_stream_t stream);
/* logit_scale.data.clamp_(0, 4.6052) (main_lavila_pretrain.py:880) without a host round trip */
int avn_clamp_scalar_f32(float* x, float lo, float hi, avn_stream_t stream);

/* ---- VideoMAE pretraining step (BASELINE.json configs[4]) ---------------------------------------
 * The encoder / decoder blocks use avn_layernorm_*, avn_gemm_bf16 and avn_attn_* above; these are the pieces around them.
 *
 * avn_tube_mask_slots: mask uint8/bool [B,N] (1 = masked; TubeMaskingGeneratorGPU, avion/data/transforms.py:279-302) ->
 * slot int32 [B,N] = position of token n in "visible ascending, then masked ascending" (the order of x[~mask] followed by
 * expand_pos_embed[mask], avion/models/model_videomae.py:383,575-577), order int32 [B,N] = its inverse, counts int32 [B]
 * = visible tokens per clip (may be NULL). */
int avn_tube_mask_slots(const uint8_t* mask, int32_t* slot, int32_t* order, int32_t* counts, int B, int N,
                        avn_stream_t stream);
/* frames uint8 [B,T,crop,crop,3] (decord layout, fused decode-crop output) ->
 *   patches bf16 [B,n_visible,3*2*p*p]: (x-mean)/std, feature order (c p0 p1 p2) of PatchEmbed (model_videomae.py:161) /
 *     Conv3d weight flattening (:153-155), rows in slot order, VISIBLE tokens only;
 *   targets fp32 [B,N-n_visible,3*2*p*p]: raw pixels standardised per patch and channel with the unbiased variance and
 *     eps 1e-6 on the std, feature order (p0 p1 p2 c) (scripts/main_videomae_pretrain.py:322-334), MASKED tokens only.
 * Either output may be NULL.  1 B read per input element; 2 B written per visible, 4 B per masked element. */
int avn_tubelet_transform_u8(const uint8_t* frames, const int32_t* slot, void* patches, float* targets, int B, int T,
                             int crop, int patch, int n_visible, const float* mean3, const float* std3, avn_stream_t stream);
/* out bf16 [B,n,D] = table[idx[b*idx_stride + j]] (+ add_row on rows j >= add_from when add_row != NULL): sinusoid position
 * rows in token order, mask_token added on the masked rows (model_videomae.py:379,574-577) */
int avn_gather_rows_f32_bf16(const float* table, const int32_t* idx, int64_t idx_stride, const float* add_row, int add_from,
                             void* out, int B, int n, int D, avn_stream_t stream);
/* out[c] (fp32) += sum over b < B, r < rows of x[b*batch_stride + r*cols + c] (bf16): d(mask_token) */
int avn_colsum_batched_bf16(const void* x, int B, int64_t batch_stride, int rows, int cols, float* out, avn_stream_t stream);
/* torch.nn.MSELoss() (scripts/main_videomae_pretrain.py:129,341): loss[0] = mean((pred - target)^2), pred bf16, target fp32,
 * deterministic two-stage sum (workspace: avn_mse_workspace_floats() floats); bwd: dpred = 2/n * gloss[0] * (pred - target). */
int avn_mse_workspace_floats(void);
int avn_mse_fwd(const void* pred, const float* target, int64_t n, float* workspace, float* loss, avn_stream_t stream);
int avn_mse_bwd(const void* pred, const float* target, const float* gloss, int64_t n, void* dpred, avn_stream_t stream);

/* ---- MaxMarginRankingLoss (avion/losses/losses.py:152-227; EK-100 MIR fine-tuning) ---------------------
 * text, image fp32 [n,E] = the all-gathered features; x = sim_matrix(text, image) (:152-160, eps 1e-8);
 * loss[0] = mean over i != j (fix_norm != 0; else over all i, j) of relu(m - x_ii + x_ij) and relu(m - x_ii + x_ji) (:197-225).
 * fwd fills `workspace` (avn_mmr_workspace_floats(n, E) floats: normalised features, norms, diagonal, row partials), which
 * bwd reads: dtext, dimage fp32 [n,E] = gloss[0] * dL/d(text, image).  Deterministic; neither x nor dL/dx is materialised.
 * (Rows with |feature| < eps take the gradient of the unclamped normalisation.) */
int avn_mmr_workspace_floats(int n, int E);
int avn_mmr_fwd(const float* text, const float* image, int n, int E, float margin, int fix_norm, float* workspace, float* loss,
                avn_stream_t stream);
int avn_mmr_bwd(const float* workspace, const float* gloss, int n, int E, float margin, int fix_norm, float* dtext,
                float* dimage, avn_stream_t stream);

#ifdef __cplusplus
}
#endif
#endif
