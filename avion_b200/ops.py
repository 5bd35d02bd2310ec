"""Tensor-level wrappers of the C-ABI kernels (one function per entry point of include/avion_b200.h).

Every launch goes through the registered torch custom operator of that entry point (`torch.ops.avion_b200.<name>`,
avion_b200/torch_ops.py), whose CUDA implementation is the ctypes call.  PyTorch is used only for device memory, streams and
operator dispatch.  Every function launches on torch's current CUDA stream.
"""
import ctypes as C
import os

import torch

from . import _lib
from ._lib import (EPI_BIAS, EPI_BIAS_GELU, EPI_BIAS_GELU_DUAL, EPI_BIAS_RESIDUAL, EPI_DGELU, EPI_F32_ADD, EPI_NONE,
                   GEMM_NN, GEMM_NT, GEMM_TN, GemmDesc)
from .torch_ops import K

bf16, f32 = torch.bfloat16, torch.float32
_checked_devices = set()

# Optional per-family device timing (bench.py): TIMED maps a family name to a list of (start, end) CUDA events recorded
# on the launching stream around each call; None disables it.
TIMED = None
TIMED_BY_EPILOGUE = False      # True: one GEMM family per (mode, epilogue) = per kernel template instantiation
_EPI_NAMES = ("none", "bias", "bias_gelu", "bias_gelu_dual", "bias_residual", "dgelu", "f32_add")


NVTX = os.environ.get("AVN_NVTX", "0") == "1"      # one NVTX range per kernel family call (nsys / ncu --nvtx filters)


class _timed:
    def __init__(self, name, t, work=0.0):
        self.on = TIMED is not None
        self.work = work
        self.name = name
        if self.on:
            self.s, self.e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            self.stream = torch.cuda.current_stream(t.device)

    def __enter__(self):
        if NVTX:
            torch.cuda.nvtx.range_push("avion_b200." + self.name)
        if self.on:
            self.s.record(self.stream)

    def __exit__(self, *a):
        if self.on:
            self.e.record(self.stream)
            TIMED.setdefault(self.name, []).append((self.s, self.e, self.work))
        if NVTX:
            torch.cuda.nvtx.range_pop()


def _cuda(t):
    """No CPU fallback: refuse anything that is not a CUDA tensor before the dispatcher does."""
    if t.device.type != "cuda":
        raise RuntimeError("avion_b200 kernels run on CUDA (sm_100a) tensors only; got a %s tensor" % t.device.type)
    return t


def _stream(t):
    """Raw stream handle for the few direct ctypes callers outside this module (graph capture helpers)."""
    from .torch_ops import _stream_of
    return _stream_of(_cuda(t))


def _p(t):
    return C.c_void_p(t.data_ptr()) if t is not None else C.c_void_p(0)


def _chk(t, dtype, name):
    if t.dtype != dtype or not t.is_contiguous():
        raise RuntimeError(f"{name}: expected contiguous {dtype}, got {t.dtype} contiguous={t.is_contiguous()}")


# ---------------------------------------------------------------------------------------------------------------
def sample_rrc_boxes(seed, n, src_w, src_h, scale=(0.5, 1.0), ratio=(3.0 / 4.0, 4.0 / 3.0), hflip_prob=0.0):
    """Host sampler (decord RRC arithmetic, explicit seed) -> int32 CPU tensor [n,5] = x, y, w, h, hflip."""
    out = torch.empty(n, 5, dtype=torch.int32)
    _lib.call("avn_sample_rrc_boxes", seed, n, src_w, src_h, scale[0], scale[1], ratio[0], ratio[1], hflip_prob, _p(out))
    return out


def patch_features(patch):
    """Row length of a patch token: 3*patch^2 rounded up to 8 elements (16-byte rows for TMA; ViT-L/14: 588 -> 592)."""
    return (3 * patch * patch + 7) // 8 * 8


def clip_transform(frames_u8, crop, patch, mean, std, boxes=None, out=None):
    """uint8 [B,T,Hs,Ws,3] -> bf16 patches [B, T*(crop/patch)^2, F], F = 3*patch^2 rounded up to a multiple of 8 (zero pad)."""
    _chk(frames_u8, torch.uint8, "frames")
    B, T, Hs, Ws, ch = frames_u8.shape
    assert ch == 3
    g = crop // patch
    if out is None:
        out = torch.empty(B, T * g * g, patch_features(patch), dtype=bf16, device=frames_u8.device)
    if boxes is not None:
        _chk(boxes, torch.int32, "boxes")
        assert boxes.shape == (B, 3) and boxes.device == frames_u8.device
    with _timed("clip_transform", _cuda(frames_u8), 3.0 * frames_u8.numel()):
        K.clip_transform_u8_bf16(frames_u8, out, B, T, Hs, Ws, crop, patch, boxes, [float(v) for v in mean], [float(v) for v in std])
    return out


def clip_resize_transform(frames_u8, boxes, crop, patch, mean, std, out=None):
    """uint8 [B,T,Hs,Ws,3] + boxes int32 [B,5] = (x, y, w, h, hflip) (the rows avn_sample_rrc_boxes writes) -> bf16 patches
    [B, T*(crop/patch)^2, F] of each clip's box resized to crop x crop (antialiased bilinear, rounded to the uint8 grid)."""
    _chk(frames_u8, torch.uint8, "frames"); _chk(boxes, torch.int32, "boxes")
    B, T, Hs, Ws, ch = frames_u8.shape
    assert ch == 3 and boxes.shape == (B, 5) and boxes.device == frames_u8.device
    g = crop // patch
    if out is None:
        out = torch.empty(B, T * g * g, patch_features(patch), dtype=bf16, device=frames_u8.device)
    with _timed("clip_resize_transform", _cuda(frames_u8), float(frames_u8.numel()) + 2.0 * out.numel()):
        K.clip_resize_transform_u8_bf16(frames_u8, out, B, T, Hs, Ws, crop, patch, boxes, [float(v) for v in mean],
                                        [float(v) for v in std])
    return out


def layernorm_fwd(x, gamma, beta, eps, out=None, save_stats=True):
    _chk(x, bf16, "x"); _chk(gamma, f32, "gamma"); _chk(beta, f32, "beta")
    D = x.shape[-1]
    rows = x.numel() // D
    y = out if out is not None else torch.empty_like(x)
    mean = rstd = None
    if save_stats:
        mean = torch.empty(rows, dtype=f32, device=x.device)
        rstd = torch.empty(rows, dtype=f32, device=x.device)
    with _timed("layernorm_fwd", _cuda(x), 4.0 * rows * D):
        K.layernorm_fwd(x, gamma, beta, y, mean, rstd, rows, D, float(eps))
    return y, mean, rstd


def layernorm_bwd(dy, x, gamma, mean, rstd, dgamma, dbeta, dres=None, dcolsum=None, out=None):
    _chk(dy, bf16, "dy"); _chk(x, bf16, "x")
    D = x.shape[-1]
    rows = x.numel() // D
    dx = out if out is not None else torch.empty_like(x)
    with _timed("layernorm_bwd", x, (8.0 if dres is not None else 6.0) * rows * D):
        K.layernorm_bwd(dy, _cuda(x), gamma, mean, rstd, dres, dx, dgamma, dbeta, dcolsum, rows, D)
    return dx


def gemm(mode, epi, M, N, Kd, A, lda, B, ldb, Cm, ldc, batch=1, a_bs=0, b_bs=0, c_bs=0, C2=None, bias=None, R=None, ldr=0,
         r_bs=0, splits=1, colsum=None):
    """Raw descriptor-level GEMM (see include/avion_b200.h); A/B/Cm/C2/R are tensors whose data_ptr is the origin."""
    name = "gemm_%s" % ("nt", "nn", "tn")[mode]
    if TIMED_BY_EPILOGUE:
        name += "_" + _EPI_NAMES[epi]
    with _timed(name, _cuda(A), 2.0 * M * N * Kd * batch):
        K.gemm_bf16(mode, epi, M, N, Kd, batch, A, lda, a_bs, B, ldb, b_bs, Cm, ldc, c_bs, C2, bias, R, ldr, r_bs, splits, colsum)


def linear_fwd(x, w, bias=None, epi=EPI_NONE, residual=None, out=None, out2=None):
    """y[M,N] = epi(x[M,K] @ w[N,K]^T); x,w bf16; bias fp32."""
    _chk(x, bf16, "x"); _chk(w, bf16, "w")
    K = x.shape[-1]
    M = x.numel() // K
    N = w.shape[0]
    y = out if out is not None else torch.empty(*x.shape[:-1], N, dtype=bf16, device=x.device)
    if epi == EPI_BIAS_GELU_DUAL and out2 is None:
        out2 = torch.empty_like(y)
    gemm(GEMM_NT, epi, M, N, K, x, K, w, K, y, N, C2=out2, bias=bias, R=residual, ldr=N)
    return (y, out2) if epi == EPI_BIAS_GELU_DUAL else y


def linear_dgrad(dy, w, epi=EPI_NONE, pre=None, out=None):
    """dx[M,K] = epi(dy[M,N] @ w[N,K]); with EPI_DGELU: dx *= gelu'(pre[M,K]), where pre = the bf16 pre-activation, the second
    output of linear_fwd(..., EPI_BIAS_GELU_DUAL)."""
    _chk(dy, bf16, "dy"); _chk(w, bf16, "w")
    N, K = w.shape
    M = dy.numel() // N
    dx = out if out is not None else torch.empty(*dy.shape[:-1], K, dtype=bf16, device=dy.device)
    gemm(GEMM_NN, epi, M, K, N, dy, N, w, K, dx, K, R=pre, ldr=K)
    return dx


def linear_wgrad(dy, x, dw, dbias=None):
    """dw[N,K] (fp32) += dy[tok,N]^T @ x[tok,K]; with dbias (fp32 [N]): dbias += dy.sum over the tokens, from the dy tiles the
    GEMM stages anyway (no separate pass over dy)."""
    _chk(dy, bf16, "dy"); _chk(x, bf16, "x"); _chk(dw, f32, "dw")
    N, K = dw.shape
    tok = dy.numel() // N
    if dbias is not None:
        _chk(dbias, f32, "dbias")
        assert dbias.numel() == N
    gemm(GEMM_TN, EPI_F32_ADD, N, K, tok, dy, N, x, K, dw, K, splits=0, colsum=dbias)      # 0: the library picks the split-K count
    return dw


def attn_fwd(qkv, heads, scale, causal=False, out=None):
    """qkv bf16 [B,S,3*H*64] -> out bf16 [B,S,H*64], lse fp32 [B,H,S]."""
    _chk(qkv, bf16, "qkv")
    B, S, W3 = qkv.shape
    assert W3 == 3 * heads * 64, "head dim must be 64"
    o = out if out is not None else torch.empty(B, S, heads * 64, dtype=bf16, device=qkv.device)
    lse = torch.empty(B, heads, S, dtype=f32, device=qkv.device)
    with _timed("attn_fwd", qkv, (2.0 if causal else 4.0) * B * heads * S * S * 64):
        K.attn_fwd(_cuda(qkv), o, lse, B, S, heads, float(scale), bool(causal))
    return o, lse


def attn_bwd(qkv, o, do, lse, heads, scale, causal=False, dqkv=None, dq_acc=None):
    _chk(qkv, bf16, "qkv"); _chk(o, bf16, "o"); _chk(do, bf16, "do"); _chk(lse, f32, "lse")
    B, S, _ = qkv.shape
    if dqkv is None:
        dqkv = torch.empty_like(qkv)
    delta = torch.empty(B, heads, S, dtype=f32, device=qkv.device)
    if dq_acc is None:
        dq_acc = torch.empty(B, S, heads * 64, dtype=f32, device=qkv.device)
    with _timed("attn_bwd", qkv, (5.0 if causal else 10.0) * B * heads * S * S * 64):
        K.attn_bwd(_cuda(qkv), o, do, lse, dqkv, delta, dq_acc, B, S, heads, float(scale), bool(causal))
    return dqkv


def cast_bf16(src, out=None):
    _chk(src, f32, "src")
    dst = out if out is not None else torch.empty(src.shape, dtype=bf16, device=src.device)
    K.cast_f32_bf16(_cuda(src), dst, src.numel())
    return dst


def transpose_cast_bf16(src):
    _chk(src, f32, "src")
    r, c = src.shape
    dst = torch.empty(c, r, dtype=bf16, device=src.device)
    K.transpose_cast_f32_bf16(_cuda(src), dst, r, c)
    return dst


def colsum(x, out):
    """out[N] (fp32) += column sums of x bf16 [rows,N]."""
    _chk(x, bf16, "x"); _chk(out, f32, "out")
    N = x.shape[-1]
    with _timed("colsum", x, 2.0 * x.numel()):
        K.colsum_bf16(_cuda(x), N, out, x.numel() // N, N)
    return out


def pos_table_fwd(cls, pos, temporal, T, G):
    D = cls.numel()
    table = torch.empty(T * G + 1, D, dtype=bf16, device=cls.device)
    K.pos_table_fwd(_cuda(cls), pos, temporal, table, T, G, D)
    return table


def broadcast_row(x, row, B, batch_stride, D):
    K.broadcast_row(_cuda(x), row, B, batch_stride, D)


def pos_table_bwd(dx, dcls, dpos, dtemporal, T, G):
    B, S, D = dx.shape
    ws = torch.empty(S, D, dtype=f32, device=dx.device)
    K.pos_table_bwd(_cuda(dx), ws, dcls, dpos, dtemporal, B, T, G, D)


def token_embed_fwd(ids, tok_emb, pos):
    _chk(ids, torch.int64, "ids"); _chk(tok_emb, f32, "tok_emb"); _chk(pos, f32, "pos")
    B, L = ids.shape
    D = tok_emb.shape[1]
    x = torch.empty(B, L, D, dtype=bf16, device=ids.device)
    K.token_embed_fwd(_cuda(ids), tok_emb, pos, x, B, L, D)
    return x


def token_embed_bwd(ids, dx, dtok, dpos):
    B, L, D = dx.shape
    K.token_embed_bwd(ids, _cuda(dx), dtok, dpos, B, L, D)


def head_fwd(x2d, row_idx, gamma, beta, proj, eps):
    """x2d bf16 [rows,D]; row_idx int64 [n] -> emb fp32 [n,E] (L2-normalised) + saved (u, mean, rstd)."""
    _chk(x2d, bf16, "x"); _chk(row_idx, torch.int64, "row_idx"); _chk(proj, f32, "proj")
    n = row_idx.numel()
    D, E = proj.shape
    dev = x2d.device
    emb = torch.empty(n, E, dtype=f32, device=dev)
    u = torch.empty(n, E, dtype=f32, device=dev)
    mean = torch.empty(n, dtype=f32, device=dev)
    rstd = torch.empty(n, dtype=f32, device=dev)
    K.head_fwd(_cuda(x2d), row_idx, gamma, beta, proj, emb, u, mean, rstd, n, D, E, float(eps))
    return emb, (u, mean, rstd)


def head_bwd(demb, x2d, row_idx, gamma, beta, proj, saved, dx2d, dgamma, dbeta, dproj, normalized=True):
    u, mean, rstd = saved
    n = row_idx.numel()
    D, E = proj.shape
    ws = torch.empty(n, D + E, dtype=f32, device=x2d.device)
    _chk(demb, f32, "demb")
    K.head_bwd(demb, _cuda(x2d), row_idx, gamma, beta, proj, u, mean, rstd, dx2d, dgamma, dbeta, dproj, ws, n, D, E, bool(normalized))


def infonce_fwd(img, txt, all_img, all_txt, scale, row_offset):
    """scale: fp32 CUDA tensor with one element (logit_scale.exp(); read on the device, no host sync)
    -> (lse [2,n], out [2] = (sum of per-row (CE_i+CE_t)/2, number of correct image->text argmaxes))."""
    _chk(scale, f32, "scale")
    assert scale.numel() == 1
    for t, nm in ((img, "img"), (txt, "txt"), (all_img, "all_img"), (all_txt, "all_txt")):
        _chk(t, f32, nm)
    n, E = img.shape
    N = all_img.shape[0]
    lse = torch.empty(2, n, dtype=f32, device=img.device)
    out = torch.zeros(2, dtype=f32, device=img.device)
    with _timed("infonce", _cuda(img), 8.0 * (n + N) * E):
        K.infonce_fwd(img, txt, all_img, all_txt, scale, n, N, E, row_offset, lse, out)
    return lse, out


def infonce_bwd(img, txt, all_img, all_txt, scale, row_offset, lse, gscale, col_side=True):
    """-> (d_all fp32 [2,N,E] (image block, text block), dscale fp32 [1]).  col_side=False: row-side gradient only (the
    gathered operands carry no gradient: the reference's local_loss=True, gather_with_grad=False mode)."""
    n, E = img.shape
    N = all_img.shape[0]
    dev = img.device
    dlogits = torch.empty(2, n, N, dtype=f32, device=dev)
    d_all = torch.zeros(2, N, E, dtype=f32, device=dev)
    dscale = torch.zeros(1, dtype=f32, device=dev)
    with _timed("infonce", _cuda(img), 8.0 * (n + N) * E + 8.0 * n * N + 8.0 * N * E):
        K.infonce_bwd(img, txt, all_img, all_txt, scale, n, N, E, row_offset, lse, float(gscale), bool(col_side), dlogits, d_all,
                      dscale)
    return d_all, dscale


# ---- VideoMAE pretraining step (csrc/videomae.cu) -----------------------------------------------------------------
def tube_mask_slots(mask):
    """mask bool/uint8 [B,N] (True = masked) -> (slot int32 [B,N], order int32 [B,N], counts int32 [B])."""
    if mask.dtype == torch.bool:
        mask = mask.view(torch.uint8)
    _chk(mask, torch.uint8, "mask")
    B, N = mask.shape
    slot = torch.empty(B, N, dtype=torch.int32, device=mask.device)
    order = torch.empty(B, N, dtype=torch.int32, device=mask.device)
    counts = torch.empty(B, dtype=torch.int32, device=mask.device)
    K.tube_mask_slots(_cuda(mask), slot, order, counts, B, N)
    return slot, order, counts


def tubelet_transform(frames_u8, slot, n_visible, patch, mean, std, want_patches=True, want_targets=True):
    """uint8 [B,T,crop,crop,3] -> (patches bf16 [B,n_visible,6*p*p] of the visible tokens, targets fp32 [B,N-n_visible,6*p*p]
    of the masked tokens); see include/avion_b200.h."""
    _chk(frames_u8, torch.uint8, "frames"); _chk(slot, torch.int32, "slot")
    B, T, Hs, Ws, ch = frames_u8.shape
    assert ch == 3 and Hs == Ws, "the VideoMAE entry takes square clips already cropped (fused decode-crop output)"
    N = (T // 2) * (Hs // patch) ** 2
    assert slot.shape == (B, N)
    F = 6 * patch * patch
    dev = frames_u8.device
    patches = torch.empty(B, n_visible, F, dtype=bf16, device=dev) if want_patches else None
    targets = torch.empty(B, N - n_visible, F, dtype=f32, device=dev) if want_targets else None
    work = float(frames_u8.numel()) * (1.0 + (2.0 * n_visible + 4.0 * (N - n_visible)) / N)
    with _timed("tubelet_transform", _cuda(frames_u8), work):
        K.tubelet_transform_u8(frames_u8, slot, patches, targets, B, T, Hs, patch, n_visible, [float(v) for v in mean],
                               [float(v) for v in std])
    return patches, targets


def gather_rows(table, order, n, add_row=None, add_from=0):
    """out bf16 [B,n,D] = table[order[b,:n]] (+ add_row on rows >= add_from)."""
    _chk(table, f32, "table"); _chk(order, torch.int32, "order")
    B, N = order.shape
    D = table.shape[1]
    out = torch.empty(B, n, D, dtype=bf16, device=table.device)
    K.gather_rows_f32_bf16(_cuda(table), order, N, add_row, add_from, out, B, n, D)
    return out


def colsum_batched(x, B, batch_stride, rows, cols, out):
    """out[cols] (fp32) += sum over batches and rows of a [rows, cols] bf16 block at x + b*batch_stride."""
    _chk(out, f32, "out")
    K.colsum_batched_bf16(_cuda(x), B, batch_stride, rows, cols, out)
    return out


def mse_fwd(pred, target):
    _chk(pred, bf16, "pred"); _chk(target, f32, "target")
    assert pred.shape == target.shape
    ws = torch.empty(_lib.load().avn_mse_workspace_floats(), dtype=f32, device=pred.device)
    loss = torch.empty((), dtype=f32, device=pred.device)
    with _timed("mse", pred, 6.0 * pred.numel()):
        K.mse_fwd(_cuda(pred), target, pred.numel(), ws, loss)
    return loss


def mse_bwd(pred, target, gloss):
    _chk(gloss, f32, "gloss")
    dpred = torch.empty_like(pred)
    with _timed("mse", pred, 8.0 * pred.numel()):
        K.mse_bwd(_cuda(pred), target, gloss, pred.numel(), dpred)
    return dpred


def mmr_fwd(text, image, margin, fix_norm):
    """MaxMarginRankingLoss forward on fp32 [n,E] features -> (loss 0-dim, workspace for mmr_bwd)."""
    _chk(text, f32, "text"); _chk(image, f32, "image")
    assert text.shape == image.shape and text.dim() == 2
    n, E = text.shape
    ws = torch.empty(_lib.load().avn_mmr_workspace_floats(n, E), dtype=f32, device=text.device)
    loss = torch.empty((), dtype=f32, device=text.device)
    K.mmr_fwd(_cuda(text), image, n, E, float(margin), bool(fix_norm), ws, loss)
    return loss, ws


def mmr_bwd(ws, gloss, n, E, margin, fix_norm):
    _chk(gloss, f32, "gloss")
    dtext = torch.empty(n, E, dtype=f32, device=ws.device)
    dimage = torch.empty(n, E, dtype=f32, device=ws.device)
    K.mmr_bwd(_cuda(ws), gloss, n, E, float(margin), bool(fix_norm), dtext, dimage)
    return dtext, dimage
