"""`torch.ops.avion_b200.*`: the C-ABI kernels of include/avion_b200.h registered as torch custom operators.

Two layers:

* kernel ops - one per C-ABI entry point, `avn_<name>` -> `torch.ops.avion_b200.<name>`.  Out-variant schemas: the buffers a
  kernel writes are annotated `Tensor(a!)`, scalars are plain ints / floats, nothing is returned.  The CUDA implementation
  is the ctypes call on torch's current stream (avion_b200/_lib.py); the fake (meta) implementation is a no-op, which is all
  FakeTensor / torch.compile tracing needs for an op that only mutates its arguments.  There is no CPU implementation: the
  dispatcher raises for CPU tensors.  Every launch of the product path (avion_b200/ops.py -> models/_engine.py) goes
  through these.
* functional ops - `layernorm`, `linear`, `attention`, `clip_transform`, `clip_loss`: allocate their results, have
  shape-propagating fake implementations and `register_autograd` formulas built from the kernel ops, so a block written
  with them traces under `torch.compile(fullgraph=True)` and differentiates under plain autograd
  (tests/test_torch_ops.py).

Registration happens at import (avion_b200/__init__.py imports this module); it does not need the shared library - the
library is loaded at the first call.
"""
import ctypes as C
import re

import torch

from . import _lib

NS = "avion_b200"
_LIBRARY = torch.library.Library(NS, "DEF")
_checked_devices = set()

# name -> schema arguments (the C argument order of include/avion_b200.h, minus the trailing stream)
KERNELS = {
    "clip_transform_u8_bf16": "Tensor frames, Tensor(a!) patches, int B, int T, int Hs, int Ws, int crop, int patch, Tensor? boxes, "
                              "float[] mean3, float[] std3",
    "clip_resize_transform_u8_bf16": "Tensor frames, Tensor(a!) patches, int B, int T, int Hs, int Ws, int crop, int patch, "
                                     "Tensor boxes, float[] mean3, float[] std3",
    "layernorm_fwd": "Tensor x, Tensor gamma, Tensor beta, Tensor(a!) y, Tensor(b!)? mean, Tensor(c!)? rstd, int rows, int D, float eps",
    "layernorm_bwd": "Tensor dy, Tensor x, Tensor gamma, Tensor mean, Tensor rstd, Tensor? dres, Tensor(a!) dx, Tensor(b!) dgamma, "
                     "Tensor(c!) dbeta, Tensor(d!)? dcolsum, int rows, int D",
    "attn_fwd": "Tensor qkv, Tensor(a!) out, Tensor(b!) lse, int B, int S, int H, float scale, bool causal",
    "attn_bwd": "Tensor qkv, Tensor out, Tensor dout, Tensor lse, Tensor(a!) dqkv, Tensor(b!) delta, Tensor(c!) dq_acc, int B, int S, "
                "int H, float scale, bool causal",
    "cast_f32_bf16": "Tensor src, Tensor(a!) dst, int n",
    "transpose_cast_f32_bf16": "Tensor src, Tensor(a!) dst, int rows, int cols",
    "colsum_bf16": "Tensor x, int ld, Tensor(a!) colsum, int rows, int cols",
    "pos_table_fwd": "Tensor cls, Tensor pos, Tensor? temporal, Tensor(a!) table, int T, int G, int D",
    "broadcast_row": "Tensor(a!) x, Tensor row, int B, int batch_stride, int D",
    "pos_table_bwd": "Tensor dx, Tensor(a!) ws, Tensor(b!) dcls, Tensor(c!) dpos, Tensor(d!)? dtemporal, int B, int T, int G, int D",
    "token_embed_fwd": "Tensor ids, Tensor tok_emb, Tensor pos, Tensor(a!) x, int B, int L, int D",
    "token_embed_bwd": "Tensor ids, Tensor dx, Tensor(a!) dtok, Tensor(b!) dpos, int B, int L, int D",
    "head_fwd": "Tensor x, Tensor row_idx, Tensor gamma, Tensor beta, Tensor proj, Tensor(a!) emb, Tensor(b!) u, Tensor(c!) mean, "
                "Tensor(d!) rstd, int n, int D, int E, float eps",
    "head_bwd": "Tensor demb, Tensor x, Tensor row_idx, Tensor gamma, Tensor beta, Tensor proj, Tensor u, Tensor mean, Tensor rstd, "
                "Tensor(a!) dx, Tensor(b!) dgamma, Tensor(c!) dbeta, Tensor(d!) dproj, Tensor(e!) ws, int n, int D, int E, bool normalized",
    "infonce_fwd": "Tensor img, Tensor txt, Tensor all_img, Tensor all_txt, Tensor scale, int n, int N, int E, int row_offset, "
                   "Tensor(a!) lse, Tensor(b!) out",
    "infonce_bwd": "Tensor img, Tensor txt, Tensor all_img, Tensor all_txt, Tensor scale, int n, int N, int E, int row_offset, "
                   "Tensor lse, float gscale, bool col_side, Tensor(a!) dlogits, Tensor(b!) d_all, Tensor(c!) dscale",
    "infonce_bwd_static": "Tensor img, Tensor txt, Tensor all_img, Tensor all_txt, Tensor scale, int n, int N, int E, "
                          "int row_offset, Tensor lse, float gscale, bool col_side, Tensor(a!) d_all, Tensor(b!) dscale",
    "adamw_step": "Tensor params, Tensor grads, Tensor exp_avg, Tensor exp_avg_sq, Tensor sizes, Tensor wd, Tensor chunk_tensor, "
                  "Tensor chunk_offset, int n_chunks, float lr, float beta1, float beta2, float eps, int step",
    "clamp_scalar_f32": "Tensor(a!) x, float lo, float hi",
    "tube_mask_slots": "Tensor mask, Tensor(a!) slot, Tensor(b!) order, Tensor(c!)? counts, int B, int N",
    "tubelet_transform_u8": "Tensor frames, Tensor slot, Tensor(a!)? patches, Tensor(b!)? targets, int B, int T, int crop, int patch, "
                            "int n_visible, float[] mean3, float[] std3",
    "gather_rows_f32_bf16": "Tensor table, Tensor idx, int idx_stride, Tensor? add_row, int add_from, Tensor(a!) out, int B, int n, int D",
    "colsum_batched_bf16": "Tensor x, int B, int batch_stride, int rows, int cols, Tensor(a!) out",
    "mse_fwd": "Tensor pred, Tensor target, int n, Tensor(a!) ws, Tensor(b!) loss",
    "mse_bwd": "Tensor pred, Tensor target, Tensor gloss, int n, Tensor(a!) dpred",
    "mmr_fwd": "Tensor text, Tensor image, int n, int E, float margin, bool fix_norm, Tensor(a!) ws, Tensor(b!) loss",
    "mmr_bwd": "Tensor ws, Tensor gloss, int n, int E, float margin, bool fix_norm, Tensor(a!) dtext, Tensor(b!) dimage",
}
GEMM_SCHEMA = ("int mode, int epi, int M, int N, int K, int batch, Tensor A, int lda, int a_bs, Tensor B, int ldb, int b_bs, "
               "Tensor(a!) C, int ldc, int c_bs, Tensor(b!)? C2, Tensor? bias, Tensor? R, int ldr, int r_bs, int splits, "
               "Tensor(c!)? colsum=None")


def _stream_of(t):
    dev = t.device
    if dev.index not in _checked_devices:
        with torch.cuda.device(dev):
            _lib.call("avn_device_check")
        _checked_devices.add(dev.index)
    return C.c_void_p(torch.cuda.current_stream(dev).cuda_stream)


def _arg_kinds(schema):
    kinds = []
    for a in schema.split(","):
        ty = a.strip().split(" ")[0]
        kinds.append("tensor" if ty.startswith("Tensor") else "floats" if ty == "float[]" else re.sub(r"\W", "", ty))
    return kinds


def _make_impl(cname, kinds):
    def impl(*args):
        cargs, first = [], None
        for k, v in zip(kinds, args):
            if k == "tensor":
                if v is None:
                    cargs.append(C.c_void_p(0))
                else:
                    first = v if first is None else first
                    cargs.append(C.c_void_p(v.data_ptr()))
            elif k == "floats":
                cargs.append((C.c_float * len(v))(*[float(x) for x in v]))
            elif k == "float":
                cargs.append(float(v))
            else:                                  # int, bool
                cargs.append(int(v))
        _lib.call(cname, *cargs, _stream_of(first))
    return impl


def _noop(*args):
    return None


def _gemm_impl(mode, epi, M, N, K, batch, A, lda, a_bs, B, ldb, b_bs, Cm, ldc, c_bs, C2, bias, R, ldr, r_bs, splits, colsum=None):
    d = _lib.GemmDesc(mode, epi, M, N, K, batch, A.data_ptr(), lda, a_bs, B.data_ptr(), ldb, b_bs, Cm.data_ptr(), ldc, c_bs,
                      C2.data_ptr() if C2 is not None else None, bias.data_ptr() if bias is not None else None,
                      R.data_ptr() if R is not None else None, ldr, r_bs, splits,
                      colsum.data_ptr() if colsum is not None else None)
    _lib.call("avn_gemm_bf16", C.byref(d), _stream_of(A))


for _name, _schema in KERNELS.items():
    _LIBRARY.define(f"{_name}({_schema}) -> ()")
    _LIBRARY.impl(_name, _make_impl("avn_" + _name, _arg_kinds(_schema)), "CUDA")
    torch.library.register_fake(f"{NS}::{_name}", _noop, lib=_LIBRARY)
_LIBRARY.define(f"gemm_bf16({GEMM_SCHEMA}) -> ()")
_LIBRARY.impl("gemm_bf16", _gemm_impl, "CUDA")
torch.library.register_fake(f"{NS}::gemm_bf16", _noop, lib=_LIBRARY)

K = getattr(torch.ops, NS)      # kernel-op namespace used by avion_b200/ops.py


# =====================================================================================================================
# functional, differentiable operators (what a user composes; each result is a fresh tensor)
# =====================================================================================================================
bf16, f32 = torch.bfloat16, torch.float32
EPI_NONE, EPI_BIAS, EPI_BIAS_GELU, EPI_BIAS_GELU_DUAL, EPI_BIAS_RESIDUAL, EPI_DGELU, EPI_F32_ADD = range(7)
GEMM_NT, GEMM_NN, GEMM_TN = 0, 1, 2


def _def(name, schema, impl, fake):
    _LIBRARY.define(f"{name}{schema}")
    _LIBRARY.impl(name, impl, "CUDA")
    torch.library.register_fake(f"{NS}::{name}", fake, lib=_LIBRARY)


# ---- clip_transform(frames u8 [B,T,Hs,Ws,3]) -> bf16 patches [B, T*g*g, F] ------------------------------------------
def _patch_features(patch):
    return (3 * patch * patch + 7) // 8 * 8


def _clip_transform(frames, crop, patch, mean, std, boxes=None):
    B, T, Hs, Ws, _ = frames.shape
    g = crop // patch
    out = torch.empty(B, T * g * g, _patch_features(patch), dtype=bf16, device=frames.device)
    K.clip_transform_u8_bf16(frames, out, B, T, Hs, Ws, crop, patch, boxes, mean, std)
    return out


def _clip_transform_fake(frames, crop, patch, mean, std, boxes=None):
    B, T = frames.shape[0], frames.shape[1]
    g = crop // patch
    return frames.new_empty((B, T * g * g, _patch_features(patch)), dtype=bf16)


_def("clip_transform", "(Tensor frames, int crop, int patch, float[] mean, float[] std, Tensor? boxes=None) -> Tensor",
     _clip_transform, _clip_transform_fake)


# ---- layernorm(x bf16 [...,D], gamma, beta fp32) -> (y bf16, mean, rstd) ---------------------------------------------
def _layernorm(x, gamma, beta, eps):
    D = x.shape[-1]
    rows = x.numel() // D
    y = torch.empty_like(x)
    mean = torch.empty(rows, dtype=f32, device=x.device)
    rstd = torch.empty(rows, dtype=f32, device=x.device)
    K.layernorm_fwd(x, gamma, beta, y, mean, rstd, rows, D, eps)
    return y, mean, rstd


def _layernorm_fake(x, gamma, beta, eps):
    rows = x.numel() // x.shape[-1]
    return torch.empty_like(x), x.new_empty((rows,), dtype=f32), x.new_empty((rows,), dtype=f32)


def _layernorm_bwd(dy, x, gamma, mean, rstd):
    D = x.shape[-1]
    dx = torch.empty_like(x)
    dgamma = torch.zeros(D, dtype=f32, device=x.device)
    dbeta = torch.zeros(D, dtype=f32, device=x.device)
    K.layernorm_bwd(dy, x, gamma, mean, rstd, None, dx, dgamma, dbeta, None, x.numel() // D, D)
    return dx, dgamma, dbeta


def _layernorm_bwd_fake(dy, x, gamma, mean, rstd):
    return torch.empty_like(x), gamma.new_empty(gamma.shape), gamma.new_empty(gamma.shape)


_def("layernorm", "(Tensor x, Tensor gamma, Tensor beta, float eps) -> (Tensor, Tensor, Tensor)", _layernorm, _layernorm_fake)
_def("layernorm_backward", "(Tensor dy, Tensor x, Tensor gamma, Tensor mean, Tensor rstd) -> (Tensor, Tensor, Tensor)",
     _layernorm_bwd, _layernorm_bwd_fake)


def _ln_setup(ctx, inputs, output):
    x, gamma, _, _ = inputs
    _, mean, rstd = output
    ctx.save_for_backward(x, gamma, mean, rstd)


def _ln_backward(ctx, dy, _dmean, _drstd):
    x, gamma, mean, rstd = ctx.saved_tensors
    dx, dgamma, dbeta = K.layernorm_backward(dy.contiguous(), x, gamma, mean, rstd)
    return dx, dgamma, dbeta, None


torch.library.register_autograd(f"{NS}::layernorm", _ln_backward, setup_context=_ln_setup, lib=_LIBRARY)


# ---- linear(x bf16 [...,K], w bf16 [N,K], bias fp32?, act) -> y bf16 [...,N]; act: 0 none, 1 gelu(erf) ---------------
def _linear(x, w, bias=None, residual=None, act=0):
    Kd = x.shape[-1]
    M, N = x.numel() // Kd, w.shape[0]
    y = torch.empty(*x.shape[:-1], N, dtype=bf16, device=x.device)
    if act == 1:
        epi = EPI_BIAS_GELU
    elif residual is not None:
        epi = EPI_BIAS_RESIDUAL
    else:
        epi = EPI_BIAS if bias is not None else EPI_NONE
    K.gemm_bf16(GEMM_NT, epi, M, N, Kd, 1, x, Kd, 0, w, Kd, 0, y, N, 0, None, bias, residual, N, 0, 1)
    return y


def _linear_fake(x, w, bias=None, residual=None, act=0):
    return x.new_empty((*x.shape[:-1], w.shape[0]), dtype=bf16)


def _linear_bwd(dy, x, w, bias, act, need_bias):
    """-> (dx bf16, dw fp32 [N,K], dbias fp32 [N]); act == 1 recomputes the pre-activation x w^T + b with the dual epilogue."""
    Kd = x.shape[-1]
    M, N = x.numel() // Kd, w.shape[0]
    dev = x.device
    if act == 1:
        g = torch.empty(M, N, dtype=bf16, device=dev)
        pre = torch.empty(M, N, dtype=bf16, device=dev)
        K.gemm_bf16(GEMM_NT, EPI_BIAS_GELU_DUAL, M, N, Kd, 1, x, Kd, 0, w, Kd, 0, g, N, 0, pre, bias, None, N, 0, 1)
        pf = pre.float()                                                # elementwise glue of the demo op only: dy * gelu'(pre)
        dgelu = 0.5 * (1.0 + torch.erf(pf * 0.7071067811865476)) + pf * torch.exp(-0.5 * pf * pf) * 0.3989422804014327
        dy = (dy.reshape(M, N).float() * dgelu).to(bf16)
    dy = dy.reshape(M, N).contiguous()
    dx = torch.empty_like(x)
    K.gemm_bf16(GEMM_NN, EPI_NONE, M, Kd, N, 1, dy, N, 0, w, Kd, 0, dx, Kd, 0, None, None, None, Kd, 0, 1)
    dw = torch.zeros(N, Kd, dtype=f32, device=dev)
    K.gemm_bf16(GEMM_TN, EPI_F32_ADD, N, Kd, M, 1, dy, N, 0, x, Kd, 0, dw, Kd, 0, None, None, None, 0, 0, 0)
    db = torch.zeros(N, dtype=f32, device=dev)
    if need_bias:
        K.colsum_bf16(dy, N, db, M, N)
    return dx, dw, db


def _linear_bwd_fake(dy, x, w, bias, act, need_bias):
    return torch.empty_like(x), w.new_empty(w.shape, dtype=f32), w.new_empty((w.shape[0],), dtype=f32)


_def("linear", "(Tensor x, Tensor w, Tensor? bias=None, Tensor? residual=None, int act=0) -> Tensor", _linear, _linear_fake)
_def("linear_backward", "(Tensor dy, Tensor x, Tensor w, Tensor? bias, int act, bool need_bias) -> (Tensor, Tensor, Tensor)",
     _linear_bwd, _linear_bwd_fake)


def _linear_setup(ctx, inputs, output):
    x, w, bias, residual, act = (tuple(inputs) + (None, None, 0))[:5] if len(inputs) < 5 else inputs
    ctx.save_for_backward(x, w, bias)
    ctx.act, ctx.has_res = act, residual is not None


def _linear_backward(ctx, dy):
    x, w, bias = ctx.saved_tensors
    dx, dw, db = K.linear_backward(dy.contiguous(), x, w, bias, ctx.act, bias is not None)
    return dx, dw.to(w.dtype), (db if bias is not None else None), (dy if ctx.has_res else None), None


torch.library.register_autograd(f"{NS}::linear", _linear_backward, setup_context=_linear_setup, lib=_LIBRARY)


# ---- attention(qkv bf16 [B,S,3*H*64]) -> (out bf16 [B,S,H*64], lse fp32 [B,H,S]) --------------------------------------
def _attention(qkv, heads, scale, causal=False):
    B, S, _ = qkv.shape
    out = torch.empty(B, S, heads * 64, dtype=bf16, device=qkv.device)
    lse = torch.empty(B, heads, S, dtype=f32, device=qkv.device)
    K.attn_fwd(qkv, out, lse, B, S, heads, scale, causal)
    return out, lse


def _attention_fake(qkv, heads, scale, causal=False):
    B, S, _ = qkv.shape
    return qkv.new_empty((B, S, heads * 64)), qkv.new_empty((B, heads, S), dtype=f32)


def _attention_bwd(qkv, out, dout, lse, heads, scale, causal):
    B, S, _ = qkv.shape
    dqkv = torch.empty_like(qkv)
    delta = torch.empty(B, heads, S, dtype=f32, device=qkv.device)
    dq_acc = torch.empty(B, S, heads * 64, dtype=f32, device=qkv.device)
    K.attn_bwd(qkv, out, dout, lse, dqkv, delta, dq_acc, B, S, heads, scale, causal)
    return dqkv


def _attention_bwd_fake(qkv, out, dout, lse, heads, scale, causal):
    return torch.empty_like(qkv)


_def("attention", "(Tensor qkv, int heads, float scale, bool causal=False) -> (Tensor, Tensor)", _attention, _attention_fake)
_def("attention_backward", "(Tensor qkv, Tensor out, Tensor dout, Tensor lse, int heads, float scale, bool causal) -> Tensor",
     _attention_bwd, _attention_bwd_fake)


def _attn_setup(ctx, inputs, output):
    qkv, heads, scale, causal = (tuple(inputs) + (False,))[:4]
    out, lse = output
    ctx.save_for_backward(qkv, out, lse)
    ctx.cfg = (heads, scale, causal)


def _attn_backward(ctx, dout, _dlse):
    qkv, out, lse = ctx.saved_tensors
    return K.attention_backward(qkv, out, dout.contiguous(), lse, *ctx.cfg), None, None, None


torch.library.register_autograd(f"{NS}::attention", _attn_backward, setup_context=_attn_setup, lib=_LIBRARY)


# ---- clip_loss(img, txt fp32 [n,E], scale fp32 []) -> (loss, acc): single-rank ClipLoss (losses.py:124-149) -----------
def _clip_loss(img, txt, scale):
    n, E = img.shape
    lse = torch.empty(2, n, dtype=f32, device=img.device)
    out = torch.zeros(2, dtype=f32, device=img.device)
    K.infonce_fwd(img, txt, img, txt, scale, n, n, E, 0, lse, out)
    return out[0] / n, 100.0 * out[1] / n, lse


def _clip_loss_fake(img, txt, scale):
    return img.new_empty(()), img.new_empty(()), img.new_empty((2, img.shape[0]))


def _clip_loss_bwd(img, txt, scale, lse):
    n, E = img.shape
    dlogits = torch.empty(2, n, n, dtype=f32, device=img.device)
    d_all = torch.zeros(2, n, E, dtype=f32, device=img.device)
    dscale = torch.zeros(1, dtype=f32, device=img.device)
    K.infonce_bwd(img, txt, img, txt, scale, n, n, E, 0, lse, 1.0, True, dlogits, d_all, dscale)
    return d_all[0], d_all[1], dscale.reshape(())


def _clip_loss_bwd_fake(img, txt, scale, lse):
    return torch.empty_like(img), torch.empty_like(txt), img.new_empty(())


_def("clip_loss", "(Tensor img, Tensor txt, Tensor scale) -> (Tensor, Tensor, Tensor)", _clip_loss, _clip_loss_fake)
_def("clip_loss_backward", "(Tensor img, Tensor txt, Tensor scale, Tensor lse) -> (Tensor, Tensor, Tensor)", _clip_loss_bwd,
     _clip_loss_bwd_fake)


def _clip_loss_setup(ctx, inputs, output):
    img, txt, scale = inputs
    ctx.save_for_backward(img, txt, scale, output[2])


def _clip_loss_backward(ctx, gloss, _gacc, _glse):
    img, txt, scale, lse = ctx.saved_tensors
    di, dt, ds = K.clip_loss_backward(img, txt, scale, lse)
    return di * gloss, dt * gloss, ds * gloss


torch.library.register_autograd(f"{NS}::clip_loss", _clip_loss_backward, setup_context=_clip_loss_setup, lib=_LIBRARY)


def block_forward(x, p, heads, eps=1e-6, causal=False):
    """One pre-LN residual block (avion/models/transformer.py:144-150) written with the functional custom ops above:
    x bf16 [B,S,D]; p: dict with fp32 ln1_w/ln1_b/ln2_w/ln2_b/biases and bf16 qkv_w/out_w/fc1_w/fc2_w.  The train step itself
    runs the hand-scheduled engine (models/_engine.py); this composition is the traceable, autograd-differentiable form."""
    h, _, _ = K.layernorm(x, p["ln1_w"], p["ln1_b"], eps)
    qkv = K.linear(h, p["qkv_w"], p["qkv_b"])
    o, _ = K.attention(qkv, heads, 64 ** -0.5, causal)
    x = K.linear(o, p["out_w"], p["out_b"], x)
    h, _, _ = K.layernorm(x, p["ln2_w"], p["ln2_b"], eps)
    a = K.linear(h, p["fc1_w"], p["fc1_b"], None, 1)
    return K.linear(a, p["fc2_w"], p["fc2_b"], x)
