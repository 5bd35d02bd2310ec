"""Hand-scheduled forward / backward of the two CLIP towers on the sm_100a kernels.

This is the host-side "runtime" of the hot path: no autograd tracing inside a tower - each tower is ONE
torch.autograd.Function whose forward launches the kernels in order and keeps exactly the activations the
hand-written backward needs (per block: block input, qkv, attention output + LSE, post-attention residual, LayerNorm
statistics).  The MLP hidden activations - and, unless HBM clearly has room for them (keep_ln_outputs), the LayerNorm
outputs - are recomputed in the backward (selective recompute:
+~4.6 % FLOPs instead of the reference's whole-block checkpoint, +33 %), which is what lets 256 clips x 3137 tokens
x 12 blocks fit in 180 GB.

Reference semantics: avion/models/transformer.py:144-150 (block), :309-378 (video tower), :455-478 (text tower).
"""
import os

import torch

from .. import ops
from ..ops import (EPI_BIAS, EPI_BIAS_GELU, EPI_BIAS_GELU_DUAL, EPI_BIAS_RESIDUAL, EPI_DGELU, EPI_F32_ADD, EPI_NONE, GEMM_NT,
                   GEMM_TN, bf16, f32)

FOLD_BIAS_GRAD = os.environ.get("AVN_FOLD_BIAS_GRAD", "1") != "0"      # "0": separate column-sum launches (A/B runs)
BLOCK_KEYS = ("ln1_w", "ln1_b", "qkv_w", "qkv_b", "out_w", "out_b", "ln2_w", "ln2_b", "fc1_w", "fc1_b", "fc2_w", "fc2_b")
_MATS = ("qkv_w", "out_w", "fc1_w", "fc2_w")


class Workspace:
    """Named scratch buffers reused across blocks and steps (nothing here is read after the op that consumes it)."""

    def __init__(self):
        self.buf = {}
        self.flags = {}

    def get(self, name, shape, dtype, device):
        t = self.buf.get(name)
        n = 1
        for s in shape:
            n *= s
        if t is None or t.numel() < n or t.dtype != dtype or t.device != device:
            t = torch.empty(n, dtype=dtype, device=device)
            self.buf[name] = t
        return t[:n].view(*shape)


def _zeros_like_f32(p):
    return torch.zeros(p.shape, dtype=f32, device=p.device)


def keep_ln_outputs(B, S, D, n_layers, dev, margin=24e9):
    """Keep the two LayerNorm outputs of every block for the backward (2 x 2 bytes x B*S*D per block: 29.6 GB for 256 clips of
    ViT-B/16 16f) instead of recomputing them, when HBM clearly has the room: free memory now, minus what this forward will
    save anyway (x, qkv, o, x_mid per block), the workspaces and the extra LayerNorm outputs, must leave `margin` bytes."""
    force = os.environ.get("AVN_KEEP_LN")       # "0" / "1" override the memory test
    if force is not None:
        return force == "1"
    bsd = float(B) * S * D
    need = (12.0 * n_layers + 32.0 + 4.0 * n_layers) * bsd
    free, _ = torch.cuda.mem_get_info(dev)
    return free + torch.cuda.memory_reserved(dev) - torch.cuda.memory_allocated(dev) - need >= margin


def blocks_forward(x, layers, heads, eps, causal, ws, save):
    """x bf16 [B,S,D] -> x_out; `layers` = list of dicts (BLOCK_KEYS; matrices already bf16).  Returns (x, saved list)."""
    B, S, D = x.shape
    dev = x.device
    scale = 64 ** -0.5
    saved = []
    key = ("keep_ln", B, S, D, len(layers))
    if save and key not in ws.flags:          # decided once per shape, at the first training forward
        ws.flags[key] = keep_ln_outputs(B, S, D, len(layers), dev)
    keep = save and ws.flags[key]
    for L in layers:
        h = torch.empty(B, S, D, dtype=bf16, device=dev) if keep else ws.get("h", (B, S, D), bf16, dev)
        h1 = h if keep else None
        _, m1, r1 = ops.layernorm_fwd(x, L["ln1_w"], L["ln1_b"], eps, out=h, save_stats=save)
        qkv = ops.linear_fwd(h, L["qkv_w"], L["qkv_b"], EPI_BIAS) if save else \
            ops.linear_fwd(h, L["qkv_w"], L["qkv_b"], EPI_BIAS, out=ws.get("qkv", (B, S, 3 * D), bf16, dev))
        o, lse = ops.attn_fwd(qkv, heads, scale, causal, out=None if save else ws.get("o", (B, S, D), bf16, dev))
        x_mid = ops.linear_fwd(o, L["out_w"], L["out_b"], EPI_BIAS_RESIDUAL, residual=x)
        if keep:
            h = torch.empty(B, S, D, dtype=bf16, device=dev)
        h2 = h if keep else None
        _, m2, r2 = ops.layernorm_fwd(x_mid, L["ln2_w"], L["ln2_b"], eps, out=h, save_stats=save)
        a = ops.linear_fwd(h, L["fc1_w"], L["fc1_b"], EPI_BIAS_GELU, out=ws.get("a", (B, S, L["fc1_w"].shape[0]), bf16, dev))
        x_out = ops.linear_fwd(a, L["fc2_w"], L["fc2_b"], EPI_BIAS_RESIDUAL, residual=x_mid)
        if save:
            saved.append((x, m1, r1, qkv, o, lse, x_mid, m2, r2, h1, h2))
        x = x_out
    return x, saved


def blocks_backward(g, saved, layers, grads, heads, eps, causal, ws, last_colsum_done=False):
    """g bf16 [B,S,D] = dL/d(x_out of the last block).  Accumulates parameter gradients (fp32) into `grads`
    (list of dicts keyed like `layers`) and returns dL/d(x_in of the first block).  Consumes `saved`."""
    B, S, D = g.shape
    dev = g.device
    scale = 64 ** -0.5
    n_layers = len(layers)
    fc2_bias_done = last_colsum_done
    for li in range(n_layers - 1, -1, -1):
        L, G = layers[li], grads[li]
        x_in, m1, r1, qkv, o, lse, x_mid, m2, r2, h1, h2 = saved.pop()
        Hd = L["fc1_w"].shape[0]
        # ---- MLP half: x_out = x_mid + fc2(gelu(fc1(ln_2(x_mid))))
        if h2 is not None:
            h = h2                                  # kept by the forward (keep_ln_outputs)
        else:
            h = ws.get("h", (B, S, D), bf16, dev)
            ops.layernorm_fwd(x_mid, L["ln2_w"], L["ln2_b"], eps, out=h, save_stats=False)
        a = ws.get("a", (B, S, Hd), bf16, dev)
        pre = ws.get("pre", (B, S, Hd), bf16, dev)      # receives the fc1 pre-activation (gelu' is applied by the dgrad epilogue)
        ops.linear_fwd(h, L["fc1_w"], L["fc1_b"], EPI_BIAS_GELU_DUAL, out=a, out2=pre)
        # bias gradients = column sums of the incoming gradient: taken by the weight-gradient GEMM from the tiles it stages
        # anyway (FOLD_BIAS_GRAD; the two that feed a residual stream come out of the LayerNorm backward)
        if not fc2_bias_done and not FOLD_BIAS_GRAD:
            ops.colsum(g, G["fc2_b"])
        ops.linear_wgrad(g, a, G["fc2_w"], dbias=G["fc2_b"] if (FOLD_BIAS_GRAD and not fc2_bias_done) else None)
        dpre = ops.linear_dgrad(g, L["fc2_w"], EPI_DGELU, pre=pre, out=a)      # a is dead after the wgrad: reuse it
        if not FOLD_BIAS_GRAD:
            ops.colsum(dpre, G["fc1_b"])
        ops.linear_wgrad(dpre, h, G["fc1_w"], dbias=G["fc1_b"] if FOLD_BIAS_GRAD else None)
        dh = ops.linear_dgrad(dpre, L["fc1_w"], out=ws.get("dh", (B, S, D), bf16, dev))
        g1 = ws.get("g1", (B, S, D), bf16, dev)
        ops.layernorm_bwd(dh, x_mid, L["ln2_w"], m2, r2, G["ln2_w"], G["ln2_b"], dres=g, dcolsum=G["out_b"], out=g1)
        # ---- attention half: x_mid = x_in + out_proj(attn(qkv(ln_1(x_in))))
        ops.linear_wgrad(g1, o, G["out_w"])
        do = ops.linear_dgrad(g1, L["out_w"], out=dh)
        dqkv = ops.attn_bwd(qkv, o, do, lse, heads, scale, causal, dqkv=ws.get("dqkv", (B, S, 3 * D), bf16, dev),
                            dq_acc=ws.get("dq_acc", (B, S, D), f32, dev))
        del qkv, o, lse
        if not FOLD_BIAS_GRAD:
            ops.colsum(dqkv, G["qkv_b"])
        del h, h2
        if h1 is not None:
            h = h1
        else:
            h = ws.get("h", (B, S, D), bf16, dev)
            ops.layernorm_fwd(x_in, L["ln1_w"], L["ln1_b"], eps, out=h, save_stats=False)
        ops.linear_wgrad(dqkv, h, G["qkv_w"], dbias=G["qkv_b"] if FOLD_BIAS_GRAD else None)
        dh1 = ops.linear_dgrad(dqkv, L["qkv_w"], out=dh)
        # dL/dx_in = g1 + LN1'(dh1); its column sum is the fc2 bias gradient of the block below
        below = grads[li - 1]["fc2_b"] if li > 0 else None
        g_new = ws.get("g", (B, S, D), bf16, dev)      # the incoming g is dead once g1 exists
        ops.layernorm_bwd(dh1, x_in, L["ln1_w"], m1, r1, G["ln1_w"], G["ln1_b"], dres=g1, dcolsum=below, out=g_new)
        g = g_new
        fc2_bias_done = True
        del x_in, x_mid, h, h1
    return g


def cast_layers(layer_params):
    """fp32 master parameters -> per-step bf16 copies of the matrices (biases and LayerNorm affine stay fp32)."""
    out = []
    for P in layer_params:
        L = dict(P)
        for k in _MATS:
            L[k] = ops.cast_bf16(P[k].detach())
        for k in BLOCK_KEYS:
            if k not in _MATS:
                L[k] = P[k].detach()
        out.append(L)
    return out


def zeros_f32_like_many(tensors):
    """One zero-filled fp32 allocation (one memset instead of one fill kernel per gradient tensor) carved into views shaped like
    `tensors`; every view starts on a 16-byte boundary (TMA reduce-add target of the wgrad GEMM)."""
    sizes = [(t.numel() + 3) // 4 * 4 for t in tensors]
    flat = torch.zeros(sum(sizes), dtype=f32, device=tensors[0].device)
    out, off = [], 0
    for t, n in zip(tensors, sizes):
        out.append(flat[off:off + t.numel()].view(t.shape))
        off += n
    return out


def zero_grads(layer_params):
    flat = zeros_f32_like_many([P[k] for P in layer_params for k in BLOCK_KEYS])
    n = len(BLOCK_KEYS)
    return [dict(zip(BLOCK_KEYS, flat[i * n:(i + 1) * n])) for i in range(len(layer_params))]
