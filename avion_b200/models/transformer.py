"""VisionTransformer / TextTransformer with the reference's constructor arguments, forward signatures and state_dict
keys (flash-attn + fast_conv1 naming, avion/models/utils.py:42-49,63-64), running on the sm_100a kernels.

Reference: avion/models/transformer.py:194-378 (VisionTransformer), :380-478 (TextTransformer).
Each tower is a single torch.autograd.Function (see _engine.py); parameters stay fp32 nn.Parameters (so optimizers, DDP
and checkpoints see what they expect), matrices are cast to bf16 once per forward.
"""
from typing import Callable, Optional

import torch
import torch.nn as nn

from .. import ops
from ..ops import EPI_BIAS_RESIDUAL, EPI_F32_ADD, GEMM_NT, GEMM_TN, bf16, f32
from . import _engine as E

MEAN_OPENAI = (108.3272985, 116.7460125, 104.09373615000001)   # scripts/main_lavila_pretrain.py:408-413
STD_OPENAI = (68.5005327, 66.6321579, 70.32316305)


def _trunc_normal_(t, std):
    return nn.init.trunc_normal_(t, std=std)


class _Affine(nn.Module):
    """Holds LayerNorm parameters under the reference's names (.weight/.bias)."""

    def __init__(self, dim, eps):
        super().__init__()
        self.weight = nn.Parameter(torch.ones(dim))
        self.bias = nn.Parameter(torch.zeros(dim))
        self.eps = eps


class _Dense(nn.Module):
    """Holds nn.Linear-shaped parameters (.weight [out,in], .bias [out])."""

    def __init__(self, in_features, out_features, bias=True):
        super().__init__()
        self.weight = nn.Parameter(torch.empty(out_features, in_features))
        self.bias = nn.Parameter(torch.zeros(out_features)) if bias else None


class _FlashAttnParams(nn.Module):      # names of flash_attn.modules.mha.MHA
    def __init__(self, d):
        super().__init__()
        self.Wqkv = _Dense(d, 3 * d)
        self.out_proj = _Dense(d, d)


class _FlashMlpParams(nn.Module):       # names of flash_attn.modules.mlp.Mlp
    def __init__(self, d, hidden):
        super().__init__()
        self.fc1 = _Dense(d, hidden)
        self.fc2 = _Dense(hidden, d)


class _MhaParams(nn.Module):            # names of nn.MultiheadAttention
    def __init__(self, d):
        super().__init__()
        self.in_proj_weight = nn.Parameter(torch.empty(3 * d, d))
        self.in_proj_bias = nn.Parameter(torch.zeros(3 * d))
        self.out_proj = _Dense(d, d)


class _SeqMlpParams(nn.Module):         # names of the reference's nn.Sequential MLP
    def __init__(self, d, hidden):
        super().__init__()
        self.c_fc = _Dense(d, hidden)
        self.c_proj = _Dense(hidden, d)


class ResidualAttentionBlock(nn.Module):
    """Parameter container for one pre-LN block (avion/models/transformer.py:94-150)."""

    def __init__(self, d_model, n_head, mlp_ratio=4.0, eps=1e-5, flash_names=True):
        super().__init__()
        hidden = int(d_model * mlp_ratio)
        self.ln_1 = _Affine(d_model, eps)
        self.attn = _FlashAttnParams(d_model) if flash_names else _MhaParams(d_model)
        self.ln_2 = _Affine(d_model, eps)
        self.mlp = _FlashMlpParams(d_model, hidden) if flash_names else _SeqMlpParams(d_model, hidden)
        self.flash_names = flash_names

    def param_dict(self):
        a, m = self.attn, self.mlp
        if self.flash_names:
            qw, qb, f1, f2 = a.Wqkv.weight, a.Wqkv.bias, m.fc1, m.fc2
        else:
            qw, qb, f1, f2 = a.in_proj_weight, a.in_proj_bias, m.c_fc, m.c_proj
        return dict(ln1_w=self.ln_1.weight, ln1_b=self.ln_1.bias, qkv_w=qw, qkv_b=qb, out_w=a.out_proj.weight,
                    out_b=a.out_proj.bias, ln2_w=self.ln_2.weight, ln2_b=self.ln_2.bias, fc1_w=f1.weight, fc1_b=f1.bias,
                    fc2_w=f2.weight, fc2_b=f2.bias)


class Transformer(nn.Module):
    def __init__(self, width, layers, heads, mlp_ratio=4.0, eps=1e-5, flash_names=True):
        super().__init__()
        assert width == heads * 64, "the sm_100a attention kernels are built for head dim 64"
        self.width, self.layers, self.heads = width, layers, heads
        self.grad_checkpointing = False     # kept for API parity; selective recompute is always on (see _engine.py)
        self.resblocks = nn.ModuleList([ResidualAttentionBlock(width, heads, mlp_ratio, eps, flash_names) for _ in range(layers)])

    def layer_params(self):
        return [b.param_dict() for b in self.resblocks]


def _flatten(layer_params):
    return [P[k] for P in layer_params for k in E.BLOCK_KEYS]


def _unflatten(flat):
    n = len(E.BLOCK_KEYS)
    return [dict(zip(E.BLOCK_KEYS, flat[i:i + n])) for i in range(0, len(flat), n)]


# =================================================================================================================
class _VisionTowerFn(torch.autograd.Function):
    """patches bf16 [B,P,K] -> embedding fp32 [B,E] (L2-normalised if `normalize`)."""

    @staticmethod
    def forward(ctx, patches, meta, cls, pos, temporal, proj, conv_w, lnpre_w, lnpre_b, lnpost_w, lnpost_b, *flat):
        T, G, heads, eps, normalize, ws = meta
        B, P, K = patches.shape
        D = conv_w.shape[0]
        S = P + 1
        dev = patches.device
        save = any(ctx.needs_input_grad)      # False under torch.no_grad() or with frozen parameters
        layers = E.cast_layers(_unflatten(flat))
        conv_bf = ops.cast_bf16(conv_w.detach())
        if conv_w.shape[1] != K:              # patch rows are padded to a multiple of 8 features (ViT-L/14: 588 -> 592)
            padded = torch.zeros(D, K, dtype=bf16, device=dev)
            padded[:, :conv_w.shape[1]] = conv_bf
            conv_bf = padded
        table = ops.pos_table_fwd(cls.detach(), pos.detach(), temporal.detach() if temporal is not None else None, T, G)
        x_raw = torch.empty(B, S, D, dtype=bf16, device=dev)
        # conv1 (fast_conv1 Linear, no bias) + positional/temporal table, written straight into rows 1.. of [B,S,D]
        ops.gemm(GEMM_NT, EPI_BIAS_RESIDUAL, P, D, K, patches, K, conv_bf, K, x_raw[:, 1:], D, batch=B, a_bs=P * K, c_bs=S * D,
                 R=table[1:], ldr=D, r_bs=0)
        ops.broadcast_row(x_raw, table, B, S * D, D)
        x0, mpre, rpre = ops.layernorm_fwd(x_raw, lnpre_w.detach(), lnpre_b.detach(), eps, save_stats=save)
        if not save:
            del x_raw
        x, saved = E.blocks_forward(x0, layers, heads, eps, False, ws, save)
        row_idx = torch.arange(B, device=dev, dtype=torch.int64) * S
        emb, hsaved = ops.head_fwd(x.view(B * S, D), row_idx, lnpost_w.detach(), lnpost_b.detach(), proj.detach(), eps)
        out = emb if normalize else hsaved[0].clone()
        if save:
            ctx.meta = meta
            ctx.layers, ctx.saved = layers, saved
            ctx.misc = (patches, x_raw, mpre, rpre, x, row_idx, hsaved, conv_bf)
            ctx.params = (cls, pos, temporal, proj, conv_w, lnpre_w, lnpre_b, lnpost_w, lnpost_b) + tuple(flat)
        return out

    @staticmethod
    def backward(ctx, dout):
        T, G, heads, eps, normalize, ws = ctx.meta
        patches, x_raw, mpre, rpre, x_last, row_idx, hsaved, conv_bf = ctx.misc
        cls, pos, temporal, proj, conv_w, lnpre_w, lnpre_b, lnpost_w, lnpost_b = ctx.params[:9]
        flat = ctx.params[9:]
        B, S, D = x_last.shape
        P, K = patches.shape[1], patches.shape[2]
        dev = x_last.device
        dout = dout.contiguous().float()
        layers, saved = ctx.layers, ctx.saved
        grads = E.zero_grads(_unflatten(flat))
        g_cls, g_pos, g_proj = E._zeros_like_f32(cls), E._zeros_like_f32(pos), E._zeros_like_f32(proj)
        g_tmp = E._zeros_like_f32(temporal) if temporal is not None else None
        g_conv = torch.zeros(conv_w.shape[0], K, dtype=f32, device=dev)
        g_lnpre_w, g_lnpre_b = E._zeros_like_f32(lnpre_w), E._zeros_like_f32(lnpre_b)
        g_lnpost_w, g_lnpost_b = E._zeros_like_f32(lnpost_w), E._zeros_like_f32(lnpost_b)
        # head: only the CLS rows of dL/dx are non-zero
        g = ws.get("g", (B, S, D), bf16, dev)
        g.zero_()
        ops.head_bwd(dout, x_last.view(B * S, D), row_idx, lnpost_w.detach(), lnpost_b.detach(), proj.detach(), hsaved,
                     g.view(B * S, D), g_lnpost_w, g_lnpost_b, g_proj, normalized=normalize)
        g = E.blocks_backward(g, saved, layers, grads, heads, eps, False, ws)
        # ln_pre, then conv1 / CLS / positional / temporal
        dx_raw = ops.layernorm_bwd(g, x_raw, lnpre_w.detach(), mpre, rpre, g_lnpre_w, g_lnpre_b,
                                   out=ws.get("dh", (B, S, D), bf16, dev))
        ops.gemm(GEMM_TN, EPI_F32_ADD, D, K, P, dx_raw[:, 1:], D, patches, K, g_conv, K, batch=B, a_bs=S * D, b_bs=P * K,
                 splits=0)
        ops.pos_table_bwd(dx_raw, g_cls, g_pos, g_tmp, T, G)
        ctx.saved = ctx.layers = ctx.misc = None
        flat_grads = [G_[k] for G_ in grads for k in E.BLOCK_KEYS]
        if conv_w.shape[1] != K:
            g_conv = g_conv[:, :conv_w.shape[1]].contiguous()
        return (None, None, g_cls, g_pos, g_tmp, g_proj, g_conv, g_lnpre_w, g_lnpre_b, g_lnpost_w, g_lnpost_b, *flat_grads)


class VisionTransformer(nn.Module):
    """Drop-in for avion.models.transformer.VisionTransformer(use_fast_conv1=True, use_flash_attn=True)."""

    def __init__(self, image_size: int, patch_size: int, width: int, layers: int, heads: int, mlp_ratio: float,
                 num_frames: int = 1, ls_init_value: float = None, global_average_pool: bool = False, output_dim: int = None,
                 patch_dropout: float = 0., drop_rate: float = 0., attn_drop_rate: float = 0., drop_path_rate: float = 0.,
                 ln_pre: bool = True, act_layer: Callable = nn.GELU, norm_layer: Optional[Callable] = None,
                 use_fast_conv1: bool = True, use_flash_attn: bool = True):
        super().__init__()
        # the B200 path implements the configuration every AVION recipe uses; anything else fails loudly
        assert ls_init_value is None, "LayerScale (ls_init_value) is not implemented on the B200 path"
        assert not global_average_pool, "global_average_pool is not implemented on the B200 path"
        assert patch_dropout == 0. and drop_rate == 0. and attn_drop_rate == 0. and drop_path_rate == 0., \
            "dropout variants are not implemented on the B200 path"
        assert ln_pre and output_dim is not None and act_layer is nn.GELU
        self.use_fast_conv1, self.use_flash_attn = True, True     # parameter naming of the flash + fast_conv1 reference
        self.image_size, self.patch_size = (image_size, image_size), (patch_size, patch_size)
        self.width, self.output_dim = width, output_dim
        self.grid_size = (image_size // patch_size, image_size // patch_size)
        self.patches_per_frame = self.grid_size[0] * self.grid_size[1]
        self.eps = 1e-6                                           # partial(nn.LayerNorm, eps=1e-6), transformer.py:213
        self.conv1 = _Dense(3 * patch_size ** 2, width, bias=False)
        scale = width ** -0.5
        self.class_embedding = nn.Parameter(scale * torch.randn(width))
        self.positional_embedding = nn.Parameter(scale * torch.randn(self.patches_per_frame + 1, width))
        assert num_frames >= 1
        self.num_frames = num_frames
        if num_frames > 1:
            self.temporal_embedding = nn.Parameter(torch.zeros(num_frames, width))
        self.ln_pre = _Affine(width, self.eps)
        self.transformer = Transformer(width, layers, heads, mlp_ratio, self.eps, flash_names=True)
        self.ln_post = _Affine(width, self.eps)
        self.image_projection = nn.Parameter(scale * torch.randn(width, output_dim))
        self.mean, self.std = MEAN_OPENAI, STD_OPENAI
        self._ws = E.Workspace()
        self._eye = None
        self.init_parameters()

    def init_parameters(self):                                     # transformer.py:269-287
        nn.init.normal_(self.class_embedding, std=0.02)
        nn.init.normal_(self.positional_embedding, std=0.02)
        nn.init.kaiming_uniform_(self.conv1.weight, a=5 ** 0.5)    # nn.Linear default (conv1 is not re-initialised there)
        for block in self.transformer.resblocks:
            for n, p in block.named_parameters():
                if 'weight' in n:
                    _trunc_normal_(p, std=0.02)
                elif 'bias' in n:
                    nn.init.zeros_(p)
                else:
                    raise NotImplementedError('Unknown parameters named {}'.format(n))
        if self.image_projection is not None:
            nn.init.normal_(self.image_projection, std=self.width ** -0.5)

    def set_grad_checkpointing(self, enable=True):                 # transformer.py:305-307
        self.transformer.grad_checkpointing = enable

    def patches_from(self, x: torch.Tensor) -> torch.Tensor:
        """Accepts (a) uint8 clips [B,T,H,W,3] in decord layout -> fused crop/normalise/patchify kernel,
        (b) bf16 patches [B,P,3*p*p] already produced by GPUClipTransform,
        (c) the reference contract: normalised float [B,3,T,H,W] (or [B,3,H,W] when num_frames == 1)."""
        p = self.patch_size[0]
        if x.dtype == torch.uint8:
            assert x.dim() == 5 and x.shape[1] == self.num_frames
            return ops.clip_transform(x.contiguous(), self.image_size[0], p, self.mean, self.std)
        if x.dim() == 3:
            assert x.shape[2] == ops.patch_features(p), "pre-computed patches must use the padded row length of GPUClipTransform"
            return x.to(bf16).contiguous()
        if x.dim() == 4:
            x = x.unsqueeze(2)
        B, C, T, H, W = x.shape        # layout glue for the float entry only (index permutation of transformer.py:319)
        gh, gw = H // p, W // p
        x = x.reshape(B, C, T, gh, p, gw, p).permute(0, 2, 3, 5, 1, 4, 6).reshape(B, T * gh * gw, C * p * p)
        if ops.patch_features(p) != C * p * p:
            x = torch.nn.functional.pad(x, (0, ops.patch_features(p) - C * p * p))
        return x.to(bf16).contiguous()

    def _run(self, x, normalize):
        patches = self.patches_from(x)
        assert patches.shape[1] == self.num_frames * self.patches_per_frame, "clip length / image size mismatch"
        meta = (self.num_frames, self.patches_per_frame, self.transformer.heads, self.eps, normalize, self._ws)
        flat = _flatten(self.transformer.layer_params())
        tmp = self.temporal_embedding if self.num_frames > 1 else None
        proj = self.image_projection
        if proj is None:      # VideoClassifier drops the projection (model_clip.py:27-28): the head kernel gets an exact identity
            if self._eye is None or self._eye.device != patches.device:
                self._eye = torch.eye(self.width, dtype=f32, device=patches.device)
            proj = self._eye
        return _VisionTowerFn.apply(patches, meta, self.class_embedding, self.positional_embedding, tmp, proj,
                                    self.conv1.weight, self.ln_pre.weight, self.ln_pre.bias, self.ln_post.weight,
                                    self.ln_post.bias, *flat)

    def forward(self, x: torch.Tensor):
        return self._run(x, normalize=False)

    def forward_normalized(self, x: torch.Tensor):
        return self._run(x, normalize=True)


# =================================================================================================================
class _TextTowerFn(torch.autograd.Function):
    """token ids int64 [B,L] -> embedding fp32 [B,E] (L2-normalised if `normalize`)."""

    @staticmethod
    def forward(ctx, text, meta, tok_emb, pos, proj, lnf_w, lnf_b, *flat):
        heads, eps, causal, normalize, ws = meta
        B, Lc = text.shape
        D = tok_emb.shape[1]
        dev = text.device
        save = any(ctx.needs_input_grad)
        layers = E.cast_layers(_unflatten(flat))
        x0 = ops.token_embed_fwd(text.contiguous(), tok_emb.detach(), pos.detach())
        x, saved = E.blocks_forward(x0, layers, heads, eps, causal, ws, save)
        # EOT = highest id in each row (transformer.py:474)
        row_idx = (torch.arange(B, device=dev, dtype=torch.int64) * Lc + text.argmax(dim=-1)).contiguous()
        emb, hsaved = ops.head_fwd(x.view(B * Lc, D), row_idx, lnf_w.detach(), lnf_b.detach(), proj.detach(), eps)
        out = emb if normalize else hsaved[0].clone()
        if save:
            ctx.meta = meta
            ctx.layers, ctx.saved = layers, saved
            ctx.misc = (text, x, row_idx, hsaved)
            ctx.params = (tok_emb, pos, proj, lnf_w, lnf_b) + tuple(flat)
        return out

    @staticmethod
    def backward(ctx, dout):
        heads, eps, causal, normalize, ws = ctx.meta
        text, x_last, row_idx, hsaved = ctx.misc
        tok_emb, pos, proj, lnf_w, lnf_b = ctx.params[:5]
        flat = ctx.params[5:]
        B, Lc, D = x_last.shape
        dev = x_last.device
        grads = E.zero_grads(_unflatten(flat))
        g_tok, g_pos, g_proj = E._zeros_like_f32(tok_emb), E._zeros_like_f32(pos), E._zeros_like_f32(proj)
        g_lnf_w, g_lnf_b = E._zeros_like_f32(lnf_w), E._zeros_like_f32(lnf_b)
        g = ws.get("g", (B, Lc, D), bf16, dev)
        g.zero_()
        ops.head_bwd(dout.contiguous().float(), x_last.view(B * Lc, D), row_idx, lnf_w.detach(), lnf_b.detach(), proj.detach(),
                     hsaved, g.view(B * Lc, D), g_lnf_w, g_lnf_b, g_proj, normalized=normalize)
        g = E.blocks_backward(g, ctx.saved, ctx.layers, grads, heads, eps, causal, ws)
        ops.token_embed_bwd(text.contiguous(), g, g_tok, g_pos)
        ctx.saved = ctx.layers = ctx.misc = None
        flat_grads = [G_[k] for G_ in grads for k in E.BLOCK_KEYS]
        return (None, None, g_tok, g_pos, g_proj, g_lnf_w, g_lnf_b, *flat_grads)


class TextTransformer(nn.Module):
    """Drop-in for avion.models.transformer.TextTransformer (nn.MultiheadAttention parameter naming)."""

    def __init__(self, context_length: int = 77, vocab_size: int = 49408, width: int = 512, heads: int = 8, layers: int = 12,
                 ls_init_value: float = None, output_dim: int = None, act_layer: Callable = nn.GELU,
                 norm_layer: Optional[Callable] = None, causal_mask: float = True, flash_attn: bool = False,
                 flash_mlp: bool = False, fused_bias_fc: bool = False):
        super().__init__()
        assert ls_init_value is None and output_dim is not None and act_layer is nn.GELU
        self.context_length, self.vocab_size, self.width, self.output_dim = context_length, vocab_size, width, output_dim
        self.eps = 1e-5                                             # LayerNorm subclass, transformer.py:35-41
        self.token_embedding = nn.Embedding(vocab_size, width)
        self.positional_embedding = nn.Parameter(torch.empty(context_length, width))
        self.transformer = Transformer(width, layers, heads, 4.0, self.eps, flash_names=False)
        self.ln_final = _Affine(width, self.eps)
        self.text_projection = nn.Parameter(torch.empty(width, output_dim))
        self.causal_mask = causal_mask
        self._ws = E.Workspace()
        self.init_parameters()

    def init_parameters(self):                                     # transformer.py:423-441
        nn.init.normal_(self.token_embedding.weight, std=0.02)
        nn.init.normal_(self.positional_embedding, std=0.01)
        proj_std = (self.width ** -0.5) * ((2 * self.transformer.layers) ** -0.5)
        attn_std = self.width ** -0.5
        fc_std = (2 * self.width) ** -0.5
        for block in self.transformer.resblocks:
            nn.init.normal_(block.attn.in_proj_weight, std=attn_std)
            nn.init.normal_(block.attn.out_proj.weight, std=proj_std)
            nn.init.normal_(block.mlp.c_fc.weight, std=fc_std)
            nn.init.normal_(block.mlp.c_proj.weight, std=proj_std)
        nn.init.normal_(self.text_projection, std=self.width ** -0.5)

    def set_grad_checkpointing(self, enable=True):
        self.transformer.grad_checkpointing = enable

    def _run(self, text, normalize):
        assert text.dtype == torch.int64 and text.shape[1] == self.context_length
        meta = (self.transformer.heads, self.eps, bool(self.causal_mask), normalize, self._ws)
        flat = _flatten(self.transformer.layer_params())
        return _TextTowerFn.apply(text, meta, self.token_embedding.weight, self.positional_embedding, self.text_projection,
                                  self.ln_final.weight, self.ln_final.bias, *flat)

    def forward(self, text, cast_dtype=None):
        return self._run(text, normalize=False)

    def forward_normalized(self, text):
        return self._run(text, normalize=True)
