"""VideoMAE pretraining model with the reference's names, constructor arguments and state_dict keys, on the sm_100a kernels.

Reference: avion/models/model_videomae.py:310-397 (encoder), :400-464 (decoder), :467-580 (PretrainVisionTransformer),
:583-604 (VIDEOMAE_VITB16); train step scripts/main_videomae_pretrain.py:316-344.

The whole model is ONE torch.autograd.Function (_VideoMAEFn): tubelet patches of the visible tokens -> patch embedding +
sinusoid positions -> encoder blocks -> norm -> encoder_to_decoder -> [visible | mask_token] + positions -> decoder blocks ->
norm -> head on the masked rows.  The reference patch-embeds all 1568 tokens and then indexes x[~mask]; patch embedding is
per token, so this path gathers first (same values, a tenth of the bytes and FLOPs).  Token order inside the encoder and
decoder is the reference's (boolean-index order: visible ascending, then masked ascending).
"""
from functools import partial

import numpy as np
import torch
import torch.nn as nn

from .. import ops
from ..ops import EPI_BIAS, EPI_BIAS_RESIDUAL, EPI_F32_ADD, EPI_NONE, GEMM_NN, GEMM_NT, GEMM_TN, bf16, f32
from . import _engine as E

MEAN_TIMM = (0.485 * 255, 0.456 * 255, 0.406 * 255)      # scripts/main_videomae_pretrain.py:198-206 (0-255 scale)
STD_TIMM = (0.229 * 255, 0.224 * 255, 0.225 * 255)


def get_sinusoid_encoding_table(n_position, d_hid):
    """model_videomae.py:177-187 (float64 numpy table cast to fp32), shape [1, n_position, d_hid]."""
    pos = np.arange(n_position, dtype=np.float64)[:, None]
    j = np.arange(d_hid)
    table = pos / np.power(10000, 2 * (j // 2) / d_hid)[None, :]
    table[:, 0::2] = np.sin(table[:, 0::2])
    table[:, 1::2] = np.cos(table[:, 1::2])
    return torch.tensor(table, dtype=torch.float, requires_grad=False).unsqueeze(0)


class _Norm(nn.Module):
    def __init__(self, dim):
        super().__init__()
        self.weight = nn.Parameter(torch.ones(dim))
        self.bias = nn.Parameter(torch.zeros(dim))


class _Linear(nn.Module):
    def __init__(self, in_features, out_features, bias=True):
        super().__init__()
        self.weight = nn.Parameter(torch.empty(out_features, in_features))
        nn.init.xavier_uniform_(self.weight)                       # _init_weights, model_videomae.py:353-360
        self.bias = nn.Parameter(torch.zeros(out_features)) if bias else None


class Attention(nn.Module):
    """Parameter names of model_videomae.py:57-79: qkv.weight (no bias), q_bias, v_bias (the k bias is identically zero), proj."""

    def __init__(self, dim):
        super().__init__()
        self.qkv = _Linear(dim, 3 * dim, bias=False)
        self.q_bias = nn.Parameter(torch.zeros(dim))
        self.v_bias = nn.Parameter(torch.zeros(dim))
        self.proj = _Linear(dim, dim)


class Mlp(nn.Module):
    def __init__(self, dim, hidden):
        super().__init__()
        self.fc1 = _Linear(dim, hidden)
        self.fc2 = _Linear(hidden, dim)


class Block(nn.Module):
    """Parameter container of model_videomae.py:104-136 (init_values = 0: no gamma_1 / gamma_2)."""

    def __init__(self, dim, mlp_ratio):
        super().__init__()
        self.norm1 = _Norm(dim)
        self.attn = Attention(dim)
        self.norm2 = _Norm(dim)
        self.mlp = Mlp(dim, int(dim * mlp_ratio))

    def param_dict(self):
        a, m = self.attn, self.mlp
        # model_videomae.py:84-87; autograd splits the bias gradient back into q_bias / v_bias
        qkv_b = torch.cat((a.q_bias, torch.zeros_like(a.v_bias, requires_grad=False), a.v_bias))
        return dict(ln1_w=self.norm1.weight, ln1_b=self.norm1.bias, qkv_w=a.qkv.weight, qkv_b=qkv_b, out_w=a.proj.weight,
                    out_b=a.proj.bias, ln2_w=self.norm2.weight, ln2_b=self.norm2.bias, fc1_w=m.fc1.weight, fc1_b=m.fc1.bias,
                    fc2_w=m.fc2.weight, fc2_b=m.fc2.bias)


class PatchEmbed(nn.Module):
    """model_videomae.py:139-174: Conv3d (2,p,p)/(2,p,p) or, with channel_last, the same weights as a Linear over
    (c p0 p1 p2) features.  Both are one GEMM here; .proj keeps the reference's parameter shapes."""

    def __init__(self, img_size=224, patch_size=16, in_chans=3, embed_dim=768, num_frames=16, tubelet_size=2, channel_last=False):
        super().__init__()
        assert in_chans == 3 and tubelet_size == 2
        self.img_size, self.patch_size = (img_size, img_size), (patch_size, patch_size)
        self.tubelet_size = int(tubelet_size)
        self.num_patches = (img_size // patch_size) ** 2 * (num_frames // self.tubelet_size)
        self.channel_last = channel_last
        feats = in_chans * tubelet_size * patch_size * patch_size
        if channel_last:
            self.proj = _Linear(feats, embed_dim)
        else:
            conv = nn.Conv3d(in_chans, embed_dim, kernel_size=(2, patch_size, patch_size), stride=(2, patch_size, patch_size))
            self.proj = conv                                       # parameter holder only (default Conv3d init, as the reference)

    def weight2d(self):
        w = self.proj.weight
        return w if w.dim() == 2 else w.view(w.shape[0], -1)


def _flatten(layer_params):
    return [P[k] for P in layer_params for k in E.BLOCK_KEYS]


def _unflatten(flat):
    n = len(E.BLOCK_KEYS)
    return [dict(zip(E.BLOCK_KEYS, flat[i:i + n])) for i in range(0, len(flat), n)]


class _VideoMAEFn(torch.autograd.Function):
    """(patches bf16 [B,Nv,F], order int32 [B,N]) -> predictions bf16 [B,N-Nv,F_out] for the masked tokens."""

    @staticmethod
    def forward(ctx, patches, order, meta, pe_w, pe_b, encn_w, encn_b, e2d_w, mask_token, decn_w, decn_b, head_w, head_b, *flat):
        enc_heads, dec_heads, n_enc, eps, pos_enc, pos_dec, ws_e, ws_d = meta
        B, Nv, F = patches.shape
        N = order.shape[1]
        Nm = N - Nv
        De, Dd, Fo = pe_w.shape[0], e2d_w.shape[0], head_w.shape[0]
        dev = patches.device
        save = any(ctx.needs_input_grad)
        layers = E.cast_layers(_unflatten(flat))
        enc_layers, dec_layers = layers[:n_enc], layers[n_enc:]
        pe_bf, e2d_bf, head_bf = ops.cast_bf16(pe_w.detach()), ops.cast_bf16(e2d_w.detach()), ops.cast_bf16(head_w.detach())
        # ---- encoder: patch_embed(x) + pos_embed on the visible tokens (model_videomae.py:376-383)
        pos_vis = ops.gather_rows(pos_enc, order, Nv)
        x0 = torch.empty(B, Nv, De, dtype=bf16, device=dev)
        ops.gemm(GEMM_NT, EPI_BIAS_RESIDUAL, B * Nv, De, F, patches, F, pe_bf, F, x0, De, bias=pe_b.detach(), R=pos_vis, ldr=De)
        del pos_vis
        x, saved_e = E.blocks_forward(x0, enc_layers, enc_heads, eps, False, ws_e, save)
        xn, m_en, r_en = ops.layernorm_fwd(x, encn_w.detach(), encn_b.detach(), eps, save_stats=save)
        # ---- decoder input: [encoder_to_decoder(x_vis) + pos_vis | mask_token + pos_mask] (model_videomae.py:569-577)
        x_full = ops.gather_rows(pos_dec, order, N, add_row=mask_token.detach().reshape(-1), add_from=Nv)
        posd_vis = ops.gather_rows(pos_dec, order, Nv)
        ops.gemm(GEMM_NT, EPI_BIAS_RESIDUAL, Nv, Dd, De, xn, De, e2d_bf, De, x_full, Dd, batch=B, a_bs=Nv * De, c_bs=N * Dd,
                 R=posd_vis, ldr=Dd, r_bs=Nv * Dd)
        del posd_vis
        y, saved_d = E.blocks_forward(x_full, dec_layers, dec_heads, eps, False, ws_d, save)
        # ---- head(norm(x[:, -Nm:])) (model_videomae.py:461): LayerNorm runs over all rows, the GEMM reads the masked ones
        yn, m_dn, r_dn = ops.layernorm_fwd(y, decn_w.detach(), decn_b.detach(), eps, save_stats=save)
        out = torch.empty(B, Nm, Fo, dtype=bf16, device=dev)
        ops.gemm(GEMM_NT, EPI_BIAS, Nm, Fo, Dd, yn[:, Nv:], Dd, head_bf, Dd, out, Fo, batch=B, a_bs=N * Dd, c_bs=Nm * Fo,
                 bias=head_b.detach())
        if save:
            ctx.meta = meta
            ctx.layers, ctx.saved_e, ctx.saved_d = layers, saved_e, saved_d
            ctx.misc = (patches, x, xn, m_en, r_en, y, yn, m_dn, r_dn, e2d_bf, head_bf)
            ctx.params = (pe_w, pe_b, encn_w, encn_b, e2d_w, mask_token, decn_w, decn_b, head_w, head_b) + tuple(flat)
            ctx.dims = (B, N, Nv, F, De, Dd, Fo)
        return out

    @staticmethod
    def backward(ctx, dout):
        enc_heads, dec_heads, n_enc, eps, pos_enc, pos_dec, ws_e, ws_d = ctx.meta
        patches, x_enc, xn, m_en, r_en, y, yn, m_dn, r_dn, e2d_bf, head_bf = ctx.misc
        pe_w, pe_b, encn_w, encn_b, e2d_w, mask_token, decn_w, decn_b, head_w, head_b = ctx.params[:10]
        flat = ctx.params[10:]
        B, N, Nv, F, De, Dd, Fo = ctx.dims
        Nm = N - Nv
        dev = patches.device
        dout = dout.to(bf16).contiguous()
        layers = ctx.layers
        grads = E.zero_grads(_unflatten(flat))
        (g_pe_w, g_pe_b, g_encn_w, g_encn_b, g_e2d, g_mask, g_decn_w, g_decn_b, g_head_w, g_head_b) = E.zeros_f32_like_many(
            [pe_w, pe_b, encn_w, encn_b, e2d_w, mask_token, decn_w, decn_b, head_w, head_b])
        # ---- head
        ops.gemm(GEMM_TN, EPI_F32_ADD, Fo, Dd, Nm, dout, Fo, yn[:, Nv:], Dd, g_head_w, Dd, batch=B, a_bs=Nm * Fo, b_bs=N * Dd,
                 splits=0, colsum=g_head_b)           # + the head's bias gradient, from the dout tiles the GEMM stages
        dyn = ws_d.get("dh", (B, N, Dd), bf16, dev)
        dyn.zero_()                                   # the visible rows of the decoder output receive no gradient
        ops.gemm(GEMM_NN, EPI_NONE, Nm, Dd, Fo, dout, Fo, head_bf, Dd, dyn[:, Nv:], Dd, batch=B, a_bs=Nm * Fo, c_bs=N * Dd)
        g = ops.layernorm_bwd(dyn, y, decn_w.detach(), m_dn, r_dn, g_decn_w, g_decn_b, out=ws_d.get("g", (B, N, Dd), bf16, dev))
        del y, yn, dyn
        # ---- decoder blocks
        dx_full = E.blocks_backward(g, ctx.saved_d, layers[n_enc:], grads[n_enc:], dec_heads, eps, False, ws_d)
        # ---- decoder input: mask_token, encoder_to_decoder
        ops.colsum_batched(dx_full[:, Nv:], B, N * Dd, Nm, Dd, g_mask.view(-1))
        ops.gemm(GEMM_TN, EPI_F32_ADD, Dd, De, Nv, dx_full, Dd, xn, De, g_e2d, De, batch=B, a_bs=N * Dd, b_bs=Nv * De, splits=0)
        dxn = ws_e.get("dh", (B, Nv, De), bf16, dev)
        ops.gemm(GEMM_NN, EPI_NONE, Nv, De, Dd, dx_full, Dd, e2d_bf, De, dxn, De, batch=B, a_bs=N * Dd, c_bs=Nv * De)
        g = ops.layernorm_bwd(dxn, x_enc, encn_w.detach(), m_en, r_en, g_encn_w, g_encn_b, out=ws_e.get("g", (B, Nv, De), bf16, dev))
        del dx_full, dxn, x_enc, xn
        # ---- encoder blocks, patch embedding (the position table is a constant)
        dx0 = E.blocks_backward(g, ctx.saved_e, layers[:n_enc], grads[:n_enc], enc_heads, eps, False, ws_e)
        ops.linear_wgrad(dx0, patches, g_pe_w, dbias=g_pe_b)
        ctx.saved_e = ctx.saved_d = ctx.layers = ctx.misc = None
        flat_grads = [G_[k] for G_ in grads for k in E.BLOCK_KEYS]
        return (None, None, None, g_pe_w, g_pe_b, g_encn_w, g_encn_b, g_e2d, g_mask, g_decn_w, g_decn_b, g_head_w, g_head_b,
                *flat_grads)


class _Blocks(nn.Module):
    def __init__(self, dim, depth, heads, mlp_ratio):
        super().__init__()
        assert dim == heads * 64, "the sm_100a attention kernels are built for head dim 64"
        self.blocks = nn.ModuleList([Block(dim, mlp_ratio) for _ in range(depth)])
        self.norm = _Norm(dim)
        self.num_heads, self.embed_dim = heads, dim

    def get_num_layers(self):
        return len(self.blocks)

    def no_weight_decay(self):
        return {'pos_embed', 'cls_token'}


class PretrainVisionTransformerEncoder(_Blocks):
    """Parameter container of model_videomae.py:310-397 (sine-cosine positions, no class head)."""

    def __init__(self, img_size=224, patch_size=16, embed_dim=768, depth=12, num_heads=12, mlp_ratio=4., tubelet_size=2,
                 channel_last=False, num_frames=16):
        super().__init__(embed_dim, depth, num_heads, mlp_ratio)
        self.patch_embed = PatchEmbed(img_size, patch_size, 3, embed_dim, num_frames, tubelet_size, channel_last)
        self.pos_embed = get_sinusoid_encoding_table(self.patch_embed.num_patches, embed_dim)   # plain tensor, as the reference


class PretrainVisionTransformerDecoder(_Blocks):
    """Parameter container of model_videomae.py:400-464."""

    def __init__(self, patch_size=16, num_classes=1536, embed_dim=384, depth=4, num_heads=6, mlp_ratio=4., tubelet_size=2):
        super().__init__(embed_dim, depth, num_heads, mlp_ratio)
        assert num_classes == 3 * tubelet_size * patch_size ** 2
        self.head = _Linear(embed_dim, num_classes)


class PretrainVisionTransformer(nn.Module):
    """Drop-in for avion.models.model_videomae.PretrainVisionTransformer (model_videomae.py:467-580).

    forward(x, mask) -> predictions [B, N_mask, 3*2*p*p] (bf16) for the masked tokens.  `x` is either uint8 clips
    [B,T,H,W,3] (decord / fused decode-crop layout: normalisation and tubelet patchify run in the transform kernel) or the
    reference contract, normalised float [B,3,T,H,W].  `mask` is bool [B,N], True = masked, the same number per clip."""

    def __init__(self, img_size=224, patch_size=16, encoder_in_chans=3, encoder_num_classes=0, encoder_embed_dim=768,
                 encoder_depth=12, encoder_num_heads=12, decoder_num_classes=1536, decoder_embed_dim=512, decoder_depth=8,
                 decoder_num_heads=8, mlp_ratio=4., qkv_bias=False, qk_scale=None, drop_rate=0., attn_drop_rate=0.,
                 drop_path_rate=0., norm_layer=nn.LayerNorm, init_values=0., use_learnable_pos_emb=False, use_checkpoint=False,
                 use_flash_attn_at_encoder=False, use_flash_attn_at_decoder=False, tubelet_size=2, channel_last=False,
                 num_classes=0, in_chans=0, num_frames=16, mean=MEAN_TIMM, std=STD_TIMM):
        super().__init__()
        # the B200 path implements the configuration of the VideoMAE recipe; anything else fails loudly
        assert qkv_bias and qk_scale is None, "the B200 path implements qkv_bias=True (q_bias / v_bias), default scale"
        assert drop_rate == 0. and attn_drop_rate == 0. and drop_path_rate == 0., "dropout variants are not implemented"
        assert not init_values and not use_learnable_pos_emb and encoder_in_chans == 3 and encoder_num_classes == 0
        eps = norm_layer.keywords.get("eps", 1e-5) if isinstance(norm_layer, partial) else 1e-5
        self.eps = eps
        self.encoder = PretrainVisionTransformerEncoder(img_size, patch_size, encoder_embed_dim, encoder_depth,
                                                        encoder_num_heads, mlp_ratio, tubelet_size, channel_last, num_frames)
        self.decoder = PretrainVisionTransformerDecoder(patch_size, decoder_num_classes, decoder_embed_dim, decoder_depth,
                                                        decoder_num_heads, mlp_ratio, tubelet_size)
        self.encoder_to_decoder = _Linear(encoder_embed_dim, decoder_embed_dim, bias=False)
        self.mask_token = nn.Parameter(torch.zeros(1, 1, decoder_embed_dim))
        self.pos_embed = get_sinusoid_encoding_table(self.encoder.patch_embed.num_patches, decoder_embed_dim)
        nn.init.trunc_normal_(self.mask_token, std=.02)
        self.mean, self.std = tuple(mean), tuple(std)
        self.num_frames, self.img_size, self.patch_size = num_frames, img_size, patch_size
        self._ws_e, self._ws_d = E.Workspace(), E.Workspace()
        self._pos = {}
        self._nv, self.check_mask = None, True

    def get_num_layers(self):
        return len(self.encoder.blocks) + len(self.decoder.blocks)

    @torch.jit.ignore
    def no_weight_decay(self):
        return {'pos_embed', 'cls_token', 'mask_token'}

    def _pos_tables(self, dev):
        if dev not in self._pos:
            self._pos[dev] = (self.encoder.pos_embed[0].to(dev).contiguous(), self.pos_embed[0].to(dev).contiguous())
        return self._pos[dev]

    def prepare(self, x, mask, want_targets=False):
        """-> (patches bf16 [B,Nv,F], order int32 [B,N], targets fp32 [B,Nm,F] or None).  With uint8 clips the transform
        kernel produces patches (and targets) directly; with the reference's float input the tubelet rearrange
        (model_videomae.py:161) is index glue followed by a row gather."""
        p = self.patch_size
        N = self.encoder.patch_embed.num_patches
        assert mask.dim() == 2 and mask.shape[1] == N, "mask must be [B, num_patches]"
        slot, order, counts = ops.tube_mask_slots(mask.to(torch.bool).contiguous())
        if self._nv is None or self.check_mask:
            c = counts.tolist()                                    # host read; x[~mask].reshape needs one count per clip
            assert min(c) == max(c), "every clip must have the same number of visible tokens"
            self._nv = c[0]
            self.check_mask = False                                # tube masks have a fixed ratio: checked on the first call
        else:
            # later calls: no host read, but a different visible count must not pass silently (device-side assert)
            torch._assert_async((counts == self._nv).all())
        Nv = self._nv
        if x.dtype == torch.uint8:
            assert x.dim() == 5 and x.shape[1] == self.num_frames and x.shape[2] == self.img_size
            patches, targets = ops.tubelet_transform(x.contiguous(), slot, Nv, p, self.mean, self.std, True, want_targets)
            return patches, order, targets
        assert not want_targets, "regression targets come from the raw uint8 clip"
        B, C, T, H, W = x.shape
        x = x.reshape(B, C, T // 2, 2, H // p, p, W // p, p).permute(0, 2, 4, 6, 1, 3, 5, 7).reshape(B, N, C * 2 * p * p)
        idx = order[:, :Nv].long().unsqueeze(-1).expand(-1, -1, x.shape[-1])
        return torch.gather(x, 1, idx).to(bf16).contiguous(), order, None

    def forward_patches(self, patches, order):
        enc, dec = self.encoder, self.decoder
        pos_enc, pos_dec = self._pos_tables(patches.device)
        meta = (enc.num_heads, dec.num_heads, len(enc.blocks), self.eps, pos_enc, pos_dec, self._ws_e, self._ws_d)
        flat = _flatten([b.param_dict() for b in enc.blocks] + [b.param_dict() for b in dec.blocks])
        return _VideoMAEFn.apply(patches, order, meta, enc.patch_embed.weight2d(), enc.patch_embed.proj.bias, enc.norm.weight,
                                 enc.norm.bias, self.encoder_to_decoder.weight, self.mask_token, dec.norm.weight, dec.norm.bias,
                                 dec.head.weight, dec.head.bias, *flat)

    def forward(self, x, mask, return_targets=False):
        """return_targets=True (uint8 clips only) is the fused entry of the train step (scripts/main_videomae_pretrain.py:
        316-340): one pass over the clip yields the encoder's patches AND the standardised regression targets."""
        patches, order, targets = self.prepare(x, mask, want_targets=return_targets)
        out = self.forward_patches(patches, order)
        return (out, targets) if return_targets else out

    def forward_with_targets(self, clips_u8, mask):
        return self.forward(clips_u8, mask, return_targets=True)


def VIDEOMAE_VITB16(pretrained=False, **kwargs):
    """model_videomae.py:583-604."""
    init_ckpt = kwargs.pop("init_ckpt", None)
    model = PretrainVisionTransformer(
        img_size=224, patch_size=16, encoder_embed_dim=768, encoder_depth=12, encoder_num_heads=12, encoder_num_classes=0,
        decoder_num_classes=1536, decoder_embed_dim=384, decoder_num_heads=6, mlp_ratio=4, qkv_bias=True,
        norm_layer=partial(nn.LayerNorm, eps=1e-6), **kwargs)
    if pretrained:
        checkpoint = torch.load(init_ckpt, map_location="cpu")
        model.load_state_dict(checkpoint["model"])
    return model
