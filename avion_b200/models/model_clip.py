"""CLIP wrapper + factories with the reference's names and signatures (avion/models/model_clip.py:41-341)."""
import numpy as np
import torch
import torch.nn as nn

from .transformer import TextTransformer, VisionTransformer


class _DenseFn(torch.autograd.Function):
    """y = x @ w^T + b on the tcgen05 GEMM for a small classifier head: x [n,K] (any float dtype), w fp32 [N,K], b fp32 [N].
    N (3806 EK-100 action classes, 97 verbs, 300 nouns) is padded to a multiple of 8 in the bf16 operand copies."""

    @staticmethod
    def forward(ctx, x, w, b):
        from .. import ops
        n, K = x.shape
        N = w.shape[0]
        Np = (N + 7) // 8 * 8
        dev = x.device
        xb = x.detach().to(torch.bfloat16).contiguous()
        wb = torch.zeros(Np, K, dtype=torch.bfloat16, device=dev)
        ops.cast_bf16(w.detach().contiguous(), out=wb[:N])
        bp = torch.zeros(Np, dtype=torch.float32, device=dev)
        bp[:N] = b.detach()
        y = ops.linear_fwd(xb, wb, bp, ops.EPI_BIAS)
        ctx.save_for_backward(xb, wb)
        ctx.dims = (N, Np, x.dtype)
        return y[:, :N].float()

    @staticmethod
    def backward(ctx, dy):
        from .. import ops
        xb, wb = ctx.saved_tensors
        N, Np, xdtype = ctx.dims
        n, K = xb.shape
        dyb = torch.zeros(n, Np, dtype=torch.bfloat16, device=xb.device)
        dyb[:, :N] = dy
        dw = torch.zeros(Np, K, dtype=torch.float32, device=xb.device)
        db = torch.zeros(Np, dtype=torch.float32, device=xb.device)
        ops.linear_wgrad(dyb, xb, dw, dbias=db)          # bias gradient from the dy tiles the GEMM stages
        dx = ops.linear_dgrad(dyb, wb)
        return dx.to(xdtype), dw[:N], db[:N]


class VideoClassifier(nn.Module):
    """avion/models/model_clip.py:15-38: video tower without its projection -> dropout -> fc_cls (the fine-tune model of
    scripts/main_lavila_finetune_cls.py).  Same attribute names / state_dict keys (`visual.*`, `fc_cls.weight`, `fc_cls.bias`)."""

    def __init__(self, vision_model: nn.Module, dropout: float, num_classes: int, **kwargs):
        super().__init__()
        self.visual = vision_model
        self.dropout = nn.Dropout(dropout)
        if hasattr(self.visual, "image_projection"):
            self.visual.image_projection = None
        self.fc_cls = nn.Linear(vision_model.width, num_classes, bias=True)     # parameter holder; the GEMM is ours
        self.fc_cls.weight.data.normal_(mean=0.0, std=0.01)
        self.fc_cls.bias.data.zero_()

    def forward(self, image):
        image_embed = self.visual(image)
        if isinstance(image_embed, list):
            assert len(image_embed) == 1
            image_embed = image_embed[0]
        return _DenseFn.apply(self.dropout(image_embed), self.fc_cls.weight, self.fc_cls.bias)


class CLIP(nn.Module):
    """avion/models/model_clip.py:41-99.  forward(image, text) -> (image_embed, text_embed, logit_scale.exp()), both
    embeddings L2-normalised.  `image` may be uint8 clips [B,T,H,W,3] (fused transform entry), bf16 patches, or the
    reference's normalised float [B,3,T,H,W]."""

    def __init__(self, embed_dim: int, vision_model: nn.Module, text_model: nn.Module, vision_width: int = None,
                 text_width: int = None, freeze_temperature=False, **kwargs):
        super().__init__()
        assert vision_width is None and text_width is None, \
            "extra CLIP-level projections are unused by every AVION factory (model_clip.py:143-148) and not implemented"
        self.visual = vision_model
        self.textual = text_model
        self.logit_scale = nn.Parameter(torch.ones([]) * np.log(1 / 0.07))
        if freeze_temperature:
            self.logit_scale.requires_grad_(False)
        self.image_projection = None
        self.text_projection = None

    def encode_image(self, image):
        return self.visual(image)

    def encode_text(self, text, cast_dtype=None):
        return self.textual(text, cast_dtype=cast_dtype)

    def forward(self, image, text):
        # F.normalize is fused into the head kernel of each tower
        return (self.visual.forward_normalized(image), self.textual.forward_normalized(text), self.logit_scale.exp())


def _build(vision_kw, text_kw, project_embed_dim, freeze_temperature, use_grad_checkpointing, pretrain_zoo, pretrain_path):
    vision_model = VisionTransformer(output_dim=project_embed_dim, **vision_kw)
    text_model = TextTransformer(output_dim=project_embed_dim, **text_kw)
    vision_model.set_grad_checkpointing(use_grad_checkpointing)
    text_model.set_grad_checkpointing(use_grad_checkpointing)
    model = CLIP(embed_dim=project_embed_dim, vision_model=vision_model, text_model=text_model,
                 freeze_temperature=freeze_temperature)
    if pretrain_zoo in (None, "none", "random"):
        return model                               # random init (this package's extension: the reference always loads)
    if pretrain_zoo not in ("openai", "open_clip"):
        raise NotImplementedError(pretrain_zoo)    # model_clip.py:175-176
    if pretrain_path is None:
        # 'openai' downloads through clip.load (model_clip.py:150-152); there is no network behind this package
        raise NotImplementedError(f"pretrain_zoo={pretrain_zoo!r} needs pretrain_path=<local OpenAI / OpenCLIP state dict> "
                                  "(no download here); pass pretrain_zoo='random' for random init")
    # model_clip.py:150-174: OpenAI / OpenCLIP key names -> the flash + fast_conv1 names of this package
    from .utils import remap_keys_from_open_clip_to_vit
    print("=> loading {} model".format(pretrain_zoo))
    state_dict = torch.load(pretrain_path, map_location="cpu")
    if hasattr(state_dict, "state_dict"):          # a pickled OpenAI CLIP module
        state_dict = state_dict.state_dict()
    remapped = remap_keys_from_open_clip_to_vit(state_dict, len(vision_model.transformer.resblocks), len(text_model.transformer.resblocks),
                                                context_length=text_model.context_length, vocab_size=text_model.vocab_size,
                                                use_fast_conv1=True, use_flash_attn=True)
    missing, unexpected = model.load_state_dict(remapped, strict=False)
    print("missing_keys: ", missing)
    print("unexpected_keys: ", unexpected)
    return model


def CLIP_VITB16(freeze_temperature=False, use_grad_checkpointing=False, use_bidirectional_lm=False, context_length=77,
                patch_dropout=0.0, drop_path_rate=0.0, num_frames=1, use_fast_conv1=True, use_flash_attn=True,
                project_embed_dim=512, pretrain_zoo="openai", pretrain_path=None, **kwargs):
    """avion/models/model_clip.py:102-177 (ViT-B/16 video tower + 12L/512 text tower)."""
    return _build(dict(image_size=224, patch_size=16, width=768, layers=12, heads=12, mlp_ratio=4, patch_dropout=patch_dropout,
                       drop_path_rate=drop_path_rate, num_frames=num_frames),
                  dict(context_length=context_length, vocab_size=49408, width=512, heads=8, layers=12,
                       causal_mask=not use_bidirectional_lm),
                  project_embed_dim, freeze_temperature, use_grad_checkpointing, pretrain_zoo, pretrain_path)


def CLIP_VITL14(freeze_temperature=False, use_grad_checkpointing=False, use_bidirectional_lm=False, context_length=77,
                vocab_size=49408, patch_dropout=0.0, drop_path_rate=0.0, num_frames=1, use_fast_conv1=True, use_flash_attn=True,
                project_embed_dim=512, pretrain_zoo="openai", pretrain_path=None, **kwargs):
    """avion/models/model_clip.py:180-259 (ViT-L/14 video tower: 24 blocks, width 1024, 16 heads; 12L/768 text tower).
    Patch 14: token rows are 588 features padded to 592 (16-byte rows); conv1.weight keeps the reference shape [1024, 588]."""
    return _build(dict(image_size=224, patch_size=14, width=1024, layers=24, heads=16, mlp_ratio=4, patch_dropout=patch_dropout,
                       drop_path_rate=drop_path_rate, num_frames=num_frames),
                  dict(context_length=context_length, vocab_size=vocab_size, width=768, heads=12, layers=12,
                       causal_mask=not use_bidirectional_lm),
                  project_embed_dim, freeze_temperature, use_grad_checkpointing, pretrain_zoo, pretrain_path)
