"""Import the UNMODIFIED reference (/root/reference/avion, or its staged copy oracle/_ref/avion).  TEST INFRASTRUCTURE ONLY.

/root/reference exists only in the build container; oracle/stage_ref.py stages the path's files byte for byte into the
git-ignored oracle/_ref/ so that the GPU box can import them too.  Two absent packages are shimmed
(SURVEY.md section 8c): ``timm.models.layers`` (3 symbols; DropPath rates are 0 in every config) and ``clip``
(only used to download OpenAI weights, avion/models/model_clip.py:150-152).  flash-attn needs CUDA to execute,
so the reference is built with use_flash_attn=False: same math, nn.MultiheadAttention naming; the bijection to
the flash naming is avion/models/utils.py:42-49.
"""
import os
import sys
import types

import torch

LIVE_ROOT = "/root/reference"
STAGED_ROOT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "_ref")
REFERENCE_ROOT = LIVE_ROOT if os.path.isdir(os.path.join(LIVE_ROOT, "avion")) else STAGED_ROOT


def available() -> bool:
    return os.path.isfile(os.path.join(REFERENCE_ROOT, "avion", "models", "transformer.py"))


def live() -> bool:
    """True in the build container (the whole reference tree), False when only the staged files exist."""
    return REFERENCE_ROOT == LIVE_ROOT


def import_reference():
    if not available():
        raise RuntimeError("reference tree not present (expected only in the build container)")
    if "timm.models.layers" not in sys.modules:
        timm, models, layers = (types.ModuleType(n) for n in ("timm", "timm.models", "timm.models.layers"))
        layers.trunc_normal_ = torch.nn.init.trunc_normal_
        layers.to_2tuple = lambda x: x if isinstance(x, tuple) else (x, x)
        layers.DropPath = torch.nn.Identity
        timm.models, models.layers = models, layers
        sys.modules.update({"timm": timm, "timm.models": models, "timm.models.layers": layers})
    sys.modules.setdefault("clip", types.ModuleType("clip"))
    if REFERENCE_ROOT not in sys.path:
        sys.path.insert(0, REFERENCE_ROOT)
    from avion.models.transformer import VisionTransformer, TextTransformer
    from avion.models.model_clip import CLIP
    from avion.losses.losses import ClipLoss
    return VisionTransformer, TextTransformer, CLIP, ClipLoss


_FLASH_TO_MHA = {"attn.Wqkv.weight": "attn.in_proj_weight", "attn.Wqkv.bias": "attn.in_proj_bias",
                 "mlp.fc1.": "mlp.c_fc.", "mlp.fc2.": "mlp.c_proj."}


def to_reference_key(k: str) -> str:
    """flash+fast_conv1 key (our boundary) -> key of the non-flash reference module (utils.py:42-49 inverted)."""
    if k.startswith("visual.transformer."):
        for a, b in _FLASH_TO_MHA.items():
            k = k.replace(a, b)
    return k


def build_reference_clip(cfg, sd, use_flash_attn=False):
    """Reference CLIP built with the factory's literal arguments (model_clip.py:118-148) and loaded with sd.
    use_flash_attn=True (CUDA only) takes sd's flash naming as is; False maps it to the nn.MultiheadAttention naming."""
    VisionTransformer, TextTransformer, CLIP, _ = import_reference()
    vis = VisionTransformer(cfg.image_size, cfg.patch, cfg.width, cfg.layers, cfg.heads, 4,
                            output_dim=cfg.embed_dim, num_frames=cfg.frames,
                            use_fast_conv1=True, use_flash_attn=use_flash_attn)
    txt = TextTransformer(context_length=cfg.ctx, vocab_size=cfg.vocab, width=cfg.t_width, heads=cfg.t_heads,
                          layers=cfg.t_layers, output_dim=cfg.embed_dim, causal_mask=True)
    model = CLIP(embed_dim=cfg.embed_dim, vision_model=vis, text_model=txt)
    keyed = sd if use_flash_attn else {to_reference_key(k): v for k, v in sd.items()}
    missing, unexpected = model.load_state_dict(keyed, strict=True)
    assert not missing and not unexpected
    return model.float()


def import_reference_videomae():
    """avion/models/model_videomae.py unmodified (adds the `drop_path` symbol it imports from timm; rates are 0)."""
    import_reference()
    layers = sys.modules["timm.models.layers"]
    if not hasattr(layers, "drop_path"):
        layers.drop_path = lambda x, drop_prob=0., training=False: x
    from avion.models import model_videomae
    return model_videomae


def build_reference_videomae(cfg, sd):
    """Reference PretrainVisionTransformer with the VIDEOMAE_VITB16 literals (model_videomae.py:583-596) scaled to cfg."""
    from functools import partial
    M = import_reference_videomae()
    assert cfg.frames == 16, "the reference's PatchEmbed hard-codes num_frames=16 (model_videomae.py:142,322-325)"
    model = M.PretrainVisionTransformer(
        img_size=cfg.image_size, patch_size=cfg.patch, encoder_embed_dim=cfg.enc_dim, encoder_depth=cfg.enc_depth,
        encoder_num_heads=cfg.enc_heads, encoder_num_classes=0, decoder_num_classes=cfg.feats, decoder_embed_dim=cfg.dec_dim,
        decoder_num_heads=cfg.dec_heads, decoder_depth=cfg.dec_depth, mlp_ratio=4, qkv_bias=True,
        norm_layer=partial(torch.nn.LayerNorm, eps=1e-6))
    missing, unexpected = model.load_state_dict(sd, strict=True)
    assert not missing and not unexpected
    return model.float()
