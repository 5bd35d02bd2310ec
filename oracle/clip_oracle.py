"""CPU oracle for the AVION video-CLIP train-step hot path.  TEST INFRASTRUCTURE ONLY.

This file restates, in plain fp32 PyTorch on the CPU, the arithmetic of the reference's hot path
(SURVEY.md section 8a).  Only ``tests/``, ``__graft_entry__.smoke()`` and the ``cpu_baseline`` /
``--impl reference`` legs of ``bench.py`` may import it; the product package ``avion_b200`` never does.

Parity status: PINNED by execution of the reference itself.  The reference has no tests or golden
vectors for this path, so ``oracle/make_golden.py`` imports the unmodified
reference modules from /root/reference (non-flash branch, fp32, CPU), runs them on seeded inputs and
commits the outputs under ``tests/golden/``; ``tests/test_oracle.py`` checks this restatement against
those vectors (and against the live reference when /root/reference is present).
The swscale *resize* that decord applies after its crop has no source under /root/reference, so pixel
values after a resize are "parity unpinned"; crop indices, patch order and the no-resize pixel path are
bit-exact.

Every function cites the reference lines it follows (paths relative to /root/reference).
"""
from __future__ import annotations

import math
from dataclasses import dataclass, asdict

import numpy as np
import torch
import torch.nn.functional as F

# scripts/main_lavila_pretrain.py:408-413 (norm_style == "openai", 0-255 scale)
MEAN_OPENAI = (108.3272985, 116.7460125, 104.09373615000001)
STD_OPENAI = (68.5005327, 66.6321579, 70.32316305)


@dataclass(frozen=True)
class ClipCfg:
    """Shapes of one CLIP dual encoder (factory literals: avion/models/model_clip.py:118-148)."""
    image_size: int = 224
    patch: int = 16
    width: int = 768
    layers: int = 12
    heads: int = 12
    frames: int = 16
    embed_dim: int = 256          # scripts/main_lavila_pretrain.py:126 (--project-embed-dim)
    ctx: int = 77
    vocab: int = 49408
    t_width: int = 512
    t_heads: int = 8
    t_layers: int = 12

    @property
    def grid(self):
        return self.image_size // self.patch

    @property
    def tokens(self):
        return self.frames * self.grid * self.grid + 1

    def asdict(self):
        return asdict(self)


TINY = ClipCfg(image_size=64, patch=16, width=128, layers=2, heads=2, frames=4, embed_dim=64,
               ctx=77, vocab=512, t_width=128, t_heads=2, t_layers=2)
TINY14 = ClipCfg(image_size=56, patch=14, width=128, layers=2, heads=2, frames=4, embed_dim=64,
                 ctx=77, vocab=512, t_width=128, t_heads=2, t_layers=2)      # patch 14 (588 features per token) like ViT-L/14
VITB16_4F = ClipCfg(frames=4)          # BASELINE.json configs[0]
VITB16_16F = ClipCfg(frames=16)        # BASELINE.json configs[1]
VITL14_16F = ClipCfg(patch=14, width=1024, layers=24, heads=16, frames=16, t_width=768, t_heads=12)


# --------------------------------------------------------------------------------------------------
# deterministic inputs and weights (portable: numpy legacy RandomState streams are frozen)
# --------------------------------------------------------------------------------------------------
def make_state_dict(cfg: ClipCfg, seed: int = 0) -> dict:
    """Random-init weights in the reference's flash+fast_conv1 key naming
    (avion/models/utils.py:42-49,63-64; shapes: avion/models/transformer.py:226-266,400-417).
    Scales follow init_parameters (transformer.py:272-287, 426-441) except that temporal_embedding and all
    biases are made non-zero so that ordering / bias bugs cannot hide."""
    rs = np.random.RandomState(seed)

    def n(*shape, std=0.02):
        return torch.from_numpy((rs.standard_normal(shape) * std).astype(np.float32))

    D, E, G = cfg.width, cfg.embed_dim, cfg.grid * cfg.grid
    sd = {"logit_scale": torch.tensor(math.log(1 / 0.07), dtype=torch.float32)}
    v = "visual."
    sd[v + "class_embedding"] = n(D)
    sd[v + "positional_embedding"] = n(G + 1, D)
    sd[v + "temporal_embedding"] = n(cfg.frames, D)
    sd[v + "image_projection"] = n(D, E, std=D ** -0.5)
    sd[v + "conv1.weight"] = n(D, 3 * cfg.patch ** 2)
    for ln in ("ln_pre", "ln_post"):
        sd[v + ln + ".weight"] = 1.0 + n(D, std=0.1)
        sd[v + ln + ".bias"] = n(D, std=0.1)
    for i in range(cfg.layers):
        p = f"{v}transformer.resblocks.{i}."
        for ln in ("ln_1", "ln_2"):
            sd[p + ln + ".weight"] = 1.0 + n(D, std=0.1)
            sd[p + ln + ".bias"] = n(D, std=0.1)
        sd[p + "attn.Wqkv.weight"] = n(3 * D, D, std=D ** -0.5)
        sd[p + "attn.Wqkv.bias"] = n(3 * D)
        sd[p + "attn.out_proj.weight"] = n(D, D)
        sd[p + "attn.out_proj.bias"] = n(D)
        sd[p + "mlp.fc1.weight"] = n(4 * D, D)
        sd[p + "mlp.fc1.bias"] = n(4 * D)
        sd[p + "mlp.fc2.weight"] = n(D, 4 * D)
        sd[p + "mlp.fc2.bias"] = n(D)
    t = "textual."
    W = cfg.t_width
    sd[t + "positional_embedding"] = n(cfg.ctx, W, std=0.01)
    sd[t + "text_projection"] = n(W, E, std=W ** -0.5)
    sd[t + "token_embedding.weight"] = n(cfg.vocab, W)
    for i in range(cfg.t_layers):
        p = f"{t}transformer.resblocks.{i}."
        for ln in ("ln_1", "ln_2"):
            sd[p + ln + ".weight"] = 1.0 + n(W, std=0.1)
            sd[p + ln + ".bias"] = n(W, std=0.1)
        sd[p + "attn.in_proj_weight"] = n(3 * W, W, std=W ** -0.5)
        sd[p + "attn.in_proj_bias"] = n(3 * W)
        sd[p + "attn.out_proj.weight"] = n(W, W, std=W ** -0.5 * (2 * cfg.t_layers) ** -0.5)
        sd[p + "attn.out_proj.bias"] = n(W)
        sd[p + "mlp.c_fc.weight"] = n(4 * W, W, std=(2 * W) ** -0.5)
        sd[p + "mlp.c_fc.bias"] = n(4 * W)
        sd[p + "mlp.c_proj.weight"] = n(W, 4 * W, std=W ** -0.5 * (2 * cfg.t_layers) ** -0.5)
        sd[p + "mlp.c_proj.bias"] = n(W)
    sd[t + "ln_final.weight"] = 1.0 + n(W, std=0.1)
    sd[t + "ln_final.bias"] = n(W, std=0.1)
    return sd


def make_inputs(cfg: ClipCfg, batch: int, seed: int = 1234, src_hw=None):
    """Synthetic uint8 clips in decord layout [B,T,H,W,3] (avion/data/clip_dataset.py:169) and token ids with
    SOT=49406-style / EOT = row maximum, zero padded (avion/data/tokenizer.py:169-178)."""
    rs = np.random.RandomState(seed)
    H, W = src_hw if src_hw is not None else (cfg.image_size, cfg.image_size)
    clips = torch.from_numpy(rs.randint(0, 256, (batch, cfg.frames, H, W, 3)).astype(np.uint8))
    sot, eot = cfg.vocab - 2, cfg.vocab - 1
    tokens = np.zeros((batch, cfg.ctx), dtype=np.int64)
    for b in range(batch):
        L = int(rs.randint(5, cfg.ctx - 1))
        tokens[b, 0] = sot
        tokens[b, 1:L] = rs.randint(1, sot, L - 1)
        tokens[b, L] = eot
    return clips, torch.from_numpy(tokens)


# --------------------------------------------------------------------------------------------------
# T1/T2: clip transform + patchify
# --------------------------------------------------------------------------------------------------
def crop_clip(frames_u8: torch.Tensor, boxes, out_hw) -> torch.Tensor:
    """Fixed-size crop per clip (no resize): one (x, y, hflip) box per clip shared by all its frames,
    the way decord samples one crop per reader (third_party/decord/src/video/ffmpeg/threaded_decoder.cc:49-93,
    filter string 'crop=w:h:x:y,[hflip,]' :92).  boxes: int tensor [B,3]."""
    B = frames_u8.shape[0]
    h, w = out_hw
    out = []
    for b in range(B):
        x, y, flip = (int(v) for v in boxes[b])
        c = frames_u8[b, :, y:y + h, x:x + w, :]
        if flip:
            c = torch.flip(c, dims=[2])
        out.append(c)
    return torch.stack(out)


def _aa_weights(in_size: int, out_size: int) -> torch.Tensor:
    """Dense [out_size, in_size] fp32 weight matrix of torch's antialiased bilinear resize
    (aten/src/ATen/native/cpu/UpSampleKernel.cpp, HelperInterpBase::_compute_indices_min_size_weights_aa with the triangle
    filter; align_corners=False): centre = scale (i + 0.5), support = max(scale, 1), taps
    [max(0, int(centre - support + 0.5)), min(in, int(centre + support + 0.5))), weight 1 - |(j - centre + 0.5) / max(scale, 1)|
    clipped at 0 and normalised by their sum.  All arithmetic in fp32 like the float kernel."""
    f32 = np.float32
    scale = f32(in_size) / f32(out_size)
    support = scale if scale >= 1 else f32(1)
    inv = f32(1) / scale if scale >= 1 else f32(1)
    W = np.zeros((out_size, in_size), dtype=np.float32)
    for i in range(out_size):
        c = scale * (f32(i) + f32(0.5))
        lo = max(int(c - support + f32(0.5)), 0)
        hi = min(int(c + support + f32(0.5)), in_size)
        j = np.arange(lo, hi, dtype=np.float32)
        w = np.maximum(f32(0), f32(1) - np.abs((j - c + f32(0.5)) * inv)).astype(np.float32)
        W[i, lo:hi] = w / w.sum(dtype=np.float32)
    return torch.from_numpy(W)


def resized_crop_clip(frames_u8: torch.Tensor, boxes, out_hw) -> torch.Tensor:
    """Random-RESIZED crop per clip: box (x, y, w, h, hflip) of each clip -> crop -> hflip -> resize to out_hw -> uint8.
    Order of decord's filter graph 'crop=w:h:x:y,[hflip,]scale=W:H' (third_party/decord/src/video/ffmpeg/threaded_decoder.cc:92);
    the offsets are clamped into the frame as ffmpeg's crop filter does (the sampler's upper bound is one too large, :66-67).
    The resize itself restates torchvision's RandomResizedCrop of the non-fused path (scripts/main_lavila_pretrain.py:437-441):
    bilinear with antialiasing (float kernel), rounded half-to-even and clamped to uint8.  swscale (what decord really
    calls) has no source in the tree: PARITY UNPINNED against decord pixels; pinned against torch.nn.functional.interpolate
    and torchvision.transforms.v2.functional.resized_crop in tests/test_oracle.py."""
    B, T, Hs, Ws, _ = frames_u8.shape
    H, W = out_hw
    out = torch.empty(B, T, H, W, 3, dtype=torch.uint8)
    for b in range(B):
        x, y, w, h, flip = (int(v) for v in boxes[b])
        w, h = min(max(w, 1), Ws), min(max(h, 1), Hs)
        x, y = min(max(x, 0), Ws - w), min(max(y, 0), Hs - h)
        c = frames_u8[b, :, y:y + h, x:x + w, :].to(torch.float32)          # [T,h,w,3]
        if flip:
            c = torch.flip(c, dims=[2])
        Wx, Wy = _aa_weights(w, W), _aa_weights(h, H)
        hz = torch.einsum("ow,thwc->thoc", Wx, c)                            # horizontal pass first, like the aten kernel
        r = torch.einsum("ph,thoc->tpoc", Wy, hz)
        out[b] = r.round().clamp_(0, 255).to(torch.uint8)
    return out


def clip_transform(frames_u8: torch.Tensor, mean=MEAN_OPENAI, std=STD_OPENAI) -> torch.Tensor:
    """uint8 [B,T,H,W,3] -> f32 [B,3,T,H,W] normalised.
    astype(float32): avion/data/clip_dataset.py:104,164; permute: scripts/main_lavila_pretrain.py:816;
    kornia Normalize (x-mean)/std per channel on the 0-255 scale: :430,817."""
    x = frames_u8.to(torch.float32).permute(0, 4, 1, 2, 3)
    m = torch.tensor(mean, dtype=torch.float32).view(1, 3, 1, 1, 1)
    s = torch.tensor(std, dtype=torch.float32).view(1, 3, 1, 1, 1)
    return (x - m) / s


def patchify(x: torch.Tensor, p: int) -> torch.Tensor:
    """'b c t (hh sh) (ww sw) -> b (t hh ww) (c sh sw)' (avion/models/transformer.py:319)."""
    B, C, T, H, W = x.shape
    gh, gw = H // p, W // p
    x = x.reshape(B, C, T, gh, p, gw, p).permute(0, 2, 3, 5, 1, 4, 6)
    return x.reshape(B, T * gh * gw, C * p * p)


# --------------------------------------------------------------------------------------------------
# V1-V8: vision tower
# --------------------------------------------------------------------------------------------------
def _block(x, sd, p, heads, eps, names, attn_mask=None):
    """Pre-LN residual block, x: [B,S,D] (avion/models/transformer.py:144-150; flash branch math ==
    nn.MultiheadAttention branch math :140-146)."""
    qkv_w, qkv_b, fc1, fc2 = names
    B, S, D = x.shape
    hd = D // heads
    h = F.layer_norm(x, (D,), sd[p + "ln_1.weight"], sd[p + "ln_1.bias"], eps)
    qkv = F.linear(h, sd[p + qkv_w], sd[p + qkv_b]).view(B, S, 3, heads, hd)
    q, k, v = (qkv[:, :, i].transpose(1, 2) for i in range(3))         # [B,H,S,hd]
    s = (q @ k.transpose(-1, -2)) * (hd ** -0.5)
    if attn_mask is not None:
        s = s + attn_mask
    o = (torch.softmax(s, dim=-1) @ v).transpose(1, 2).reshape(B, S, D)
    x = x + F.linear(o, sd[p + "attn.out_proj.weight"], sd[p + "attn.out_proj.bias"])
    h = F.layer_norm(x, (D,), sd[p + "ln_2.weight"], sd[p + "ln_2.bias"], eps)
    h = F.gelu(F.linear(h, sd[p + fc1 + ".weight"], sd[p + fc1 + ".bias"]))   # exact erf (transformer.py:212)
    return x + F.linear(h, sd[p + fc2 + ".weight"], sd[p + fc2 + ".bias"])


def vision_tokens(sd, patches: torch.Tensor, cfg: ClipCfg) -> torch.Tensor:
    """conv1 + CLS + positional/temporal table (avion/models/transformer.py:319-329) -> [B,S,D] before ln_pre."""
    v = "visual."
    x = F.linear(patches, sd[v + "conv1.weight"])
    B = x.shape[0]
    G = cfg.grid * cfg.grid
    cls = sd[v + "class_embedding"].view(1, 1, -1).expand(B, 1, -1)
    x = torch.cat([cls, x], dim=1)
    pos = sd[v + "positional_embedding"]
    if cfg.frames > 1:
        tile = pos[1:].repeat(cfg.frames, 1) + sd[v + "temporal_embedding"].repeat_interleave(G, 0)
        pos = torch.cat([pos[:1], tile], dim=0)
    return x + pos.unsqueeze(0)


def vision_forward(sd, image: torch.Tensor, cfg: ClipCfg) -> torch.Tensor:
    """VisionTransformer.forward (avion/models/transformer.py:309-378), image f32 [B,3,T,H,W] -> [B,E]."""
    v = "visual."
    if cfg.frames == 1 and image.dim() == 4:
        image = image.unsqueeze(2)
    x = vision_tokens(sd, patchify(image, cfg.patch), cfg)
    D = cfg.width
    x = F.layer_norm(x, (D,), sd[v + "ln_pre.weight"], sd[v + "ln_pre.bias"], 1e-6)     # eps: transformer.py:213
    for i in range(cfg.layers):
        x = _block(x, sd, f"{v}transformer.resblocks.{i}.", cfg.heads, 1e-6,
                   ("attn.Wqkv.weight", "attn.Wqkv.bias", "mlp.fc1", "mlp.fc2"))
    x = F.layer_norm(x[:, 0], (D,), sd[v + "ln_post.weight"], sd[v + "ln_post.bias"], 1e-6)
    if sd.get(v + "image_projection") is None:          # VideoClassifier drops it (model_clip.py:27-28; transformer.py:374)
        return x
    return x @ sd[v + "image_projection"]


def classifier_forward(sd, image: torch.Tensor, cfg: ClipCfg) -> torch.Tensor:
    """VideoClassifier.forward (avion/models/model_clip.py:33-38) with dropout 0: fc_cls(visual(image)) -> [B, classes].
    sd holds visual.* (without image_projection), fc_cls.weight, fc_cls.bias."""
    return F.linear(vision_forward(sd, image, cfg), sd["fc_cls.weight"], sd["fc_cls.bias"])


def make_classifier_state_dict(cfg: ClipCfg, num_classes: int, seed: int = 0) -> dict:
    sd = {k: v for k, v in make_state_dict(cfg, seed).items() if k.startswith("visual.") and k != "visual.image_projection"}
    rs = np.random.RandomState(seed + 1000)
    sd["fc_cls.weight"] = torch.from_numpy((rs.standard_normal((num_classes, cfg.width)) * 0.05).astype(np.float32))
    sd["fc_cls.bias"] = torch.from_numpy((rs.standard_normal(num_classes) * 0.05).astype(np.float32))
    return sd


# --------------------------------------------------------------------------------------------------
# X1: text tower
# --------------------------------------------------------------------------------------------------
def text_forward(sd, text: torch.Tensor, cfg: ClipCfg) -> torch.Tensor:
    """TextTransformer.forward (avion/models/transformer.py:455-478): causal mask :447-453, LayerNorm eps 1e-5
    (:35-41), EOT gather = argmax of the ids (:474)."""
    t = "textual."
    W = cfg.t_width
    x = sd[t + "token_embedding.weight"][text] + sd[t + "positional_embedding"]
    mask = torch.full((cfg.ctx, cfg.ctx), float("-inf")).triu_(1)
    for i in range(cfg.t_layers):
        x = _block(x, sd, f"{t}transformer.resblocks.{i}.", cfg.t_heads, 1e-5,
                   ("attn.in_proj_weight", "attn.in_proj_bias", "mlp.c_fc", "mlp.c_proj"), attn_mask=mask)
    x = F.layer_norm(x, (W,), sd[t + "ln_final.weight"], sd[t + "ln_final.bias"], 1e-5)
    x = x[torch.arange(x.shape[0]), text.argmax(dim=-1)]
    return x @ sd[t + "text_projection"]


# --------------------------------------------------------------------------------------------------
# C1 + L1/L2: CLIP wrapper and ClipLoss
# --------------------------------------------------------------------------------------------------
def clip_forward(sd, image, text, cfg: ClipCfg):
    """CLIP.forward (avion/models/model_clip.py:91-99)."""
    return (F.normalize(vision_forward(sd, image, cfg), dim=-1),
            F.normalize(text_forward(sd, text, cfg), dim=-1),
            sd["logit_scale"].exp())


def clip_loss(img, txt, scale):
    """ClipLoss.forward with world_size == 1 (avion/losses/losses.py:124-149).  For W ranks the default mode
    equals this on the rank-concatenated batch (losses.py:119-123)."""
    logits_i = scale * img @ txt.T
    logits_t = scale * txt @ img.T
    labels = torch.arange(img.shape[0])
    loss = (F.cross_entropy(logits_i, labels) + F.cross_entropy(logits_t, labels)) / 2
    acc = 100 * (logits_i.argmax(-1) == labels).sum() / img.shape[0]
    return {"loss": loss, "clip_acc": acc, "logits_per_image": logits_i}


def clip_loss_local(img, txt, all_img, all_txt, scale, rank):
    """local_loss=True branch (avion/losses/losses.py:116-118,132-133): rank-local rows against all columns."""
    logits_i = scale * img @ all_txt.T
    logits_t = scale * txt @ all_img.T
    n = img.shape[0]
    labels = torch.arange(n) + n * rank
    loss = (F.cross_entropy(logits_i, labels) + F.cross_entropy(logits_t, labels)) / 2
    acc = 100 * (logits_i.argmax(-1) == labels).sum() / n
    return {"loss": loss, "clip_acc": acc}


def clip_loss_world(imgs, txts, scale, local_loss: bool, gather_with_grad: bool):
    """Single-process restatement of ClipLoss over W ranks for every (local_loss, gather_with_grad) mode
    (gather_features, avion/losses/losses.py:53-75, non-horovod branch; loss :103-149).

    imgs, txts: lists (one entry per rank) of leaf tensors [n,E]; scale: leaf scalar.  Returns one dict per rank with that
    rank's `loss`, `clip_acc` and the gradients its backward() leaves on ITS OWN inputs before DDP averages parameters:
    `dimg`, `dtxt` (what reaches image_features / text_features of rank r, including - with gather_with_grad - the
    reduce-scatter of every other rank's contribution, torch.distributed.nn.all_gather backward) and `dscale`
    (rank r's own d loss_r / d scale).  Pinned against the live reference under gloo in tests/test_oracle.py."""
    W = len(imgs)
    n = imgs[0].shape[0]
    outs, losses = [], []
    feat_grads = [[torch.zeros_like(imgs[r]), torch.zeros_like(txts[r])] for r in range(W)]
    for r in range(W):
        if gather_with_grad:                      # :55-60: differentiable gather, gradient flows to every rank's features
            g_img, g_txt = list(imgs), list(txts)
        else:                                     # :62-72: plain all_gather, copies carry no gradient ...
            g_img = [t.detach() for t in imgs]
            g_txt = [t.detach() for t in txts]
            if not local_loss:                    # ... unless the own slice is put back (:68-72)
                g_img[r], g_txt[r] = imgs[r], txts[r]
        all_img, all_txt = torch.cat(g_img), torch.cat(g_txt)
        if local_loss:                            # :116-118
            logits_i = scale * imgs[r] @ all_txt.T
            logits_t = scale * txts[r] @ all_img.T
            labels = torch.arange(n) + n * r      # :132-133
        else:                                     # :119-123
            logits_i = scale * all_img @ all_txt.T
            logits_t = logits_i.T
            labels = torch.arange(W * n)
        loss = (F.cross_entropy(logits_i, labels) + F.cross_entropy(logits_t, labels)) / 2
        acc = 100 * (logits_i.argmax(-1) == labels).sum() / logits_i.shape[0]
        leaves = list(imgs) + list(txts) + [scale]
        grads = torch.autograd.grad(loss, leaves, allow_unused=True)
        for q in range(W):                        # rank r's backward sends these to rank q (all_gather backward sums them)
            if grads[q] is not None:
                feat_grads[q][0] += grads[q]
            if grads[W + q] is not None:
                feat_grads[q][1] += grads[W + q]
        outs.append({"loss": loss.detach(), "clip_acc": acc, "dscale": grads[-1]})
    for r in range(W):
        outs[r]["dimg"], outs[r]["dtxt"] = feat_grads[r]
    return outs


def train_step_forward(sd, clips_u8, tokens, cfg: ClipCfg):
    """The step of scripts/main_lavila_pretrain.py:814-819 (transform -> model -> criterion), fp32."""
    img, txt, scale = clip_forward(sd, clip_transform(clips_u8), tokens, cfg)
    out = clip_loss(img, txt, scale)
    out.update(image_embed=img, text_embed=txt, logit_scale=scale)
    return out


def loss_and_grads(cfg: ClipCfg, sd, clips_u8, tokens, names=None):
    """Forward + backward through the restatement; returns (outputs, {name: grad})."""
    sd = {k: v.clone().requires_grad_(True) for k, v in sd.items()}
    out = train_step_forward(sd, clips_u8, tokens, cfg)
    out["loss"].backward()
    names = names if names is not None else list(sd)
    return out, {k: sd[k].grad for k in names}
