"""CPU oracle for the VideoMAE pretraining step (BASELINE.json configs[4]).  TEST INFRASTRUCTURE ONLY.

Plain fp32 PyTorch restatement of avion/models/model_videomae.py and the train step of
scripts/main_videomae_pretrain.py:316-341.  Only ``tests/``, ``__graft_entry__.smoke()`` and the ``cpu_baseline`` /
``--impl reference`` legs of ``bench.py`` may import it; the product package ``avion_b200`` never does.

Parity status: PINNED by execution of the reference itself - ``oracle/make_golden_videomae.py`` imports the unmodified
``PretrainVisionTransformer`` from /root/reference (non-flash branch, fp32, CPU), runs it with ``torch.nn.MSELoss`` on seeded
inputs and commits outputs, loss and gradients as ``tests/golden/videomae_tiny_b4.npz``; ``tests/test_oracle_videomae.py``
checks this restatement against that fixture (and against the live reference when /root/reference is present).
kornia is absent: ``K.enhance.Normalize`` is restated as the per-channel (x - mean) / std.
"""
from __future__ import annotations

from dataclasses import dataclass, asdict

import numpy as np
import torch
import torch.nn.functional as F

MEAN_TIMM = (0.485 * 255, 0.456 * 255, 0.406 * 255)      # scripts/main_videomae_pretrain.py:205
STD_TIMM = (0.229 * 255, 0.224 * 255, 0.225 * 255)


@dataclass(frozen=True)
class MaeCfg:
    """Shapes of VIDEOMAE_VITB16 (model_videomae.py:583-596) with --decoder-depth 4 (main_videomae_pretrain.py:51)."""
    image_size: int = 224
    patch: int = 16
    frames: int = 16
    enc_dim: int = 768
    enc_depth: int = 12
    enc_heads: int = 12
    dec_dim: int = 384
    dec_depth: int = 4
    dec_heads: int = 6
    masked_per_frame: int = 176          # int(0.9 * 196), transforms.py:285

    @property
    def grid(self):
        return self.image_size // self.patch

    @property
    def tokens(self):
        return (self.frames // 2) * self.grid * self.grid

    @property
    def feats(self):
        return 3 * 2 * self.patch * self.patch

    def asdict(self):
        return asdict(self)


TINY = MaeCfg(image_size=48, patch=16, frames=16, enc_dim=128, enc_depth=2, enc_heads=2, dec_dim=64, dec_depth=1, dec_heads=1,
              masked_per_frame=6)
VITB16 = MaeCfg()


def sinusoid_table(n_position, d_hid):
    """model_videomae.py:177-187."""
    table = np.array([[pos / np.power(10000, 2 * (j // 2) / d_hid) for j in range(d_hid)] for pos in range(n_position)])
    table[:, 0::2] = np.sin(table[:, 0::2])
    table[:, 1::2] = np.cos(table[:, 1::2])
    return torch.tensor(table, dtype=torch.float)


def make_state_dict(cfg: MaeCfg, seed: int = 0) -> dict:
    """Seeded weights under the reference's keys (model_videomae.py: Block :104-127, Attention :57-79, PatchEmbed :139-157,
    PretrainVisionTransformer :509-548).  All biases (incl. q_bias / v_bias) and mask_token are non-zero so that bias and
    ordering bugs cannot hide."""
    rs = np.random.RandomState(seed)

    def n(*shape, std=0.02):
        return torch.from_numpy((rs.standard_normal(shape) * std).astype(np.float32))

    sd = {}

    def blocks(prefix, D, depth):
        for i in range(depth):
            p = f"{prefix}blocks.{i}."
            for ln in ("norm1", "norm2"):
                sd[p + ln + ".weight"] = 1.0 + n(D, std=0.1)
                sd[p + ln + ".bias"] = n(D, std=0.1)
            sd[p + "attn.q_bias"] = n(D)
            sd[p + "attn.v_bias"] = n(D)
            sd[p + "attn.qkv.weight"] = n(3 * D, D, std=D ** -0.5)
            sd[p + "attn.proj.weight"] = n(D, D, std=D ** -0.5)
            sd[p + "attn.proj.bias"] = n(D)
            sd[p + "mlp.fc1.weight"] = n(4 * D, D, std=D ** -0.5)
            sd[p + "mlp.fc1.bias"] = n(4 * D)
            sd[p + "mlp.fc2.weight"] = n(D, 4 * D, std=(4 * D) ** -0.5)
            sd[p + "mlp.fc2.bias"] = n(D)
        sd[prefix + "norm.weight"] = 1.0 + n(D, std=0.1)
        sd[prefix + "norm.bias"] = n(D, std=0.1)

    sd["mask_token"] = n(1, 1, cfg.dec_dim)
    sd["encoder.patch_embed.proj.weight"] = n(cfg.enc_dim, 3, 2, cfg.patch, cfg.patch, std=cfg.feats ** -0.5)
    sd["encoder.patch_embed.proj.bias"] = n(cfg.enc_dim)
    blocks("encoder.", cfg.enc_dim, cfg.enc_depth)
    blocks("decoder.", cfg.dec_dim, cfg.dec_depth)
    sd["decoder.head.weight"] = n(cfg.feats, cfg.dec_dim, std=cfg.dec_dim ** -0.5)
    sd["decoder.head.bias"] = n(cfg.feats)
    sd["encoder_to_decoder.weight"] = n(cfg.dec_dim, cfg.enc_dim, std=cfg.enc_dim ** -0.5)
    return sd


def tube_mask(batch, cfg: MaeCfg, generator=None):
    """TubeMaskingGeneratorGPU.__call__ (avion/data/transforms.py:296-302) -> bool [B, N]."""
    ppf = cfg.grid * cfg.grid
    tensors = torch.zeros((batch, ppf))
    indexes = torch.argsort(torch.rand(batch, ppf, generator=generator))[:, :cfg.masked_per_frame]
    tensors[torch.arange(batch).unsqueeze(1), indexes] = 1
    return tensors.repeat(1, cfg.frames // 2).to(torch.bool)


def make_inputs(cfg: MaeCfg, batch: int, seed: int = 1234):
    """uint8 clips [B,T,H,W,3] (decord layout) with smooth + noisy content (non-trivial patch statistics), and a tube mask."""
    rs = np.random.RandomState(seed)
    base = rs.randint(0, 256, (batch, cfg.frames, 1, 1, 3))
    noise = rs.randint(-60, 61, (batch, cfg.frames, cfg.image_size, cfg.image_size, 3))
    clips = torch.from_numpy(np.clip(base + noise, 0, 255).astype(np.uint8))
    g = torch.Generator().manual_seed(seed)
    return clips, tube_mask(batch, cfg, g)


def raw_video(clips_u8):
    """inputs.permute(0, 4, 1, 2, 3) of the fused-decode branch (main_videomae_pretrain.py:317-318) as fp32."""
    return clips_u8.permute(0, 4, 1, 2, 3).float()


def normalize(video, mean=MEAN_TIMM, std=STD_TIMM):
    """kornia Normalize (main_videomae_pretrain.py:206,336): per-channel (x - mean) / std on [B,3,T,H,W]."""
    m = torch.tensor(mean, dtype=torch.float32).view(1, 3, 1, 1, 1)
    s = torch.tensor(std, dtype=torch.float32).view(1, 3, 1, 1, 1)
    return (video - m) / s


def _split(video, p):
    B, C, T, H, W = video.shape
    return video.reshape(B, C, T // 2, 2, H // p, p, W // p, p)      # b c t p0 h p1 w p2


def tubelet_patches(video, p):
    """'b c (t p0) (h p1) (w p2) -> b (t h w) (c p0 p1 p2)' (model_videomae.py:161) = Conv3d weight flattening (:153-155)."""
    x = _split(video, p).permute(0, 2, 4, 6, 1, 3, 5, 7)
    B = x.shape[0]
    return x.reshape(B, x.shape[1] * x.shape[2] * x.shape[3], -1)


def targets(video_raw, mask, p):
    """main_videomae_pretrain.py:322-334 (normalize_target=True): per patch and channel standardisation of the RAW pixels
    with the unbiased variance, eps on the std, feature order (p0 p1 p2 c); masked tokens only."""
    x = _split(video_raw, p).permute(0, 2, 4, 6, 3, 5, 7, 1)          # b t h w p0 p1 p2 c
    B, C = x.shape[0], x.shape[-1]
    sq = x.reshape(B, x.shape[1] * x.shape[2] * x.shape[3], -1, C)    # b n (p0 p1 p2) c
    norm = (sq - sq.mean(dim=-2, keepdim=True)) / (sq.var(dim=-2, unbiased=True, keepdim=True).sqrt() + 1e-6)
    patch = norm.reshape(B, sq.shape[1], -1)                          # b n (p c)
    return patch[mask].reshape(B, -1, patch.shape[-1])


def _block(x, sd, p, heads, eps):
    """Block.forward / Attention.forward / Mlp.forward (model_videomae.py:129-136, 81-101, 47-54)."""
    B, N, D = x.shape
    h = F.layer_norm(x, (D,), sd[p + "norm1.weight"], sd[p + "norm1.bias"], eps)
    bias = torch.cat((sd[p + "attn.q_bias"], torch.zeros_like(sd[p + "attn.v_bias"]), sd[p + "attn.v_bias"]))
    qkv = F.linear(h, sd[p + "attn.qkv.weight"], bias).reshape(B, N, 3, heads, -1).permute(2, 0, 3, 1, 4)
    q, k, v = qkv[0] * (D // heads) ** -0.5, qkv[1], qkv[2]
    a = (q @ k.transpose(-2, -1)).softmax(dim=-1)
    o = (a @ v).transpose(1, 2).reshape(B, N, -1)
    x = x + F.linear(o, sd[p + "attn.proj.weight"], sd[p + "attn.proj.bias"])
    h = F.layer_norm(x, (D,), sd[p + "norm2.weight"], sd[p + "norm2.bias"], eps)
    h = F.gelu(F.linear(h, sd[p + "mlp.fc1.weight"], sd[p + "mlp.fc1.bias"]))
    return x + F.linear(h, sd[p + "mlp.fc2.weight"], sd[p + "mlp.fc2.bias"])


def forward(sd, video, mask, cfg: MaeCfg, eps=1e-6):
    """PretrainVisionTransformer.forward (model_videomae.py:567-580) on a normalised video [B,3,T,H,W] -> [B, N_mask, F]."""
    B = video.shape[0]
    w = sd["encoder.patch_embed.proj.weight"]
    x = F.linear(tubelet_patches(video, cfg.patch), w.reshape(w.shape[0], -1), sd["encoder.patch_embed.proj.bias"])
    x = x + sinusoid_table(cfg.tokens, cfg.enc_dim)[None]                     # :379
    x = x[~mask].reshape(B, -1, cfg.enc_dim)                                  # :383
    for i in range(cfg.enc_depth):
        x = _block(x, sd, f"encoder.blocks.{i}.", cfg.enc_heads, eps)
    x = F.layer_norm(x, (cfg.enc_dim,), sd["encoder.norm.weight"], sd["encoder.norm.bias"], eps)
    x = F.linear(x, sd["encoder_to_decoder.weight"])                          # :570
    pos = sinusoid_table(cfg.tokens, cfg.dec_dim)[None].expand(B, -1, -1)
    pos_vis = pos[~mask].reshape(B, -1, cfg.dec_dim)
    pos_mask = pos[mask].reshape(B, -1, cfg.dec_dim)
    x = torch.cat([x + pos_vis, sd["mask_token"] + pos_mask], dim=1)          # :577
    for i in range(cfg.dec_depth):
        x = _block(x, sd, f"decoder.blocks.{i}.", cfg.dec_heads, eps)
    x = x[:, -pos_mask.shape[1]:]                                             # :461
    x = F.layer_norm(x, (cfg.dec_dim,), sd["decoder.norm.weight"], sd["decoder.norm.bias"], eps)
    return F.linear(x, sd["decoder.head.weight"], sd["decoder.head.bias"])


def train_step_forward(sd, clips_u8, mask, cfg: MaeCfg):
    """main_videomae_pretrain.py:316-341 -> dict(outputs, targets, loss)."""
    raw = raw_video(clips_u8)
    tgt = targets(raw, mask, cfg.patch)
    out = forward(sd, normalize(raw), mask, cfg)
    return dict(outputs=out, targets=tgt, loss=F.mse_loss(out, tgt))


def loss_and_grads(cfg: MaeCfg, sd, clips_u8, mask, names=None):
    leaf = {k: v.clone().requires_grad_(True) for k, v in sd.items()}
    out = train_step_forward(leaf, clips_u8, mask, cfg)
    out["loss"].backward()
    names = names if names is not None else list(leaf)
    return out, {k: leaf[k].grad.detach() for k in names}
