"""One tiny train step (transform -> towers -> ClipLoss -> backward) on cuda:0, checked against the CPU oracle."""
import os
import sys

import torch


def run():
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    if root not in sys.path:
        sys.path.insert(0, root)
    from oracle import clip_oracle as O            # checker only
    from avion_b200.models.model_clip import CLIP
    from avion_b200.models.transformer import TextTransformer, VisionTransformer
    from avion_b200.losses import ClipLoss

    assert torch.cuda.is_available(), "smoke() needs a CUDA device"
    cfg, B = O.TINY, 6
    sd = O.make_state_dict(cfg, 0)
    clips, tokens = O.make_inputs(cfg, B, 1234)
    model = build_model(cfg).cuda()
    model.load_state_dict(sd, strict=True)
    crit = ClipLoss(world_size=1)
    img, txt, scale = model(clips.cuda(), tokens.cuda())
    out = crit(img, txt, scale)
    out["loss"].backward()
    torch.cuda.synchronize()
    ref = O.train_step_forward(sd, clips, tokens, cfg)
    cos_i = torch.nn.functional.cosine_similarity(img.detach().cpu(), ref["image_embed"], dim=-1).min().item()
    cos_t = torch.nn.functional.cosine_similarity(txt.detach().cpu(), ref["text_embed"], dim=-1).min().item()
    rel = abs(out["loss"].item() - ref["loss"].item()) / abs(ref["loss"].item())
    gn = model.visual.conv1.weight.grad.norm().item()
    print(f"smoke: loss={out['loss'].item():.5f} oracle={ref['loss'].item():.5f} rel={rel:.2e} cos_img={cos_i:.5f} cos_txt={cos_t:.5f} "
          f"|dconv1|={gn:.3e}")
    assert cos_i >= 0.999 and cos_t >= 0.999, "embedding cosine below 0.999"
    assert rel < 1e-3 * 5, "loss deviates from the oracle"
    assert gn > 0 and gn == gn


def build_model(cfg):
    from avion_b200.models.model_clip import CLIP
    from avion_b200.models.transformer import TextTransformer, VisionTransformer
    vis = VisionTransformer(cfg.image_size, cfg.patch, cfg.width, cfg.layers, cfg.heads, 4, num_frames=cfg.frames,
                            output_dim=cfg.embed_dim)
    txt = TextTransformer(context_length=cfg.ctx, vocab_size=cfg.vocab, width=cfg.t_width, heads=cfg.t_heads,
                          layers=cfg.t_layers, output_dim=cfg.embed_dim)
    return CLIP(cfg.embed_dim, vis, txt)
