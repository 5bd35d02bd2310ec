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
    cfg, B = O.TINY, 32
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
    assert rel < 1e-3, "loss deviates from the oracle by more than 1e-3 (north-star tolerance)"
    assert gn > 0 and gn == gn
    run_videomae()


def run_videomae():
    """One tiny VideoMAE pretraining step (BASELINE.json configs[4]) on cuda:0, checked against the CPU oracle."""
    from functools import partial
    from oracle import videomae_oracle as V        # checker only
    from avion_b200.losses import MSELoss
    from avion_b200.models.model_videomae import PretrainVisionTransformer
    cfg = V.TINY
    sd = V.make_state_dict(cfg, 0)
    clips, mask = V.make_inputs(cfg, 4, 1234)
    model = PretrainVisionTransformer(
        img_size=cfg.image_size, patch_size=cfg.patch, encoder_embed_dim=cfg.enc_dim, encoder_depth=cfg.enc_depth,
        encoder_num_heads=cfg.enc_heads, decoder_num_classes=cfg.feats, decoder_embed_dim=cfg.dec_dim,
        decoder_num_heads=cfg.dec_heads, decoder_depth=cfg.dec_depth, mlp_ratio=4, qkv_bias=True,
        norm_layer=partial(torch.nn.LayerNorm, eps=1e-6), num_frames=cfg.frames).cuda()
    model.load_state_dict(sd, strict=True)
    out, tgt = model(clips.cuda(), mask.cuda(), return_targets=True)
    loss = MSELoss()(out, tgt)
    loss.backward()
    torch.cuda.synchronize()
    ref = V.train_step_forward(sd, clips, mask, cfg)
    cos = torch.nn.functional.cosine_similarity(out.detach().float().cpu(), ref["outputs"], dim=-1).min().item()
    rel = abs(loss.item() - ref["loss"].item()) / abs(ref["loss"].item())
    terr = (tgt.cpu() - ref["targets"]).abs().max().item()
    print(f"smoke videomae: loss={loss.item():.5f} oracle={ref['loss'].item():.5f} rel={rel:.2e} cos_pred={cos:.5f} target_err={terr:.1e}")
    assert cos >= 0.999 and rel < 1e-3 and terr <= 2e-5


def build_model(cfg):
    from avion_b200.models.model_clip import CLIP
    from avion_b200.models.transformer import TextTransformer, VisionTransformer
    vis = VisionTransformer(cfg.image_size, cfg.patch, cfg.width, cfg.layers, cfg.heads, 4, num_frames=cfg.frames,
                            output_dim=cfg.embed_dim)
    txt = TextTransformer(context_length=cfg.ctx, vocab_size=cfg.vocab, width=cfg.t_width, heads=cfg.t_heads,
                          layers=cfg.t_layers, output_dim=cfg.embed_dim)
    return CLIP(cfg.embed_dim, vis, txt)
