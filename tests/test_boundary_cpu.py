"""Drop-in boundary on CPU: constructors, state_dict keys and shapes equal the reference's flash + fast_conv1 naming."""
import pytest
import torch

from oracle import clip_oracle as O
from oracle import ref_import as R


def test_state_dict_contract_matches_oracle_naming():
    from avion_b200.smoke import build_model
    cfg = O.TINY
    model = build_model(cfg)
    sd = O.make_state_dict(cfg, 0)
    ours = model.state_dict()
    assert set(ours) == set(sd)
    for k in sd:
        assert tuple(ours[k].shape) == tuple(sd[k].shape), k
    model.load_state_dict(sd, strict=True)


@pytest.mark.skipif(not R.available(), reason="reference tree only exists in the build container")
def test_state_dict_contract_matches_live_reference():
    from avion_b200.smoke import build_model
    cfg = O.TINY
    ref = R.build_reference_clip(cfg, O.make_state_dict(cfg, 0))
    ref_keys = {k: tuple(v.shape) for k, v in ref.state_dict().items()}
    ours = {R.to_reference_key(k): tuple(v.shape) for k, v in build_model(cfg).state_dict().items()}
    assert ours == ref_keys


def test_factory_signature_and_param_count():
    from avion_b200.models import CLIP_VITB16
    m = CLIP_VITB16(num_frames=16, project_embed_dim=256, use_fast_conv1=True, use_flash_attn=True, freeze_temperature=False,
                    pretrain_zoo="random", use_quick_gelu=True)          # swallowed by **kwargs like the reference (model_clip.py:115)
    n = sum(p.numel() for p in m.parameters())
    assert abs(n - 149.3e6) < 0.2e6
    assert m.visual.temporal_embedding.shape == (16, 768)
    assert float(m.logit_scale.exp()) == pytest.approx(1 / 0.07, rel=1e-5)


def test_factory_loads_open_clip_checkpoint_through_the_remap(tmp_path):
    """avion/models/model_clip.py:150-174: pretrain_zoo='open_clip' / 'openai' run the checkpoint through
    remap_keys_from_open_clip_to_vit (Conv2d conv1 -> Linear, in_proj -> Wqkv, text keys under textual.) before load_state_dict;
    the default zoo is 'openai' like the reference's, which without a local path has nothing to load from (no network)."""
    from avion_b200.models import model_clip as MC
    from avion_b200.models.transformer import TextTransformer, VisionTransformer
    D, TW, E, L = 64, 64, 32, 2
    g = torch.Generator().manual_seed(0)
    r = lambda *s: torch.randn(*s, generator=g)      # noqa: E731
    sd = {"logit_scale": r(()), "visual.proj": r(D, E), "positional_embedding": r(77, TW), "text_projection": r(TW, E),
          "token_embedding.weight": r(100, TW), "ln_final.weight": r(TW), "ln_final.bias": r(TW),
          "visual.conv1.weight": r(D, 3, 16, 16), "visual.class_embedding": r(D), "visual.positional_embedding": r(5, D),
          "visual.ln_pre.weight": r(D), "visual.ln_pre.bias": r(D), "visual.ln_post.weight": r(D), "visual.ln_post.bias": r(D)}
    for i in range(L):
        for tower, W in (("visual.", D), ("", TW)):
            for n, shape in (("attn.in_proj_weight", (3 * W, W)), ("attn.in_proj_bias", (3 * W,)), ("attn.out_proj.weight", (W, W)),
                             ("attn.out_proj.bias", (W,)), ("ln_1.weight", (W,)), ("ln_1.bias", (W,)), ("ln_2.weight", (W,)),
                             ("ln_2.bias", (W,)), ("mlp.c_fc.weight", (4 * W, W)), ("mlp.c_fc.bias", (4 * W,)),
                             ("mlp.c_proj.weight", (W, 4 * W)), ("mlp.c_proj.bias", (W,))):
                sd[f"{tower}transformer.resblocks.{i}.{n}"] = r(*shape)
    path = tmp_path / "open_clip.pt"
    torch.save({"state_dict": {"module." + k: v for k, v in sd.items()}}, path)
    vis = dict(image_size=32, patch_size=16, width=D, layers=L, heads=1, mlp_ratio=4, num_frames=1)
    txt = dict(context_length=77, vocab_size=100, width=TW, heads=1, layers=L)
    model = MC._build(vis, txt, E, False, False, "open_clip", str(path))
    assert torch.equal(model.visual.conv1.weight, sd["visual.conv1.weight"].flatten(1))
    assert torch.equal(model.visual.transformer.resblocks[1].attn.Wqkv.weight, sd["visual.transformer.resblocks.1.attn.in_proj_weight"])
    assert torch.equal(model.visual.transformer.resblocks[0].mlp.fc2.bias, sd["visual.transformer.resblocks.0.mlp.c_proj.bias"])
    assert torch.equal(model.textual.transformer.resblocks[1].attn.in_proj_weight, sd["transformer.resblocks.1.attn.in_proj_weight"])
    assert torch.equal(model.textual.token_embedding.weight, sd["token_embedding.weight"])
    assert not torch.equal(model.visual.image_projection, sd["visual.proj"])      # projections are re-initialised (utils.py:62-63)
    with pytest.raises(NotImplementedError):
        MC._build(vis, txt, E, False, False, "openai", None)
    import inspect
    assert inspect.signature(MC.CLIP_VITB16).parameters["pretrain_zoo"].default == "openai"
    assert inspect.signature(MC.CLIP_VITL14).parameters["pretrain_zoo"].default == "openai"
