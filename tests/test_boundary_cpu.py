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
                    use_quick_gelu=True)          # swallowed by **kwargs like the reference (model_clip.py:115)
    n = sum(p.numel() for p in m.parameters())
    assert abs(n - 149.3e6) < 0.2e6
    assert m.visual.temporal_embedding.shape == (16, 768)
    assert float(m.logit_scale.exp()) == pytest.approx(1 / 0.07, rel=1e-5)
