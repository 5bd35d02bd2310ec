"""VideoMAE oracle restatement vs the golden vectors of the unmodified reference (oracle/make_golden_videomae.py), and the
host-side pieces of avion_b200.models.model_videomae that need no GPU (state_dict keys, position table, mask generator)."""
import os

import numpy as np
import pytest
import torch

from oracle import ref_import as R
from oracle import videomae_oracle as O

GOLDEN = os.path.join(os.path.dirname(__file__), "golden")


def load_golden(name="videomae_tiny_b4"):
    g = np.load(os.path.join(GOLDEN, name + ".npz"))
    return g, O.MaeCfg(*[int(v) for v in g["cfg"]])


def test_oracle_matches_golden_fwd_bwd():
    g, cfg = load_golden()
    sd = O.make_state_dict(cfg, int(g["wseed"]))
    clips, mask = O.make_inputs(cfg, int(g["batch"]), int(g["iseed"]))
    assert np.array_equal(mask.numpy(), g["mask"])
    keys = [k[5:] for k in g.files if k.startswith("grad:")]
    out, grads = O.loss_and_grads(cfg, sd, clips, mask, keys)
    np.testing.assert_allclose(out["targets"].numpy()[:, :12], g["targets"], rtol=1e-5, atol=1e-6)
    np.testing.assert_allclose(out["outputs"].detach().numpy()[:, :12], g["outputs"], rtol=1e-4, atol=1e-5)
    np.testing.assert_allclose(float(out["loss"]), float(g["loss"]), rtol=1e-5)
    assert len(keys) >= 16
    for k in keys:
        ref = g["grad:" + k]
        got = grads[k].numpy().reshape(-1)[:ref.size].reshape(ref.shape)
        np.testing.assert_allclose(got, ref, rtol=2e-3, atol=1e-6 + 1e-4 * np.abs(ref).max(), err_msg=k)


def test_mask_is_a_tube_with_fixed_ratio():
    cfg = O.VITB16
    m = O.tube_mask(3, cfg, torch.Generator().manual_seed(5)).view(3, cfg.frames // 2, -1)
    assert (m == m[:, :1]).all()                       # the same spatial mask at every tubelet time step
    assert (m[:, 0].sum(-1) == 176).all()              # int(0.9 * 196)


@pytest.mark.skipif(not R.available(), reason="reference tree only exists in the build container")
def test_oracle_matches_live_reference():
    cfg = O.TINY
    sd = O.make_state_dict(cfg, 11)
    clips, mask = O.make_inputs(cfg, 3, 99)
    model = R.build_reference_videomae(cfg, sd)
    with torch.no_grad():
        ref = model(O.normalize(O.raw_video(clips)), mask)
        out = O.train_step_forward(sd, clips, mask, cfg)
    np.testing.assert_allclose(out["outputs"].numpy(), ref.numpy(), rtol=1e-4, atol=1e-5)


def test_host_model_matches_reference_contract():
    """state_dict keys / shapes of the B200 model = the oracle's (reference-named) state dict; same position table; the mask
    generator reproduces the reference's RNG stream."""
    from avion_b200.data.transforms import TubeMaskingGeneratorGPU
    from avion_b200.models import model_videomae as M
    from functools import partial
    cfg = O.TINY
    model = M.PretrainVisionTransformer(
        img_size=cfg.image_size, patch_size=cfg.patch, encoder_embed_dim=cfg.enc_dim, encoder_depth=cfg.enc_depth,
        encoder_num_heads=cfg.enc_heads, decoder_num_classes=cfg.feats, decoder_embed_dim=cfg.dec_dim,
        decoder_num_heads=cfg.dec_heads, decoder_depth=cfg.dec_depth, mlp_ratio=4, qkv_bias=True,
        norm_layer=partial(torch.nn.LayerNorm, eps=1e-6))
    sd = O.make_state_dict(cfg, 0)
    ours = model.state_dict()
    assert set(ours) == set(sd)
    assert all(ours[k].shape == sd[k].shape for k in sd)
    model.load_state_dict(sd, strict=True)
    assert model.eps == 1e-6
    assert torch.equal(model.pos_embed[0], O.sinusoid_table(cfg.tokens, cfg.dec_dim))
    assert torch.equal(model.encoder.pos_embed[0], O.sinusoid_table(cfg.tokens, cfg.enc_dim))
    big = M.VIDEOMAE_VITB16(decoder_depth=4)
    n = sum(p.numel() for p in big.parameters())
    assert len(big.decoder.blocks) == 4 and len(big.encoder.blocks) == 12 and 93e6 < n < 96e6
    torch.manual_seed(3)
    a = TubeMaskingGeneratorGPU(5, (8, 14, 14), 0.9, device="cpu")()
    b = O.tube_mask(5, O.VITB16, torch.Generator().manual_seed(3))      # same stream as the global generator at seed 3
    assert a.dtype == torch.bool and a.shape == (5, 1568) and torch.equal(a, b)
    with pytest.raises(RuntimeError):                                   # no CPU fallback
        model(torch.zeros(1, 16, 48, 48, 3, dtype=torch.uint8), a[:1, :72])
