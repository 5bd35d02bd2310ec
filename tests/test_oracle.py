"""The oracle restatement vs the golden vectors produced by the unmodified reference (oracle/make_golden.py),
and vs the live reference when /root/reference is present."""
import os

import numpy as np
import pytest
import torch

from oracle import clip_oracle as O
from oracle import ref_import as R

GOLDEN = os.path.join(os.path.dirname(__file__), "golden")


def load_golden(name):
    g = np.load(os.path.join(GOLDEN, name + ".npz"))
    cfg = O.ClipCfg(*[int(v) for v in g["cfg"]])
    return g, cfg


@pytest.mark.parametrize("name", ["tiny_b6", "tiny_b32", "tiny14_b32"])
def test_oracle_matches_golden_fwd_bwd(name):
    g, cfg = load_golden(name)
    sd = O.make_state_dict(cfg, int(g["wseed"]))
    clips, tokens = O.make_inputs(cfg, int(g["batch"]), int(g["iseed"]))
    keys = [k[5:] for k in g.files if k.startswith("grad:")]
    out, grads = O.loss_and_grads(cfg, sd, clips, tokens, keys)
    np.testing.assert_allclose(out["image_embed"].detach().numpy(), g["image_embed"], rtol=1e-4, atol=2e-6)
    np.testing.assert_allclose(out["text_embed"].detach().numpy(), g["text_embed"], rtol=1e-4, atol=2e-6)
    np.testing.assert_allclose(float(out["loss"]), float(g["loss"]), rtol=1e-5)
    assert float(out["clip_acc"]) == float(g["clip_acc"])
    assert keys
    for k in keys:
        ref = g["grad:" + k]
        got = grads[k].numpy().reshape(-1)[:ref.size].reshape(ref.shape)
        np.testing.assert_allclose(got, ref, rtol=2e-3, atol=1e-6 + 1e-4 * np.abs(ref).max(), err_msg=k)


def test_oracle_patch_order_bit_exact():
    g, cfg = load_golden("tiny_b6")
    clips, _ = O.make_inputs(cfg, int(g["batch"]), int(g["iseed"]))
    p = O.patchify(O.clip_transform(clips), cfg.patch)[:2, :5, :40].numpy()
    assert np.array_equal(p, g["patches_probe"])


@pytest.mark.slow
def test_oracle_matches_golden_cfg1():
    """BASELINE.json configs[0]: ViT-B/16 4f + 12L/512 text, batch 8, forward + ClipLoss."""
    g, cfg = load_golden("vitb16_4f_b8")
    sd = O.make_state_dict(cfg, int(g["wseed"]))
    clips, tokens = O.make_inputs(cfg, int(g["batch"]), int(g["iseed"]))
    with torch.no_grad():
        out = O.train_step_forward(sd, clips, tokens, cfg)
    np.testing.assert_allclose(out["image_embed"].numpy(), g["image_embed"], rtol=1e-3, atol=1e-5)
    np.testing.assert_allclose(out["text_embed"].numpy(), g["text_embed"], rtol=1e-3, atol=1e-5)
    np.testing.assert_allclose(float(out["loss"]), float(g["loss"]), rtol=1e-5)


@pytest.mark.skipif(not R.available(), reason="reference tree only exists in the build container")
def test_oracle_matches_live_reference():
    cfg = O.TINY
    sd = O.make_state_dict(cfg, 7)
    clips, tokens = O.make_inputs(cfg, 5, 77)
    model = R.build_reference_clip(cfg, sd)
    _, _, _, ClipLoss = R.import_reference()
    with torch.no_grad():
        img, txt, scale = model(O.clip_transform(clips), tokens)
        ref = ClipLoss(world_size=1)(img, txt, scale)
        out = O.train_step_forward(sd, clips, tokens, cfg)
    np.testing.assert_allclose(out["image_embed"].numpy(), img.numpy(), rtol=1e-4, atol=2e-6)
    np.testing.assert_allclose(out["text_embed"].numpy(), txt.numpy(), rtol=1e-4, atol=2e-6)
    np.testing.assert_allclose(float(out["loss"]), float(ref["loss"]), rtol=1e-5)
