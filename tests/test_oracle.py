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


# ---- ClipLoss over ranks: every (local_loss, gather_with_grad) mode of gather_features (losses.py:53-75) -------------------
def test_clip_loss_world_matches_reference_golden():
    """oracle.clip_loss_world vs tests/golden/cliploss_modes_w2.npz = the UNMODIFIED reference ClipLoss run in two gloo
    processes (oracle/make_golden_cliploss.py): loss, accuracy, feature gradients and logit-scale gradient per rank and mode."""
    from oracle import cliploss_modes as M
    g, n, E, world, seed, scale = M.load_golden()
    img, txt = M.features(n, E, world, seed)
    for ll, gwg in M.MODES:
        outs = M.oracle_world(img, txt, scale, n, world, ll, gwg)
        for r in range(world):
            for k in ("loss", "clip_acc", "dimg", "dtxt", "dscale"):
                np.testing.assert_allclose(outs[r][k].detach().numpy(), g[f"r{r}:{int(ll)}{int(gwg)}:{k}"], rtol=1e-5, atol=1e-6,
                                           err_msg=f"mode local_loss={ll} gather_with_grad={gwg} rank {r} {k}")
    # the two local_loss=False modes return the same loss on both ranks; the local_loss modes do not
    assert float(g["r0:00:loss"]) == float(g["r1:00:loss"]) and float(g["r0:10:loss"]) != float(g["r1:10:loss"])
    # local_loss without gather_with_grad: only the row side flows, so the gradient differs from the gather_with_grad one
    assert not np.allclose(g["r0:10:dimg"], g["r0:11:dimg"])


@pytest.mark.skipif(not R.available(), reason="reference tree only exists in the build container")
def test_clip_loss_world_matches_live_reference_gloo():
    """Same comparison against the reference executed now (two gloo processes), not against the committed vectors."""
    from oracle import make_golden_cliploss as G
    from oracle import cliploss_modes as M
    got = G.run_reference_world(port=29933 + os.getpid() % 50)
    img, txt = M.features(G.N_LOCAL, G.E, G.WORLD, G.SEED)
    for ll, gwg in M.MODES:
        outs = M.oracle_world(img, txt, G.SCALE, G.N_LOCAL, G.WORLD, ll, gwg)
        for r in range(G.WORLD):
            for k in ("loss", "clip_acc", "dimg", "dtxt", "dscale"):
                np.testing.assert_allclose(outs[r][k].detach().numpy(), got[r][f"{int(ll)}{int(gwg)}:{k}"], rtol=1e-5, atol=1e-6)


# ---- T0/T1 with resize: the "Resized" half of RandomResizedCrop ----------------------------------------------------------
def test_resized_crop_restatement_matches_torch_and_torchvision():
    """oracle.resized_crop_clip (own restatement of the antialiased-bilinear weights) vs torch's float kernel (equal up to fp32
    summation order: identical after rounding except at exact .5 ties) and vs torchvision's uint8 resized_crop (its Pillow-style
    fixed-point path differs from the float path by at most one grey level)."""
    import torch.nn.functional as F
    from torchvision.transforms.v2 import functional as TF
    rs = np.random.RandomState(3)
    clips = torch.from_numpy(rs.randint(0, 256, (3, 2, 96, 120, 3)).astype(np.uint8))
    boxes = torch.tensor([[7, 5, 100, 80, 0], [0, 0, 120, 96, 1], [30, 20, 41, 37, 0]], dtype=torch.int32)   # down, down+flip, up
    got = O.resized_crop_clip(clips, boxes, (64, 64))
    for b in range(3):
        x, y, w, h, flip = (int(v) for v in boxes[b])
        crop = clips[b, :, y:y + h, x:x + w, :].permute(0, 3, 1, 2)          # [T,3,h,w]
        if flip:
            crop = torch.flip(crop, dims=[3])
        ref_f = F.interpolate(crop.float(), size=(64, 64), mode="bilinear", antialias=True, align_corners=False)
        ref_u8 = ref_f.round().clamp(0, 255).to(torch.uint8).permute(0, 2, 3, 1)
        d = (got[b].int() - ref_u8.int()).abs()
        assert d.max() <= 1 and (d > 0).float().mean() < 1e-3, (b, int(d.max()), float((d > 0).float().mean()))
        tv = TF.resized_crop(clips[b].permute(0, 3, 1, 2), y, x, h, w, [64, 64], antialias=True)
        if flip:
            tv = torch.flip(tv, dims=[3])
        assert (got[b].int() - tv.permute(0, 2, 3, 1).int()).abs().max() <= 1
    # clamping of the sampler's off-by-one offsets (x + w == Ws + 1)
    far = O.resized_crop_clip(clips[:1], torch.tensor([[21, 17, 100, 80, 0]]), (64, 64))
    assert torch.equal(far, O.resized_crop_clip(clips[:1], torch.tensor([[20, 16, 100, 80, 0]]), (64, 64)))


def test_bf16_point_oracle_without_points_is_the_fp32_oracle():
    """oracle.bf16_points.Emu with no rounding point enabled is the fp32 oracle (to round-off): the rounding points are the only
    thing it adds.  With all points it moves by the bf16 noise of the pipeline - for the tiny batch-6 fixture by more than the
    1e-3 loss tolerance, which is the measured noise floor any bf16 implementation of this fixture lives with."""
    from oracle.bf16_points import ALL_POINTS, Emu
    g, cfg = load_golden("tiny_b6")
    sd = O.make_state_dict(cfg, int(g["wseed"]))
    clips, tokens = O.make_inputs(cfg, int(g["batch"]), int(g["iseed"]))
    with torch.no_grad():
        loss, img, txt = Emu(()).loss(sd, clips, tokens, cfg)
        np.testing.assert_allclose(loss, float(g["loss"]), rtol=2e-6)
        np.testing.assert_allclose(img.numpy(), g["image_embed"], rtol=1e-4, atol=2e-6)
        noisy, img16, _ = Emu(ALL_POINTS).loss(sd, clips, tokens, cfg)
    assert 1e-4 < abs(noisy - float(g["loss"])) / float(g["loss"]) < 1e-2
    assert torch.nn.functional.cosine_similarity(img16, torch.from_numpy(g["image_embed"]), dim=-1).min() > 0.9995
