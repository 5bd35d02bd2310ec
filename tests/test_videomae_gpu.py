"""VideoMAE pretraining step (BASELINE.json configs[4]) on the CUDA path vs the committed golden vectors of the unmodified
reference and vs the oracle on the same seeded inputs.

Bit-exact: token order (slot / order tables), visible-patch bytes (bf16 of the reference's fp32 (x-mean)/std, feature order
c p0 p1 p2).  Float: regression targets within 2e-5 absolute of the fp32 reference (values are O(1)); loss within 1e-3
relative (north_star); predictions cosine >= 0.999 per token in bf16; gradients cosine > 0.99."""
import os
from functools import partial

import numpy as np
import pytest
import torch
import torch.nn.functional as F

from oracle import videomae_oracle as O

pytestmark = pytest.mark.gpu
GOLDEN = os.path.join(os.path.dirname(__file__), "golden")


def build(cfg):
    from avion_b200.models.model_videomae import PretrainVisionTransformer
    return PretrainVisionTransformer(
        img_size=cfg.image_size, patch_size=cfg.patch, encoder_embed_dim=cfg.enc_dim, encoder_depth=cfg.enc_depth,
        encoder_num_heads=cfg.enc_heads, decoder_num_classes=cfg.feats, decoder_embed_dim=cfg.dec_dim,
        decoder_num_heads=cfg.dec_heads, decoder_depth=cfg.dec_depth, mlp_ratio=4, qkv_bias=True,
        norm_layer=partial(torch.nn.LayerNorm, eps=1e-6), num_frames=cfg.frames).cuda()


def load_golden(name="videomae_tiny_b4"):
    g = np.load(os.path.join(GOLDEN, name + ".npz"))
    return g, O.MaeCfg(*[int(v) for v in g["cfg"]])


@pytest.mark.parametrize("B,N,ratio", [(1, 72, 0.7), (5, 1568, 0.9), (3, 33, 0.5), (2, 64, 0.0), (2, 64, 1.0)])
def test_mask_slots_bit_exact(B, N, ratio):
    from avion_b200 import ops
    g = torch.Generator().manual_seed(N)
    mask = torch.rand(B, N, generator=g) < ratio
    slot, order, counts = ops.tube_mask_slots(mask.cuda())
    for b in range(B):
        vis = torch.nonzero(~mask[b]).flatten()
        msk = torch.nonzero(mask[b]).flatten()
        ref_order = torch.cat([vis, msk]).int()                     # order of x[~mask] followed by x[mask]
        assert torch.equal(order[b].cpu(), ref_order)
        assert torch.equal(slot[b].cpu()[ref_order.long()], torch.arange(N, dtype=torch.int32))
        assert int(counts[b]) == vis.numel()


@pytest.mark.parametrize("cfg,batch", [(O.TINY, 4), (O.MaeCfg(), 2)])
def test_tubelet_transform_patches_bit_exact_targets_close(cfg, batch):
    from avion_b200 import ops
    clips, mask = O.make_inputs(cfg, batch, 17)
    slot, order, counts = ops.tube_mask_slots(mask.cuda())
    Nv = int(counts[0])
    patches, targets = ops.tubelet_transform(clips.cuda(), slot, Nv, cfg.patch, O.MEAN_TIMM, O.STD_TIMM)
    raw = O.raw_video(clips)
    ref_p = O.tubelet_patches(O.normalize(raw), cfg.patch)[~mask].reshape(batch, Nv, -1).to(torch.bfloat16)
    assert torch.equal(patches.cpu().view(torch.int16), ref_p.view(torch.int16))
    ref_t = O.targets(raw, mask, cfg.patch)
    assert targets.shape == ref_t.shape
    assert (targets.cpu() - ref_t).abs().max() <= 2e-5
    # either output alone
    p2, t2 = ops.tubelet_transform(clips.cuda(), slot, Nv, cfg.patch, O.MEAN_TIMM, O.STD_TIMM, want_targets=False)
    assert t2 is None and torch.equal(p2, patches)
    p3, t3 = ops.tubelet_transform(clips.cuda(), slot, Nv, cfg.patch, O.MEAN_TIMM, O.STD_TIMM, want_patches=False)
    assert p3 is None and torch.equal(t3, targets)


def test_constant_patch_target_is_zero():
    """A flat patch has zero variance: the reference yields 0 / (0 + 1e-6) = 0."""
    from avion_b200 import ops
    cfg = O.TINY
    clips = torch.full((1, cfg.frames, cfg.image_size, cfg.image_size, 3), 200, dtype=torch.uint8)
    mask = torch.ones(1, cfg.tokens, dtype=torch.bool)
    slot, _, _ = ops.tube_mask_slots(mask.cuda())
    _, t = ops.tubelet_transform(clips.cuda(), slot, 0, cfg.patch, O.MEAN_TIMM, O.STD_TIMM, want_patches=False)
    assert torch.count_nonzero(t) == 0


def test_gather_rows_and_batched_colsum():
    from avion_b200 import ops
    g = torch.Generator().manual_seed(0)
    table = torch.randn(72, 64, generator=g)
    add = torch.randn(64, generator=g)
    order = torch.stack([torch.randperm(72, generator=g) for _ in range(3)]).int()
    out = ops.gather_rows(table.cuda(), order.cuda(), 72, add_row=add.cuda(), add_from=24).cpu()
    ref = table[order.long()]
    ref[:, 24:] += add
    assert torch.equal(out.view(torch.int16), ref.to(torch.bfloat16).view(torch.int16))
    x = torch.randn(3, 72, 64, generator=g).to(torch.bfloat16)
    acc = torch.zeros(64, device="cuda")
    ops.colsum_batched(x.cuda()[:, 24:], 3, 72 * 64, 48, 64, acc)
    np.testing.assert_allclose(acc.cpu().numpy(), x[:, 24:].float().sum((0, 1)).numpy(), rtol=1e-5, atol=1e-4)


@pytest.mark.parametrize("n", [4 * 48 * 1536, 1024, 64 * 1408 * 1536])
def test_mse_loss_and_gradient(n):
    from avion_b200.losses import MSELoss
    g = torch.Generator(device="cuda").manual_seed(1)
    rows = n // 512
    pred = torch.randn(rows, 512, generator=g, device="cuda").to(torch.bfloat16).requires_grad_(True)
    tgt = torch.randn(rows, 512, generator=g, device="cuda")
    loss = MSELoss()(pred, tgt)
    (loss * 3.0).backward()
    p32 = pred.detach().float().requires_grad_(True)
    ref = F.mse_loss(p32, tgt)
    (ref * 3.0).backward()
    assert abs(float(loss) - float(ref)) <= 1e-5 * float(ref)
    assert ((pred.grad.float() - p32.grad).abs() <= 2.0 ** -8 * p32.grad.abs() + 1e-30).all()      # one bf16 rounding
    assert float(MSELoss()(pred.detach(), tgt)) == float(loss)      # deterministic reduction


def run_step(cfg, sd, clips, mask):
    from avion_b200.losses import MSELoss
    model = build(cfg)
    model.load_state_dict(sd, strict=True)
    out, tgt = model.forward_with_targets(clips.cuda(), mask.cuda())
    loss = MSELoss()(out, tgt)
    loss.backward()
    return model, out.detach().float().cpu(), tgt.cpu(), loss.detach().cpu()


def test_tiny_step_matches_reference_golden():
    g, cfg = load_golden()
    sd = O.make_state_dict(cfg, int(g["wseed"]))
    clips, mask = O.make_inputs(cfg, int(g["batch"]), int(g["iseed"]))
    model, out, tgt, loss = run_step(cfg, sd, clips, mask)
    assert (tgt[:, :12] - torch.from_numpy(g["targets"])).abs().max() <= 2e-5
    ref_out = torch.from_numpy(g["outputs"])
    assert F.cosine_similarity(out[:, :12], ref_out, dim=-1).min() >= 0.999
    assert abs(float(loss) - float(g["loss"])) <= 1e-3 * abs(float(g["loss"]))
    params = dict(model.named_parameters())
    checked = 0
    for k in g.files:
        if not k.startswith("grad:"):
            continue
        ref = torch.from_numpy(g[k]).float().reshape(-1)
        got = params[k[5:]].grad.detach().float().cpu().reshape(-1)[:ref.numel()]
        cos = F.cosine_similarity(got, ref, dim=0).item()
        rel = ((got - ref).norm() / (ref.norm() + 1e-20)).item()
        assert cos > 0.99 and rel < 0.1, (k, cos, rel)
        checked += 1
    assert checked >= 16


def test_step_matches_oracle_other_seed_and_float_entry():
    cfg = O.TINY
    sd = O.make_state_dict(cfg, 21)
    clips, mask = O.make_inputs(cfg, 6, 5)
    ref, ref_grads = O.loss_and_grads(cfg, sd, clips, mask, ["mask_token", "encoder_to_decoder.weight"])
    model, out, tgt, loss = run_step(cfg, sd, clips, mask)
    assert F.cosine_similarity(out, ref["outputs"].detach(), dim=-1).min() >= 0.999
    assert abs(float(loss) - float(ref["loss"])) <= 1e-3 * float(ref["loss"])
    for k, r in ref_grads.items():
        got = dict(model.named_parameters())[k].grad.cpu()
        assert F.cosine_similarity(got.reshape(-1), r.reshape(-1), dim=0) > 0.99, k
    with torch.no_grad():          # the reference's input contract: normalised float [B,3,T,H,W] + mask
        out_f = model(O.normalize(O.raw_video(clips)).cuda(), mask.cuda()).float().cpu()
    assert F.cosine_similarity(out_f, out, dim=-1).min() > 0.9999


def test_train_steps_reduce_loss():
    from avion_b200.losses import MSELoss
    from avion_b200.optim import FusedAdamW
    cfg = O.TINY
    model = build(cfg)
    model.load_state_dict(O.make_state_dict(cfg, 0), strict=True)
    clips, mask = O.make_inputs(cfg, 8, 7)
    clips, mask = clips.cuda(), mask.cuda()
    opt = FusedAdamW(model.parameters(), lr=1e-3)
    losses = []
    for _ in range(10):
        opt.zero_grad(set_to_none=True)
        out, tgt = model.forward_with_targets(clips, mask)
        loss = MSELoss()(out, tgt)
        loss.backward()
        opt.step()
        losses.append(float(loss))
    assert losses[-1] < losses[0] - 0.05, losses


def test_full_size_forward_backward_properties():
    """VIDEOMAE_VITB16 at BASELINE sizes (8 clips): shapes, finite loss near the variance of the targets, and the gradient of
    mask_token equals the sum of the decoder-input gradients (checked through a second run with the token perturbed)."""
    from avion_b200.data.transforms import TubeMaskingGeneratorGPU
    from avion_b200.losses import MSELoss
    from avion_b200.models import VIDEOMAE_VITB16
    torch.manual_seed(0)
    model = VIDEOMAE_VITB16(decoder_depth=4).cuda()
    clips = torch.randint(0, 256, (8, 16, 224, 224, 3), dtype=torch.uint8, device="cuda")
    mask = TubeMaskingGeneratorGPU(8, (8, 14, 14), 0.9, device="cuda")()
    out, tgt = model.forward_with_targets(clips, mask)
    assert out.shape == (8, 1408, 1536) and tgt.shape == (8, 1408, 1536)
    loss = MSELoss()(out, tgt)
    loss.backward()
    assert torch.isfinite(loss) and 0.5 < float(loss) < 3.0
    assert all(p.grad is not None and torch.isfinite(p.grad).all() for p in model.parameters())
    g = model.mask_token.grad.clone()
    eps = 0.05 * g / g.norm()
    with torch.no_grad():
        l0 = float(MSELoss()(model.forward_with_targets(clips, mask)[0], tgt))
        model.mask_token += eps
        l1 = float(MSELoss()(model.forward_with_targets(clips, mask)[0], tgt))
    assert l1 > l0                                         # stepping along +grad increases the loss


def test_graphed_step_matches_eager():
    """avion_b200.graph.GraphedStep: one CUDA-graph replay of zero_grad+forward+loss+backward gives the eager step's loss and
    gradients (split-K reductions are order-dependent, hence allclose), follows new inputs, and feeds FusedAdamW."""
    from avion_b200.graph import GraphedStep
    from avion_b200.losses import MSELoss
    from avion_b200.optim import FusedAdamW
    cfg = O.TINY
    model = build(cfg)
    model.load_state_dict(O.make_state_dict(cfg, 0), strict=True)
    crit = MSELoss()
    opt = FusedAdamW(model.parameters(), lr=1e-3)
    batches = [tuple(t.cuda() for t in O.make_inputs(cfg, 4, s)) for s in (1, 2)]

    def fwd_bwd(clips, mask):
        opt.zero_grad(set_to_none=True)
        out, tgt = model(clips, mask, return_targets=True)
        loss = crit(out, tgt)
        loss.backward()
        return loss.detach()

    eager = []
    for clips, mask in batches:
        loss = fwd_bwd(clips, mask)
        eager.append((float(loss), {k: p.grad.clone() for k, p in model.named_parameters()}))
    step = GraphedStep(fwd_bwd, list(batches[0]))
    for (clips, mask), (l_ref, g_ref) in zip(batches, eager):
        loss = step(clips, mask)
        assert abs(float(loss) - l_ref) <= 1e-5 * abs(l_ref)
        for k, p in model.named_parameters():
            assert torch.allclose(p.grad, g_ref[k], rtol=1e-3, atol=1e-5 * float(g_ref[k].abs().max()) + 1e-12), k
    before = model.mask_token.detach().clone()
    losses = []
    for _ in range(8):
        losses.append(float(step(*batches[0])))
        opt.step()
    assert not torch.equal(before, model.mask_token.detach()) and losses[-1] < losses[0], losses
