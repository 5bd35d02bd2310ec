"""End-to-end parity of the CUDA path (through the avion.models / avion.losses API) against the committed golden
vectors of the unmodified reference and against the oracle, on the same seeded inputs.

Tolerances (BASELINE.json north_star): loss within 1e-3 relative, embedding cosine >= 0.999 in bf16."""
import os

import numpy as np
import pytest
import torch
import torch.nn.functional as F

from oracle import clip_oracle as O

pytestmark = pytest.mark.gpu
GOLDEN = os.path.join(os.path.dirname(__file__), "golden")
LOSS_TOL = 1e-3          # BASELINE.json north_star: loss within 1e-3 relative of the reference PyTorch path
COS_TOL = 0.999          # embeddings: cosine similarity >= 0.999 in bf16


def load_golden(name):
    g = np.load(os.path.join(GOLDEN, name + ".npz"))
    return g, O.ClipCfg(*[int(v) for v in g["cfg"]])


def run_step(cfg, sd, clips, tokens, backward=True):
    from avion_b200.smoke import build_model
    from avion_b200.losses import ClipLoss
    model = build_model(cfg).cuda()
    model.load_state_dict(sd, strict=True)
    with torch.set_grad_enabled(backward):
        img, txt, scale = model(clips.cuda(), tokens.cuda())
        out = ClipLoss(world_size=1)(img, txt, scale)
    if backward:
        out["loss"].backward()
    return model, img.detach().cpu(), txt.detach().cpu(), out


@pytest.mark.parametrize("name", ["tiny_b6", "tiny_b32", "tiny14_b32"])      # tiny14: patch 14, the ViT-L/14 patchify path
def test_tiny_step_matches_reference_golden(name):
    g, cfg = load_golden(name)
    sd = O.make_state_dict(cfg, int(g["wseed"]))
    clips, tokens = O.make_inputs(cfg, int(g["batch"]), int(g["iseed"]))
    model, img, txt, out = run_step(cfg, sd, clips, tokens)
    assert F.cosine_similarity(img, torch.from_numpy(g["image_embed"]), dim=-1).min() >= 0.999
    assert F.cosine_similarity(txt, torch.from_numpy(g["text_embed"]), dim=-1).min() >= 0.999
    assert abs(float(out["loss"]) - float(g["loss"])) <= LOSS_TOL * abs(float(g["loss"]))      # north star: 1e-3, every fixture
    assert float(out["clip_acc"]) == pytest.approx(float(g["clip_acc"]), rel=1e-6)
    params = dict(model.named_parameters())
    checked = 0
    for k in g.files:
        if not k.startswith("grad:"):
            continue
        ref = torch.from_numpy(g[k]).float().reshape(-1)
        got = params[k[5:]].grad.detach().float().cpu().reshape(-1)[:ref.numel()]
        cos = F.cosine_similarity(got, ref, dim=0).item()
        rel = ((got - ref).norm() / (ref.norm() + 1e-20)).item()
        assert cos > 0.99 and rel < 0.1, (k, cos, rel)     # bf16 activations/gradients vs the fp32 reference
        checked += 1
    assert checked >= 15


@pytest.mark.parametrize("name", ["tiny_b6", "tiny_b32", "tiny14_b32"])
def test_tiny_step_matches_bf16_point_oracle(name):
    """The sharper comparison: the same fixtures against the oracle evaluated WITH the pipeline's bf16 rounding points
    (oracle/bf16_points.py).  The rounding noise every bf16 implementation has against fp32 (rms 1.4e-3 of the loss at batch 6,
    tools/parity_noise.py) cancels, so a biased kernel (exponential, accumulation order, a missing rounding) shows up here even
    where the fp32 comparison could not tell it from noise."""
    from oracle.bf16_points import ALL_POINTS, Emu
    g, cfg = load_golden(name)
    sd = O.make_state_dict(cfg, int(g["wseed"]))
    clips, tokens = O.make_inputs(cfg, int(g["batch"]), int(g["iseed"]))
    _, img, txt, out = run_step(cfg, sd, clips, tokens, backward=False)
    with torch.no_grad():
        loss, e_img, e_txt = Emu(ALL_POINTS).loss(sd, clips, tokens, cfg)
    assert F.cosine_similarity(img, e_img, dim=-1).min() >= 0.9999
    assert F.cosine_similarity(txt, e_txt, dim=-1).min() >= 0.9999
    # At batch 32 the north-star tolerance holds against the emulation as well.  At batch 6 the loss hangs on 12 pooled rows of
    # 128 values: ONE bf16 rounding that falls the other way on such a row (fp32 summation order is all it takes) moves this
    # loss by ~4e-4, so kernel and emulation - both valid bf16 evaluations - sit a few such flips apart (measured: 1.7e-3; the
    # emulation itself is 2.5e-3 from the fp32 reference on this fixture, the kernel 0.8e-3: DESIGN.md section 2).
    tol = LOSS_TOL if int(g["batch"]) >= 32 else 3e-3
    assert abs(float(out["loss"]) - loss) <= tol * abs(loss)


def test_full_size_block_matches_oracle():
    """One ViT-B/16 16-frame block at full size (S = 3137 tokens, 12 heads of 64, width 768; 2 clips) through the engine's
    kernels against the oracle's fp32 block (avion/models/transformer.py:144-150): every token of the output, not just the
    pooled embedding.  (The tiny fixtures never leave one K/V tile; this one runs 25 x 25 tiles per head.)"""
    from avion_b200.models import _engine as E
    torch.manual_seed(0)
    B, S, D, H = 2, 3137, 768, 12
    rs = np.random.RandomState(5)
    p = "b."
    sd = {p + "ln_1.weight": 1 + 0.1 * rs.standard_normal(D), p + "ln_1.bias": 0.1 * rs.standard_normal(D),
          p + "ln_2.weight": 1 + 0.1 * rs.standard_normal(D), p + "ln_2.bias": 0.1 * rs.standard_normal(D),
          p + "attn.Wqkv.weight": rs.standard_normal((3 * D, D)) * D ** -0.5 * 2.0, p + "attn.Wqkv.bias": 0.02 * rs.standard_normal(3 * D),
          p + "attn.out_proj.weight": 0.02 * rs.standard_normal((D, D)), p + "attn.out_proj.bias": 0.02 * rs.standard_normal(D),
          p + "mlp.fc1.weight": 0.02 * rs.standard_normal((4 * D, D)), p + "mlp.fc1.bias": 0.02 * rs.standard_normal(4 * D),
          p + "mlp.fc2.weight": 0.02 * rs.standard_normal((D, 4 * D)), p + "mlp.fc2.bias": 0.02 * rs.standard_normal(D)}
    sd = {k: torch.from_numpy(np.asarray(v, dtype=np.float32)) for k, v in sd.items()}
    x = torch.from_numpy(rs.standard_normal((B, S, D)).astype(np.float32)).bfloat16()
    with torch.no_grad():
        ref = O._block(x.float(), sd, p, H, 1e-6, ("attn.Wqkv.weight", "attn.Wqkv.bias", "mlp.fc1", "mlp.fc2"))
    names = dict(ln1_w="ln_1.weight", ln1_b="ln_1.bias", qkv_w="attn.Wqkv.weight", qkv_b="attn.Wqkv.bias", out_w="attn.out_proj.weight",
                 out_b="attn.out_proj.bias", ln2_w="ln_2.weight", ln2_b="ln_2.bias", fc1_w="mlp.fc1.weight", fc1_b="mlp.fc1.bias",
                 fc2_w="mlp.fc2.weight", fc2_b="mlp.fc2.bias")
    L = {k: sd[p + v].cuda() for k, v in names.items()}
    for k in ("qkv_w", "out_w", "fc1_w", "fc2_w"):
        L[k] = L[k].bfloat16()
    y, _ = E.blocks_forward(x.cuda(), [L], H, 1e-6, False, E.Workspace(), save=False)
    y = y.float().cpu()
    cos = F.cosine_similarity(y, ref, dim=-1)
    assert cos.min() >= COS_TOL, float(cos.min())
    # the block's own contribution (output minus the residual input), where an attention error cannot hide behind x
    dy, dref = y - x.float(), ref - x.float()
    rel = ((dy - dref).norm() / dref.norm()).item()
    assert rel < 2e-2, rel
    assert F.cosine_similarity(dy.reshape(B * S, D), dref.reshape(B * S, D), dim=-1).min() > 0.99


def test_vitb16_4f_forward_matches_reference_golden():
    """BASELINE.json configs[0]: ViT-B/16 4-frame + 12L/512 text, batch 8, forward + ClipLoss."""
    g, cfg = load_golden("vitb16_4f_b8")
    sd = O.make_state_dict(cfg, int(g["wseed"]))
    clips, tokens = O.make_inputs(cfg, int(g["batch"]), int(g["iseed"]))
    _, img, txt, out = run_step(cfg, sd, clips, tokens, backward=False)
    assert F.cosine_similarity(img, torch.from_numpy(g["image_embed"]), dim=-1).min() >= COS_TOL
    assert F.cosine_similarity(txt, torch.from_numpy(g["text_embed"]), dim=-1).min() >= COS_TOL
    assert abs(float(out["loss"]) - float(g["loss"])) <= LOSS_TOL * abs(float(g["loss"]))


def test_input_forms_agree():
    """uint8 clips (fused transform), precomputed bf16 patches and the reference's normalised float layout."""
    from avion_b200.smoke import build_model
    from avion_b200.data import GPUClipTransform
    cfg = O.TINY
    model = build_model(cfg).cuda()
    model.load_state_dict(O.make_state_dict(cfg, 3), strict=True)
    clips, _ = O.make_inputs(cfg, 4, 5)
    with torch.no_grad():
        a = model.encode_image(clips.cuda())
        b = model.encode_image(GPUClipTransform(cfg.image_size, cfg.patch)(clips.cuda()))
        c = model.encode_image(O.clip_transform(clips).cuda())
    assert torch.equal(a, b)
    assert F.cosine_similarity(a, c, dim=-1).min() > 0.9999


def test_train_steps_reduce_loss():
    """A few AdamW steps on one batch: the loss must go down (gradients point the right way end to end)."""
    from avion_b200.smoke import build_model
    from avion_b200.losses import ClipLoss
    cfg = O.TINY
    model = build_model(cfg).cuda()
    model.load_state_dict(O.make_state_dict(cfg, 0), strict=True)
    clips, tokens = O.make_inputs(cfg, 16, 7)
    clips, tokens = clips.cuda(), tokens.cuda()
    opt = torch.optim.AdamW(model.parameters(), lr=3e-4)
    crit = ClipLoss(world_size=1)
    losses = []
    for _ in range(8):
        opt.zero_grad(set_to_none=True)
        out = crit(*model(clips, tokens))
        out["loss"].backward()
        opt.step()
        losses.append(float(out["loss"]))
    assert losses[-1] < losses[0] - 0.2, losses


def test_clip_step_replays_as_cuda_graph():
    """The whole CLIP step (fused transform -> towers -> ClipLoss -> backward) is free of host synchronisation - ClipLoss reads
    logit_scale.exp() on the device - so it captures into ONE CUDA graph.  GraphedStep(params=...)
    keeps every gradient as a view of one flat buffer (what a multi-rank run all-reduces after the replay): loss, accuracy and
    gradients of a replay equal the eager step's, and the replay follows new inputs."""
    from avion_b200.graph import GraphedStep
    from avion_b200.losses import ClipLoss
    from avion_b200.smoke import build_model
    cfg = O.TINY
    model = build_model(cfg).cuda()
    model.load_state_dict(O.make_state_dict(cfg, 0), strict=True)
    crit = ClipLoss(world_size=1)
    batches = [tuple(t.cuda() for t in O.make_inputs(cfg, 8, s)) for s in (11, 12)]

    def fwd_bwd(clips, tokens):
        out = crit(*model(clips, tokens))
        out["loss"].backward()
        return out["loss"].detach(), out["clip_acc"]

    eager = []
    for clips, tokens in batches:
        model.zero_grad(set_to_none=True)
        loss, acc = fwd_bwd(clips, tokens)
        eager.append((float(loss), float(acc), {k: p.grad.clone() for k, p in model.named_parameters()}))
    model.zero_grad(set_to_none=True)
    step = GraphedStep(fwd_bwd, list(batches[0]), params=list(model.parameters()))
    assert all(p.grad.untyped_storage().data_ptr() == step.flat.untyped_storage().data_ptr() for p in model.parameters())
    for (clips, tokens), (l_ref, a_ref, g_ref) in zip(batches, eager):
        loss, acc = step(clips, tokens)
        assert abs(float(loss) - l_ref) <= 1e-5 * abs(l_ref) and float(acc) == a_ref
        for k, p in model.named_parameters():
            assert torch.allclose(p.grad, g_ref[k], rtol=2e-3, atol=1e-5 * float(g_ref[k].abs().max()) + 1e-12), k
