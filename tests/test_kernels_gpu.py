"""Per-kernel parity on the B200: every C-ABI kernel against a plain fp32 PyTorch statement of the same arithmetic
(and the oracle where one exists).  Integer/byte work is checked bit-exactly; floating point within stated tolerances."""
import math

import numpy as np
import pytest
import torch
import torch.nn.functional as F

from oracle import clip_oracle as O

pytestmark = pytest.mark.gpu

bf16, f32 = torch.bfloat16, torch.float32


@pytest.fixture(scope="module")
def ops():
    from avion_b200 import ops as _ops
    return _ops


def rnd(*shape, scale=1.0, dtype=bf16, seed=None):
    g = torch.Generator(device="cuda")
    g.manual_seed(seed if seed is not None else (hash(shape) & 0xffff))
    return (torch.randn(*shape, generator=g, device="cuda", dtype=f32) * scale).to(dtype)


def rel_err(a, b):
    a, b = a.float(), b.float()
    return float((a - b).norm() / (b.norm() + 1e-12))


# ---------------------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("shape", [(2, 4, 64, 64, 16), (3, 2, 224, 224, 16), (1, 1, 96, 96, 32)])
def test_clip_transform_bit_exact(ops, shape):
    B, T, H, W, p = shape
    clips = torch.from_numpy(np.random.RandomState(1).randint(0, 256, (B, T, H, W, 3)).astype(np.uint8))
    ref = O.patchify(O.clip_transform(clips), p).to(bf16)
    got = ops.clip_transform(clips.cuda(), H, p, O.MEAN_OPENAI, O.STD_OPENAI)
    assert torch.equal(got.cpu(), ref)


@pytest.mark.parametrize("shape", [(2, 3, 56, 56, 14), (1, 2, 224, 224, 14)])
def test_clip_transform_patch14_bit_exact(ops, shape):
    """ViT-L/14 patchify: 588 features per token, rows zero-padded to 592 (16-byte rows for TMA)."""
    B, T, H, W, p = shape
    clips = torch.from_numpy(np.random.RandomState(3).randint(0, 256, (B, T, H, W, 3)).astype(np.uint8))
    ref = O.patchify(O.clip_transform(clips), p).to(bf16)
    got = ops.clip_transform(clips.cuda(), H, p, O.MEAN_OPENAI, O.STD_OPENAI).cpu()
    assert got.shape[-1] == 592
    assert torch.equal(got[..., :588], ref)
    assert torch.count_nonzero(got[..., 588:]) == 0
    boxes = torch.tensor([[0, 0, 1]] * B, dtype=torch.int32)
    ref_f = O.patchify(O.clip_transform(O.crop_clip(clips, boxes, (H, W))), p).to(bf16)
    got_f = ops.clip_transform(clips.cuda(), H, p, O.MEAN_OPENAI, O.STD_OPENAI, boxes=boxes.cuda()).cpu()
    assert torch.equal(got_f[..., :588], ref_f)


def test_clip_transform_crop_flip_bit_exact(ops):
    B, T, Hs, Ws, crop, p = 4, 3, 101, 157, 64, 16
    clips = torch.from_numpy(np.random.RandomState(2).randint(0, 256, (B, T, Hs, Ws, 3)).astype(np.uint8))
    boxes = torch.tensor([[0, 0, 0], [93, 37, 1], [5, 11, 0], [92, 36, 1]], dtype=torch.int32)
    ref = O.patchify(O.clip_transform(O.crop_clip(clips, boxes, (crop, crop))), p).to(bf16)
    got = ops.clip_transform(clips.cuda(), crop, p, O.MEAN_OPENAI, O.STD_OPENAI, boxes=boxes.cuda())
    assert torch.equal(got.cpu(), ref)


def test_clip_transform_clamps_boxes_and_takes_any_alignment(ops):
    """Offsets one past the last valid position (decord's sampler draws x in [0, W-w+1], threaded_decoder.cc:66-67) are clamped
    into the frame like ffmpeg's crop filter does - the kernel never reads a neighbouring row.  A frames pointer that is not
    16-byte aligned exercises the bulk-load alignment logic and the guarded copy of the buffer's first / last row."""
    from avion_b200.data.transforms import MEAN_TIMM, STD_TIMM
    B, T, Hs, Ws, crop, p = 2, 2, 70, 90, 64, 16
    clips = torch.from_numpy(np.random.RandomState(5).randint(0, 256, (B, T, Hs, Ws, 3)).astype(np.uint8))
    edge = torch.tensor([[26, 6, 0], [26, 6, 1]], dtype=torch.int32)
    past = torch.tensor([[27, 7, 0], [400, 99, 1]], dtype=torch.int32)
    ref = O.patchify(O.clip_transform(O.crop_clip(clips, edge, (crop, crop))), p).to(bf16)
    assert torch.equal(ops.clip_transform(clips.cuda(), crop, p, O.MEAN_OPENAI, O.STD_OPENAI, boxes=past.cuda()).cpu(), ref)
    for shift in (1, 5, 15):
        buf = torch.zeros(clips.numel() + 16, dtype=torch.uint8, device="cuda")
        view = buf[shift:shift + clips.numel()].view(clips.shape)
        view.copy_(clips)
        assert view.data_ptr() % 16 == shift and view.is_contiguous()
        assert torch.equal(ops.clip_transform(view, crop, p, O.MEAN_OPENAI, O.STD_OPENAI, boxes=edge.cuda()).cpu(), ref)
    # another mean / std pair (timm style): still bit-exact bf16((x - mean) / std)
    ref_t = O.patchify(O.clip_transform(O.crop_clip(clips, edge, (crop, crop)), MEAN_TIMM, STD_TIMM), p).to(bf16)
    assert torch.equal(ops.clip_transform(clips.cuda(), crop, p, MEAN_TIMM, STD_TIMM, boxes=edge.cuda()).cpu(), ref_t)


@pytest.mark.parametrize("case", ["down16", "mixed16", "patch14"])
def test_clip_resize_transform_matches_oracle(ops, case):
    """Fused crop -> hflip -> resize -> normalise -> patchify (decord 'crop=..,hflip,scale=..', threaded_decoder.cc:92) against the
    oracle's restatement of torchvision's antialiased-bilinear resized_crop: box arithmetic and patch order exact, pixel values
    equal to bf16(normalise(oracle uint8)) except where an fp32 summation-order difference crosses a rounding tie (then one
    grey level apart).  swscale parity is unpinned (no source in the tree)."""
    if case == "patch14":
        B, T, Hs, Ws, crop, p = 2, 2, 90, 130, 56, 14
        boxes = torch.tensor([[3, 4, 120, 80, 0], [40, 10, 33, 47, 1]], dtype=torch.int32)
    elif case == "down16":
        B, T, Hs, Ws, crop, p = 2, 3, 288, 384, 224, 16
        boxes = torch.tensor([[10, 20, 300, 260, 0], [0, 0, 384, 288, 1]], dtype=torch.int32)
    else:
        B, T, Hs, Ws, crop, p = 4, 2, 101, 157, 64, 16
        boxes = torch.tensor([[0, 0, 157, 101, 0], [93, 37, 64, 64, 1], [5, 11, 40, 30, 0], [58, 2, 100, 100, 1]], dtype=torch.int32)
    clips = torch.from_numpy(np.random.RandomState(7).randint(0, 256, (B, T, Hs, Ws, 3)).astype(np.uint8))
    u8 = O.resized_crop_clip(clips, boxes, (crop, crop))
    ref = O.patchify(O.clip_transform(u8), p)
    got = ops.clip_resize_transform(clips.cuda(), boxes.cuda(), crop, p, O.MEAN_OPENAI, O.STD_OPENAI).cpu()
    feat = 3 * p * p
    assert got.shape[-1] == (feat + 7) // 8 * 8 and torch.count_nonzero(got[..., feat:]) == 0
    got = got[..., :feat].float()
    same = got == ref.to(bf16).float()
    assert same.float().mean() > 0.999, float(same.float().mean())
    one_level = 1.0 / min(O.STD_OPENAI) + 2.0 ** -6
    assert float((got - ref).abs().max()) <= one_level
    if case == "mixed16":      # a box of exactly the output size is a plain crop (+ flip): bit-exact with the crop kernel
        plain = ops.clip_transform(clips.cuda(), crop, p, O.MEAN_OPENAI, O.STD_OPENAI, boxes=torch.tensor([[0, 0, 0], [93, 37, 1], [0, 0, 0], [0, 0, 0]], dtype=torch.int32).cuda())
        assert torch.equal(plain[1].cpu().float(), got[1])


# the last four give every persistent CTA several trips round its TMA ring (296 CTAs x 8 rows x up to 8 stages; 148 CTAs x 4
# stages at D = 1024) and end in a ragged tile; (3, 384) has fewer rows than one tile
@pytest.mark.parametrize("rows,D", [(1000, 768), (77 * 3, 512), (65 * 6, 128), (50, 1024), (3, 384), (25003, 768), (9001, 1024),
                                    (40005, 384), (60001, 128)])
def test_layernorm_fwd_bwd(ops, rows, D):
    x = rnd(rows, D, seed=1)
    g = (1 + 0.1 * torch.randn(D, device="cuda")).float()
    b = (0.1 * torch.randn(D, device="cuda")).float()
    y, mean, rstd = ops.layernorm_fwd(x, g, b, 1e-6)
    xr = x.float().requires_grad_(True)
    gr, br = g.clone().requires_grad_(True), b.clone().requires_grad_(True)
    yr = F.layer_norm(xr, (D,), gr, br, 1e-6)
    assert rel_err(y, yr) < 4e-3
    dy = rnd(rows, D, seed=2)
    dres = rnd(rows, D, seed=3)
    yr.backward(dy.float())
    dg, db, dc = (torch.zeros(D, device="cuda") for _ in range(3))
    dx = ops.layernorm_bwd(dy, x, g, mean, rstd, dg, db, dres=dres, dcolsum=dc)
    assert rel_err(dx, xr.grad + dres.float()) < 5e-3
    assert rel_err(dg, gr.grad) < 2e-3 and rel_err(db, br.grad) < 2e-3
    assert rel_err(dc, dx.float().sum(0)) < 1e-4
    assert rel_err(mean, x.float().mean(-1)) < 1e-5 and rel_err(rstd, (x.float().var(-1, unbiased=False) + 1e-6).rsqrt()) < 1e-5
    # without the residual gradient / column sum (ln_pre's backward)
    dg2, db2 = torch.zeros(D, device="cuda"), torch.zeros(D, device="cuda")
    dx2 = ops.layernorm_bwd(dy, x, g, mean, rstd, dg2, db2)
    assert rel_err(dx2, xr.grad) < 5e-3 and rel_err(dg2, gr.grad) < 2e-3


# the last two are large enough for the CTA-pair (cta_group::2) kernels, with an odd number of ragged 128-row tiles
GEMM_SHAPES = [(256, 256, 128), (1000, 768, 768), (130, 384, 128), (2304, 3072, 768), (128 * 150, 768, 3072), (300, 200, 72),
               (128 * 201 + 37, 768, 256), (128 * 201 + 37, 256, 768)]


@pytest.mark.parametrize("M,N,K", GEMM_SHAPES)
def test_gemm_nt_epilogues(ops, M, N, K):
    x, w = rnd(M, K, seed=1), rnd(N, K, scale=K ** -0.5, seed=2)
    bias = torch.randn(N, device="cuda")
    res = rnd(M, N, seed=3)
    acc = x.float() @ w.float().T
    assert rel_err(ops.linear_fwd(x, w), acc) < 4e-3
    assert rel_err(ops.linear_fwd(x, w, bias, ops.EPI_BIAS), acc + bias) < 4e-3
    assert rel_err(ops.linear_fwd(x, w, bias, ops.EPI_BIAS_GELU), F.gelu(acc + bias)) < 5e-3
    h, pre = ops.linear_fwd(x, w, bias, ops.EPI_BIAS_GELU_DUAL)        # gelu(pre) and pre, what the MLP backward needs
    assert rel_err(pre, acc + bias) < 4e-3 and rel_err(h, F.gelu(acc + bias)) < 5e-3
    assert rel_err(ops.linear_fwd(x, w, bias, ops.EPI_BIAS_RESIDUAL, residual=res), acc + bias + res.float()) < 4e-3


@pytest.mark.parametrize("M,N,K", GEMM_SHAPES)
def test_gemm_nn_dgrad(ops, M, N, K):
    dy, w = rnd(M, N, seed=4), rnd(N, K, scale=N ** -0.5, seed=5)
    ref = dy.float() @ w.float()
    assert rel_err(ops.linear_dgrad(dy, w), ref) < 4e-3
    pre = rnd(M, K, scale=1.5, seed=6)   # stands for the pre-activation, the second output of EPI_BIAS_GELU_DUAL
    pre32 = pre.float().requires_grad_(True)
    F.gelu(pre32).sum().backward()       # exact-erf gelu'
    assert rel_err(ops.linear_dgrad(dy, w, ops.EPI_DGELU, pre=pre), ref * pre32.grad) < 5e-3


@pytest.mark.parametrize("tok,N,K", [(512, 256, 128), (1000, 768, 768), (77 * 9, 384, 128), (128 * 150, 2304, 768), (333, 200, 72)])
def test_gemm_tn_wgrad(ops, tok, N, K):
    dy, x = rnd(tok, N, seed=7), rnd(tok, K, seed=8)
    dw = torch.ones(N, K, device="cuda")
    ops.linear_wgrad(dy, x, dw)
    assert rel_err(dw, 1 + dy.float().T @ x.float()) < 3e-3


@pytest.mark.parametrize("tok,N,K", [(4096, 768, 768), (3137 * 4, 2304, 768), (8200, 3072, 768), (1000, 512, 2048), (520, 200, 96),
                                     (77 * 16, 1536, 512), (64, 256, 256), (200000, 256, 256)])
def test_gemm_tn_wgrad_with_bias_gradient(ops, tok, N, K):
    """The weight-gradient GEMM also returns dY.sum(0) (the bias gradient) from the dY tiles it stages: every tile geometry
    (CTA pairs / single CTA, 256- and 128-wide, ragged M and token counts, several tiles per CTA, split-K), accumulating."""
    dy, x = rnd(tok, N, seed=17), rnd(tok, K, seed=18)
    dw = torch.ones(N, K, device="cuda")
    db = torch.full((N,), 2.0, device="cuda")
    ops.linear_wgrad(dy, x, dw, dbias=db)
    assert rel_err(dw, 1 + dy.float().T @ x.float()) < 3e-3
    ref = 2 + dy.float().sum(0)
    assert rel_err(db, ref) < 5e-5, rel_err(db, ref)
    ops.linear_wgrad(dy, x, dw, dbias=db)                 # += on both outputs
    assert rel_err(db, 2 + 2 * dy.float().sum(0)) < 5e-5
    dw2 = torch.ones(N, K, device="cuda")
    ops.linear_wgrad(dy, x, dw2)                          # without the bias gradient: same weight gradient path as before
    assert rel_err(dw2, 1 + dy.float().T @ x.float()) < 3e-3


def test_gemm_batched_patch_embed_layout(ops):
    """Patch-embed geometry: A [B][P][K] -> C rows 1.. of [B][P+1][D] with a shared residual table."""
    B, P, K, D = 3, 196, 256, 384
    a, w = rnd(B, P, K, seed=1), rnd(D, K, scale=K ** -0.5, seed=2)
    table = rnd(P + 1, D, seed=3)
    x = torch.zeros(B, P + 1, D, dtype=bf16, device="cuda")
    ops.gemm(ops.GEMM_NT, ops.EPI_BIAS_RESIDUAL, P, D, K, a, K, w, K, x[:, 1:], D, batch=B, a_bs=P * K, c_bs=(P + 1) * D,
             R=table[1:], ldr=D, r_bs=0)
    ref = a.float() @ w.float().T + table[1:].float()
    assert rel_err(x[:, 1:], ref) < 4e-3
    assert float(x[:, 0].abs().max()) == 0.0
    dw = torch.zeros(D, K, device="cuda")
    ops.gemm(ops.GEMM_TN, ops.EPI_F32_ADD, D, K, P, x[:, 1:], D, a, K, dw, K, batch=B, a_bs=(P + 1) * D, b_bs=P * K, splits=5)
    assert rel_err(dw, torch.einsum("bpd,bpk->dk", x[:, 1:].float(), a.float())) < 3e-3


def attn_ref(qkv, H, scale, causal):
    B, S, _ = qkv.shape
    q, k, v = (t.transpose(1, 2) for t in qkv.float().view(B, S, 3, H, 64).unbind(2))
    s = (q @ k.transpose(-1, -2)) * scale
    if causal:
        s = s + torch.full((S, S), float("-inf"), device=qkv.device).triu_(1)
    lse = torch.logsumexp(s, -1)
    o = (torch.softmax(s, -1) @ v).transpose(1, 2).reshape(B, S, H * 64)
    return o, lse


@pytest.mark.parametrize("B,S,H,causal", [(2, 128, 2, False), (2, 65, 2, False), (3, 785, 3, False), (1, 3137, 2, False),
                                          (4, 77, 2, True), (2, 300, 2, True), (2, 256, 1, False),
                                          # more (clip, head, K/V tile) items than SMs: the persistent backward walks several
                                          # items per CTA (causal: a different number of query tiles per item)
                                          (16, 300, 6, False), (20, 400, 4, True), (40, 77, 8, True)])
def test_attention_fwd_bwd(ops, B, S, H, causal):
    qkv = rnd(B, S, 3 * H * 64, seed=11)
    scale = 0.125
    o, lse = ops.attn_fwd(qkv, H, scale, causal)
    q32 = qkv.float().requires_grad_(True)
    o_ref, lse_ref = attn_ref(q32, H, scale, causal)
    assert rel_err(o, o_ref) < 6e-3
    assert float((lse - lse_ref).abs().max()) < 2e-3
    do = rnd(B, S, H * 64, seed=12)
    o_ref.backward(do.float())
    dqkv = ops.attn_bwd(qkv, o, do, lse, H, scale, causal)
    g = q32.grad.view(B, S, 3, H * 64)
    d = dqkv.view(B, S, 3, H * 64)
    for i, name in enumerate("qkv"):
        assert rel_err(d[:, :, i], g[:, :, i]) < 1.5e-2, name


@pytest.mark.parametrize("B,S,H,causal,mag", [(4, 200, 2, False, 1.0), (4, 200, 2, False, 3.0), (2, 787, 2, False, 2.0), (4, 77, 2, True, 1.0),
                                              (4, 32, 2, True, 2.0)])
def test_attention_fwd_accuracy_vs_fp64(ops, B, S, H, causal, mag):
    """A third of the exponentials of the forward kernel run as a polynomial on the FMA pipe: the output error against an
    fp64 softmax must stay at the level of the bf16 rounding of the result itself (measured 1.03-1.33x), the LSE within 2e-4."""
    torch.manual_seed(3)
    qkv = (torch.randn(B, S, 3 * H * 64, device="cuda") * mag).to(bf16)
    o, lse = ops.attn_fwd(qkv, H, 0.125, causal)
    q, k, v = qkv.double().view(B, S, 3, H, 64).permute(2, 0, 3, 1, 4)
    s = (q @ k.transpose(-1, -2)) * 0.125
    if causal:
        s = s.masked_fill(torch.ones(S, S, device="cuda", dtype=torch.bool).triu(1), float("-inf"))
    ref = torch.softmax(s, -1) @ v
    got = o.double().view(B, S, H, 64).permute(0, 2, 1, 3)
    err = (got - ref).pow(2).mean().sqrt()
    ideal = (ref.to(bf16).double() - ref).pow(2).mean().sqrt()      # rounding the exact result to bf16
    assert float(err) < 1.5 * float(ideal)
    assert float((lse.double() - torch.logsumexp(s, -1)).abs().max()) < 2e-4


def test_cast_transpose_colsum(ops):
    w = torch.randn(300, 136, device="cuda")
    assert torch.equal(ops.cast_bf16(w), w.to(bf16))
    assert torch.equal(ops.transpose_cast_bf16(w), w.T.contiguous().to(bf16))
    x = rnd(5000, 384, seed=1)
    out = torch.ones(384, device="cuda")
    ops.colsum(x, out)
    assert rel_err(out, 1 + x.float().sum(0)) < 1e-5


def test_pos_table_and_token_embed(ops):
    T, G, D, B = 4, 16, 128, 3
    cls, pos, tmp = torch.randn(D, device="cuda"), torch.randn(G + 1, D, device="cuda"), torch.randn(T, D, device="cuda")
    table = ops.pos_table_fwd(cls, pos, tmp, T, G)
    ref = torch.cat([(cls + pos[0])[None], pos[1:].repeat(T, 1) + tmp.repeat_interleave(G, 0)])
    assert torch.equal(table, ref.to(bf16))
    dx = rnd(B, T * G + 1, D, seed=5)
    dcls, dpos, dtmp = torch.zeros(D, device="cuda"), torch.zeros(G + 1, D, device="cuda"), torch.zeros(T, D, device="cuda")
    ops.pos_table_bwd(dx, dcls, dpos, dtmp, T, G)
    s = dx.float().sum(0)
    assert rel_err(dcls, s[0]) < 1e-5 and rel_err(dpos[0], s[0]) < 1e-5
    assert rel_err(dpos[1:], s[1:].view(T, G, D).sum(0)) < 1e-5
    assert rel_err(dtmp, s[1:].view(T, G, D).sum(1)) < 1e-5
    ids = torch.randint(0, 50, (B, 77), device="cuda")
    emb, p = torch.randn(50, D, device="cuda"), torch.randn(77, D, device="cuda")
    x = ops.token_embed_fwd(ids, emb, p)
    assert torch.equal(x, (emb[ids] + p).to(bf16))
    dxe = rnd(B, 77, D, seed=6)
    demb, dp = torch.zeros_like(emb), torch.zeros_like(p)
    ops.token_embed_bwd(ids, dxe, demb, dp)
    ref = torch.zeros_like(emb).index_add_(0, ids.view(-1), dxe.float().view(-1, D))
    assert rel_err(demb, ref) < 1e-5 and rel_err(dp, dxe.float().sum(0)) < 1e-5


def test_head_fwd_bwd(ops):
    rows, D, E, n = 500, 768, 256, 37
    x = rnd(rows, D, seed=1)
    idx = torch.randperm(rows, device="cuda")[:n].contiguous()
    g, b = (1 + 0.1 * torch.randn(D, device="cuda")), 0.1 * torch.randn(D, device="cuda")
    proj = torch.randn(D, E, device="cuda") * D ** -0.5
    emb, saved = ops.head_fwd(x, idx, g, b, proj, 1e-6)
    xr = x.float().requires_grad_(True)
    gr, br, pr = (t.clone().requires_grad_(True) for t in (g, b, proj))
    ref = F.normalize(F.layer_norm(xr[idx], (D,), gr, br, 1e-6) @ pr, dim=-1)
    assert rel_err(emb, ref) < 1e-5
    demb = torch.randn(n, E, device="cuda")
    ref.backward(demb)
    dx = torch.zeros(rows, D, dtype=bf16, device="cuda")
    dg, db, dp = torch.zeros(D, device="cuda"), torch.zeros(D, device="cuda"), torch.zeros(D, E, device="cuda")
    ops.head_bwd(demb, x, idx, g, b, proj, saved, dx, dg, db, dp)
    assert rel_err(dx, xr.grad) < 5e-3
    assert rel_err(dg, gr.grad) < 1e-4 and rel_err(db, br.grad) < 1e-4 and rel_err(dp, pr.grad) < 1e-4


@pytest.mark.parametrize("n,N,E,off", [(6, 6, 64, 0), (32, 128, 64, 64), (256, 2048, 256, 512), (5, 20, 64, 15)])
def test_infonce_fwd_bwd(ops, n, N, E, off):
    torch.manual_seed(0)
    all_img = F.normalize(torch.randn(N, E, device="cuda"), dim=-1)
    all_txt = F.normalize(all_img + 0.7 * torch.randn(N, E, device="cuda"), dim=-1)
    scale = 14.3
    ai, at = all_img.clone().requires_grad_(True), all_txt.clone().requires_grad_(True)
    sc = torch.tensor(scale, device="cuda", requires_grad=True)
    img, txt = ai[off:off + n], at[off:off + n]
    li, lt = sc * img @ at.T, sc * txt @ ai.T
    labels = torch.arange(n, device="cuda") + off
    loss = (F.cross_entropy(li, labels) + F.cross_entropy(lt, labels)) / 2
    correct = (li.argmax(-1) == labels).sum()
    loss.backward()
    scale_dev = torch.full((1,), scale, device="cuda")          # the kernels read logit_scale.exp() from device memory
    lse, out = ops.infonce_fwd(all_img[off:off + n].contiguous(), all_txt[off:off + n].contiguous(), all_img, all_txt, scale_dev, off)
    assert abs(float(out[0]) / n - float(loss)) < 1e-4 * max(1.0, abs(float(loss)))
    assert int(out[1]) == int(correct)
    d_all, dscale = ops.infonce_bwd(all_img[off:off + n].contiguous(), all_txt[off:off + n].contiguous(), all_img, all_txt, scale_dev,
                                    off, lse, 1.0)
    assert rel_err(d_all[0], ai.grad) < 1e-4 and rel_err(d_all[1], at.grad) < 1e-4
    assert abs(float(dscale) - float(sc.grad)) < 1e-4 * max(1.0, abs(float(sc.grad)))
    # col_side=False (local_loss without gather_with_grad, losses.py:61-72): the gathered operands are constants, so only the
    # direct use of the local rows is differentiated
    ci, ct = all_img[off:off + n].clone().requires_grad_(True), all_txt[off:off + n].clone().requires_grad_(True)
    l2 = (F.cross_entropy(scale * ci @ all_txt.T, labels) + F.cross_entropy(scale * ct @ all_img.T, labels)) / 2
    l2.backward()
    d_rows, _ = ops.infonce_bwd(all_img[off:off + n].contiguous(), all_txt[off:off + n].contiguous(), all_img, all_txt, scale_dev,
                                off, lse, 1.0, col_side=False)
    assert rel_err(d_rows[0, off:off + n], ci.grad) < 1e-4 and rel_err(d_rows[1, off:off + n], ct.grad) < 1e-4
    mask = torch.ones(N, dtype=torch.bool, device="cuda")
    mask[off:off + n] = False
    assert float(d_rows[:, mask].abs().max()) == 0.0 if N > n else True


def test_fused_adamw_matches_torch(ops):
    """O1: one-launch multi-tensor AdamW == torch.optim.AdamW (two groups as in main_lavila_pretrain.py:296-320)."""
    from avion_b200.optim import FusedAdamW, clamp_logit_scale
    torch.manual_seed(5)
    shapes = [(1,), (7,), (1000, 33), (16384 * 2 + 5,), (768, 768), ()]
    ref = [torch.nn.Parameter(torch.randn(s, device="cuda")) for s in shapes]
    # an odd-offset view: exercises the unaligned (scalar) path
    base = torch.randn(4099, device="cuda")
    ref.append(torch.nn.Parameter(base[3:]))
    ours = [torch.nn.Parameter(p.detach().clone()) if i < len(shapes) else torch.nn.Parameter(base.clone()[3:])
            for i, p in enumerate(ref)]
    groups = lambda ps: [{"params": ps[:3], "weight_decay": 0.01}, {"params": ps[3:], "weight_decay": 0.0}]   # noqa: E731
    o_ref = torch.optim.AdamW(groups(ref), lr=3e-3, betas=(0.9, 0.98), eps=1e-8)
    o_our = FusedAdamW(groups(ours), lr=3e-3, betas=(0.9, 0.98), eps=1e-8)
    for step in range(5):
        for a, b in zip(ref, ours):
            g = torch.randn_like(a)
            a.grad = g.clone()
            b.grad = g.clone()
        o_ref.step()
        o_our.step()
    for a, b in zip(ref, ours):
        assert torch.allclose(a, b, rtol=2e-6, atol=1e-7), float((a - b).abs().max())
    sd = o_our.state_dict()
    assert set(sd["state"][0].keys()) == {"step", "exp_avg", "exp_avg_sq"} and float(sd["state"][0]["step"]) == 5
    import copy
    o_ref.load_state_dict(copy.deepcopy(sd))         # checkpoints are interchangeable (deep copy: load_state_dict keeps same-device tensors by reference)
    ls = torch.nn.Parameter(torch.tensor(5.0, device="cuda"))
    clamp_logit_scale(ls)
    assert abs(float(ls) - 4.6052) < 1e-6
    # an lr schedule changes lr every iteration (main_lavila_pretrain.py:805-808): the device tables must not be rebuilt
    built = []
    orig = o_our._build_tables
    o_our._build_tables = lambda *a: (built.append(1), orig(*a))[1]
    for step in range(3):
        for grp_r, grp_o in zip(o_ref.param_groups, o_our.param_groups):
            grp_r["lr"] = grp_o["lr"] = 3e-3 * (0.5 + 0.1 * step)
        for a, b in zip(ref, ours):
            b.grad.copy_(torch.randn_like(a))
            a.grad.copy_(b.grad)
        o_ref.step()
        o_our.step()
    assert not built
    for a, b in zip(ref, ours):
        assert torch.allclose(a, b, rtol=4e-6, atol=2e-7), float((a - b).abs().max())
