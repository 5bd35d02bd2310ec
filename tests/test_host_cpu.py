"""CPU-side checks: the C-ABI library loads and exports every symbol include/avion_b200.h declares, the host RRC
sampler equals the literal restatement of decord's sampler bit for bit, and the product path refuses CPU tensors."""
import ctypes as C
import os
import re

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def built():
    import __graft_entry__ as g
    g.build()
    return g


def test_library_exports_every_declared_symbol(built):
    from avion_b200 import _lib
    lib = _lib.load()
    header = open(os.path.join(ROOT, "include", "avion_b200.h")).read()
    declared = set(re.findall(r"\b(avn_[a-z0-9_]+)\s*\(", header))
    declared -= {"avn_gemm_desc", "avn_stream_t"}
    assert len(declared) >= 20
    for name in sorted(declared):
        assert hasattr(lib, name), name
    assert set(_lib.SIGNATURES) <= declared
    assert lib.avn_version() >= 100
    import ctypes
    assert lib.avn_sizeof_gemm_desc() == ctypes.sizeof(_lib.GemmDesc)      # the ctypes mirror matches the header's struct


def test_rrc_sampler_matches_restatement(built):
    """Bit-exact crop boxes vs the literal C++ restatement of decord's sampler (oracle/rrc_ref.cpp)."""
    from avion_b200 import ops
    lib = C.CDLL(os.path.join(ROOT, "oracle", "_build", "librrc_oracle.so"))
    lib.rrc_ref.argtypes = [C.c_uint64, C.c_int, C.c_int, C.c_double, C.c_double, C.c_double, C.c_double, C.c_double, C.c_void_p]
    n_fallback = 0
    for (W, H, hp) in [(384, 288, 0.0), (320, 240, 0.5), (100, 1000, 0.5), (1000, 90, 0.0), (224, 224, 1.0)]:
        got = ops.sample_rrc_boxes(42, 64, W, H, hflip_prob=hp)
        for i in range(64):
            out = (C.c_int32 * 5)()
            lib.rrc_ref(42 + i, W, H, 0.5, 1.0, 3 / 4, 4 / 3, hp, out)
            assert list(out) == got[i].tolist()
            x, y, w, h, _ = got[i].tolist()
            assert 0 < w <= W and 0 < h <= H and x >= 0 and y >= 0
            n_fallback += int(w == W or h == H)
    assert n_fallback > 0          # the extreme aspect ratios exercise the centre-crop fallback


def test_rrc_sampler_rejects_bad_arguments(built):
    from avion_b200 import ops
    with pytest.raises(RuntimeError):
        ops.sample_rrc_boxes(0, 4, 0, 10)


def test_ops_refuse_cpu_tensors(built):
    """No CPU fallback: the product path raises instead of computing on the host."""
    from avion_b200 import ops
    x = torch.zeros(4, 128, dtype=torch.bfloat16)
    g = torch.ones(128)
    with pytest.raises(RuntimeError, match="CUDA"):
        ops.layernorm_fwd(x, g, g, 1e-6)


def _fake_open_clip_sd(layers=2, width=16, t_width=8, ctx=5, vocab=11, prefix=""):
    import torch
    g = torch.Generator().manual_seed(0)
    r = lambda *s: torch.randn(*s, generator=g)      # noqa: E731
    sd = {"logit_scale": r(()), "visual.proj": r(width, 4), "positional_embedding": r(ctx, t_width), "text_projection": r(t_width, 4),
          "token_embedding.weight": r(vocab, t_width), "ln_final.weight": r(t_width), "ln_final.bias": r(t_width),
          "visual.conv1.weight": r(width, 3, 2, 2), "visual.class_embedding": r(width), "visual.positional_embedding": r(5, width),
          "visual.ln_pre.weight": r(width)}
    for i in range(layers):
        for n, shape in (("attn.in_proj_weight", (3 * width, width)), ("attn.in_proj_bias", (3 * width,)),
                         ("attn.out_proj.weight", (width, width)), ("attn.out_proj.bias", (width,)), ("ln_1.weight", (width,)),
                         ("ln_1.bias", (width,)), ("ln_2.weight", (width,)), ("ln_2.bias", (width,)),
                         ("mlp.c_fc.weight", (4 * width, width)), ("mlp.c_fc.bias", (4 * width,)),
                         ("mlp.c_proj.weight", (width, 4 * width)), ("mlp.c_proj.bias", (width,))):
            sd[f"visual.transformer.resblocks.{i}.{n}"] = r(*shape)
            sd[f"transformer.resblocks.{i}.{n}"] = r(*[s // 2 for s in shape])
    return {prefix + k: v for k, v in sd.items()}


def test_checkpoint_remap_matches_reference():
    """avion_b200.models.utils vs avion/models/utils.py on a synthetic OpenCLIP-style state dict (live reference)."""
    import pytest
    import torch
    from oracle import ref_import as R
    from avion_b200.models import utils as U
    # properties that hold without the reference
    out = U.remap_keys_from_open_clip_to_vit({"state_dict": _fake_open_clip_sd(prefix="module.")}, 2, 2, context_length=7,
                                             vocab_size=13, use_fast_conv1=True, use_flash_attn=True)
    assert out["visual.conv1.weight"].shape == (16, 12) and out["textual.positional_embedding"].shape == (7, 8)
    assert out["textual.token_embedding.weight"].shape == (13, 8) and "visual.transformer.resblocks.1.attn.Wqkv.weight" in out
    assert "visual.image_projection" not in out and "logit_scale" not in out
    if not R.available():
        pytest.skip("reference tree only exists in the build container")
    R.import_reference()
    from avion.models import utils as RU
    for kw in (dict(use_fast_conv1=True, use_flash_attn=True, context_length=7, vocab_size=13),
               dict(use_fast_conv1=False, use_flash_attn=False, context_length=3, vocab_size=11)):
        a = U.remap_keys_from_open_clip_to_vit(_fake_open_clip_sd(prefix="module."), 2, 2, **kw)
        b = RU.remap_keys_from_open_clip_to_vit(_fake_open_clip_sd(prefix="module."), 2, 2, **kw)
        assert list(a) == list(b) and all(torch.equal(a[k], b[k]) for k in a)
    cur = {"visual.temporal_embedding": torch.zeros(8, 6), "visual.positional_embedding": torch.zeros(5, 6)}
    for frames, fix in ((8, "bilinear"), (8, "interp"), (2, "bilinear")):
        new = lambda: {"visual.temporal_embedding": torch.arange(24.).view(4, 6), "visual.positional_embedding": torch.ones(5, 6)}  # noqa: E731
        a = U.inflate_positional_embeds(cur, new(), num_frames=frames, load_temporal_fix=fix)
        b = RU.inflate_positional_embeds(cur, new(), num_frames=frames, load_temporal_fix=fix)
        assert all(torch.equal(a[k], b[k]) for k in b)
