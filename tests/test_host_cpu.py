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
