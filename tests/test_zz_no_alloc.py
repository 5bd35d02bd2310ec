"""avn_infonce_bwd_static: the InfoNCE backward whose d(logit) scratch is a file-scope `static __device__` array, for callers
that may not allocate device memory (include/avion_b200.h; no counterpart in the reference).

Called through the C-ABI (ctypes) and compared with avn_infonce_bwd on the same inputs, twice in a row: the second call is
SMALLER than the first, so the shared scratch still holds the first call's values around the region it rewrites.
Collected last on purpose (a failure here hides nothing)."""
import ctypes as C

import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu

TOL = 1e-4      # the tolerance of tests/test_kernels_gpu.py::test_infonce_fwd_bwd


def _p(t):
    return C.c_void_p(t.data_ptr())


def _rel(a, b):
    return float((a.double() - b.double()).norm() / b.double().norm().clamp_min(1e-30))


def _case(lib_call, n, N, E, off, col_side, seed):
    g = torch.Generator(device="cuda").manual_seed(seed)
    all_img = F.normalize(torch.randn(N, E, device="cuda", generator=g), dim=-1)
    all_txt = F.normalize(all_img + 0.7 * torch.randn(N, E, device="cuda", generator=g), dim=-1)
    img, txt = all_img[off:off + n].contiguous(), all_txt[off:off + n].contiguous()
    scale = torch.full((1,), 14.3, device="cuda")
    stream = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    lse = torch.empty(2, n, device="cuda")
    out = torch.zeros(2, device="cuda")
    lib_call("avn_infonce_fwd", _p(img), _p(txt), _p(all_img), _p(all_txt), _p(scale), n, N, E, off, _p(lse), _p(out), stream)
    # existing entry point: caller-provided workspace
    dlogits = torch.empty(2, n, N, device="cuda")
    d_ref, ds_ref = torch.zeros(2, N, E, device="cuda"), torch.zeros(1, device="cuda")
    lib_call("avn_infonce_bwd", _p(img), _p(txt), _p(all_img), _p(all_txt), _p(scale), n, N, E, off, _p(lse), 1.0, col_side,
             _p(dlogits), _p(d_ref), _p(ds_ref), stream)
    # new entry point: outputs only; nothing may be allocated on the device during the call
    d_new, ds_new = torch.zeros(2, N, E, device="cuda"), torch.zeros(1, device="cuda")
    torch.cuda.synchronize()
    free_before = torch.cuda.mem_get_info()[0]
    lib_call("avn_infonce_bwd_static", _p(img), _p(txt), _p(all_img), _p(all_txt), _p(scale), n, N, E, off, _p(lse), 1.0, col_side,
             _p(d_new), _p(ds_new), stream)
    torch.cuda.synchronize()
    free_after = torch.cuda.mem_get_info()[0]
    return d_ref, ds_ref, d_new, ds_new, free_before, free_after


def test_infonce_bwd_static_matches_infonce_bwd_twice_in_a_row():
    from avion_b200 import _lib
    lib = _lib.load()
    max_floats = lib.avn_infonce_static_max_floats()
    assert max_floats == 1 << 20
    # first call fills the whole scratch (2*256*2048 floats = the maximum); the second, smaller one runs over its leftovers
    for i, (n, N, E, off, col_side) in enumerate([(256, 2048, 256, 512, 1), (40, 320, 64, 80, 1), (40, 320, 64, 80, 0)]):
        assert 2 * n * N <= max_floats
        d_ref, ds_ref, d_new, ds_new, free_before, free_after = _case(_lib.call, n, N, E, off, col_side, seed=10 + i)
        assert float(d_ref.abs().max()) > 0
        assert _rel(d_new, d_ref) < TOL, (i, _rel(d_new, d_ref))
        assert abs(float(ds_new) - float(ds_ref)) < TOL * max(1.0, abs(float(ds_ref))), (i, float(ds_new), float(ds_ref))
        if i > 0:       # (the first call on a device may load the module lazily; later calls must leave device memory alone)
            assert free_after >= free_before, (i, free_before, free_after)


def test_infonce_bwd_static_rejects_sizes_beyond_the_scratch():
    from avion_b200 import _lib
    lib = _lib.load()
    n, N, E = 512, 2048, 64                                     # 2*n*N = 2^21 floats > 2^20
    z = torch.zeros(N, E, device="cuda")
    lse, scale = torch.zeros(2, n, device="cuda"), torch.ones(1, device="cuda")
    d_all, ds = torch.zeros(2, N, E, device="cuda"), torch.zeros(1, device="cuda")
    stream = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    with pytest.raises(RuntimeError, match="exceeds the static scratch"):
        _lib.call("avn_infonce_bwd_static", _p(z), _p(z), _p(z), _p(z), _p(scale), n, N, E, 0, _p(lse), 1.0, 1, _p(d_all), _p(ds),
                  stream)
    torch.cuda.synchronize()
    assert float(d_all.abs().max()) == 0.0 and float(ds) == 0.0      # nothing was launched
