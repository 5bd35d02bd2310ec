"""The C-ABI kernels as torch custom operators (`torch.ops.avion_b200.*`, avion_b200/torch_ops.py).

CPU part (no GPU needed): every kernel entry point of include/avion_b200.h has a registered operator with an out-variant
schema and a fake implementation; a residual block composed of the functional operators propagates shapes through
FakeTensorMode (forward and backward) on fake CUDA tensors - which is what torch.compile's tracer needs.
GPU part: torch.compile(fullgraph=True) of that block runs the kernels and agrees with the hand-scheduled engine."""
import os
import re

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
bf16, f32 = torch.bfloat16, torch.float32


def _header_kernels():
    src = open(os.path.join(ROOT, "include", "avion_b200.h")).read()
    names = set(re.findall(r"\bint\s+avn_(\w+)\s*\(", src))
    host_only = {"version", "device_check", "sample_rrc_boxes", "adamw_chunk_elems", "mse_workspace_floats", "mmr_workspace_floats",
                 "infonce_static_max_floats", "sizeof_gemm_desc"}
    return names - host_only


def test_every_kernel_entry_point_is_a_registered_op():
    from avion_b200 import torch_ops as T
    ns = torch.ops.avion_b200
    declared = _header_kernels()
    registered = set(T.KERNELS) | {"gemm_bf16"}
    assert declared == registered, (declared - registered, registered - declared)
    for name in registered:
        op = getattr(ns, name).default
        schema = str(op._schema)
        assert schema.endswith("-> ()"), schema
        # out-variant: mutates its result buffers (adamw_step updates its tensors through device POINTER TABLES, which a
        # schema cannot annotate; it is only ever called eagerly from FusedAdamW.step)
        assert "!" in schema or name == "adamw_step", schema
    for name in ("layernorm", "linear", "attention", "clip_transform", "clip_loss"):
        assert hasattr(ns, name)


def _fake_block_params(D, dev):
    p = {}
    for k, shape, dt in (("ln1_w", (D,), f32), ("ln1_b", (D,), f32), ("qkv_w", (3 * D, D), bf16), ("qkv_b", (3 * D,), f32),
                         ("out_w", (D, D), bf16), ("out_b", (D,), f32), ("ln2_w", (D,), f32), ("ln2_b", (D,), f32),
                         ("fc1_w", (4 * D, D), bf16), ("fc1_b", (4 * D,), f32), ("fc2_w", (D, 4 * D), bf16), ("fc2_b", (D,), f32)):
        p[k] = torch.empty(shape, dtype=dt, device=dev, requires_grad=True)
    return p


def test_block_traces_with_fake_cuda_tensors():
    """Fake implementations propagate shapes/dtypes through a whole block and its backward without a GPU or the library."""
    from torch._subclasses.fake_tensor import FakeTensorMode
    from avion_b200 import torch_ops as T
    with FakeTensorMode():
        x = torch.empty(2, 197, 128, dtype=bf16, device="cuda", requires_grad=True)
        p = _fake_block_params(128, "cuda")
        y = T.block_forward(x, p, heads=2)
        assert y.shape == x.shape and y.dtype == bf16 and y.device.type == "cuda"
        # (the autograd engine needs a real CUDA context to run a backward on CUDA tensors, so on a GPU-less box the backward
        # operators are traced directly; the GPU test below differentiates through the registered formulas for real)
        ns = torch.ops.avion_b200
        dx, dg, db = ns.layernorm_backward(y.detach(), x.detach(), p["ln1_w"].detach(), torch.empty(394, device="cuda"), torch.empty(394, device="cuda"))
        assert dx.shape == x.shape and dg.shape == (128,) and db.dtype == f32
        dxl, dw, dbias = ns.linear_backward(torch.empty(2, 197, 512, dtype=bf16, device="cuda"), x.detach(), p["fc1_w"].detach(),
                                            p["fc1_b"].detach(), 1, True)
        assert dxl.shape == x.shape and dw.shape == (512, 128) and dw.dtype == f32 and dbias.shape == (512,)
        qkv = torch.empty(2, 197, 384, dtype=bf16, device="cuda")
        o, lse = ns.attention(qkv, 2, 0.125)
        assert ns.attention_backward(qkv, o, o, lse, 2, 0.125, False).shape == qkv.shape and lse.shape == (2, 2, 197)
        frames = torch.empty(2, 4, 64, 64, 3, dtype=torch.uint8, device="cuda")
        patches = torch.ops.avion_b200.clip_transform(frames, 64, 16, [1., 2., 3.], [1., 1., 1.])
        assert patches.shape == (2, 64, 768) and patches.dtype == bf16
        img = torch.empty(6, 64, device="cuda", requires_grad=True)
        loss, acc, _ = torch.ops.avion_b200.clip_loss(img, img, torch.empty((), device="cuda"))
        assert loss.shape == () and acc.shape == ()


def test_kernel_ops_have_no_cpu_implementation():
    """No CPU fallback behind the operators either: the dispatcher refuses CPU tensors."""
    x = torch.zeros(4, 128, dtype=bf16)
    g = torch.ones(128)
    with pytest.raises((NotImplementedError, RuntimeError)):
        torch.ops.avion_b200.layernorm_fwd(x, g, g, torch.empty_like(x), None, None, 4, 128, 1e-6)


# ---------------------------------------------------------------------------------------------------------------------
@pytest.mark.gpu
def test_compiled_block_matches_engine_and_reference():
    from avion_b200 import torch_ops as T
    from avion_b200.models import _engine as E
    torch.manual_seed(0)
    B, S, D, H = 2, 197, 128, 2
    dev = "cuda"
    x = (torch.randn(B, S, D, device=dev) * 0.5).to(bf16)
    shapes = dict(ln1_w=(D,), ln1_b=(D,), qkv_w=(3 * D, D), qkv_b=(3 * D,), out_w=(D, D), out_b=(D,), ln2_w=(D,), ln2_b=(D,),
                  fc1_w=(4 * D, D), fc1_b=(4 * D,), fc2_w=(D, 4 * D), fc2_b=(D,))
    master = {k: (1.0 + 0.1 * torch.randn(s, device=dev) if k in ("ln1_w", "ln2_w") else torch.randn(s, device=dev) * (0.05 if len(s) == 2 else 0.02))
              for k, s in shapes.items()}
    p = {k: (v.to(bf16) if v.dim() == 2 else v.clone()).requires_grad_(True) for k, v in master.items()}
    xg = x.clone().requires_grad_(True)
    eager = T.block_forward(xg, p, H)
    compiled = torch.compile(T.block_forward, fullgraph=True)(xg, p, H)
    assert torch.equal(eager, compiled)                                   # same kernels, same order: bit-identical
    # the engine's hand-scheduled block on the same weights
    L = {k: v.detach() for k, v in p.items()}
    y_engine, _ = E.blocks_forward(x, [L], H, 1e-6, False, E.Workspace(), save=False)
    assert torch.equal(eager.detach(), y_engine)
    # fp32 reference + gradients through the registered autograd formulas
    import torch.nn.functional as F
    xr = x.float().requires_grad_(True)
    pr = {k: v.detach().float().requires_grad_(True) for k, v in p.items()}
    h = F.layer_norm(xr, (D,), pr["ln1_w"], pr["ln1_b"], 1e-6)
    qkv = F.linear(h, pr["qkv_w"], pr["qkv_b"]).view(B, S, 3, H, 64)
    q, k_, v = (qkv[:, :, i].transpose(1, 2) for i in range(3))
    o = (torch.softmax(q @ k_.transpose(-1, -2) * 0.125, -1) @ v).transpose(1, 2).reshape(B, S, D)
    xm = xr + F.linear(o, pr["out_w"], pr["out_b"])
    h2 = F.layer_norm(xm, (D,), pr["ln2_w"], pr["ln2_b"], 1e-6)
    ref = xm + F.linear(F.gelu(F.linear(h2, pr["fc1_w"], pr["fc1_b"])), pr["fc2_w"], pr["fc2_b"])
    cos = F.cosine_similarity(eager.detach().float().reshape(-1), ref.detach().reshape(-1), dim=0).item()
    assert cos > 0.9999
    w = torch.randn_like(ref)
    (ref * w).sum().backward()
    (eager.float() * w).sum().backward()
    for name, got, want in [("x", xg.grad, xr.grad)] + [(k, p[k].grad, pr[k].grad) for k in shapes]:
        c = F.cosine_similarity(got.float().reshape(-1), want.reshape(-1), dim=0).item()
        assert c > 0.99, (name, c)


@pytest.mark.gpu
def test_opcheck_functional_ops():
    """torch.library.opcheck: schema, fake-tensor and autograd registration of the functional operators are consistent."""
    torch.manual_seed(1)
    x = (torch.randn(2, 64, 128, device="cuda") * 0.5).to(bf16).requires_grad_(True)
    g = torch.ones(128, device="cuda", requires_grad=True)
    b = torch.zeros(128, device="cuda", requires_grad=True)
    tests = ("test_schema", "test_faketensor", "test_autograd_registration")
    torch.library.opcheck(torch.ops.avion_b200.layernorm.default, (x, g, b, 1e-6), test_utils=tests)
    qkv = (torch.randn(2, 64, 3 * 128, device="cuda") * 0.5).to(bf16).requires_grad_(True)
    torch.library.opcheck(torch.ops.avion_b200.attention.default, (qkv, 2, 0.125, False), test_utils=tests)
    y = torch.empty_like(x)
    torch.library.opcheck(torch.ops.avion_b200.layernorm_fwd.default, (x.detach(), g.detach(), b.detach(), y, None, None, 128, 128, 1e-6),
                          test_utils=("test_schema", "test_faketensor"))
