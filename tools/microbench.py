"""Per-kernel timing at the ViT-B/16 16-frame shapes (CUDA events, inputs larger than L2).  python tools/microbench.py [B] [which]"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from avion_b200 import ops  # noqa: E402

B = int(sys.argv[1]) if len(sys.argv) > 1 else 32
which = sys.argv[2] if len(sys.argv) > 2 else "all"
S, H, D = 3137, 12, 768
dev = "cuda"
bf = torch.bfloat16


def timeit(fn, reps=5, warm=2):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


def report(name, ms, flops=None, bytes_=None):
    s = f"{name:28s} {ms:8.3f} ms"
    if flops:
        s += f"  {flops / ms / 1e9:8.1f} TFLOP/s"
    if bytes_:
        s += f"  {bytes_ / ms / 1e6:8.1f} GB/s"
    print(s, flush=True)


if which in ("all", "attn", "attn_fwd", "attn_bwd"):
    qkv = (torch.randn(B, S, 3 * D, device=dev) * 0.5).to(bf)
    o, lse = ops.attn_fwd(qkv, H, 0.125)
    if which != "attn_bwd":
        report("attn_fwd", timeit(lambda: ops.attn_fwd(qkv, H, 0.125, out=o)), flops=4.0 * B * H * S * S * 64)
    if which != "attn_fwd":
        do = torch.randn_like(o)
        dqkv = torch.empty_like(qkv)
        dq_acc = torch.empty(B, S, D, device=dev)
        report("attn_bwd", timeit(lambda: ops.attn_bwd(qkv, o, do, lse, H, 0.125, dqkv=dqkv, dq_acc=dq_acc)), flops=10.0 * B * H * S * S * 64)

if which in ("all", "gemm"):
    M = B * S
    x = torch.randn(M, D, device=dev).to(bf)
    w1 = (torch.randn(4 * D, D, device=dev) * 0.03).to(bf)
    b1 = torch.randn(4 * D, device=dev)
    h = torch.empty(M, 4 * D, device=dev, dtype=bf)
    pre = torch.empty_like(h)
    report("gemm_nt fc1 bias+gelu", timeit(lambda: ops.linear_fwd(x, w1, b1, ops.EPI_BIAS_GELU, out=h)), flops=2.0 * M * D * 4 * D)
    report("gemm_nt fc1 dual", timeit(lambda: ops.linear_fwd(x, w1, b1, ops.EPI_BIAS_GELU_DUAL, out=h, out2=pre)), flops=2.0 * M * D * 4 * D)
    wq = (torch.randn(3 * D, D, device=dev) * 0.03).to(bf)
    bq = torch.randn(3 * D, device=dev)
    q = torch.empty(M, 3 * D, device=dev, dtype=bf)
    report("gemm_nt qkv bias", timeit(lambda: ops.linear_fwd(x, wq, bq, ops.EPI_BIAS, out=q)), flops=2.0 * M * D * 3 * D)
    w2 = (torch.randn(D, 4 * D, device=dev) * 0.03).to(bf)
    b2 = torch.randn(D, device=dev)
    y = torch.empty(M, D, device=dev, dtype=bf)
    report("gemm_nt fc2 bias+res", timeit(lambda: ops.linear_fwd(h, w2, b2, ops.EPI_BIAS_RESIDUAL, residual=x, out=y)), flops=2.0 * M * D * 4 * D)
    wo = (torch.randn(D, D, device=dev) * 0.03).to(bf)
    report("gemm_nt proj bias+res", timeit(lambda: ops.linear_fwd(x, wo, b2, ops.EPI_BIAS_RESIDUAL, residual=x, out=y)), flops=2.0 * M * D * D)
    report("gemm_nn fc2 dgrad dgelu", timeit(lambda: ops.linear_dgrad(y, w2, ops.EPI_DGELU, pre=pre, out=h)), flops=2.0 * M * D * 4 * D)
    report("gemm_nn fc1 dgrad", timeit(lambda: ops.linear_dgrad(h, w1, out=y)), flops=2.0 * M * D * 4 * D)
    report("gemm_nn qkv dgrad", timeit(lambda: ops.linear_dgrad(q, wq, out=y)), flops=2.0 * M * D * 3 * D)
    dw = torch.zeros(4 * D, D, device=dev)
    report("gemm_tn fc1 wgrad", timeit(lambda: ops.linear_wgrad(h, x, dw)), flops=2.0 * M * D * 4 * D)
    dwo = torch.zeros(D, D, device=dev)
    report("gemm_tn proj wgrad", timeit(lambda: ops.linear_wgrad(y, x, dwo)), flops=2.0 * M * D * D)

if which in ("all", "ln"):
    M = B * S
    x = torch.randn(M, D, device=dev).to(bf)
    g, b = torch.ones(D, device=dev), torch.zeros(D, device=dev)
    y = torch.empty_like(x)
    _, mean, rstd = ops.layernorm_fwd(x, g, b, 1e-6, out=y)
    report("layernorm_fwd", timeit(lambda: ops.layernorm_fwd(x, g, b, 1e-6, out=y, save_stats=False)), bytes_=4.0 * M * D)
    dg, db, dc = (torch.zeros(D, device=dev) for _ in range(3))
    dx = torch.empty_like(x)
    report("layernorm_bwd (+res,+colsum)", timeit(lambda: ops.layernorm_bwd(y, x, g, mean, rstd, dg, db, dres=x, dcolsum=dc, out=dx)), bytes_=8.0 * M * D)
    report("colsum 4D", timeit(lambda: ops.colsum(torch.empty(0, device=dev, dtype=bf) if False else x, dc)), bytes_=2.0 * M * D)

if which in ("all", "transform"):
    Hs, Ws = 288, 384
    clips = torch.randint(0, 256, (B, 16, 224, 224, 3), dtype=torch.uint8, device=dev)
    mean, std = (108.3272985, 116.7460125, 104.09373615000001), (68.5005327, 66.6321579, 70.32316305)
    for p in (16, 14):
        out = ops.clip_transform(clips, 224, p, mean, std)
        report(f"clip_transform p={p}", timeit(lambda: ops.clip_transform(clips, 224, p, mean, std, out=out)), bytes_=float(clips.numel()) + 2.0 * out.numel())
    nb = max(1, B // 2)
    src = torch.randint(0, 256, (nb, 16, Hs, Ws, 3), dtype=torch.uint8, device=dev)
    boxes = ops.sample_rrc_boxes(7, nb, Ws, Hs).to(dev)
    out = ops.clip_resize_transform(src, boxes, 224, 16, mean, std)
    area = float((boxes[:, 2] * boxes[:, 3]).sum()) * 16 * 3
    report("clip_resize_transform p=16 (RRC 0.5-1.0 of 288x384)", timeit(lambda: ops.clip_resize_transform(src, boxes, 224, 16, mean, std, out=out)),
           bytes_=area + 2.0 * out.numel())
