#!/usr/bin/env python
"""Time / profile one GEMM shape of the train step in isolation: python tools/gemm_probe.py MODE M N K [reps]
MODE = nt | nn | tn; shapes in the C = A x B convention of avn_gemm_bf16 (tn: C[M,N] += A[K,M]^T B[K,N], fp32).
Environment: AVN_GEMM_CG=1 forces single-CTA tiles, AVN_GEMM_SPLITS=n forces the split-K count (tn), AVN_GEMM_COLSUM=1 also
asks the tn kernel for the column sums of A (the bias gradient), =2 times a separate colsum launch after the GEMM instead."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from avion_b200 import ops  # noqa: E402

mode, M, N, K = sys.argv[1], int(sys.argv[2]), int(sys.argv[3]), int(sys.argv[4])
reps = int(sys.argv[5]) if len(sys.argv) > 5 else 5
bf = torch.bfloat16
torch.manual_seed(0)
splits = int(os.environ.get("AVN_GEMM_SPLITS", "0"))
if mode == "tn":
    a = (torch.randn(K, M, device="cuda") * 0.1).to(bf)
    b = (torch.randn(K, N, device="cuda") * 0.1).to(bf)
    c = torch.zeros(M, N, device="cuda")
    cs_mode = int(os.environ.get("AVN_GEMM_COLSUM", "0"))
    cs = torch.zeros(M, device="cuda")
    if cs_mode == 2:
        def fn():
            ops.gemm(ops.GEMM_TN, ops.EPI_F32_ADD, M, N, K, a, M, b, N, c, N, splits=splits)
            ops.colsum(a, cs)
    else:
        fn = lambda: ops.gemm(ops.GEMM_TN, ops.EPI_F32_ADD, M, N, K, a, M, b, N, c, N, splits=splits,      # noqa: E731
                              colsum=cs if cs_mode == 1 else None)
elif mode == "nt":
    a = (torch.randn(M, K, device="cuda") * 0.1).to(bf)
    b = (torch.randn(N, K, device="cuda") * 0.1).to(bf)
    c = torch.empty(M, N, dtype=bf, device="cuda")
    epi = os.environ.get("AVN_GEMM_EPI", "none")
    bias = torch.randn(N, device="cuda")
    if epi == "dual":
        c2 = torch.empty(M, N, dtype=bf, device="cuda")
        fn = lambda: ops.gemm(ops.GEMM_NT, ops.EPI_BIAS_GELU_DUAL, M, N, K, a, K, b, K, c, N, C2=c2, bias=bias)      # noqa: E731
    elif epi == "gelu":
        fn = lambda: ops.gemm(ops.GEMM_NT, ops.EPI_BIAS_GELU, M, N, K, a, K, b, K, c, N, bias=bias)      # noqa: E731
    else:
        fn = lambda: ops.gemm(ops.GEMM_NT, ops.EPI_NONE, M, N, K, a, K, b, K, c, N)      # noqa: E731
else:
    a = (torch.randn(M, K, device="cuda") * 0.1).to(bf)
    b = (torch.randn(K, N, device="cuda") * 0.1).to(bf)
    c = torch.empty(M, N, dtype=bf, device="cuda")
    if os.environ.get("AVN_GEMM_EPI") == "dgelu":      # C = (A B) * R, R = gelu'(pre-activation) [M,N]
        r = torch.rand(M, N, device="cuda").to(bf)
        fn = lambda: ops.gemm(ops.GEMM_NN, ops.EPI_DGELU, M, N, K, a, K, b, N, c, N, R=r, ldr=N)      # noqa: E731
    else:
        fn = lambda: ops.gemm(ops.GEMM_NN, ops.EPI_NONE, M, N, K, a, K, b, N, c, N)      # noqa: E731
for _ in range(2):
    fn()
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
best = 1e9
for _ in range(reps):
    e0.record(); fn(); e1.record(); torch.cuda.synchronize()
    best = min(best, e0.elapsed_time(e1))
print(f"gemm {mode} M={M} N={N} K={K} cg={os.environ.get('AVN_GEMM_CG', 'auto')} splits={splits or 'auto'} "
      f"colsum={os.environ.get('AVN_GEMM_COLSUM', '0')}: {best:.3f} ms  "
      f"{2.0 * M * N * K / best / 1e9:.0f} TFLOP/s", flush=True)
