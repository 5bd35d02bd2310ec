#!/usr/bin/env python
"""Same-box comparators: this package's kernels against what the reference runs on the SAME B200, interleaved in one process.

SURVEY.md 2b K6 / 8(d): the reference's attention is flash_attn_qkvpacked_func (flash-attn 2.8.3, sm_100 SASS = FA2 mma.sync
kernels; avion/models/transformer.py:114-120,148), its GEMMs are cuBLAS(Lt) through nn.Linear (:135,227,320) with separate
ATen bias / GELU / residual passes, its LayerNorm is ATen's.  Arms are timed with CUDA events, alternating A/B/A/B inside each
repetition so that clock drift hits both alike; every operand set is larger than L2 or rotated.  nvidia-smi clocks are sampled
during the run.  Prints a table and writes gpurun_out/same_box.json.

  python tools/compare_same_box.py [--quick] [--skip-block]
"""
import argparse
import json
import os
import subprocess
import sys
import threading
from functools import partial

import torch
import torch.nn.functional as F

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from avion_b200 import ops  # noqa: E402
from avion_b200.models import _engine as E  # noqa: E402

bf16, f32 = torch.bfloat16, torch.float32
DEV = "cuda"


def time_arms(arms, reps=5, inner=3, warm=2):
    """arms: {name: fn}.  Returns {name: best-of-reps mean ms over `inner` calls}, arms interleaved within every repetition."""
    for fn in arms.values():
        for _ in range(warm):
            fn()
    torch.cuda.synchronize()
    best = {k: float("inf") for k in arms}
    for _ in range(reps):
        for name, fn in arms.items():
            s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            s.record()
            for _ in range(inner):
                fn()
            e.record()
            torch.cuda.synchronize()
            best[name] = min(best[name], s.elapsed_time(e) / inner)
    return best


class Clocks:
    def __init__(self):
        self.lines = []
        try:
            self.p = subprocess.Popen(["nvidia-smi", "-i", "0", "--query-gpu=clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active",
                                       "--format=csv,noheader,nounits", "-lms", "500"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=lambda: [self.lines.append(ln.strip()) for ln in self.p.stdout], daemon=True).start()
        except Exception:
            self.p = None

    def stop(self):
        if self.p is None:
            return {}
        self.p.terminate()
        sm = sorted(float(ln.split(",")[0]) for ln in self.lines if ln and ln.split(",")[0].strip().replace(".", "").isdigit())
        return {"sm_mhz_median": sm[len(sm) // 2] if sm else None, "samples": len(sm)}


def attention_rows(quick):
    from flash_attn import flash_attn_qkvpacked_func
    shapes = [("ViT-B/16 16f  B=32", 32, 3137, 12, False), ("ViT-B/16 16f  B=256", 256, 3137, 12, False),
              ("ViT-L/14 16f  B=16", 16, 4097, 16, False), ("VideoMAE dec  B=64", 64, 1568, 6, False),
              ("VideoMAE enc  B=64", 64, 160, 12, False), ("text causal   B=256", 256, 77, 8, True)]
    if quick:
        shapes = shapes[:1] + shapes[3:]
    rows = []
    for name, B, S, H, causal in shapes:
        torch.manual_seed(0)
        qkv = (torch.randn(B, S, 3 * H * 64, device=DEV) * 0.5).to(bf16)
        dout = (torch.randn(B, S, H * 64, device=DEV) * 0.1).to(bf16)
        o = torch.empty(B, S, H * 64, dtype=bf16, device=DEV)
        dqkv = torch.empty_like(qkv)
        dq_acc = torch.empty(B, S, H * 64, dtype=f32, device=DEV)
        q5 = qkv.view(B, S, 3, H, 64).detach().requires_grad_(True)
        state = {}

        def ours_fwd():
            state["o"], state["lse"] = ops.attn_fwd(qkv, H, 0.125, causal, out=o)

        def ours_fwdbwd():
            oo, lse = ops.attn_fwd(qkv, H, 0.125, causal, out=o)
            ops.attn_bwd(qkv, oo, dout, lse, H, 0.125, causal, dqkv=dqkv, dq_acc=dq_acc)

        def fa_fwd():
            with torch.no_grad():
                flash_attn_qkvpacked_func(q5, 0.0, 0.125, causal)

        def fa_fwdbwd():
            out = flash_attn_qkvpacked_func(q5, 0.0, 0.125, causal)
            out.backward(dout.view(B, S, H, 64))
            q5.grad = None

        t = time_arms({"ours_fwd": ours_fwd, "fa2_fwd": fa_fwd, "ours_fwdbwd": ours_fwdbwd, "fa2_fwdbwd": fa_fwdbwd})
        fl = (2.0 if causal else 4.0) * B * H * S * S * 64
        rows.append({"case": "attention " + name, "S": S, "H": H, "B": B, "causal": causal,
                     "ours_fwd_ms": t["ours_fwd"], "ref_fwd_ms": t["fa2_fwd"], "ours_fwdbwd_ms": t["ours_fwdbwd"], "ref_fwdbwd_ms": t["fa2_fwdbwd"],
                     "ours_fwd_tflops": fl / t["ours_fwd"] / 1e9, "ref_fwd_tflops": fl / t["fa2_fwd"] / 1e9,
                     "ours_fwdbwd_tflops": 3.5 * fl / t["ours_fwdbwd"] / 1e9, "ref_fwdbwd_tflops": 3.5 * fl / t["fa2_fwdbwd"] / 1e9,
                     "ratio_fwd": t["fa2_fwd"] / t["ours_fwd"], "ratio_fwdbwd": t["fa2_fwdbwd"] / t["ours_fwdbwd"]})
        del qkv, dout, o, dqkv, dq_acc, q5
        torch.cuda.empty_cache()
    return rows


def gemm_rows(quick):
    M = 803072 if not quick else 100352          # 256 (32) clips x 3137 tokens
    D, Hd = 768, 3072
    torch.manual_seed(1)
    x = (torch.randn(M, D, device=DEV) * 0.5).to(bf16)
    xh = (torch.randn(M, Hd, device=DEV) * 0.5).to(bf16)
    res = (torch.randn(M, D, device=DEV) * 0.5).to(bf16)
    rows = []

    def w_(n, k):
        return (torch.randn(n, k, device=DEV) * 0.02).to(bf16)

    def add(case, ours, ref, flops):
        t = time_arms({"ours": ours, "ref": ref})
        rows.append({"case": case, "ours_ms": t["ours"], "ref_ms": t["ref"], "ours_tflops": flops / t["ours"] / 1e9,
                     "ref_tflops": flops / t["ref"] / 1e9, "ratio": t["ref"] / t["ours"]})

    # ---- forward (NT) with the epilogue the block needs; the reference pays a separate ATen pass for GELU / residual
    wq, bq = w_(3 * D, D), torch.randn(3 * D, device=DEV) * 0.02
    yq = torch.empty(M, 3 * D, dtype=bf16, device=DEV)
    add("gemm NT qkv   [M,768]x[2304,768] +bias", lambda: ops.linear_fwd(x, wq, bq, ops.EPI_BIAS, out=yq),
        lambda: F.linear(x, wq, bq.to(bf16)), 2.0 * M * 3 * D * D)
    wo, bo = w_(D, D), torch.randn(D, device=DEV) * 0.02
    yo = torch.empty(M, D, dtype=bf16, device=DEV)
    add("gemm NT out   [M,768]x[768,768] +bias+residual", lambda: ops.linear_fwd(x, wo, bo, ops.EPI_BIAS_RESIDUAL, residual=res, out=yo),
        lambda: res + F.linear(x, wo, bo.to(bf16)), 2.0 * M * D * D)
    w1, b1 = w_(Hd, D), torch.randn(Hd, device=DEV) * 0.02
    y1 = torch.empty(M, Hd, dtype=bf16, device=DEV)
    add("gemm NT fc1   [M,768]x[3072,768] +bias+GELU", lambda: ops.linear_fwd(x, w1, b1, ops.EPI_BIAS_GELU, out=y1),
        lambda: F.gelu(F.linear(x, w1, b1.to(bf16))), 2.0 * M * Hd * D)
    w2, b2 = w_(D, Hd), torch.randn(D, device=DEV) * 0.02
    add("gemm NT fc2   [M,3072]x[768,3072] +bias+residual", lambda: ops.linear_fwd(xh, w2, b2, ops.EPI_BIAS_RESIDUAL, residual=res, out=yo),
        lambda: res + F.linear(xh, w2, b2.to(bf16)), 2.0 * M * Hd * D)
    # ---- dgrad (NN)
    dyq = yq
    dx = torch.empty(M, D, dtype=bf16, device=DEV)
    add("gemm NN dqkv  [M,2304]x[2304,768]", lambda: ops.linear_dgrad(dyq, wq, out=dx), lambda: torch.matmul(dyq, wq), 2.0 * M * 3 * D * D)
    add("gemm NN dfc1  [M,3072]x[3072,768]", lambda: ops.linear_dgrad(xh, w1, out=dx), lambda: torch.matmul(xh, w1), 2.0 * M * Hd * D)
    pre = y1
    dxh = torch.empty(M, Hd, dtype=bf16, device=DEV)
    add("gemm NN dfc2  [M,768]x[768,3072] * gelu'", lambda: ops.linear_dgrad(x, w2, ops.EPI_DGELU, pre=pre, out=dxh),
        lambda: torch.ops.aten.gelu_backward(torch.matmul(x, w2), pre), 2.0 * M * Hd * D)      # pre = the bf16 pre-activation
    # ---- wgrad (TN), fp32 accumulate into the master gradient
    dwq = torch.zeros(3 * D, D, device=DEV)
    add("gemm TN wqkv  [2304,M]x[M,768] -> fp32 +=", lambda: ops.linear_wgrad(dyq, x, dwq), lambda: dwq.add_(torch.matmul(dyq.t(), x)), 2.0 * M * 3 * D * D)
    dw1 = torch.zeros(Hd, D, device=DEV)
    add("gemm TN wfc1  [3072,M]x[M,768] -> fp32 +=", lambda: ops.linear_wgrad(xh, x, dw1), lambda: dw1.add_(torch.matmul(xh.t(), x)), 2.0 * M * Hd * D)
    dw2 = torch.zeros(D, Hd, device=DEV)
    add("gemm TN wfc2  [768,M]x[M,3072] -> fp32 +=", lambda: ops.linear_wgrad(x, xh, dw2), lambda: dw2.add_(torch.matmul(x.t(), xh)), 2.0 * M * Hd * D)
    return rows


def layernorm_rows(quick):
    rows = []
    for D, M in ((768, 803072 if not quick else 100352), (1024, 262208)):
        torch.manual_seed(2)
        x = (torch.randn(M, D, device=DEV)).to(bf16)
        g, b = torch.ones(D, device=DEV), torch.zeros(D, device=DEV)
        dy = (torch.randn(M, D, device=DEV)).to(bf16)
        y = torch.empty_like(x)
        dxo = torch.empty_like(x)
        dg, db = torch.zeros(D, device=DEV), torch.zeros(D, device=DEV)
        st = {}

        def ours_fwd():
            _, st["m"], st["r"] = ops.layernorm_fwd(x, g, b, 1e-6, out=y)
        ours_fwd()

        def ours_bwd():
            ops.layernorm_bwd(dy, x, g, st["m"], st["r"], dg, db, out=dxo)
        xr = x.detach().requires_grad_(True)
        gr, br = g.detach().to(bf16).requires_grad_(True), b.detach().to(bf16).requires_grad_(True)

        def ref_fwd():
            with torch.no_grad():
                F.layer_norm(x, (D,), gr, br, 1e-6)

        def ref_fwdbwd():
            out = F.layer_norm(xr, (D,), gr, br, 1e-6)
            out.backward(dy)
            xr.grad = gr.grad = br.grad = None

        def ours_fwdbwd():
            ours_fwd()
            ours_bwd()
        t = time_arms({"ours_fwd": ours_fwd, "ref_fwd": ref_fwd, "ours_fwdbwd": ours_fwdbwd, "ref_fwdbwd": ref_fwdbwd})
        rows.append({"case": f"layernorm [M={M},{D}] bf16 (ATen F.layer_norm, bf16 in/out: torch's best case)",
                     "ours_fwd_ms": t["ours_fwd"], "ref_fwd_ms": t["ref_fwd"], "ours_fwdbwd_ms": t["ours_fwdbwd"], "ref_fwdbwd_ms": t["ref_fwdbwd"],
                     "ours_fwd_gbs": 4.0 * M * D / t["ours_fwd"] / 1e6, "ref_fwd_gbs": 4.0 * M * D / t["ref_fwd"] / 1e6,
                     "ratio_fwd": t["ref_fwd"] / t["ours_fwd"], "ratio_fwdbwd": t["ref_fwdbwd"] / t["ours_fwdbwd"]})
    return rows


def block_rows(quick):
    """The reference's own ResidualAttentionBlock (use_flash_attn=True: FlashMHA + FlashMlp, avion/models/transformer.py:94-150)
    forward + backward under torch.autocast(bfloat16), against the engine's block on the same shape."""
    from oracle import ref_import as R
    if not R.available():
        return [{"case": "reference block", "unavailable": "oracle/_ref not staged"}]
    R.import_reference()
    from avion.models.transformer import ResidualAttentionBlock
    import torch.nn as nn
    B, S, D, H = (32 if not quick else 8), 3137, 768, 12
    torch.manual_seed(3)
    blk = ResidualAttentionBlock(D, H, 4.0, norm_layer=partial(nn.LayerNorm, eps=1e-6), use_flash_attn=True).to(DEV)
    x32 = torch.randn(B, S, D, device=DEV)
    xr = x32.clone().requires_grad_(True)
    g32 = torch.randn(B, S, D, device=DEV) * 0.1

    def ref_step():
        with torch.autocast("cuda", dtype=bf16):
            y = blk(xr)
        y.backward(g32)
        xr.grad = None
        for p_ in blk.parameters():
            p_.grad = None

    def ref_step_ckpt():      # the recipe trains with --grad-checkpointing (transformer.py:185-191): the block runs twice forward
        from torch.utils.checkpoint import checkpoint
        with torch.autocast("cuda", dtype=bf16):
            y = checkpoint(blk, xr, use_reentrant=False)
        y.backward(g32)
        xr.grad = None
        for p_ in blk.parameters():
            p_.grad = None

    sd = blk.state_dict()
    P = dict(ln1_w=sd["ln_1.weight"], ln1_b=sd["ln_1.bias"], qkv_w=sd["attn.Wqkv.weight"], qkv_b=sd["attn.Wqkv.bias"],
             out_w=sd["attn.out_proj.weight"], out_b=sd["attn.out_proj.bias"], ln2_w=sd["ln_2.weight"], ln2_b=sd["ln_2.bias"],
             fc1_w=sd["mlp.fc1.weight"], fc1_b=sd["mlp.fc1.bias"], fc2_w=sd["mlp.fc2.weight"], fc2_b=sd["mlp.fc2.bias"])
    P = {k: v.detach().float().contiguous() for k, v in P.items()}
    ws = E.Workspace()
    xb, gb = x32.to(bf16), g32.to(bf16)

    def ours_step():
        layers = E.cast_layers([P])
        grads = E.zero_grads([P])
        y, saved = E.blocks_forward(xb, layers, H, 1e-6, False, ws, save=True)
        E.blocks_backward(gb, saved, layers, grads, H, 1e-6, False, ws)

    t = time_arms({"ours": ours_step, "ref": ref_step, "ref_ckpt": ref_step_ckpt}, reps=4, inner=2)
    with torch.no_grad():
        with torch.autocast("cuda", dtype=bf16):
            yr = blk(x32)
        yo, _ = E.blocks_forward(xb, E.cast_layers([P]), H, 1e-6, False, ws, save=False)
    cos = F.cosine_similarity(yo.float().reshape(-1), yr.float().reshape(-1), dim=0).item()
    fl = 3 * B * (2.0 * S * D * (3 * D + D + 8 * D) + 4.0 * S * S * D)
    return [{"case": f"block fwd+bwd ViT-B/16 16f B={B} (reference ResidualAttentionBlock, flash-attn 2.8.3 + cuBLAS, bf16 autocast)",
             "ours_ms": t["ours"], "ref_ms": t["ref"], "ref_checkpointed_ms": t["ref_ckpt"], "ours_tflops": fl / t["ours"] / 1e9,
             "ref_tflops": fl / t["ref"] / 1e9, "ratio": t["ref"] / t["ours"], "ratio_vs_recipe_checkpointing": t["ref_ckpt"] / t["ours"],
             "output_cosine": cos}]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--quick", action="store_true")
    ap.add_argument("--skip-block", action="store_true")
    ap.add_argument("--out", default=os.path.join(ROOT, "gpurun_out", "same_box.json"))
    args = ap.parse_args()
    clocks = Clocks()
    rows = []
    for fn in (attention_rows, gemm_rows, layernorm_rows) + (() if args.skip_block else (block_rows,)):
        try:
            rows += fn(args.quick)
        except Exception as ex:      # one comparator failing (e.g. flash-attn refusing a shape) must not lose the others
            rows.append({"case": fn.__name__, "error": repr(ex)[:300]})
        torch.cuda.empty_cache()
    res = {"gpu": torch.cuda.get_device_name(0), "clocks": clocks.stop(), "rows": rows}
    os.makedirs(os.path.dirname(args.out), exist_ok=True)
    json.dump(res, open(args.out, "w"), indent=1)
    print(f"same-box comparison on {res['gpu']}  clocks {res['clocks']}")
    for r in rows:
        if "error" in r or "unavailable" in r:
            print(f"  {r['case']}: {r.get('error') or r.get('unavailable')}")
        elif "ours_fwd_ms" in r:
            print(f"  {r['case']:58s} fwd {r['ours_fwd_ms']:8.3f} ms vs {r['ref_fwd_ms']:8.3f} ms (x{r['ratio_fwd']:.2f})   "
                  f"fwd+bwd {r['ours_fwdbwd_ms']:8.3f} vs {r['ref_fwdbwd_ms']:8.3f} ms (x{r['ratio_fwdbwd']:.2f})")
        else:
            extra = f"  ckpt x{r['ratio_vs_recipe_checkpointing']:.2f} cos {r['output_cosine']:.5f}" if "output_cosine" in r else ""
            print(f"  {r['case']:58s} {r['ours_ms']:8.3f} ms ({r['ours_tflops']:6.0f} TF/s) vs {r['ref_ms']:8.3f} ms ({r['ref_tflops']:6.0f} TF/s)  x{r['ratio']:.2f}{extra}")


if __name__ == "__main__":
    main()
