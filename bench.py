#!/usr/bin/env python
"""Benchmark of the video-CLIP train step (BASELINE.json: clips/sec, ViT-B/16 16-frame CLIP train step on B200).

  python bench.py --gpus 1 --steps K --warmup W              # our arm (N>1: launched by torch.distributed.run)
  python bench.py --impl reference --steps K --warmup W      # the reference algorithm on the host CPU cores (oracle port)

One step = fused clip transform -> video tower + text tower -> ClipLoss (all-gather + InfoNCE) -> backward ->
gradient all-reduce (DDP) -> AdamW step + logit_scale clamp (scripts/main_lavila_pretrain.py:793-880).
Prints ONE JSON line (rank 0).
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import torch

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
# stdout carries exactly ONE JSON line.  NCCL prints its version banner to file descriptor 1 whatever NCCL_DEBUG_FILE says, so
# fd 1 is pointed at stderr for the whole process and the JSON line goes out through a private duplicate of the real stdout.
_REAL_STDOUT = os.fdopen(os.dup(1), "w")
os.dup2(2, 1)


def emit(line):
    _REAL_STDOUT.write(json.dumps(line) + "\n")
    _REAL_STDOUT.flush()

# algorithmic FLOPs per clip-text pair, forward; train step = 3x (no recompute counted)
FWD_GFLOP_PER_PAIR = {"vitb16_16f": 905.3, "vitb16_4f": 162.95, "vitl14_16f": 4142.8,
                      "videomae_16f": 70.9}      # per clip: encoder 31.82 + enc->dec 0.09 + decoder (4 L) 37.30 + head 1.66


def peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        p = json.load(open(path))
        return dict(hbm=p["hbm_gbs"], tf_burst=p["bf16_tflops"], tf_sust=p.get("bf16_tflops_sustained", p["bf16_tflops"]),
                    source="measured")
    return dict(hbm=6650.0, tf_burst=1590.0, tf_sust=1400.0, source="fallback")


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled during the timed region."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.proc, self.lines = index, None, []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q, "--format=csv,noheader,nounits",
                                          "-lms", "200"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.th = threading.Thread(target=self._read, daemon=True)
            self.th.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 8:
                continue
            try:
                sm.append(float(f[1])); mx.append(float(f[2]))
            except ValueError:
                continue
            for nm, v in zip(names, f[4:8]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None, "reasons": sorted(reasons),
                "samples": len(sm)}


def param_groups(model):
    """Weight-decay grouping of scripts/main_lavila_pretrain.py:296-320."""
    wd, no_wd = [], []
    for n, p in model.named_parameters():
        if not p.requires_grad:
            continue
        if p.ndim < 2 or "bias" in n or "ln" in n or "bn" in n or "pos_embed" in n or "positional_embedding" in n:
            no_wd.append(p)
        else:
            wd.append(p)
    return [{"params": wd, "weight_decay": 0.01}, {"params": no_wd, "weight_decay": 0.0}]


def synth_batch(cfg, batch, seed, pin):
    g = torch.Generator().manual_seed(seed)
    clips = torch.randint(0, 256, (batch, cfg.frames, cfg.image_size, cfg.image_size, 3), dtype=torch.uint8, generator=g)
    sot, eot = cfg.vocab - 2, cfg.vocab - 1
    tokens = torch.zeros(batch, cfg.ctx, dtype=torch.int64)
    lens = torch.randint(5, cfg.ctx - 1, (batch,), generator=g)
    body = torch.randint(1, sot, (batch, cfg.ctx), generator=g)
    for b in range(batch):
        L = int(lens[b])
        tokens[b, 0] = sot
        tokens[b, 1:L] = body[b, 1:L]
        tokens[b, L] = eot
    if pin:
        clips, tokens = clips.pin_memory(), tokens.pin_memory()
    return clips, tokens


# =====================================================================================================================
def cpu_reference_clips_per_s(cfg, pairs, steps, warmup, threads):
    """The reference's CPU path on the host cores: forward + ClipLoss + backward + AdamW + clamp on `pairs` clip-text pairs per
    step (scripts/main_lavila_pretrain.py:814-880), fp32.  Runs the UNMODIFIED reference modules (avion.models.model_clip.CLIP with
    use_flash_attn=False - flash-attn has no CPU kernels -, avion.losses.ClipLoss) from oracle/_ref when they were staged
    (oracle/stage_ref.py; kind "reference"), else the oracle port (kind "port").  Returns (clips/s, ms/step, kind)."""
    from oracle import clip_oracle as O
    from oracle import ref_import as R
    torch.set_num_threads(threads)
    clips, tokens = O.make_inputs(cfg, pairs, 1234)
    if R.available():
        model = R.build_reference_clip(cfg, O.make_state_dict(cfg, 0))
        crit = R.import_reference()[3](world_size=1)
        params = list(model.parameters())

        def fwd():
            img, txt, scale = model(O.clip_transform(clips), tokens)      # Permute + Normalize (kornia, absent) restated; rest is the reference
            return crit(img, txt, scale)["loss"]

        def clamp():
            model.logit_scale.data.clamp_(0, 4.6052)
        kind = "reference"
    else:
        sd = {k: v.clone().requires_grad_(True) for k, v in O.make_state_dict(cfg, 0).items()}
        params = list(sd.values())

        def fwd():
            return O.train_step_forward(sd, clips, tokens, cfg)["loss"]

        def clamp():
            sd["logit_scale"].data.clamp_(0, 4.6052)
        kind = "port"
    opt = torch.optim.AdamW(params, lr=3e-5, betas=(0.9, 0.999), eps=1e-8, weight_decay=0.01)
    times = []
    for it in range(warmup + steps):
        t0 = time.perf_counter()
        opt.zero_grad(set_to_none=True)
        fwd().backward()
        opt.step()
        clamp()
        dt = time.perf_counter() - t0
        if it >= warmup:
            times.append(dt)
    mean = sum(times) / len(times)
    return pairs / mean, mean * 1e3, kind


def cpu_reference_videomae_clips_per_s(clips_n, steps, warmup, threads):
    """VideoMAE pretraining step (oracle port, fp32) on the host cores: targets + forward + MSE + backward + AdamW."""
    from oracle import videomae_oracle as V
    torch.set_num_threads(threads)
    cfg = V.VITB16
    sd = {k: v.clone().requires_grad_(True) for k, v in V.make_state_dict(cfg, 0).items()}
    opt = torch.optim.AdamW(list(sd.values()), lr=1.5e-4, betas=(0.9, 0.95), eps=1e-8, weight_decay=0.05)
    clips, mask = V.make_inputs(cfg, clips_n, 1234)
    times = []
    for it in range(warmup + steps):
        t0 = time.perf_counter()
        opt.zero_grad(set_to_none=True)
        out = V.train_step_forward(sd, clips, mask, cfg)
        out["loss"].backward()
        opt.step()
        dt = time.perf_counter() - t0
        if it >= warmup:
            times.append(dt)
    mean = sum(times) / len(times)
    return clips_n / mean, mean * 1e3


def run_reference(args):
    from oracle import clip_oracle as O
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    cfg = O.VITB16_16F if args.frames == 16 else O.ClipCfg(frames=args.frames)
    threads = os.cpu_count() or 1
    pairs = args.cpu_pairs
    if args.model == "videomae":
        steps, warmup = min(args.steps, 2), min(args.warmup, 1)
        n = max(pairs, 48)
        v, ms = cpu_reference_videomae_clips_per_s(n, steps, warmup, threads)
        sample = f"{n} clips/step, fp32, targets+fwd+MSE+bwd+AdamW, {steps} timed steps after {warmup} warm-up"
        emit({"impl": "reference", "metric": "clips/sec VideoMAE ViT-B/16 16f pretrain step", "value": v, "unit": "clips/s",
                          "n_gpus": args.gpus, "steps": steps, "warmup": warmup, "ms_per_step": ms, "higher_is_better": True,
                          "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
                          "config": {"workload": "VideoMAE ViT-B/16 16f 224^2 pretrain step, 90% tube masking (oracle port on host CPU)",
                                     "clips_per_step": n},
                          "cpu_baseline": {"value": v, "unit": "clips/s", "cores": threads, "kind": "port", "sample": sample},
                          "e2e": {"value": v, "unit": "clips/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}})
        return
    # bound the whole run to a few minutes: at most 2 timed + 1 warm-up CPU steps of `pairs` pairs (each ~10-40 s)
    steps, warmup = min(args.steps, 2), min(args.warmup, 1)
    v, ms, kind = cpu_reference_clips_per_s(cfg, pairs, steps, warmup, threads)
    what = "the reference's own avion.models CLIP + avion.losses ClipLoss (oracle/_ref)" if kind == "reference" else "oracle port"
    sample = (f"{what}: {pairs} clip-text pairs/step of the arm's config (ViT-B/16, {args.frames} frames, 224^2), fp32, "
              f"fwd+ClipLoss+bwd+AdamW+clamp, {steps} timed steps after {warmup} warm-up")
    line = {"impl": "reference", "metric": "clips/sec ViT-B/16 16f CLIP train step", "value": v, "unit": "clips/s", "n_gpus": args.gpus,
            "steps": steps, "warmup": warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic",
            "config": {"workload": f"AVION CLIP ViT-B/16 {args.frames}f 224^2 full train step on the host CPU ({what}), bounded sample",
                       "pairs_per_step": pairs, "frames": args.frames, "embed_dim": 256},
            "cpu_baseline": {"value": v, "unit": "clips/s", "cores": threads, "kind": kind, "sample": sample},
            "e2e": {"value": v, "unit": "clips/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    emit(line)


# =====================================================================================================================
def run_ours(args):
    import torch.distributed as dist
    from torch.nn.parallel import DistributedDataParallel as DDP
    from avion_b200 import _lib, ops
    from avion_b200.losses import ClipLoss
    from avion_b200.models import CLIP_VITB16, CLIP_VITL14

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    assert world == args.gpus, f"--gpus {args.gpus} but WORLD_SIZE={world}"
    _lib.load()                                     # fail loudly if the CUDA library is missing

    from types import SimpleNamespace
    cfg = SimpleNamespace(frames=args.frames, image_size=224, vocab=49408, ctx=77)
    B = args.batch
    torch.manual_seed(0)
    mae = args.model == "videomae"               # BASELINE.json configs[4]
    if mae:
        return run_videomae(args, world, rank, local_rank, dev)
    factory = CLIP_VITL14 if args.model == "vitl14" else CLIP_VITB16      # vitl14 = BASELINE.json configs[3]
    model = factory(num_frames=args.frames, project_embed_dim=256, pretrain_zoo="random").to(dev)
    with torch.no_grad():
        model.visual.temporal_embedding.normal_(std=0.02)
    use_graph = args.graph == "on"
    # eager: torch DDP around the module (scripts/main_lavila_pretrain.py:283-286).  graph: the bare module, captured once;
    # GraphedStep averages one flat gradient buffer over the ranks after each replay (same reduction, one NCCL call per step)
    net = DDP(model, device_ids=[local_rank], bucket_cap_mb=200, gradient_as_bucket_view=True) if (world > 1 and not use_graph) else model
    crit = ClipLoss(local_loss=False, gather_with_grad=True, cache_labels=True, rank=rank, world_size=world)
    if args.torch_optimizer:
        opt = torch.optim.AdamW(param_groups(model), lr=3e-5, betas=(0.9, 0.999), eps=1e-8, fused=True)
    else:
        from avion_b200.optim import FusedAdamW, clamp_logit_scale
        opt = FusedAdamW(param_groups(model), lr=3e-5, betas=(0.9, 0.999), eps=1e-8)

    host = [synth_batch(cfg, B, 1234 + rank + 1000 * i, pin=True) for i in range(2)]
    resident = [(c.to(dev), t.to(dev)) for c, t in host]
    h2d_bytes = host[0][0].numel() + host[0][1].numel() * 8

    def finish_step():
        opt.step()
        if args.torch_optimizer:
            with torch.no_grad():
                model.logit_scale.clamp_(0, 4.6052)
        else:
            clamp_logit_scale(model.logit_scale)

    def eager_step(clips, tokens):
        opt.zero_grad(set_to_none=True)
        img, txt, scale = net(clips, tokens)
        out = crit(img, txt, scale)
        out["loss"].backward()
        if world > 1 and use_graph:                    # bare module: average the gradients by hand in the eager (warm-up) steps
            from avion_b200.graph import average_module_gradients
            average_module_gradients(model)
        finish_step()
        return out

    step = eager_step

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(ms):
        if world == 1:
            return ms
        t = torch.tensor([ms], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t)

    # ---------------- phase A: inputs resident in HBM (value) ----------------
    for i in range(args.warmup):
        out = step(*resident[i % 2])
    barrier()
    timed = None
    if use_graph:
        # per-family device time and the launch count come from eager steps (events cannot be recorded inside a replay; kernel
        # durations do not depend on how the kernels were launched); then the step is captured and the timed loop replays it
        from avion_b200.graph import GraphedStep
        ops.TIMED = {}
        launches0 = _lib.launch_count
        n_fam = min(args.steps, 2)
        for i in range(n_fam):
            out = eager_step(*resident[i % 2])
        barrier()
        launches = (_lib.launch_count - launches0) * args.steps // n_fam      # the replayed graph launches the same kernels
        timed, ops.TIMED = ops.TIMED, None
        fam_steps = n_fam
        del out
        opt.zero_grad(set_to_none=True)

        def fwd_bwd(clips, tokens):
            img, txt, scale = model(clips, tokens)
            o = crit(img, txt, scale)
            o["loss"].backward()
            return o["loss"].detach(), o["clip_acc"]
        graphed = GraphedStep(fwd_bwd, list(resident[0]), params=list(model.parameters()))

        def step(clips, tokens):      # noqa: F811
            loss, acc = graphed(clips, tokens)
            finish_step()
            return {"loss": loss, "clip_acc": acc}
        for i in range(args.warmup):
            out = step(*resident[i % 2])
        barrier()
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    if not use_graph:
        ops.TIMED = {}
        launches0 = _lib.launch_count
        fam_steps = args.steps
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(args.steps):
        out = step(*resident[i % 2])
    e1.record()
    barrier()
    if not use_graph:
        launches = _lib.launch_count - launches0
        timed, ops.TIMED = ops.TIMED, None
    ms_total = max_over_ranks(e0.elapsed_time(e1))
    clocks = sampler.stop() if rank == 0 else None
    loss_val = float(out["loss"].detach())
    ms_step = ms_total / args.steps
    value = world * B / (ms_step / 1e3)

    # ---------------- phase B: end to end through the public API with host buffers ----------------
    copy_stream = torch.cuda.Stream(dev)
    dbuf = [(torch.empty_like(c, device=dev), torch.empty_like(t, device=dev)) for c, t in host]
    ready = [torch.cuda.Event(), torch.cuda.Event()]
    consumed = [torch.cuda.Event(), torch.cuda.Event()]

    def prefetch(i):
        s = i % 2
        with torch.cuda.stream(copy_stream):
            copy_stream.wait_event(consumed[s])
            dbuf[s][0].copy_(host[s][0], non_blocking=True)
            dbuf[s][1].copy_(host[s][1], non_blocking=True)
            ready[s].record(copy_stream)

    for s in range(2):
        consumed[s].record()
    barrier()
    e2, e3 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e2.record()
    prefetch(0)
    d2h = 0
    for i in range(args.steps):
        if i + 1 < args.steps:
            prefetch(i + 1)                          # next step's H2D overlaps this step's compute
        torch.cuda.current_stream().wait_event(ready[i % 2])
        out = step(*dbuf[i % 2])
        consumed[i % 2].record()
        _ = out["loss"].item()                       # device -> host read of the step's result
        d2h += 4
    e3.record()
    barrier()
    ms_e2e = max_over_ranks(e2.elapsed_time(e3)) / args.steps
    e2e_value = world * B / (ms_e2e / 1e3)

    def finish():
        """Leave the process group.  A captured graph holds NCCL kernels of this communicator: release it first, and do not let
        a straggling communicator teardown hold the process after the result line is out."""
        if world > 1:
            if use_graph:
                torch.cuda.synchronize()
                dist.barrier()
                _REAL_STDOUT.flush()
                os._exit(0)
            dist.destroy_process_group()

    if rank != 0:
        finish()
        return

    # ---------------- per-family device time and roofline ----------------
    pk = peaks()
    fam = []
    for name, evs in timed.items():
        ms = sum(s.elapsed_time(e) for s, e, _ in evs) / fam_steps
        work = sum(w for _, _, w in evs) / fam_steps
        n = len(evs) / fam_steps
        tensor = name.startswith("gemm") or name.startswith("attn")
        ach = work / (ms * 1e-3) / (1e12 if tensor else 1e9) if ms > 0 else 0.0
        peak = pk["tf_sust"] if tensor else pk["hbm"]
        fam.append({"kernel": name, "launches_per_step": n, "ms_per_step": ms, "share_of_step": ms / ms_step,
                    "bound": "tensor" if tensor else "hbm", "achieved": ach, "unit": "TFLOP/s" if tensor else "GB/s",
                    "peak": peak, "frac": ach / peak})
    fam.sort(key=lambda d: -d["ms_per_step"])
    traffic = None
    tpath = os.path.join(ROOT, "profiles", "traffic.json")
    top = fam[0] if fam else None
    if top and os.path.exists(tpath):
        traffic = json.load(open(tpath)).get(top["kernel"])
    roofline = None
    if top:
        roofline = {"kernel": top["kernel"], "bound": top["bound"], "achieved": top["achieved"], "peak": top["peak"], "unit": top["unit"],
                    "frac": top["frac"], "traffic": traffic, "peak_source": pk["source"] + " (sustained: timed inside the step)"}
    key = {("vitb16", 16): "vitb16_16f", ("vitb16", 4): "vitb16_4f", ("vitl14", 16): "vitl14_16f"}.get((args.model, args.frames))
    step_tf = None
    if key:
        step_tf = 3 * FWD_GFLOP_PER_PAIR[key] * 1e9 * B / (ms_step * 1e-3) / 1e12

    cpu_baseline = None
    if world == 1 and not args.no_cpu_baseline:
        import psutil
        from oracle import clip_oracle as O
        pairs = args.cpu_pairs if psutil.virtual_memory().available > 48e9 else 1
        threads = os.cpu_count() or 1
        v, ms, kind = cpu_reference_clips_per_s(O.ClipCfg(frames=args.frames), pairs, 1, 0, threads)
        what = "the reference's own CLIP + ClipLoss modules (oracle/_ref, fp32, non-flash branch)" if kind == "reference" else \
            "oracle port (fp32 PyTorch restatement of the reference path)"
        cpu_baseline = {"value": v, "unit": "clips/s", "cores": threads, "kind": kind,
                        "sample": f"{what}, {pairs} clip-text pairs of this workload, one fwd+ClipLoss+bwd+AdamW step, no warm-up "
                                  f"({ms / 1e3:.1f} s)"}

    arch = "ViT-L/14" if args.model == "vitl14" else "ViT-B/16"
    which = "configs[3]" if args.model == "vitl14" else "configs[1]"
    line = {"metric": f"clips/sec {arch} 16f CLIP train step", "value": value, "unit": "clips/s", "n_gpus": world,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
            "config": {"workload": f"AVION CLIP {arch} {args.frames}f 224^2 full train step, {B} clips/GPU (BASELINE.json {which})",
                       "clips_per_gpu": B, "global_batch": world * B, "frames": args.frames, "embed_dim": 256, "optimizer": "AdamW (torch fused)" if args.torch_optimizer else "AdamW (avion_b200 multi-tensor kernel, one launch)",
                       "parallelism": f"dp{world}", "launch": "cuda_graph(zero grads+fwd+ClipLoss+bwd) + flat gradient all-reduce + eager optimizer" if use_graph else "eager",
                       "l2": "inputs_exceed_l2 (617 MB uint8 clips + GB-scale activations per step)"},
            "loss": loss_val, "gpu_launches": launches,
            "e2e": {"value": e2e_value, "unit": "clips/s", "ms_per_step": ms_e2e, "h2d_bytes_per_step": h2d_bytes,
                    "d2h_bytes_per_step": d2h // args.steps},
            "clocks": clocks, "roofline": roofline, "kernels": fam[:10],
            "step_tensor": {"algorithmic_tflops": step_tf, "peak": pk["tf_sust"], "frac": step_tf / pk["tf_sust"] if step_tf else None},
            "cpu_baseline": cpu_baseline}
    emit(line)
    finish()


# =====================================================================================================================
def run_videomae(args, world, rank, local_rank, dev):
    """BASELINE.json configs[4]: VideoMAE ViT-B/16 pretrain step (scripts/main_videomae_pretrain.py:316-352): tube mask ->
    fused transform (visible patches + standardised targets) -> encoder on visible tokens -> decoder -> MSE -> backward ->
    gradient all-reduce (DDP) -> AdamW."""
    import torch.distributed as dist
    from torch.nn.parallel import DistributedDataParallel as DDP
    from avion_b200 import _lib, ops
    from avion_b200.data.transforms import TubeMaskingGeneratorGPU
    from avion_b200.losses import MSELoss
    from avion_b200.models import VIDEOMAE_VITB16
    from avion_b200.optim import FusedAdamW

    B = args.batch
    model = VIDEOMAE_VITB16(decoder_depth=4).to(dev)
    # The step is replayed as ONE CUDA graph (the eager step is bound by its ~1800 Python/ctypes launches, not by the GPU); with
    # several ranks GraphedStep averages one flat gradient buffer over NCCL after each replay (DDP's reduction, one call).
    # --no-graph: eager launches, torch DDP around the module.
    use_graph = not args.no_graph
    net = DDP(model, device_ids=[local_rank], bucket_cap_mb=200, gradient_as_bucket_view=True) if (world > 1 and not use_graph) else model
    crit = MSELoss()
    groups = param_groups(model)
    groups[0]["weight_decay"] = 0.05                                   # main_videomae_pretrain.py --wd default
    opt = (torch.optim.AdamW(groups, lr=1.5e-4, betas=(0.9, 0.95), fused=True) if args.torch_optimizer
           else FusedAdamW(groups, lr=1.5e-4, betas=(0.9, 0.95), eps=1e-8))
    gen = TubeMaskingGeneratorGPU(B, (8, 14, 14), 0.9, device=dev)
    host = []
    for i in range(2):
        g = torch.Generator().manual_seed(1234 + rank + 1000 * i)
        host.append(torch.randint(0, 256, (B, 16, 224, 224, 3), dtype=torch.uint8, generator=g).pin_memory())
    resident = [c.to(dev) for c in host]
    h2d_bytes = host[0].numel() + B * 1568                             # clips + the bool mask made on the host by the generator

    def fwd_bwd(clips, mask):
        opt.zero_grad(set_to_none=True)
        out, tgt = net(clips, mask, return_targets=True)
        loss = crit(out, tgt)
        loss.backward()
        return loss.detach()

    def eager_step(clips):
        mask = gen()                                                   # :319 (host RNG + argsort, copied to the device)
        loss = fwd_bwd(clips, mask)
        if world > 1 and use_graph:                                    # bare module: average the gradients by hand
            from avion_b200.graph import average_module_gradients
            average_module_gradients(model)
        opt.step()
        return loss

    # the mask (host RNG) and the optimizer step stay outside the graph

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(ms):
        if world == 1:
            return ms
        t = torch.tensor([ms], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t)

    for i in range(args.warmup):
        eager_step(resident[i % 2])
    barrier()
    # per-family device time: an EAGER pass (events cannot be timed inside a graph replay); kernel durations do not depend on
    # how the kernels were launched.  Also counts the launches of one step.
    ops.TIMED = {}
    ops.TIMED_BY_EPILOGUE = True      # a "kernel" = one template instantiation; the GEMM modes here mix 2-4 epilogues each
    launches0 = _lib.launch_count
    for i in range(args.steps):
        eager_step(resident[i % 2])
    barrier()
    launches = _lib.launch_count - launches0
    timed, ops.TIMED = ops.TIMED, None
    if use_graph:          # capture AFTER the eager passes: from here on p.grad are views of the graph's flat gradient buffer
        from avion_b200.graph import GraphedStep
        opt.zero_grad(set_to_none=True)

        def fwd_bwd_nozero(clips, mask):
            out, tgt = model(clips, mask, return_targets=True)
            loss = crit(out, tgt)
            loss.backward()
            return loss.detach()
        graphed = GraphedStep(fwd_bwd_nozero, [resident[0], gen()], params=list(model.parameters()))

        def step(clips):
            loss = graphed(clips, gen())
            opt.step()
            return loss
        for i in range(args.warmup):
            step(resident[i % 2])
        barrier()
    else:
        step = eager_step
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(args.steps):
        loss = step(resident[i % 2])
    e1.record()
    barrier()
    ms_step = max_over_ranks(e0.elapsed_time(e1)) / args.steps
    clocks = sampler.stop() if rank == 0 else None
    loss_val = float(loss)
    value = world * B / (ms_step / 1e3)

    # end to end: pinned uint8 clips copied H2D every step (double-buffered on a copy stream), loss read back
    copy_stream = torch.cuda.Stream(dev)
    dbuf = [torch.empty_like(c, device=dev) for c in host]
    ready, consumed = [torch.cuda.Event(), torch.cuda.Event()], [torch.cuda.Event(), torch.cuda.Event()]

    def prefetch(i):
        s = i % 2
        with torch.cuda.stream(copy_stream):
            copy_stream.wait_event(consumed[s])
            dbuf[s].copy_(host[s], non_blocking=True)
            ready[s].record(copy_stream)

    for s in range(2):
        consumed[s].record()
    barrier()
    e2, e3 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e2.record()
    prefetch(0)
    for i in range(args.steps):
        if i + 1 < args.steps:
            prefetch(i + 1)
        torch.cuda.current_stream().wait_event(ready[i % 2])
        loss = step(dbuf[i % 2])
        consumed[i % 2].record()
        _ = loss.item()
    e3.record()
    barrier()
    ms_e2e = max_over_ranks(e2.elapsed_time(e3)) / args.steps
    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return
    pk = peaks()
    fam = []
    for name, evs in timed.items():
        ms = sum(s.elapsed_time(e) for s, e, _ in evs) / args.steps
        work = sum(w for _, _, w in evs) / args.steps
        tensor = name.startswith("gemm") or name.startswith("attn")
        ach = work / (ms * 1e-3) / (1e12 if tensor else 1e9) if ms > 0 else 0.0
        peak = pk["tf_sust"] if tensor else pk["hbm"]
        fam.append({"kernel": name, "launches_per_step": len(evs) / args.steps, "ms_per_step": ms, "share_of_step": ms / ms_step,
                    "bound": "tensor" if tensor else "hbm", "achieved": ach, "unit": "TFLOP/s" if tensor else "GB/s",
                    "peak": peak, "frac": ach / peak})
    fam.sort(key=lambda d: -d["ms_per_step"])
    top = fam[0]
    traffic = None
    tpath = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(tpath):
        traffic = json.load(open(tpath)).get("videomae:" + top["kernel"])
    roofline = {"kernel": top["kernel"], "bound": top["bound"], "achieved": top["achieved"], "peak": top["peak"], "unit": top["unit"],
                "frac": top["frac"], "traffic": traffic, "peak_source": pk["source"] + " (sustained: timed inside the step)"}
    step_tf = 3 * FWD_GFLOP_PER_PAIR["videomae_16f"] * 1e9 * B / (ms_step * 1e-3) / 1e12
    cpu_baseline = None
    if world == 1 and not args.no_cpu_baseline:
        threads = os.cpu_count() or 1
        v, ms = cpu_reference_videomae_clips_per_s(48, 1, 0, threads)
        cpu_baseline = {"value": v, "unit": "clips/s", "cores": threads, "kind": "port",
                        "sample": f"oracle port (fp32 PyTorch restatement), 48 clips, one targets+fwd+MSE+bwd+AdamW step, no warm-up "
                                  f"({ms / 1e3:.1f} s)"}
    line = {"metric": "clips/sec VideoMAE ViT-B/16 16f pretrain step", "value": value, "unit": "clips/s", "n_gpus": world,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
            "config": {"workload": f"VideoMAE ViT-B/16 16f 224^2 pretrain step, 90% tube masking, 4-layer decoder, {B} clips/GPU "
                                   "(BASELINE.json configs[4])", "clips_per_gpu": B, "global_batch": world * B,
                       "optimizer": "AdamW (torch fused)" if args.torch_optimizer else "AdamW (avion_b200 multi-tensor kernel)",
                       "parallelism": f"dp{world}", "launch": "cuda_graph(zero grads+fwd+loss+bwd) + flat gradient all-reduce + eager mask/optimizer" if use_graph else "eager (torch DDP)",
                       "l2": f"inputs_exceed_l2 ({host[0].numel() / 1e6:.0f} MB uint8 clips per step)"},
            "loss": loss_val, "gpu_launches": launches,
            "e2e": {"value": world * B / (ms_e2e / 1e3), "unit": "clips/s", "ms_per_step": ms_e2e, "h2d_bytes_per_step": h2d_bytes,
                    "d2h_bytes_per_step": 4},
            "clocks": clocks, "roofline": roofline, "kernels": fam[:14],
            "step_tensor": {"algorithmic_tflops": step_tf, "peak": pk["tf_sust"], "frac": step_tf / pk["tf_sust"]},
            "cpu_baseline": cpu_baseline}
    emit(line)
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=4)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--batch", type=int, default=None, help="clips per GPU (default 256; 64 for vitl14 / videomae)")
    ap.add_argument("--frames", type=int, default=16)
    ap.add_argument("--model", default="vitb16", choices=["vitb16", "vitl14", "videomae"],
                    help="vitl14: BASELINE.json configs[3]; videomae: configs[4] (both default to 64 clips/GPU)")
    ap.add_argument("--cpu-pairs", type=int, default=2)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-graph", action="store_true", help="videomae: launch the step eagerly instead of replaying a CUDA graph")
    ap.add_argument("--graph", default="off", choices=["on", "off"], help="CLIP: replay the step as one CUDA graph (default: eager launches)")
    ap.add_argument("--torch-optimizer", action="store_true", help="torch.optim.AdamW(fused=True) instead of avion_b200.optim.FusedAdamW")
    args = ap.parse_args()
    if args.batch is None:
        args.batch = 256 if args.model == "vitb16" else 64
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
