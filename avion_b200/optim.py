"""Fused multi-tensor AdamW on the sm_100a library: ONE kernel launch per step for all parameters of all groups.

Drop-in for the optimizer of scripts/main_lavila_pretrain.py:296-341 (`torch.optim.AdamW(param_groups, lr, betas, eps,
weight_decay)` with a weight-decay and a no-weight-decay group); same update, same `state_dict()` layout
(`state[p] = {"step", "exp_avg", "exp_avg_sq"}`), so checkpoints written by either optimizer load into the other.
`clamp_logit_scale` is the `logit_scale.data.clamp_(0, 4.6052)` of :880 as a device-side kernel.
"""
import torch

from . import _lib
from .torch_ops import K


class FusedAdamW(torch.optim.Optimizer):
    def __init__(self, params, lr=1e-3, betas=(0.9, 0.999), eps=1e-8, weight_decay=1e-2):
        if lr < 0 or eps < 0 or not (0 <= betas[0] < 1) or not (0 <= betas[1] < 1) or weight_decay < 0:
            raise ValueError("invalid AdamW hyper-parameters")
        super().__init__(params, dict(lr=lr, betas=betas, eps=eps, weight_decay=weight_decay))
        self._tables = None      # {(device, pointer signature): device tables}; rebuilt only when the set of tensors changes

    def load_state_dict(self, state_dict):
        """Checkpoints of torch's fused / capturable AdamW keep `step` as a CUDA tensor; reading it every step would cost one
        host synchronisation per parameter, so the step counters come back to the host once, here."""
        super().load_state_dict(state_dict)
        for st in self.state.values():
            if torch.is_tensor(st.get("step")) and st["step"].is_cuda:
                st["step"] = st["step"].detach().float().cpu()
        self._tables = None

    def _build_tables(self, entries, dev):
        chunk = _lib.load().avn_adamw_chunk_elems()
        ct, co = [], []
        for i, (p, _, _, _, _) in enumerate(entries):
            n = p.numel()
            for off in range(0, n, chunk):
                ct.append(i)
                co.append(off)
        i64 = lambda v: torch.tensor(v, dtype=torch.int64, device=dev)      # noqa: E731
        return dict(params=i64([e[0].data_ptr() for e in entries]), grads=i64([e[1].data_ptr() for e in entries]),
                    exp_avg=i64([e[2].data_ptr() for e in entries]), exp_avg_sq=i64([e[3].data_ptr() for e in entries]),
                    sizes=i64([e[0].numel() for e in entries]),
                    wd=torch.tensor([e[4] for e in entries], dtype=torch.float32, device=dev),
                    chunk_tensor=torch.tensor(ct, dtype=torch.int32, device=dev), chunk_offset=i64(co), n_chunks=len(ct))

    @torch.no_grad()
    def step(self, closure=None):
        loss = None
        if closure is not None:
            with torch.enable_grad():
                loss = closure()
        # all groups that share (lr, betas, eps) go out in one launch; weight decay is per tensor
        launches = {}
        for group in self.param_groups:
            for p in group["params"]:
                if p.grad is None:
                    continue
                if p.dtype != torch.float32 or not p.is_cuda or p.grad.is_sparse:
                    raise RuntimeError("FusedAdamW: fp32 CUDA parameters with dense gradients only (the masters of the B200 path)")
                if not p.is_contiguous() or not p.grad.is_contiguous():
                    raise RuntimeError("FusedAdamW: parameters and gradients must be contiguous")
                st = self.state[p]
                if len(st) == 0:
                    st["step"] = torch.tensor(0.0)
                    st["exp_avg"] = torch.zeros_like(p, memory_format=torch.preserve_format)
                    st["exp_avg_sq"] = torch.zeros_like(p, memory_format=torch.preserve_format)
                st["step"] += 1
                key = (float(group["lr"]), float(group["betas"][0]), float(group["betas"][1]), float(group["eps"]),
                       int(st["step"]), p.device)
                launches.setdefault(key, []).append((p, p.grad, st["exp_avg"], st["exp_avg_sq"], float(group["weight_decay"])))
        # Device tables depend only on WHICH tensors go out together (pointers, weight decay) - not on lr / betas / eps / step,
        # which are launch arguments: a per-iteration lr schedule (main_lavila_pretrain.py:805-808) must not rebuild them.
        if self._tables is None:
            self._tables = {}
        tables = {}
        for k, v in launches.items():
            sig = (k[5], tuple((e[0].data_ptr(), e[1].data_ptr(), e[2].data_ptr(), e[3].data_ptr(), e[4]) for e in v))
            # (gradient buffers are stable under DDP's gradient_as_bucket_view and zero_grad(set_to_none=False); with
            # set_to_none=True PyTorch's caching allocator normally hands the same blocks back - otherwise this rebuilds)
            t = self._tables.get(sig)
            if t is None:
                if len(self._tables) > 8:
                    self._tables.clear()
                t = self._tables[sig] = self._build_tables(v, k[5])
            tables[k] = t
        for k, v in launches.items():
            t = tables[k]
            lr, b1, b2, eps, step, dev = k
            K.adamw_step(t["params"], t["grads"], t["exp_avg"], t["exp_avg_sq"], t["sizes"], t["wd"], t["chunk_tensor"],
                         t["chunk_offset"], t["n_chunks"], lr, b1, b2, eps, step)
        return loss


def clamp_logit_scale(logit_scale, lo=0.0, hi=4.6052):
    """scripts/main_lavila_pretrain.py:880 on the device (no host synchronisation)."""
    K.clamp_scalar_f32(logit_scale.data, float(lo), float(hi))
