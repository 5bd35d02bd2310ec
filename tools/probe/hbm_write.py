"""HBM bandwidth by access mix (B200): pure write (fill), pure read (sum), copy.  python tools/probe/hbm_write.py"""
import torch

n = 4 << 30          # bytes
a = torch.empty(n // 2, dtype=torch.bfloat16, device="cuda")
b = torch.empty(n // 2, dtype=torch.bfloat16, device="cuda")


def t(fn, reps=10):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    best = 1e9
    for _ in range(reps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1))
    return best


ms = t(lambda: a.zero_())
print(f"fill  (write {n / 2**30:.0f} GiB)            {ms:7.3f} ms  {n / ms / 1e6:8.1f} GB/s written")
ms = t(lambda: torch.cuda.memset if False else a.fill_(1.0))
print(f"fill_ kernel                      {ms:7.3f} ms  {n / ms / 1e6:8.1f} GB/s written")
ms = t(lambda: b.copy_(a))
print(f"copy  (read + write)              {ms:7.3f} ms  {2 * n / ms / 1e6:8.1f} GB/s total, {n / ms / 1e6:8.1f} GB/s written")
ms = t(lambda: a.float().sum() if False else torch.sum(a.view(torch.int16)[: n // 4].to(torch.int32)) if False else a.max())
print(f"max   (read only)                 {ms:7.3f} ms  {n / ms / 1e6:8.1f} GB/s read")
