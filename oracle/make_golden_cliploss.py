"""Golden vectors for the four (local_loss, gather_with_grad) modes of the reference ClipLoss over 2 ranks.

Runs the UNMODIFIED reference `avion.losses.losses.ClipLoss` (gather_features :19-77, loss :103-149) in two gloo processes
on CPU (fp32) and records, per mode and rank: loss, clip_acc, the gradients left on that rank's image/text features and
on its logit scale.  Build-container only (needs /root/reference); writes tests/golden/cliploss_modes_w2.npz.

  python oracle/make_golden_cliploss.py
"""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
N_LOCAL, E, WORLD, SCALE, SEED = 5, 16, 2, 9.0, 11
MODES = [(False, False), (False, True), (True, False), (True, True)]


def make_features():
    g = torch.Generator().manual_seed(SEED)
    img = torch.nn.functional.normalize(torch.randn(WORLD * N_LOCAL, E, generator=g), dim=-1)
    txt = torch.nn.functional.normalize(img + 0.6 * torch.randn(WORLD * N_LOCAL, E, generator=g), dim=-1)
    return img, txt


def run_reference_rank(rank, world, ClipLoss):
    """-> {mode: dict(loss, clip_acc, dimg, dtxt, dscale)} for this rank (process group must be initialised)."""
    g_img, g_txt = make_features()
    off = rank * N_LOCAL
    res = {}
    for ll, gwg in MODES:
        img = g_img[off:off + N_LOCAL].clone().requires_grad_(True)
        txt = g_txt[off:off + N_LOCAL].clone().requires_grad_(True)
        sc = torch.tensor(SCALE, requires_grad=True)
        out = ClipLoss(local_loss=ll, gather_with_grad=gwg, rank=rank, world_size=world)(img, txt, sc)
        out["loss"].backward()
        z = torch.zeros_like(img)
        res[(ll, gwg)] = dict(loss=out["loss"].detach(), clip_acc=out["clip_acc"].float(),
                              dimg=img.grad if img.grad is not None else z, dtxt=txt.grad if txt.grad is not None else z,
                              dscale=sc.grad)
    return res


def _worker(rank, world, port, q):
    sys.path.insert(0, ROOT)
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from oracle import ref_import as R
        ClipLoss = R.import_reference()[3]
        res = run_reference_rank(rank, world, ClipLoss)
        q.put((rank, {f"{int(ll)}{int(gwg)}:{k}": v.numpy() for (ll, gwg), d in res.items() for k, v in d.items()}))
    finally:
        dist.destroy_process_group()


def run_reference_world(port=29911):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, WORLD, port, q)) for r in range(WORLD)]
    for p in procs:
        p.start()
    got = dict(q.get(timeout=120) for _ in range(WORLD))
    for p in procs:
        p.join(timeout=60)
    return got


if __name__ == "__main__":
    got = run_reference_world()
    flat = {f"r{r}:{k}": v for r, d in got.items() for k, v in d.items()}
    flat["meta"] = np.array([N_LOCAL, E, WORLD, SEED], dtype=np.int64)
    flat["scale"] = np.float32(SCALE)
    out = os.path.join(ROOT, "tests", "golden", "cliploss_modes_w2.npz")
    np.savez(out, **flat)
    print("wrote", out, len(flat), "arrays")
