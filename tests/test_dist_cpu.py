"""world_size-2 gloo test (CPU) of the sharded-loss plumbing: packed all-gather layout, reduce-scatter of the feature
gradients and the scalar all-reduce reproduce the reference's single-process ClipLoss on the concatenated batch
(avion/losses/losses.py:119-123) - loss, accuracy and gradients.  The per-rank arithmetic is the oracle's here; on the
GPU box tests/test_dist_gpu.py runs the same check through the CUDA kernels over NCCL."""
import os
import sys

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _worker(rank, world, port, n, E, q):
    sys.path.insert(0, ROOT)
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from avion_b200.losses import _dist
        from oracle import clip_oracle as O
        torch.manual_seed(0)
        N = world * n
        g_img = torch.nn.functional.normalize(torch.randn(N, E), dim=-1)
        g_txt = torch.nn.functional.normalize(g_img + 0.5 * torch.randn(N, E), dim=-1)
        scale = torch.tensor(9.0)
        off = rank * n
        both = _dist.gather_packed(g_img[off:off + n].contiguous(), g_txt[off:off + n].contiguous(), world)
        assert torch.equal(both[0], g_img) and torch.equal(both[1], g_txt)
        # rank-local rows against all columns; local features are rows off.. of the gathered tensors
        ai, at = both[0].clone().requires_grad_(True), both[1].clone().requires_grad_(True)
        out = O.clip_loss_local(ai[off:off + n], at[off:off + n], ai, at, scale, rank)
        out["loss"].backward()
        local = _dist.scatter_sum(torch.stack([ai.grad, at.grad]), world)
        stats = _dist.mean_over_ranks(torch.stack([out["loss"].detach(), out["clip_acc"].float()]), world)
        # single-process reference on the concatenated batch
        fi, ft = g_img.clone().requires_grad_(True), g_txt.clone().requires_grad_(True)
        full = O.clip_loss(fi, ft, scale)
        full["loss"].backward()
        ok = (torch.allclose(stats[0], full["loss"].detach(), atol=1e-6) and torch.allclose(stats[1], full["clip_acc"].float(), atol=1e-4)
              and torch.allclose(local[0], world * fi.grad[off:off + n], atol=1e-6)
              and torch.allclose(local[1], world * ft.grad[off:off + n], atol=1e-6))
        q.put((rank, bool(ok)))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("n,E", [(4, 16), (7, 32)])
def test_sharded_loss_plumbing_gloo_world2(n, E):
    world = 2
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29600 + (os.getpid() % 200) + n
    procs = [ctx.Process(target=_worker, args=(r, world, port, n, E, q)) for r in range(world)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(timeout=120)
        assert p.exitcode == 0
    res = dict(q.get(timeout=5) for _ in range(world))
    assert res == {0: True, 1: True}


def _worker_mmr(rank, world, port, n, E, q):
    """MaxMarginRankingLoss's rank routing (packed all-gather; local-row gradient by default, reduce-scatter with
    gather_with_grad) under gloo.  The two kernel calls are replaced by the oracle's arithmetic (CPU has no kernels); the GPU
    box runs the same check through the kernels over NCCL (tests/test_dist_gpu.py)."""
    sys.path.insert(0, ROOT)
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from avion_b200 import ops
        from avion_b200.losses import MaxMarginRankingLoss
        from oracle import ranking_oracle as R

        def mmr_fwd(text, image, margin, fix_norm):                 # workspace = the features themselves
            return R.max_margin_loss(image, text, margin, fix_norm).detach(), torch.stack([text, image])

        def mmr_bwd(ws, gloss, n_all, E_, margin, fix_norm):
            with torch.enable_grad():
                _, di, dt = R.loss_and_grads(ws[1], ws[0], margin, fix_norm)
            return dt * gloss, di * gloss

        ops.mmr_fwd, ops.mmr_bwd = mmr_fwd, mmr_bwd
        torch.manual_seed(0)
        N = world * n
        g_img = torch.randn(N, E)
        g_txt = 0.25 * g_img + torch.randn(N, E)
        off = rank * n
        full, fdi, fdt = R.loss_and_grads(g_img, g_txt, 0.2, True)
        ok = True
        for gwg, factor in ((False, 1.0), (True, float(world))):
            img = g_img[off:off + n].clone().requires_grad_(True)
            txt = g_txt[off:off + n].clone().requires_grad_(True)
            loss = MaxMarginRankingLoss(gather_with_grad=gwg, rank=rank, world_size=world)(img, txt)["loss"]
            loss.backward()
            ok = ok and torch.allclose(loss, full, rtol=1e-6) \
                and torch.allclose(img.grad, factor * fdi[off:off + n], rtol=1e-5, atol=1e-8) \
                and torch.allclose(txt.grad, factor * fdt[off:off + n], rtol=1e-5, atol=1e-8)
        q.put((rank, bool(ok)))
    finally:
        dist.destroy_process_group()


def test_max_margin_rank_routing_gloo_world2():
    world = 2
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29850 + (os.getpid() % 100)
    procs = [ctx.Process(target=_worker_mmr, args=(r, world, port, 5, 16, q)) for r in range(world)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(timeout=120)
        assert p.exitcode == 0
    assert dict(q.get(timeout=5) for _ in range(world)) == {0: True, 1: True}


# ---- all four gather_features modes through avion_b200.losses.ClipLoss (host logic) -------------------------------------
def _cpu_infonce_kernels():
    """CPU stand-ins with the SEMANTICS of avn_infonce_fwd / avn_infonce_bwd (include/avion_b200.h): rank-local rows against
    the gathered columns; d_all [2,N,E] = row-side gradient at rows row_offset.. plus, if col_side, the column-side gradient
    of every gathered row.  Only the routing around them (avion_b200/losses/losses.py, _dist.py) is under test here."""
    import torch.nn.functional as F

    def fwd(img, txt, all_img, all_txt, scale, off):
        n = img.shape[0]
        li, lt = scale * img @ all_txt.T, scale * txt @ all_img.T
        labels = torch.arange(n) + off
        lse = torch.stack([torch.logsumexp(li, -1), torch.logsumexp(lt, -1)])
        loss_sum = 0.5 * (F.cross_entropy(li, labels, reduction="sum") + F.cross_entropy(lt, labels, reduction="sum"))
        correct = (li.argmax(-1) == labels).sum().float()
        return lse, torch.stack([loss_sum, correct])

    def bwd(img, txt, all_img, all_txt, scale, off, lse, gscale, col_side=True):
        n, N = img.shape[0], all_img.shape[0]
        i_, t_ = img.clone().requires_grad_(True), txt.clone().requires_grad_(True)
        ai, at = all_img.clone().requires_grad_(True), all_txt.clone().requires_grad_(True)
        sc = scale.clone().requires_grad_(True)
        labels = torch.arange(n) + off
        with torch.enable_grad():
            loss = gscale * 0.5 * (F.cross_entropy(sc * i_ @ at.T, labels) + F.cross_entropy(sc * t_ @ ai.T, labels))
            gi, gt, gai, gat, gs = torch.autograd.grad(loss, [i_, t_, ai, at, sc])
        d_all = torch.zeros(2, N, img.shape[1])
        d_all[0, off:off + n] += gi
        d_all[1, off:off + n] += gt
        if col_side:
            d_all[0] += gai
            d_all[1] += gat
        return d_all, gs.reshape(1)

    return fwd, bwd


def _worker_modes(rank, world, port, q):
    sys.path.insert(0, ROOT)
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from avion_b200 import ops
        from avion_b200.losses import ClipLoss
        from oracle import cliploss_modes as M
        ops.infonce_fwd, ops.infonce_bwd = _cpu_infonce_kernels()
        n, E, scale = 5, 16, 9.0
        g_img, g_txt = M.features(n, E, world, 11)
        off = rank * n
        bad = []
        for ll, gwg in M.MODES:
            img = g_img[off:off + n].clone().requires_grad_(True)
            txt = g_txt[off:off + n].clone().requires_grad_(True)
            sc = torch.tensor(scale, requires_grad=True)
            out = ClipLoss(local_loss=ll, gather_with_grad=gwg, rank=rank, world_size=world)(img, txt, sc)
            out["loss"].backward()
            ref = M.oracle_world(g_img, g_txt, scale, n, world, ll, gwg)[rank]
            bad += [f"local_loss={ll} gather_with_grad={gwg}: {b}" for b in M.check_rank(out, (img.grad, txt.grad, sc.grad), ref)]
        q.put((rank, bad))
    finally:
        dist.destroy_process_group()


def test_clip_loss_all_gather_modes_gloo_world2():
    """ClipLoss(local_loss, gather_with_grad) for all four combinations == the oracle's restatement of gather_features +
    ClipLoss (pinned to the reference by tests/golden/cliploss_modes_w2.npz): loss, accuracy, feature gradients and the
    logit-scale gradient of each rank."""
    world = 2
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29700 + (os.getpid() % 100)
    procs = [ctx.Process(target=_worker_modes, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(timeout=120)
        assert p.exitcode == 0
    res = dict(q.get(timeout=5) for _ in range(world))
    assert res == {0: [], 1: []}, res
