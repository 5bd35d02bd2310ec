"""2-GPU NCCL test: ClipLoss through the CUDA kernels, sharded over ranks, equals the single-GPU result on the
concatenated batch (loss, accuracy, feature gradients x world, logit-scale gradient after the DDP mean)."""
import os
import sys

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _worker(rank, world, port, n, E, q):
    sys.path.insert(0, ROOT)
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    try:
        from avion_b200.losses import ClipLoss
        torch.manual_seed(0)
        N = world * n
        g_img = torch.nn.functional.normalize(torch.randn(N, E), dim=-1).cuda()
        g_txt = torch.nn.functional.normalize(g_img.cpu() + 0.5 * torch.randn(N, E), dim=-1).cuda()
        off = rank * n
        img = g_img[off:off + n].clone().requires_grad_(True)
        txt = g_txt[off:off + n].clone().requires_grad_(True)
        sc = torch.tensor(9.0, device="cuda", requires_grad=True)
        out = ClipLoss(local_loss=False, gather_with_grad=True, rank=rank, world_size=world)(img, txt, sc)
        out["loss"].backward()
        dsc = sc.grad.clone()
        dist.all_reduce(dsc)
        dsc /= world                                    # what DDP does to parameter gradients
        fi, ft = g_img.clone().requires_grad_(True), g_txt.clone().requires_grad_(True)
        fs = torch.tensor(9.0, device="cuda", requires_grad=True)
        full = ClipLoss(world_size=1)(fi, ft, fs)
        full["loss"].backward()
        ok = (torch.allclose(out["loss"], full["loss"], atol=1e-5) and torch.allclose(out["clip_acc"], full["clip_acc"], atol=1e-3)
              and torch.allclose(img.grad, world * fi.grad[off:off + n], atol=1e-6)
              and torch.allclose(txt.grad, world * ft.grad[off:off + n], atol=1e-6)
              and torch.allclose(dsc, fs.grad, atol=1e-5))
        q.put((rank, bool(ok)))
    finally:
        dist.destroy_process_group()


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs 2 GPUs")
def test_sharded_clip_loss_nccl_world2():
    world = 2
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, world, 29731, 48, 64, q)) for r in range(world)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(timeout=300)
        assert p.exitcode == 0
    assert dict(q.get(timeout=5) for _ in range(world)) == {0: True, 1: True}


def _worker_mmr(rank, world, port, n, E, q):
    """MaxMarginRankingLoss sharded over ranks == the single-process loss on the concatenated batch: default mode gives every
    rank the full loss and the gradient of its own rows; gather_with_grad sums every rank's contribution (= world x)."""
    sys.path.insert(0, ROOT)
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    try:
        from avion_b200.losses import MaxMarginRankingLoss
        torch.manual_seed(0)
        N = world * n
        g_img = torch.randn(N, E).cuda()
        g_txt = (0.25 * g_img.cpu() + torch.randn(N, E)).cuda()
        off = rank * n
        fi, ft = g_img.clone().requires_grad_(True), g_txt.clone().requires_grad_(True)
        full = MaxMarginRankingLoss(world_size=1)(fi, ft)["loss"]
        full.backward()
        ok = True
        for gwg, factor in ((False, 1.0), (True, float(world))):
            img = g_img[off:off + n].clone().requires_grad_(True)
            txt = g_txt[off:off + n].clone().requires_grad_(True)
            loss = MaxMarginRankingLoss(gather_with_grad=gwg, rank=rank, world_size=world)(img, txt)["loss"]
            loss.backward()
            ok = ok and torch.allclose(loss, full, rtol=1e-5) \
                and torch.allclose(img.grad, factor * fi.grad[off:off + n], rtol=1e-4, atol=1e-7) \
                and torch.allclose(txt.grad, factor * ft.grad[off:off + n], rtol=1e-4, atol=1e-7)
        q.put((rank, bool(ok)))
    finally:
        dist.destroy_process_group()


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs 2 GPUs")
def test_sharded_max_margin_loss_nccl_world2():
    world = 2
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker_mmr, args=(r, world, 29741, 40, 64, q)) for r in range(world)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(timeout=300)
        assert p.exitcode == 0
    assert dict(q.get(timeout=5) for _ in range(world)) == {0: True, 1: True}
