"""Multi-rank tests of the sharded losses through the CUDA kernels: every gather_features mode of ClipLoss against the oracle
(two ranks on one GPU over gloo, and over NCCL when two GPUs are present), MaxMarginRankingLoss over NCCL."""
import os
import sys

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _worker(rank, world, port, n, E, q, backend, share_gpu):
    """ClipLoss on this rank through the CUDA kernels (avn_infonce_fwd / bwd with row offsets, packed all-gather, reduce-scatter)
    for all four (local_loss, gather_with_grad) modes, checked against the ORACLE's restatement of gather_features + ClipLoss
    (avion/losses/losses.py:53-75,103-149; oracle.clip_oracle.clip_loss_world, pinned to the reference by
    tests/golden/cliploss_modes_w2.npz) computed on the CPU in this worker - not against the kernels themselves."""
    sys.path.insert(0, ROOT)
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dev = torch.device("cuda", 0 if share_gpu else rank)
    torch.cuda.set_device(dev)
    if backend == "nccl":
        dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    else:
        dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from avion_b200.losses import ClipLoss
        from oracle import cliploss_modes as M
        scale = 9.0
        g_img, g_txt = M.features(n, E, world, 11)
        off = rank * n
        bad = []
        for ll, gwg in M.MODES:
            img = g_img[off:off + n].to(dev).requires_grad_(True)
            txt = g_txt[off:off + n].to(dev).requires_grad_(True)
            sc = torch.tensor(scale, device=dev, requires_grad=True)
            out = ClipLoss(local_loss=ll, gather_with_grad=gwg, rank=rank, world_size=world)(img, txt, sc)
            out["loss"].backward()
            ref = M.oracle_world(g_img, g_txt, scale, n, world, ll, gwg)[rank]
            bad += [f"local_loss={ll} gather_with_grad={gwg}: {b}"
                    for b in M.check_rank(out, (img.grad, txt.grad, sc.grad), ref, atol=3e-6, rtol=2e-4)]
        q.put((rank, bad))
    finally:
        dist.destroy_process_group()


def _run_world(backend, share_gpu, n, E, port):
    world = 2
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, world, port, n, E, q, backend, share_gpu)) for r in range(world)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(timeout=300)
        assert p.exitcode == 0
    res = dict(q.get(timeout=5) for _ in range(world))
    assert res == {0: [], 1: []}, res


@pytest.mark.parametrize("n,E", [(5, 16), (48, 64)])
def test_clip_loss_modes_two_ranks_one_gpu(n, E):
    """World size 2 on ONE GPU: both ranks run the kernels on cuda:0 and exchange through gloo (staged via the host,
    avion_b200/losses/_dist.py), so the driver's single-GPU test run exercises row offsets, the packed gather and the
    reduce-scatter of every mode instead of skipping them."""
    _run_world("gloo", True, n, E, 29721 + n)


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs 2 GPUs")
def test_clip_loss_modes_nccl_world2():
    _run_world("nccl", False, 48, 64, 29731)


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
