"""Collective plumbing of the sharded loss (torch.distributed; NCCL on the GPUs, gloo in the CPU tests).

One packed all-gather in the forward and one reduce-scatter in the backward replace the reference's two list-form
all_gathers + torch.cat and their two autograd reduce_scatters (avion/losses/losses.py:54-60)."""
import torch
import torch.distributed as dist


def _via_host(group):
    """gloo moves CUDA tensors through host memory itself for some collectives and refuses others; staging them explicitly
    keeps every mode of the loss testable with two ranks on ONE GPU (tests/test_dist_gpu.py) - NCCL never takes this path."""
    return dist.get_backend(group) == "gloo"


def gather_packed(img, txt, world, group=None):
    """img, txt [n,E] -> [2, world*n, E] (image block, text block), rank-major rows, in ONE collective."""
    n, E = img.shape
    packed = torch.stack([img, txt])                                        # [2,n,E]
    if packed.is_cuda and _via_host(group):
        host = torch.empty(world * 2, n, E, dtype=packed.dtype)
        dist.all_gather_into_tensor(host, packed.cpu(), group=group)
        out = host.to(packed.device)
    else:
        out = torch.empty(world * 2, n, E, dtype=packed.dtype, device=packed.device)
        dist.all_gather_into_tensor(out, packed, group=group)               # [W*2,n,E]: rank r at rows 2r, 2r+1
    return out.view(world, 2, n, E).permute(1, 0, 2, 3).reshape(2, world * n, E).contiguous()


def scatter_sum(d_all, world, group=None, extra=None):
    """d_all [2, world*n, E] (this rank's contribution to every rank's features) -> [2,n,E] summed over ranks.
    `extra` (optional, fp32 [k]): rank-local scalars that ride in the same collective; their sum over ranks is returned as a
    second result (the logit-scale gradient of the full-matrix modes: one reduce-scatter instead of that plus an all-reduce)."""
    _, N, E = d_all.shape
    n = N // world
    inp = d_all.view(2, world, n, E).permute(1, 0, 2, 3).reshape(world, 2 * n * E)
    k = 0 if extra is None else extra.numel()
    if k:
        inp = torch.cat([inp, extra.reshape(1, k).expand(world, k)], dim=1)
    inp = inp.contiguous().view(-1)            # [world * (2nE + k)]: chunk q goes to rank q
    if d_all.is_cuda and _via_host(group):
        host = torch.empty(2 * n * E + k, dtype=d_all.dtype)
        dist.reduce_scatter_tensor(host, inp.cpu(), op=dist.ReduceOp.SUM, group=group)
        out = host.to(d_all.device)
    else:
        out = torch.empty(2 * n * E + k, dtype=d_all.dtype, device=d_all.device)
        dist.reduce_scatter_tensor(out, inp, op=dist.ReduceOp.SUM, group=group)
    local = out[:2 * n * E].view(2, n, E)
    return (local, out[2 * n * E:]) if k else local


def mean_over_ranks(stats, world, group=None):
    if stats.is_cuda and _via_host(group):
        host = stats.cpu()
        dist.all_reduce(host, op=dist.ReduceOp.SUM, group=group)
        return host.to(stats.device) / world
    dist.all_reduce(stats, op=dist.ReduceOp.SUM, group=group)
    return stats / world
