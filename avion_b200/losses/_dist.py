"""Collective plumbing of the sharded loss (torch.distributed; NCCL on the GPUs, gloo in the CPU tests).

One packed all-gather in the forward and one reduce-scatter in the backward replace the reference's two list-form
all_gathers + torch.cat and their two autograd reduce_scatters (avion/losses/losses.py:54-60)."""
import torch
import torch.distributed as dist


def gather_packed(img, txt, world, group=None):
    """img, txt [n,E] -> [2, world*n, E] (image block, text block), rank-major rows, in ONE collective."""
    n, E = img.shape
    packed = torch.stack([img, txt])                                        # [2,n,E]
    out = torch.empty(world * 2, n, E, dtype=packed.dtype, device=packed.device)
    dist.all_gather_into_tensor(out, packed, group=group)                   # [W*2,n,E]: rank r at rows 2r, 2r+1
    return out.view(world, 2, n, E).permute(1, 0, 2, 3).reshape(2, world * n, E).contiguous()


def scatter_sum(d_all, world, group=None):
    """d_all [2, world*n, E] (this rank's contribution to every rank's features) -> [2,n,E] summed over ranks."""
    _, N, E = d_all.shape
    n = N // world
    inp = d_all.view(2, world, n, E).permute(1, 0, 2, 3).reshape(world * 2, n, E).contiguous()
    out = torch.empty(2, n, E, dtype=d_all.dtype, device=d_all.device)
    dist.reduce_scatter_tensor(out, inp, op=dist.ReduceOp.SUM, group=group)
    return out


def mean_over_ranks(stats, world, group=None):
    dist.all_reduce(stats, op=dist.ReduceOp.SUM, group=group)
    return stats / world
