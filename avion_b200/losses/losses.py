"""ClipLoss with the reference's constructor and result dict (avion/losses/losses.py:80-149), computed by the sharded
InfoNCE kernels: each rank evaluates its local rows of both logit matrices against the all-gathered embeddings
(one packed NCCL all-gather; one reduce-scatter in the backward, as torch.distributed.nn.all_gather does, :55-60)."""
import torch
import torch.distributed as dist
import torch.nn as nn

from .. import ops
from . import _dist


class _InfoNCEFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, img, txt, scale, rank, world, local_loss, gather_with_grad, group):
        img32, txt32 = img.detach().float().contiguous(), txt.detach().float().contiguous()
        n, E = img32.shape
        if world > 1:
            both = _dist.gather_packed(img32, txt32, world, group)           # ONE collective for both towers
            all_img, all_txt = both[0], both[1]
        else:
            all_img, all_txt = img32, txt32
        s = scale.detach().float().reshape(1).contiguous()      # stays on the device: no host read of the parameter per step
        off = rank * n
        lse, out = ops.infonce_fwd(img32, txt32, all_img, all_txt, s, off)
        stats = torch.stack([out[0] / n, 100.0 * out[1] / n])                  # rank-local mean loss, top-1 accuracy
        if world > 1 and not local_loss:
            # the reference's default mode evaluates the full NxN problem on every rank (:119-123); its scalar is the
            # mean over all N rows = the mean of the rank-local means
            stats = _dist.mean_over_ranks(stats, world, group)
        ctx.save_for_backward(img32, txt32, all_img, all_txt, lse, s)
        ctx.cfg = (off, rank, world, local_loss, gather_with_grad, group, img.dtype, txt.dtype)
        loss, acc = stats[0].clone(), stats[1].clone()
        ctx.mark_non_differentiable(acc)
        return loss, acc

    @staticmethod
    def backward(ctx, gloss, _gacc):
        img32, txt32, all_img, all_txt, lse, s = ctx.saved_tensors
        off, rank, world, local_loss, gather_with_grad, group, dt_i, dt_t = ctx.cfg
        n, E = img32.shape
        # The four modes of gather_features (avion/losses/losses.py:53-75) with the loss of :116-133:
        #   gather_with_grad             : every rank's loss reaches every rank's features (all_gather backward = sum over ranks)
        #   neither                      : full-matrix loss, remote features constant, own slice re-inserted (:68-72)
        #                                  = 1/W of the sum of the rank-local losses' gradients
        #   local_loss only              : the gathered copies carry NO gradient and the own slice is NOT re-inserted, so only
        #                                  the row side (the direct use of image_features / text_features, :117-118) flows
        rows_only = world > 1 and local_loss and not gather_with_grad
        d_all, dscale = ops.infonce_bwd(img32, txt32, all_img, all_txt, s, off, lse, 1.0, col_side=not rows_only)
        if world > 1 and not rows_only:
            # d_all holds this rank's contribution to every rank's features: sum them where they belong
            if local_loss:
                local = _dist.scatter_sum(d_all, world, group)
            else:
                # full-matrix loss: every rank's reference backward holds d(full loss)/d scale = the mean of the rank-local
                # ones; it rides in the same reduce-scatter
                local, dsum = _dist.scatter_sum(d_all, world, group, extra=dscale)
                dscale = dsum / world
            if not gather_with_grad:
                local = local / world
        else:
            local = d_all[:, off:off + n]
        g = gloss.float()
        return ((local[0] * g).to(dt_i), (local[1] * g).to(dt_t), (dscale[0] * g).to(gloss.dtype).reshape(()),
                None, None, None, None, None)


class ClipLoss(nn.Module):
    def __init__(self, local_loss=False, gather_with_grad=False, cache_labels=False, rank=0, world_size=1, use_horovod=False,
                 group=None):
        super().__init__()
        assert not use_horovod, "horovod is not supported (the reference never enables it either)"
        self.local_loss = local_loss
        self.gather_with_grad = gather_with_grad
        self.cache_labels = cache_labels          # labels are implicit (row_offset + r) in the kernel
        self.rank = rank
        self.world_size = world_size
        self.group = group

    def forward(self, image_features, text_features, logit_scale):
        loss, acc = _InfoNCEFn.apply(image_features, text_features, logit_scale, self.rank, self.world_size, self.local_loss,
                                     self.gather_with_grad, self.group)
        return {"loss": loss, "clip_acc": acc}


class _MSEFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, pred, target):
        loss = ops.mse_fwd(pred, target)
        ctx.save_for_backward(pred, target)
        return loss

    @staticmethod
    def backward(ctx, gloss):
        pred, target = ctx.saved_tensors
        return ops.mse_bwd(pred, target, gloss.to(torch.float32).reshape(1).contiguous()), None


class MSELoss(nn.Module):
    """torch.nn.MSELoss() of the VideoMAE step (scripts/main_videomae_pretrain.py:129,341): mean((input - target)^2) with bf16
    predictions and fp32 targets, deterministic two-stage reduction, fused (pred - target) backward."""

    def forward(self, input, target):
        return _MSEFn.apply(input.contiguous(), target.contiguous())


class _MaxMarginFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, img, txt, margin, fix_norm, rank, world, local_loss, gather_with_grad, group):
        img32, txt32 = img.detach().float().contiguous(), txt.detach().float().contiguous()
        n, E = img32.shape
        if world > 1:
            both = _dist.gather_packed(img32, txt32, world, group)
            all_img, all_txt = both[0], both[1]
        else:
            all_img, all_txt = img32, txt32
        loss, ws = ops.mmr_fwd(all_txt, all_img, margin, fix_norm)
        ctx.save_for_backward(ws)
        ctx.cfg = (margin, fix_norm, rank, world, local_loss, gather_with_grad, group, n, E, img.dtype, txt.dtype)
        return loss

    @staticmethod
    def backward(ctx, gloss):
        (ws,) = ctx.saved_tensors
        margin, fix_norm, rank, world, local_loss, gather_with_grad, group, n, E, dt_i, dt_t = ctx.cfg
        dtxt, dimg = ops.mmr_bwd(ws, gloss.float().reshape(1).contiguous(), n * world, E, margin, fix_norm)
        if world > 1 and gather_with_grad:
            # torch.distributed.nn.all_gather backward: every rank's loss sends gradient to every rank's features (:55-60)
            local = _dist.scatter_sum(torch.stack([dimg, dtxt]), world, group)
            dimg, dtxt = local[0], local[1]
        elif world > 1:
            if local_loss:                        # gathered copies carry no gradient and the local slice is not re-inserted (:61-72)
                return (torch.zeros_like(dimg[:n]).to(dt_i), torch.zeros_like(dtxt[:n]).to(dt_t)) + (None,) * 7
            dimg, dtxt = dimg[rank * n:(rank + 1) * n], dtxt[rank * n:(rank + 1) * n]
        return (dimg.to(dt_i), dtxt.to(dt_t), None, None, None, None, None, None, None)


class MaxMarginRankingLoss(nn.Module):
    """avion/losses/losses.py:163-227 (EK-100 multi-instance retrieval fine-tuning): same constructor and result dict; the loss
    over the all-gathered features is evaluated by the `avn_mmr_*` kernels without materialising the similarity matrix."""

    def __init__(self, margin=0.2, fix_norm=True, local_loss=False, gather_with_grad=False, rank=0, world_size=1,
                 use_horovod=False, group=None):
        super().__init__()
        assert not use_horovod, "horovod is not supported"
        self.fix_norm, self.margin = fix_norm, margin
        self.local_loss, self.gather_with_grad = local_loss, gather_with_grad
        self.rank, self.world_size, self.group = rank, world_size, group

    def forward(self, image_features, text_features, weight=None):
        loss = _MaxMarginFn.apply(image_features, text_features, self.margin, self.fix_norm, self.rank, self.world_size,
                                  self.local_loss, self.gather_with_grad, self.group)
        return {"loss": loss, "max_margin_loss": loss}
