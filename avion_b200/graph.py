"""CUDA-graph replay of a train step.

A VideoMAE step at the recipe's 64 clips/GPU is ~1800 kernel launches of ~13 us each: in eager mode the Python/ctypes launch
path (~16 us per launch) is slower than the GPU.  `GraphedStep` captures `zero grads -> forward -> loss -> backward` once (every
kernel of the C-ABI library launches on torch's current stream, so a stream capture records them; activations, workspaces and
gradients live at fixed addresses) and replays it with one launch.  The optimizer step stays outside the graph (its bias
correction depends on the host-side step count); it reads the gradients the replay left in place.

Data parallelism (one process per GPU): DDP's reducer hooks do not run inside a replay, so with `params=` the gradients of
all parameters are made views of ONE flat fp32 buffer and, after every replay, that buffer is averaged over the ranks with a
single NCCL all-reduce - the same reduction DDP performs bucket by bucket (scripts/main_lavila_pretrain.py:283-286), issued
once per step over NVLink.  Collectives INSIDE the step (ClipLoss's embedding all-gather and its reduce-scatter,
avion_b200/losses/_dist.py) are captured with the kernels: the step function must be free of host synchronisation
(ClipLoss reads logit_scale on the device for that reason).
"""
import torch
import torch.distributed as dist


class GraphedStep:
    """fn(*static_inputs) -> loss tensor (or tuple / dict of tensors); must not synchronise with the host after its warm-up calls.

    step = GraphedStep(fn, example_inputs)                  # fn zeroes its own gradients (single process)
    step = GraphedStep(fn, example_inputs, params=model.parameters(), group=None)
                                                            # fn = forward + loss + backward only; gradients are zeroed inside
                                                            # the graph and averaged over the ranks after each replay
    loss = step(*new_inputs)                                # inputs are copied into the static buffers"""

    def __init__(self, fn, example_inputs, warmup=3, params=None, group=None):
        self.static_inputs = [x.clone() for x in example_inputs]
        self.group = group
        self.world = dist.get_world_size(group) if dist.is_available() and dist.is_initialized() else 1
        self.flat = None
        if params is not None:
            params = [p for p in params if p.requires_grad]
            dev = params[0].device
            sizes = [(p.numel() + 3) // 4 * 4 for p in params]          # every view starts on a 16-byte boundary
            self.flat = torch.zeros(sum(sizes), dtype=torch.float32, device=dev)
            off = 0
            for p, n in zip(params, sizes):
                assert p.dtype == torch.float32, "fp32 master parameters"
                p.grad = self.flat[off:off + p.numel()].view_as(p)
                off += n
            inner = fn

            def fn(*inputs):      # noqa: F811
                self.flat.zero_()                                       # one memset node; autograd then accumulates into the views
                return inner(*inputs)
        cur = torch.cuda.current_stream()
        side = torch.cuda.Stream()
        side.wait_stream(cur)
        with torch.cuda.stream(side):
            for _ in range(warmup):
                fn(*self.static_inputs)
        cur.wait_stream(side)
        torch.cuda.synchronize()
        self.graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(self.graph):
            self.static_out = fn(*self.static_inputs)
        self.static_loss = self.static_out

    def __call__(self, *inputs):
        for dst, src in zip(self.static_inputs, inputs):
            if dst.data_ptr() != src.data_ptr():
                dst.copy_(src, non_blocking=True)
        self.graph.replay()
        if self.flat is not None and self.world > 1:
            average_gradients(self.flat, self.world, self.group)
        return self.static_out


def average_gradients(flat, world, group=None):
    """Mean over the ranks of one flat gradient buffer, in place (ReduceOp.AVG on NCCL; sum and scale elsewhere)."""
    if dist.get_backend(group) == "nccl":
        dist.all_reduce(flat, op=dist.ReduceOp.AVG, group=group)
    else:
        dist.all_reduce(flat, op=dist.ReduceOp.SUM, group=group)
        flat.div_(world)


def average_module_gradients(module, group=None):
    """Eager counterpart (warm-up steps of a graphed run, tests): average every parameter gradient over the ranks."""
    world = dist.get_world_size(group)
    grads = [p.grad for p in module.parameters() if p.grad is not None]
    if world > 1 and grads:
        flat = torch._utils._flatten_dense_tensors(grads)
        average_gradients(flat, world, group)
        for g, f in zip(grads, torch._utils._unflatten_dense_tensors(flat, grads)):
            g.copy_(f)
