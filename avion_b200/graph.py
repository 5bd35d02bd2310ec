"""CUDA-graph replay of a launch-bound train step.

A VideoMAE step at the recipe's 64 clips/GPU is ~1800 kernel launches of ~13 us each: in eager mode the Python/ctypes launch
path (~16 us per launch) is slower than the GPU.  `GraphedStep` captures `zero_grad -> forward -> loss -> backward` once (every
kernel of the C-ABI library launches on torch's current stream, so a stream capture records them; activations, workspaces and
gradients live in the graph's private pool at fixed addresses) and replays it with one launch.  The optimizer step stays
outside the graph (its bias correction depends on the host-side step count); it reads the gradients the replay left in place.
"""
import torch


class GraphedStep:
    """fn(*static_inputs) -> loss tensor; must be free of host synchronisation after its first (warm-up) calls.

    step = GraphedStep(fn, example_inputs); loss = step(*new_inputs)   # inputs are copied into the static buffers"""

    def __init__(self, fn, example_inputs, warmup=3):
        self.static_inputs = [x.clone() for x in example_inputs]
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
            self.static_loss = fn(*self.static_inputs)

    def __call__(self, *inputs):
        for dst, src in zip(self.static_inputs, inputs):
            if dst.data_ptr() != src.data_ptr():
                dst.copy_(src, non_blocking=True)
        self.graph.replay()
        return self.static_loss
