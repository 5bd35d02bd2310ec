"""avion_b200: B200-native (sm_100a) implementation of AVION's video-CLIP train-step hot path.

API mirrors avion.models / avion.losses / avion.data for that path; everything runs on the hand-written CUDA kernels in
avion_b200/csrc (C-ABI: include/avion_b200.h).  There is no CPU or PyTorch fallback.
"""
from . import torch_ops  # noqa: F401  (registers torch.ops.avion_b200.*)
from . import data, losses, models  # noqa: F401

__version__ = "0.1.0"
