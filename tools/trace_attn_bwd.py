"""Phase timeline of attn_bwd for CTA (0,0,0): python tools/trace_attn_bwd.py [B].  Prints clock64 deltas per iteration."""
import ctypes
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from avion_b200 import _lib, ops  # noqa: E402

B = int(sys.argv[1]) if len(sys.argv) > 1 else 8
S, H, D = 3137, 12, 768
dev = "cuda"
lib = _lib.load()
qkv = (torch.randn(B, S, 3 * D, device=dev) * 0.5).to(torch.bfloat16)
o, lse = ops.attn_fwd(qkv, H, 0.125)
do = torch.randn_like(o)
dqkv = torch.empty_like(qkv)
dq_acc = torch.empty(B, S, D, device=dev)
for _ in range(2):
    ops.attn_bwd(qkv, o, do, lse, H, 0.125, dqkv=dqkv, dq_acc=dq_acc)
trace = torch.zeros(3 * 32 * 16, dtype=torch.int64, device=dev)
lib.avn_debug_set_attn_bwd_trace.argtypes = [ctypes.c_void_p]
assert lib.avn_debug_set_attn_bwd_trace(ctypes.c_void_p(trace.data_ptr())) == 0
ops.attn_bwd(qkv, o, do, lse, H, 0.125, dqkv=dqkv, dq_acc=dq_acc)
torch.cuda.synchronize()
lib.avn_debug_set_attn_bwd_trace(ctypes.c_void_p(0))
tr = trace.cpu().view(3, 32, 16)
t0 = int(tr[0, 0, 0])
names = {0: ["iter_start", "sdp_full", "ld0", "math0", "ld1", "math1", "ds_free", "sts", "p_free", "p_arrived", "ds_arrived"],
         1: ["iter_start", "sdp_issued", "p_ready", "ds_ready", "dq_free"],
         2: ["iter_start", "dq_full", "reduce_issued"]}
for role, label in ((0, "compute warp 0"), (1, "MMA warp"), (2, "drain warp 8")):
    print(label, names[role])
    for it in range(25):
        row = [int(tr[role, it, s]) - t0 for s in range(len(names[role]))]
        print(f"  t={it:2d} " + " ".join(f"{v:7d}" for v in row))
