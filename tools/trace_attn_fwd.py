"""Phase timeline of attn_fwd for CTA (0,0,0): python tools/trace_attn_fwd.py [B].  Prints clock64 offsets per K/V tile.
Needs a library built with tracing: NVCC_APPEND_FLAGS=-DAVN_TRACE python avion_b200/csrc/build.py --force"""
import ctypes
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from avion_b200 import _lib, ops  # noqa: E402

B = int(sys.argv[1]) if len(sys.argv) > 1 else 8
S, H, D = 3137, 12, 768
lib = _lib.load()
qkv = (torch.randn(B, S, 3 * D, device="cuda") * 0.5).to(torch.bfloat16)
o, lse = ops.attn_fwd(qkv, H, 0.125)
ops.attn_fwd(qkv, H, 0.125, out=o)
trace = torch.zeros(2 * 32 * 16, dtype=torch.int64, device="cuda")
lib.avn_debug_set_attn_fwd_trace.argtypes = [ctypes.c_void_p]
assert lib.avn_debug_set_attn_fwd_trace(ctypes.c_void_p(trace.data_ptr())) == 0
ops.attn_fwd(qkv, H, 0.125, out=o)
torch.cuda.synchronize()
lib.avn_debug_set_attn_fwd_trace(ctypes.c_void_p(0))
tr = trace.cpu().view(2, 32, 16)
t0 = int(tr[0, 0, 0])
names = ["iter_start", "s_full", "loaded", "max", "exp", "o_full(+fix)", "p_stored"]
for role in (0, 1):
    print("softmax group", role, names)
    for it in range(25):
        print(f"  j={it:2d} " + " ".join(f"{int(tr[role, it, s]) - t0:7d}" for s in range(len(names))))
