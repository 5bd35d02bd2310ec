import sys, torch
sys.path.insert(0, __import__('os').path.dirname(__import__('os').path.dirname(__import__('os').path.dirname(__import__('os').path.abspath(__file__)))))
from avion_b200 import ops
B, S, H = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3])
D = H * 64
torch.manual_seed(0)
qkv = (torch.randn(B, S, 3 * D, device='cuda') * 0.5).to(torch.bfloat16)
o, lse = ops.attn_fwd(qkv, H, 0.125)
do = torch.randn_like(o)
print('fwd ok', flush=True)
dqkv = ops.attn_bwd(qkv, o, do, lse, H, 0.125)
torch.cuda.synchronize()
print('bwd ok', float(dqkv.float().abs().mean()), flush=True)
