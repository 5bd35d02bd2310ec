"""Generate tests/golden/videomae_tiny_b4.npz by running the UNMODIFIED reference VideoMAE model on CPU (fp32).

Run in the build container only:  python oracle/make_golden_videomae.py
The step follows scripts/main_videomae_pretrain.py:316-341 with the reference's own rearrange / MSELoss calls; weights,
clips and masks come from oracle.videomae_oracle (frozen seeds), so the fixture stores seeds, config and reference outputs.
"""
import os
import sys

import numpy as np
import torch
from einops import rearrange

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle import videomae_oracle as O        # noqa: E402
from oracle import ref_import as R             # noqa: E402

GOLDEN = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden")
GRAD_KEYS = ["mask_token", "encoder.patch_embed.proj.weight", "encoder.patch_embed.proj.bias", "encoder.blocks.0.attn.q_bias",
             "encoder.blocks.0.attn.v_bias", "encoder.blocks.0.attn.qkv.weight", "encoder.blocks.1.mlp.fc1.weight",
             "encoder.blocks.1.norm2.weight", "encoder.norm.bias", "encoder_to_decoder.weight", "decoder.blocks.0.attn.qkv.weight",
             "decoder.blocks.0.attn.proj.bias", "decoder.blocks.0.mlp.fc2.weight", "decoder.norm.weight", "decoder.head.weight",
             "decoder.head.bias"]


def run(name, cfg, batch, wseed, iseed):
    sd = O.make_state_dict(cfg, wseed)
    clips, mask = O.make_inputs(cfg, batch, iseed)
    model = R.build_reference_videomae(cfg, sd)
    model.train()
    p = cfg.patch
    videos = clips.permute(0, 4, 1, 2, 3).float()                                          # :317-318
    sq = rearrange(videos, 'b c (t p0) (h p1) (w p2) -> b (t h w) (p0 p1 p2) c', p0=2, p1=p, p2=p)
    norm = (sq - sq.mean(dim=-2, keepdim=True)) / (sq.var(dim=-2, unbiased=True, keepdim=True).sqrt() + 1e-6)
    patch = rearrange(norm, 'b n p c -> b n (p c)')
    B, _, C = patch.shape
    targets = patch[mask].reshape(B, -1, C)                                                # :333-334
    outputs = model(O.normalize(videos), mask)                                             # :336,340
    loss = torch.nn.MSELoss()(outputs, target=targets)                                     # :129,341
    loss.backward()
    res = dict(cfg=np.array(list(cfg.asdict().values()), dtype=np.int64), batch=batch, wseed=wseed, iseed=iseed,
               mask=mask.numpy(), loss=loss.detach().numpy(), outputs=outputs.detach().numpy()[:, :12].copy(),
               targets=targets.numpy()[:, :12].copy())
    params = dict(model.named_parameters())
    for k in GRAD_KEYS:
        g = params[k].grad.detach().numpy()
        res["grad:" + k] = g.reshape(-1)[:8192] if g.size > 8192 else g
    np.savez_compressed(os.path.join(GOLDEN, name + ".npz"), **res)
    print(f"{name}: loss={float(loss):.6f} outputs {tuple(outputs.shape)}")


if __name__ == "__main__":
    torch.manual_seed(0)
    run("videomae_tiny_b4", O.TINY, 4, 3, 4321)
