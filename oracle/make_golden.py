"""Generate tests/golden/*.npz by running the UNMODIFIED reference on CPU (fp32, non-flash branch).

Run in the build container only:  python oracle/make_golden.py
Weights and inputs come from oracle.clip_oracle.make_state_dict / make_inputs (frozen numpy RandomState streams),
so a fixture stores only seeds, config and the reference's outputs.
"""
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle import clip_oracle as O            # noqa: E402
from oracle import ref_import as R             # noqa: E402

GOLDEN = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden")

GRAD_KEYS = ["logit_scale", "visual.conv1.weight", "visual.class_embedding", "visual.temporal_embedding",
             "visual.positional_embedding", "visual.image_projection", "visual.ln_pre.weight",
             "visual.transformer.resblocks.0.attn.Wqkv.weight", "visual.transformer.resblocks.0.attn.Wqkv.bias",
             "visual.transformer.resblocks.1.mlp.fc1.weight", "visual.transformer.resblocks.1.mlp.fc2.bias",
             "visual.transformer.resblocks.1.ln_2.weight", "visual.transformer.resblocks.0.attn.out_proj.weight",
             "textual.text_projection", "textual.positional_embedding",
             "textual.transformer.resblocks.0.attn.in_proj_weight", "textual.transformer.resblocks.1.mlp.c_fc.bias",
             "textual.ln_final.weight"]


def run(name, cfg, batch, wseed, iseed, backward):
    _, _, _, ClipLoss = R.import_reference()
    sd = O.make_state_dict(cfg, wseed)
    clips, tokens = O.make_inputs(cfg, batch, iseed)
    model = R.build_reference_clip(cfg, sd)
    model.train()
    crit = ClipLoss(world_size=1)
    t0 = time.perf_counter()
    # the step of scripts/main_lavila_pretrain.py:814-819; kornia Normalize restated per SURVEY.md 8c
    image = O.clip_transform(clips)
    with torch.set_grad_enabled(backward):
        img, txt, scale = model(image, tokens)
        out = crit(img, txt, scale)
    if backward:
        out["loss"].backward()
    dt = time.perf_counter() - t0
    res = dict(cfg=np.array(list(cfg.asdict().values()), dtype=np.int64), batch=batch, wseed=wseed, iseed=iseed,
               image_embed=img.detach().numpy(), text_embed=txt.detach().numpy(),
               logit_scale=scale.detach().numpy(), loss=out["loss"].detach().numpy(),
               clip_acc=out["clip_acc"].detach().numpy(),
               patches_probe=O.patchify(image, cfg.patch)[:2, :5, :40].numpy())
    if backward:
        params = dict(model.named_parameters())
        for k in GRAD_KEYS:
            rk = R.to_reference_key(k)
            if rk in params and params[rk].grad is not None:
                g = params[rk].grad.detach().numpy()
                if g.size > 8192:                      # keep fixtures small: first rows only
                    g = g.reshape(-1)[:8192]
                res["grad:" + k] = g
    np.savez_compressed(os.path.join(GOLDEN, name + ".npz"), **res)
    print(f"{name}: loss={float(out['loss']):.6f} acc={float(out['clip_acc']):.2f} ({dt:.1f}s)")


if __name__ == "__main__":
    os.makedirs(GOLDEN, exist_ok=True)
    torch.manual_seed(0)
    run("tiny_b6", O.TINY, 6, 0, 1234, backward=True)
    run("tiny_b32", O.TINY, 32, 1, 99, backward=True)
    run("tiny14_b32", O.TINY14, 32, 2, 77, backward=True)          # patch 14: the ViT-L/14 patchify / padded conv1 path
    run("vitb16_4f_b8", O.VITB16_4F, 8, 0, 1234, backward=False)   # BASELINE.json configs[0]
