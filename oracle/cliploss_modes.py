"""Shared pieces of the ClipLoss mode tests: the (local_loss, gather_with_grad) matrix of gather_features
(avion/losses/losses.py:53-75), features seeded like the golden fixture, and the comparison with the oracle."""
import os

import numpy as np
import torch

from . import clip_oracle as O

MODES = [(False, False), (False, True), (True, False), (True, True)]
GOLDEN = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden", "cliploss_modes_w2.npz")


def features(n_local, E, world, seed):
    """Same generator calls as oracle/make_golden_cliploss.py:make_features."""
    g = torch.Generator().manual_seed(seed)
    img = torch.nn.functional.normalize(torch.randn(world * n_local, E, generator=g), dim=-1)
    txt = torch.nn.functional.normalize(img + 0.6 * torch.randn(world * n_local, E, generator=g), dim=-1)
    return img, txt


def oracle_world(img, txt, scale, n_local, world, local_loss, gather_with_grad):
    imgs = [img[r * n_local:(r + 1) * n_local].clone().requires_grad_(True) for r in range(world)]
    txts = [txt[r * n_local:(r + 1) * n_local].clone().requires_grad_(True) for r in range(world)]
    sc = torch.tensor(float(scale), requires_grad=True)
    return O.clip_loss_world(imgs, txts, sc, local_loss, gather_with_grad)


def check_rank(out, grads, ref, atol=2e-6, rtol=1e-5):
    """out: {'loss','clip_acc'} of our ClipLoss on this rank; grads: (dimg, dtxt, dscale) left by backward(); ref: the oracle's
    dict for this rank.  Returns a list of mismatch descriptions (empty = parity)."""
    bad = []
    pairs = [("loss", out["loss"]), ("clip_acc", out["clip_acc"]), ("dimg", grads[0]), ("dtxt", grads[1]), ("dscale", grads[2])]
    for name, got in pairs:
        a = torch.as_tensor(got).detach().float().cpu().reshape(-1)
        b = torch.as_tensor(ref[name]).detach().float().cpu().reshape(-1)
        if not torch.allclose(a, b, atol=atol, rtol=rtol):
            bad.append(f"{name}: max |diff| {(a - b).abs().max().item():.3e}")
    return bad


def load_golden():
    g = np.load(GOLDEN)
    n_local, E, world, seed = (int(v) for v in g["meta"])
    return g, n_local, E, world, seed, float(g["scale"])
