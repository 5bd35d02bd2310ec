#!/usr/bin/env python
"""Where does the loss error of the bf16 towers come from?  CPU emulation of the kernel pipeline's rounding points.

The CUDA path rounds to bf16 at fixed places (patches, weights, every activation written to HBM, P inside attention).
This tool replays the oracle's arithmetic (oracle/clip_oracle.py) with a switchable bf16 rounding at each of those places
and reports the distribution of |loss - fp32 loss| / loss over input seeds, so that a tolerance failure can be attributed
to a rounding point (or to a biased kernel, if the emulation passes and the kernel does not).  Analysis tool only: nothing
in the product imports it.

  python tools/parity_noise.py --cfg tiny --batch 6 --seeds 20
"""
import argparse
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import clip_oracle as O  # noqa: E402

from oracle.bf16_points import ALL_POINTS, Emu  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--cfg", default="tiny", choices=["tiny", "tiny14"])
    ap.add_argument("--batch", type=int, default=6)
    ap.add_argument("--seeds", type=int, default=16)
    args = ap.parse_args()
    cfg = O.TINY if args.cfg == "tiny" else O.TINY14
    torch.set_grad_enabled(False)
    variants = {"all points (= kernel pipeline)": ALL_POINTS}
    for p in ALL_POINTS:
        variants[f"all but {p}"] = tuple(q for q in ALL_POINTS if q != p)
    for p in ALL_POINTS:
        variants[f"only {p}"] = (p,)
    variants["all but resid+table (fp32 residual stream)"] = tuple(q for q in ALL_POINTS if q not in ("resid", "table"))
    rows = {k: [] for k in variants}
    for s in range(args.seeds):
        sd = O.make_state_dict(cfg, 100 + s)
        clips, tokens = O.make_inputs(cfg, args.batch, 2000 + s)
        ref, _, _ = Emu(()).loss(sd, clips, tokens, cfg)
        for name, pts in variants.items():
            got, _, _ = Emu(pts).loss(sd, clips, tokens, cfg)
            rows[name].append((got - ref) / ref)
    print(f"cfg={args.cfg} batch={args.batch} seeds={args.seeds}: relative loss error vs fp32 (signed mean, rms, max |.|, share > 1e-3)")
    for name, v in rows.items():
        t = torch.tensor(v, dtype=torch.float64)
        print(f"  {name:48s} mean {t.mean():+.2e}  rms {t.pow(2).mean().sqrt():.2e}  max {t.abs().max():.2e}  "
              f">1e-3: {(t.abs() > 1e-3).float().mean():.2f}")


if __name__ == "__main__":
    main()
