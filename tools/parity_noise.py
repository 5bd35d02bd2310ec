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
import torch.nn.functional as F

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import clip_oracle as O  # noqa: E402

ALL_POINTS = ("patches", "weights", "table", "resid", "ln_out", "qkv", "p", "attn_out", "mlp_hidden")


class Emu:
    def __init__(self, points):
        self.points = set(points)

    def r(self, x, point):
        return x.bfloat16().float() if point in self.points else x

    def block(self, x, sd, p, heads, eps, names, mask=None):
        qkv_w, qkv_b, fc1, fc2 = names
        B, S, D = x.shape
        hd = D // heads
        h = self.r(F.layer_norm(x, (D,), sd[p + "ln_1.weight"], sd[p + "ln_1.bias"], eps), "ln_out")
        qkv = self.r(F.linear(h, self.r(sd[p + qkv_w], "weights"), sd[p + qkv_b]), "qkv").view(B, S, 3, heads, hd)
        q, k, v = (qkv[:, :, i].transpose(1, 2) for i in range(3))
        s = (q @ k.transpose(-1, -2)) * (hd ** -0.5)
        if mask is not None:
            s = s + mask
        m = s.max(dim=-1, keepdim=True).values
        e = torch.exp(s - m)
        l = e.sum(-1, keepdim=True)                       # the kernel sums the unrounded exponentials
        o = (self.r(e, "p") @ v) / l
        o = self.r(o.transpose(1, 2).reshape(B, S, D), "attn_out")
        x = self.r(x + F.linear(o, self.r(sd[p + "attn.out_proj.weight"], "weights"), sd[p + "attn.out_proj.bias"]), "resid")
        h = self.r(F.layer_norm(x, (D,), sd[p + "ln_2.weight"], sd[p + "ln_2.bias"], eps), "ln_out")
        a = self.r(F.gelu(F.linear(h, self.r(sd[p + fc1 + ".weight"], "weights"), sd[p + fc1 + ".bias"])), "mlp_hidden")
        return self.r(x + F.linear(a, self.r(sd[p + fc2 + ".weight"], "weights"), sd[p + fc2 + ".bias"]), "resid")

    def vision(self, sd, clips, cfg):
        v = "visual."
        patches = self.r(O.patchify(O.clip_transform(clips), cfg.patch), "patches")
        x = F.linear(patches, self.r(sd[v + "conv1.weight"], "weights"))
        G = cfg.grid * cfg.grid
        pos = sd[v + "positional_embedding"]
        if cfg.frames > 1:
            tile = pos[1:].repeat(cfg.frames, 1) + sd[v + "temporal_embedding"].repeat_interleave(G, 0)
            pos_all = torch.cat([pos[:1] + sd[v + "class_embedding"], tile], dim=0)
        else:
            pos_all = torch.cat([pos[:1] + sd[v + "class_embedding"], pos[1:]], dim=0)
        table = self.r(pos_all, "table")
        B = x.shape[0]
        x = torch.cat([table[:1].expand(B, 1, -1), x + table[1:]], dim=1)
        x = self.r(x, "resid")
        D = cfg.width
        x = self.r(F.layer_norm(x, (D,), sd[v + "ln_pre.weight"], sd[v + "ln_pre.bias"], 1e-6), "resid")
        for i in range(cfg.layers):
            x = self.block(x, sd, f"{v}transformer.resblocks.{i}.", cfg.heads, 1e-6,
                           ("attn.Wqkv.weight", "attn.Wqkv.bias", "mlp.fc1", "mlp.fc2"))
        x = F.layer_norm(x[:, 0], (D,), sd[v + "ln_post.weight"], sd[v + "ln_post.bias"], 1e-6)
        return x @ sd[v + "image_projection"]

    def text(self, sd, text, cfg):
        t = "textual."
        W = cfg.t_width
        x = self.r(sd[t + "token_embedding.weight"][text] + sd[t + "positional_embedding"], "resid")
        mask = torch.full((cfg.ctx, cfg.ctx), float("-inf")).triu_(1)
        for i in range(cfg.t_layers):
            x = self.block(x, sd, f"{t}transformer.resblocks.{i}.", cfg.t_heads, 1e-5,
                           ("attn.in_proj_weight", "attn.in_proj_bias", "mlp.c_fc", "mlp.c_proj"), mask=mask)
        x = F.layer_norm(x, (W,), sd[t + "ln_final.weight"], sd[t + "ln_final.bias"], 1e-5)
        x = x[torch.arange(x.shape[0]), text.argmax(dim=-1)]
        return x @ sd[t + "text_projection"]

    def loss(self, sd, clips, tokens, cfg):
        img = F.normalize(self.vision(sd, clips, cfg), dim=-1)
        txt = F.normalize(self.text(sd, tokens, cfg), dim=-1)
        return O.clip_loss(img, txt, sd["logit_scale"].exp())["loss"].item(), img, txt


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
