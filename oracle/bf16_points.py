"""bf16 rounding-point variant of the CPU oracle.  TEST INFRASTRUCTURE ONLY (same rules as oracle/clip_oracle.py).

`Emu(points)` replays oracle/clip_oracle.py's arithmetic (every function there cites the reference lines it follows) in fp32
with a switchable round-to-bf16 at each place where the CUDA path stores bf16: patches, weight matrices, the positional table,
the residual stream, LayerNorm outputs, qkv, the attention probabilities P, the attention output and the MLP hidden
activation.  With ALL_POINTS it is what an ideal (infinitely precise between roundings) implementation of the kernel pipeline
computes - the reference's own bf16-autocast training path rounds at the same places except the residual stream and the table.

Two uses:
  * tests/test_e2e_gpu.py compares the kernels with it at the north-star tolerance: the emulation removes the rounding noise
    that any bf16 implementation has against the fp32 reference, so what remains is kernel error (a biased exponential, a wrong
    accumulation) - a much sharper detector than the fp32 comparison;
  * tools/parity_noise.py measures how far an ideal bf16 pipeline is from the fp32 oracle (the noise floor of the loss tolerance).

Parity status: derived from the pinned fp32 oracle (Emu(()) equals clip_oracle to fp32 round-off, tests/test_oracle.py); the rounding
points themselves restate avion_b200's storage formats, not the reference.
"""
import torch
import torch.nn.functional as F

from . import clip_oracle as O

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
