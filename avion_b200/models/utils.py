"""Checkpoint glue of the fine-tune / pretrain entry scripts: OpenCLIP / OpenAI key names -> this package's (= the reference's
flash-attn + fast_conv1) names, and temporal-embedding inflation.  Host-side dict work, same results as
avion/models/utils.py:7-147 (remap_keys_from_open_clip_to_vit :15-98, inflate_positional_embeds :101-147).
"""
from collections import OrderedDict

import torch
import torch.nn.functional as F

# top-level OpenCLIP keys that move under textual./visual. (utils.py:31-39)
_TOP = {"logit_scale": "logit_scale", "visual.proj": "visual.image_projection",
        "positional_embedding": "textual.positional_embedding", "text_projection": "textual.text_projection",
        "token_embedding.weight": "textual.token_embedding.weight", "ln_final.weight": "textual.ln_final.weight",
        "ln_final.bias": "textual.ln_final.bias"}
# nn.MultiheadAttention / Sequential-MLP names -> flash_attn MHA / Mlp names (utils.py:43-48)
_FLASH = {"attn.in_proj_weight": "attn.Wqkv.weight", "attn.in_proj_bias": "attn.Wqkv.bias",
          "attn.out_proj.weight": "attn.out_proj.weight", "attn.out_proj.bias": "attn.out_proj.bias",
          "mlp.c_fc.weight": "mlp.fc1.weight", "mlp.c_fc.bias": "mlp.fc1.bias",
          "mlp.c_proj.weight": "mlp.fc2.weight", "mlp.c_proj.bias": "mlp.fc2.bias"}
_TEXT_BLOCK = ("attn.in_proj_weight", "attn.in_proj_bias", "attn.out_proj.weight", "attn.out_proj.bias", "ln_1.weight",
               "ln_1.bias", "ln_2.weight", "ln_2.bias", "mlp.c_fc.weight", "mlp.c_fc.bias", "mlp.c_proj.weight", "mlp.c_proj.bias")
_SKIPPED = ("visual.proj", "text_projection", "logit_scale")       # re-initialised by the factories (utils.py:62-63)


def enable_grad_checkpointing(model, enable: bool):
    if hasattr(model, "set_grad_checkpointing"):
        model.set_grad_checkpointing(enable=enable)
    else:
        print("{} has no attribute named 'set_grad_checkpointing'".format(model._get_name()))


def _resize_rows(t, rows):
    """Truncate or zero-pad the leading dimension."""
    if rows <= t.shape[0]:
        return t[:rows, :]
    return torch.cat((t, torch.zeros((rows - t.shape[0], t.shape[1]), dtype=t.dtype)), dim=0)


def remap_keys_from_open_clip_to_vit(clip_state_dict, visual_transformer_layers=12, textual_transformer_layers=12,
                                     context_length=77, vocab_size=49408, use_fast_conv1=False, use_flash_attn=False):
    sd = clip_state_dict.get("state_dict", clip_state_dict)
    if next(iter(sd)).startswith("module."):
        sd = OrderedDict((k.replace("module.", ""), v) for k, v in sd.items())
    rename = dict(_TOP)
    if use_flash_attn:
        for i in range(visual_transformer_layers):
            for src, dst in _FLASH.items():
                rename[f"visual.transformer.resblocks.{i}.{src}"] = f"visual.transformer.resblocks.{i}.{dst}"
    for i in range(textual_transformer_layers):
        for name in _TEXT_BLOCK:
            rename[f"transformer.resblocks.{i}.{name}"] = f"textual.transformer.resblocks.{i}.{name}"
    out = OrderedDict()
    for key, value in sd.items():
        if key in _SKIPPED:
            continue
        if use_fast_conv1 and key == "visual.conv1.weight":
            out[key] = value.flatten(1)                               # Conv2d [D,3,p,p] -> Linear [D, 3*p*p]
        elif key not in rename:
            out[key] = value
        elif key == "positional_embedding":
            out[rename[key]] = _resize_rows(value, context_length)
        elif key == "token_embedding.weight":
            assert vocab_size >= value.shape[0]
            out[rename[key]] = _resize_rows(value, vocab_size)
        else:
            out[rename[key]] = value
    return out


def inflate_positional_embeds(current_model_state_dict, new_state_dict, num_frames=4, load_temporal_fix="bilinear"):
    key = "visual.temporal_embedding"
    if key in new_state_dict and key in current_model_state_dict:
        loaded = new_state_dict[key]
        have, dim = loaded.shape
        if have > num_frames:
            new_state_dict[key] = loaded[:num_frames, :]
        elif have < num_frames:
            if load_temporal_fix == "zeros":
                grown = torch.zeros([have, num_frames, dim])          # shape quirk of the reference kept (utils.py:123)
                grown[:have] = loaded
            elif load_temporal_fix in ("interp", "bilinear"):
                mode = "bilinear" if load_temporal_fix == "bilinear" else "nearest"
                grown = F.interpolate(loaded[None, None], (num_frames, dim), mode=mode)[0, 0]
            else:
                raise NotImplementedError
            new_state_dict[key] = grown
    key = "visual.positional_embedding"
    if key in new_state_dict and key in current_model_state_dict:
        if new_state_dict[key].shape[0] != current_model_state_dict[key].shape[0]:
            raise NotImplementedError(
                "Loading models with different spatial resolution / patch number not yet implemented, sorry.")
    return new_state_dict
