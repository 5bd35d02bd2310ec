"""GPU-side clip transform: the fused replacement of Permute + kornia Normalize (+ the model's patchify)
(scripts/main_lavila_pretrain.py:425-435,816-817; avion/data/transforms.py:10-26; avion/models/transformer.py:319)."""
import torch
import torch.nn as nn

from .. import ops

MEAN_OPENAI = (108.3272985, 116.7460125, 104.09373615000001)
STD_OPENAI = (68.5005327, 66.6321579, 70.32316305)
MEAN_TIMM = (0.485 * 255, 0.456 * 255, 0.406 * 255)
STD_TIMM = (0.229 * 255, 0.224 * 255, 0.225 * 255)


class GPUClipTransform(nn.Module):
    """uint8 decord frames [B,T,Hs,Ws,3] -> bf16 patches [B, T*(crop/patch)^2, 3*patch^2] in one kernel.

    boxes (optional, int32 on the clips' device), one row per clip:
      [B,3] = (x, y, hflip)       crop a crop x crop window at (x, y) - no resize;
      [B,5] = (x, y, w, h, hflip) decord's random-RESIZED crop: the w x h box at (x, y) is resized to crop x crop
                                  (antialiased bilinear, rounded to uint8 like decoded frames) - rows of `sample_boxes`."""

    def __init__(self, crop=224, patch=16, mean=MEAN_OPENAI, std=STD_OPENAI):
        super().__init__()
        self.crop, self.patch, self.mean, self.std = crop, patch, tuple(mean), tuple(std)

    def forward(self, frames_u8, boxes=None):
        if boxes is not None and boxes.shape[-1] == 5:
            return ops.clip_resize_transform(frames_u8, boxes, self.crop, self.patch, self.mean, self.std)
        return ops.clip_transform(frames_u8, self.crop, self.patch, self.mean, self.std, boxes=boxes)

    @staticmethod
    def sample_boxes(seed, n, src_w, src_h, **kw):
        return ops.sample_rrc_boxes(seed, n, src_w, src_h, **kw)


class TubeMaskingGeneratorGPU:
    """avion/data/transforms.py:279-302: one random spatial mask per clip, repeated over the tubelet time steps.  Same torch
    RNG calls as the reference (so the same seed gives the same masks); the result is returned on `device` as bool [B, N]
    already flattened (the train loop's `.flatten(1).to(torch.bool)`, scripts/main_videomae_pretrain.py:319, is a no-op on it)."""

    def __init__(self, batch, input_size, mask_ratio, device=0):
        self.batch = batch
        self.frames, self.height, self.width = input_size
        self.num_patches_per_frame = self.height * self.width
        self.total_patches = self.frames * self.num_patches_per_frame
        self.num_masks_per_frame = int(mask_ratio * self.num_patches_per_frame)
        self.total_masks = self.frames * self.num_masks_per_frame
        self.device = device

    def __repr__(self):
        return "Maks: total patches {}, mask patches {}".format(self.total_patches, self.total_masks)

    def __call__(self):
        tensors = torch.zeros((self.batch, self.num_patches_per_frame))
        indexes = torch.argsort(torch.rand(self.batch, self.num_patches_per_frame))[:, :self.num_masks_per_frame]
        tensors[torch.arange(self.batch).unsqueeze(1), indexes] = 1
        masks = tensors.repeat(1, self.frames)
        dev = self.device if isinstance(self.device, (str, torch.device)) else torch.device("cuda", self.device)
        return masks.to(dev, non_blocking=True).to(torch.bool)
