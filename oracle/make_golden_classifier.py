"""tests/golden/classifier_tiny_b6.npz from the UNMODIFIED reference VideoClassifier (avion/models/model_clip.py:15-38) over
the reference VisionTransformer (non-flash branch, fp32, CPU), cross-entropy on seeded labels as the fine-tune loop does
(scripts/main_lavila_finetune_cls.py).  Build container only."""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle import clip_oracle as O            # noqa: E402
from oracle import ref_import as R             # noqa: E402

GOLDEN = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden")
CLASSES = 37      # not a multiple of 8, like EK-100's 97 verbs / 3806 actions

if __name__ == "__main__":
    VisionTransformer, _, _, _ = R.import_reference()
    from avion.models.model_clip import VideoClassifier
    cfg, B = O.TINY, 6
    sd = O.make_classifier_state_dict(cfg, CLASSES, 5)
    vis = VisionTransformer(cfg.image_size, cfg.patch, cfg.width, cfg.layers, cfg.heads, 4, output_dim=cfg.embed_dim,
                            num_frames=cfg.frames, use_fast_conv1=True, use_flash_attn=False)
    model = VideoClassifier(vis, dropout=0.0, num_classes=CLASSES).float()
    missing, unexpected = model.load_state_dict({R.to_reference_key(k): v for k, v in sd.items()}, strict=True)
    model.train()
    clips, _ = O.make_inputs(cfg, B, 4242)
    labels = torch.from_numpy(np.random.RandomState(1).randint(0, CLASSES, B))
    logits = model(O.clip_transform(clips))
    loss = torch.nn.functional.cross_entropy(logits, labels)
    loss.backward()
    p = dict(model.named_parameters())
    np.savez_compressed(os.path.join(GOLDEN, "classifier_tiny_b6.npz"), logits=logits.detach().numpy(), loss=loss.detach().numpy(),
                        labels=labels.numpy(), **{"grad:" + k: p[R.to_reference_key(k)].grad.numpy().reshape(-1)[:8192]
                                                  for k in ("fc_cls.weight", "fc_cls.bias", "visual.conv1.weight",
                                                            "visual.ln_post.weight", "visual.transformer.resblocks.1.mlp.fc2.weight")})
    print("loss", float(loss), logits.shape)
