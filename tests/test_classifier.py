"""VideoClassifier (fine-tune head): oracle vs the reference's golden vectors (CPU) and the CUDA path vs
both (GPU): logits cosine >= 0.999 per clip, cross-entropy within 1e-3 relative, gradients cos > 0.99."""
import os

import numpy as np
import pytest
import torch
import torch.nn.functional as F

from oracle import clip_oracle as O

G = np.load(os.path.join(os.path.dirname(__file__), "golden", "classifier_tiny_b6.npz"))
CLASSES = 37


def test_oracle_matches_golden():
    cfg = O.TINY
    sd = O.make_classifier_state_dict(cfg, CLASSES, 5)
    clips, _ = O.make_inputs(cfg, 6, 4242)
    with torch.no_grad():
        logits = O.classifier_forward(sd, O.clip_transform(clips), cfg)
    np.testing.assert_allclose(logits.numpy(), G["logits"], rtol=1e-4, atol=1e-5)
    np.testing.assert_allclose(float(F.cross_entropy(logits, torch.from_numpy(G["labels"]))), float(G["loss"]), rtol=1e-5)


@pytest.mark.gpu
def test_cuda_classifier_matches_reference_golden():
    from avion_b200.models import VideoClassifier, VisionTransformer
    cfg = O.TINY
    sd = O.make_classifier_state_dict(cfg, CLASSES, 5)
    vis = VisionTransformer(cfg.image_size, cfg.patch, cfg.width, cfg.layers, cfg.heads, 4, num_frames=cfg.frames,
                            output_dim=cfg.embed_dim)
    model = VideoClassifier(vis, dropout=0.0, num_classes=CLASSES).cuda()
    assert set(model.state_dict()) == set(sd)
    model.load_state_dict(sd, strict=True)
    clips, _ = O.make_inputs(cfg, 6, 4242)
    logits = model(clips.cuda())
    assert logits.shape == (6, CLASSES) and logits.dtype == torch.float32
    loss = F.cross_entropy(logits, torch.from_numpy(G["labels"]).cuda())
    loss.backward()
    assert F.cosine_similarity(logits.detach().cpu(), torch.from_numpy(G["logits"]), dim=-1).min() >= 0.999
    assert abs(float(loss) - float(G["loss"])) <= 1e-3 * float(G["loss"])
    params = dict(model.named_parameters())
    for k in G.files:
        if k.startswith("grad:"):
            ref = torch.from_numpy(G[k]).float()
            got = params[k[5:]].grad.detach().float().cpu().reshape(-1)[:ref.numel()]
            assert F.cosine_similarity(got, ref, dim=0) > 0.99, k
