"""MaxMarginRankingLoss: oracle restatement vs the golden vectors of the unmodified reference."""
import os

import numpy as np

from oracle import ranking_oracle as O

G = np.load(os.path.join(os.path.dirname(__file__), "golden", "ranking_n12.npz"))
TAGS = ("a", "b", "c")


def case(tag):
    n, E, seed, fix = (int(v) for v in G[f"{tag}:cfg"])
    return n, E, seed, bool(fix), float(G[f"{tag}:margin"])


def test_oracle_matches_golden():
    for tag in TAGS:
        n, E, seed, fix, margin = case(tag)
        img, txt = O.make_features(n, E, seed)
        loss, di, dt = O.loss_and_grads(img, txt, margin, fix)
        np.testing.assert_allclose(float(loss), float(G[f"{tag}:loss"]), rtol=1e-6)
        np.testing.assert_allclose(di.numpy(), G[f"{tag}:dimage"], rtol=1e-4, atol=1e-7)
        np.testing.assert_allclose(dt.numpy(), G[f"{tag}:dtext"], rtol=1e-4, atol=1e-7)
