"""tests/golden/ranking_n12.npz from the UNMODIFIED reference MaxMarginRankingLoss (avion/losses/losses.py:163-227), run in the
build container only.  The reference's gather_features calls dist.all_gather even at world_size 1, hence the 1-rank gloo group."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle import ranking_oracle as O         # noqa: E402
from oracle import ref_import as R             # noqa: E402

GOLDEN = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden")

if __name__ == "__main__":
    R.import_reference()
    from avion.losses.losses import MaxMarginRankingLoss
    dist.init_process_group("gloo", init_method="tcp://127.0.0.1:29517", rank=0, world_size=1)
    res = {}
    for tag, n, E, seed, margin, fix in (("a", 12, 32, 0, 0.2, True), ("b", 12, 32, 1, 0.4, False), ("c", 7, 40, 2, 0.05, True)):
        img, txt = O.make_features(n, E, seed)
        i, t = img.clone().requires_grad_(True), txt.clone().requires_grad_(True)
        out = MaxMarginRankingLoss(margin=margin, fix_norm=fix)(i, t)
        out["loss"].backward()
        res.update({f"{tag}:cfg": np.array([n, E, seed, int(fix)]), f"{tag}:margin": np.float64(margin),
                    f"{tag}:loss": out["loss"].detach().numpy(), f"{tag}:dimage": i.grad.numpy(), f"{tag}:dtext": t.grad.numpy()})
        print(tag, float(out["loss"]))
    np.savez_compressed(os.path.join(GOLDEN, "ranking_n12.npz"), **res)
    dist.destroy_process_group()
