"""CPU oracle for MaxMarginRankingLoss (avion/losses/losses.py:152-227).  TEST INFRASTRUCTURE ONLY (see clip_oracle.py).

Parity status: PINNED - `oracle/make_golden_ranking.py` runs the unmodified reference class (single-rank gloo group) and
commits loss + feature gradients as tests/golden/ranking_n12.npz; tests/test_oracle_ranking.py checks this restatement.
"""
import numpy as np
import torch


def make_features(n, E, seed):
    rs = np.random.RandomState(seed)
    img = torch.from_numpy(rs.standard_normal((n, E)).astype(np.float32))
    txt = torch.from_numpy((0.25 * img.numpy() + 0.97 * rs.standard_normal((n, E))).astype(np.float32))   # weakly correlated pairs
    return img, txt


def sim_matrix(a, b, eps=1e-8):
    """losses.py:152-160."""
    a_n = a.norm(dim=1, keepdim=True).clamp_min(eps)
    b_n = b.norm(dim=1, keepdim=True).clamp_min(eps)
    return (a / a_n) @ (b / b_n).t()


def max_margin_loss(image, text, margin=0.2, fix_norm=True):
    """losses.py:197-225 written per pair: for every (i, j): relu(m - x_ii + x_ij) and relu(m - x_ii + x_ji); the mean runs
    over the 2 n (n-1) off-diagonal terms with fix_norm, over all 2 n^2 terms otherwise."""
    x = sim_matrix(text, image)
    n = x.shape[0]
    d = torch.diag(x).unsqueeze(1)                       # x_ii, broadcast along j
    terms = torch.relu(margin - d + x) + torch.relu(margin - d + x.t())
    if fix_norm:
        off = ~torch.eye(n, dtype=torch.bool)
        return terms[off].sum() / (2 * n * (n - 1))
    return terms.sum() / (2 * n * n)


def loss_and_grads(image, text, margin=0.2, fix_norm=True):
    i, t = image.clone().requires_grad_(True), text.clone().requires_grad_(True)
    loss = max_margin_loss(i, t, margin, fix_norm)
    loss.backward()
    return loss.detach(), i.grad, t.grad
