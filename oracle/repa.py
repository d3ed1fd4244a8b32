"""REPA projector and alignment losses (oracle; test infrastructure only)."""
import math

import torch
import torch.nn.functional as F


def repa_mlp(h, w1, b1, w2, b2):
    """common/repa.py:43-57 — ``fc2(relu(fc1(h)))``."""
    return F.linear(F.relu(F.linear(h, w1, b1)), w2, b2)


def reference_calculate_loss(zs_list, features):
    """common/repa.py:137-154 verbatim semantics: per target module, token-mean of the
    projected states vs the teacher features, MSE, averaged over modules (no grad upstream)."""
    loss = torch.zeros(1, dtype=zs_list[0].dtype)
    for zs in zs_list:
        z = zs.mean(dim=1)
        loss = loss + torch.mean((z - features) ** 2)
    return loss / len(zs_list)


def adaptive_windows(n_in, n_out):
    """adaptive_avg_pool index rule: start=floor(i*n_in/n_out), end=ceil((i+1)*n_in/n_out)."""
    return [(int(math.floor(i * n_in / n_out)), int(math.ceil((i + 1) * n_in / n_out))) for i in range(n_out)]


def repa_cosine_loss(z, y, grid_hw, teacher_hw=(16, 16), eps=1e-8):
    """REPA alignment as re-specified in SURVEY.md §8 A8 (arXiv 2410.06940, cited at
    utils/patch_pixart_sigma_transformer.py:27): parity unpinned by the reference.

    z: [B, h*w, E] projected student tokens (any float dtype), y: [B, th*tw, E] teacher tokens.
    The student grid is average-pooled to the teacher grid with adaptive_avg_pool2d windows,
    then L = -mean_{b,n} cos(z_pool, y); all math in fp32.
    """
    B, N, E = z.shape
    h, w = grid_hw
    th, tw = teacher_hw
    assert N == h * w and y.shape[1] == th * tw
    zf = z.float().reshape(B, h, w, E).permute(0, 3, 1, 2)
    zp = F.adaptive_avg_pool2d(zf, (th, tw)).permute(0, 2, 3, 1).reshape(B, th * tw, E)
    cos = F.cosine_similarity(zp, y.float(), dim=-1, eps=eps)
    return -cos.mean()
