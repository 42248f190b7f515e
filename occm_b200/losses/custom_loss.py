"""``losses/custom_loss.py`` surface.

``OCSoftmax`` is the class BASELINE.json's north_star names.  It is NOT in the reference
(``/root/reference/losses/custom_loss.py`` defines only four training-loss functions):
**parity unpinned** -- the definition follows the published OC-Softmax (Zhang, Jiang & Duan 2021):
``score = <x/||x||, w/||w||>``; loss ``softplus(alpha*(m_real - s))`` for bona fide (label 0) and
``softplus(alpha*(s - m_fake))`` for spoof (label 1), alpha=20, m_real=0.9, m_fake=0.2.  The cosine
scores are computed by the ``occ_score_cos`` kernel; the scalar loss on top of them is plumbing.

The reference's four loss functions are training-only (``oc_training.py:380-382``), outside the
scoring hot path; they are kept here with the reference's names and behaviour
(``custom_loss.py:4-30, 32-57, 59-74, 78-99``) so that ``from losses.custom_loss import *`` still works.
"""
from __future__ import annotations

import torch
import torch.nn as nn
import torch.nn.functional as F

from .. import _lib


def cosine_scores(x: torch.Tensor, center: torch.Tensor) -> torch.Tensor:
    """OC-Softmax score per row of ``x`` [B,D] against ``center`` [D] (CUDA kernel ``occ_score_cos``)."""
    if not x.is_cuda:
        raise ValueError("occm_b200: OC-Softmax scoring needs CUDA tensors (no CPU path)")
    lib = _lib.load()
    xe = x.detach().to(torch.float32).contiguous()
    c = center.detach().to(device=x.device, dtype=torch.float32).reshape(-1).contiguous()
    B, D = xe.shape
    if c.numel() != D:
        raise ValueError("occm_b200: centre and embedding dimensions differ")
    out = torch.empty(B, device=x.device, dtype=torch.float32)
    with torch.cuda.device(x.device):
        _lib.check(lib.occ_score_cos(xe.data_ptr(), c.data_ptr(), B, D, out.data_ptr(),
                                     torch.cuda.current_stream(x.device).cuda_stream), None, "occ_score_cos")
    return out


class OCSoftmax(nn.Module):
    def __init__(self, feat_dim: int = 160, m_real: float = 0.9, m_fake: float = 0.2, alpha: float = 20.0):
        super().__init__()
        self.feat_dim, self.m_real, self.m_fake, self.alpha = feat_dim, m_real, m_fake, alpha
        self.center = nn.Parameter(torch.randn(1, feat_dim))
        nn.init.kaiming_uniform_(self.center, 0.25)

    def score(self, x: torch.Tensor) -> torch.Tensor:
        return cosine_scores(x, self.center)

    def forward(self, x: torch.Tensor, labels: torch.Tensor):
        scores = self.score(x)
        margin = torch.where(labels.to(scores.device) == 0, self.m_real - scores, scores - self.m_fake)
        return F.softplus(self.alpha * margin).mean(), scores


# ---------------------------------------------------------------------------------------------
# training-only losses of the reference (not accelerated; plain torch on whatever device x is on)
def compactness_loss(batch_embeddings):
    """custom_loss.py:4-30: mean distance of each of the first 6 (bona fide) embeddings to the mean of the other 5."""
    e = batch_embeddings[:6]
    d = []
    for i in range(len(e)):
        rest = torch.cat((e[:i], e[i + 1:]), dim=0).mean(dim=0, keepdim=True)
        d.append(F.pairwise_distance(e[i:i + 1], rest, p=2))
    return torch.cat(d).mean()


def triplet_loss(batch_embeddings, margin=9.0):
    """custom_loss.py:32-57: relu(d(bona1,bona2) - d(bona1,spoof1) + margin)."""
    a, p, n = (batch_embeddings[i].unsqueeze(0) for i in range(3))
    return F.relu(F.pairwise_distance(a, p, p=2) - F.pairwise_distance(a, n, p=2) + margin)


def euclidean_distance_loss(batch_embeddings):
    """custom_loss.py:59-74: mean L2 distance over the pairs (0,1),(0,2),(0,3),(2,1),(2,3)."""
    pairs = [(0, 1), (0, 2), (0, 3), (2, 1), (2, 3)]
    tot = 0.0
    for i, j in pairs:
        tot = tot + F.pairwise_distance(batch_embeddings[i].unsqueeze(0), batch_embeddings[j].unsqueeze(0), p=2)
    return tot / len(pairs)


def descriptiveness_loss(batch_embeddings, labels):
    """custom_loss.py:78-99: summed cross entropy / batch size."""
    return F.cross_entropy(batch_embeddings, labels, reduction="none").sum() / len(batch_embeddings)
