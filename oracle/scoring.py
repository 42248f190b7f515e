"""ORACLE (test infrastructure only) -- CPU restatement of the reference's
one-class scoring (`/root/reference/oc_classifier.py`) and EER
(`/root/reference/evaluate_metrics.py`), plus the composed model forward.

Pinning: `compute_eer` here is checked against the reference's own
`evaluate_metrics.compute_eer` (imported unmodified in this container by
`oracle/gen_golden.py`) on the KAT vectors of SURVEY.md 8c and on random
tie-heavy vectors; the outputs are committed in `tests/golden/eer_kat.json`.
`oc_classifier.py` cannot be imported (needs librosa/torchattacks), so its three
scoring functions are restated from the cited lines.

Only tests / smoke / bench's cpu_baseline leg may import this.
"""
from __future__ import annotations

from typing import Dict, Iterable, List, Optional, Sequence, Tuple

import numpy as np
import torch
import torch.nn.functional as F

from . import backend as _be
from . import frontend as _fe


def amodel_forward(sd: Dict[str, torch.Tensor], wav: torch.Tensor) -> Tuple[torch.Tensor, torch.Tensor]:
    """`AModel.forward` (sslassist.py:506-597) on CPU fp32: wav [B,N] or [B,N,1] -> (emb, out)."""
    if wav.ndim == 3:
        wav = wav[:, :, 0]                      # sslassist.py:42-43
    with torch.no_grad():
        feat = _fe.xlsr_forward(sd, wav)
        return _be.aasist_backend(sd, feat)


def reference_embedding(embs: Sequence[torch.Tensor]) -> Tuple[torch.Tensor, torch.Tensor]:
    """`create_reference_embedding2` (oc_classifier.py:182-197): embs is a list of [1,160]
    batch-1 embeddings; reference = mean over the stacked list, threshold = max distance."""
    ref = torch.mean(torch.stack(list(embs)), dim=0)
    d = [F.pairwise_distance(ref, e, p=2) for e in embs]
    return ref, torch.max(torch.stack(d))


def one_class_scores(ref: torch.Tensor, emb: torch.Tensor) -> torch.Tensor:
    """oc_classifier.py:261 -- ||ref - emb + 1e-6||_2 per row (eps is added to the difference)."""
    return F.pairwise_distance(ref, emb, p=2)


def score_lines_1c(dist: Iterable[float], threshold: float) -> List[str]:
    """oc_classifier.py:262-265: '<repr(float)>, 1 \\n' if distance > threshold else '..., 0 \\n'."""
    return [f"{float(d)}, {1 if float(d) > threshold else 0} \n" for d in dist]


def score_lines_2c(out: torch.Tensor) -> List[str]:
    """oc_classifier.py:309-312: logit of class 0, one bare float per line."""
    return [f"{float(o)}\n" for o in out[:, 0]]


def ocsoftmax_scores(x: torch.Tensor, center: torch.Tensor) -> torch.Tensor:
    """OC-Softmax cosine score (Zhang, Jiang & Duan 2021), not in the reference:
    score = <x/||x||, w/||w||>.  PARITY UNPINNED: no reference code or test exists for it."""
    return F.normalize(x, p=2, dim=1) @ F.normalize(center.reshape(-1), p=2, dim=0)


def ocsoftmax_loss(scores: torch.Tensor, labels: torch.Tensor, m_real=0.9, m_fake=0.2, alpha=20.0) -> torch.Tensor:
    """labels: 0 = bona fide, 1 = spoof.  softplus(alpha*(m_real - s)) / softplus(alpha*(s - m_fake))."""
    s = scores.clone()
    s = torch.where(labels == 0, m_real - s, s - m_fake)
    return F.softplus(alpha * s).mean()


# ----------------------------------------------------------------------------- EER
def compute_det_curve(target_scores: np.ndarray, nontarget_scores: np.ndarray):
    """evaluate_metrics.py:3-21.  Stable ascending sort of [target, nontarget]; cumulative
    target count c_i; FRR_i = c_i/N_t, FAR_i = (N_n - (i - c_i))/N_n, each with a leading
    (0, 1) point; thresholds = [min - 0.001, sorted scores]."""
    t = np.asarray(target_scores)
    n = np.asarray(nontarget_scores)
    allv = np.concatenate((t, n))
    is_t = np.concatenate((np.ones(t.size), np.zeros(n.size)))
    order = np.argsort(allv, kind="mergesort")
    c = np.cumsum(is_t[order])
    rejected_non = np.arange(1, allv.size + 1) - c
    frr = np.concatenate(([0.0], c / t.size))
    far = np.concatenate(([1.0], (n.size - rejected_non) / n.size))
    thr = np.concatenate(([allv[order[0]] - 0.001], allv[order]))
    return frr, far, thr


def compute_eer(target_scores: np.ndarray, nontarget_scores: np.ndarray) -> Tuple[float, float]:
    """evaluate_metrics.py:34-40: first argmin of |FRR-FAR|; EER = mean of the two."""
    frr, far, thr = compute_det_curve(target_scores, nontarget_scores)
    i = int(np.argmin(np.abs(frr - far)))
    return float(np.mean((frr[i], far[i]))), float(thr[i])


def eer_from_labels(scores: Sequence[float], labels: Sequence[str]) -> Tuple[float, float]:
    """evaluate.py:127-145: spoof scores are the *target* class."""
    s = np.asarray(scores, dtype=np.float64)
    is_spoof = np.asarray([l == "spoof" for l in labels])
    return compute_eer(s[is_spoof], s[~is_spoof])


def pad(x: np.ndarray, max_len: int = 64600) -> np.ndarray:
    """/root/reference/data_utils_SSL.py:47-54, restated: truncate, or tile the utterance until max_len samples."""
    x_len = x.shape[0]
    if x_len >= max_len:
        return x[:max_len]
    num_repeats = int(max_len / x_len) + 1
    return np.tile(x, (1, num_repeats))[:, :max_len][0]


def pcm16_to_float(x: np.ndarray) -> np.ndarray:
    """What librosa.load / soundfile return for a PCM16 file (oc_classifier.py:93-95): int16 / 32768 as fp32."""
    return (x.astype(np.float32) / np.float32(32768.0)).astype(np.float32)


def calculate_confusion_matrix(target_scores, nontarget_scores, threshold):
    """/root/reference/evaluate_metrics.py:23-32 (tp, tn, fp, fn)."""
    target_scores, nontarget_scores = np.asarray(target_scores), np.asarray(nontarget_scores)
    tp = np.sum(target_scores > threshold)
    tn = np.sum(nontarget_scores <= threshold)
    fn = np.sum(target_scores <= threshold)
    fp = np.sum(nontarget_scores > threshold)
    return int(tp), int(tn), int(fp), int(fn)
