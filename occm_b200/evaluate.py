"""``python -m occm_b200.evaluate --score_file S --protocol_file P --metadata_file M --threshold T``

Drop-in for the reference's ``evaluate.py`` command line (``/root/reference/evaluate.py:148-199``): same arguments, same
printed lines (label counts, the sklearn-ordered confusion matrix, TP/TN/FP/FN as the reference indexes them, and
``EER = ..., threshold = ...``).  The counting and the EER run on the GPU (``occ_confusion_counts``, ``occ_eer``); the
label join is the O(N) ``scoring.load_metadata_from_proto``.  No CPU compute path: without a CUDA device it raises.
"""
from __future__ import annotations

import argparse
from typing import List, Optional, Sequence

import numpy as np
import torch

from .evaluate_metrics import calculate_confusion_matrix
from .scoring import (calculate_EER, eval_dict, load_metadata, load_metadata_from_proto, load_proto_file,  # noqa: F401
                      load_score_file)


def confusion_matrix(labels: Sequence[str], scores, threshold: float) -> np.ndarray:
    """``sklearn.metrics.confusion_matrix(labels, predictions)`` for ``predictions = "spoof" if score > threshold else
    "bonafide"`` (evaluate.py:167-188): rows = true label, columns = prediction, both over the SORTED union of the label
    strings that occur (so an unmatched ``""`` label adds a row/column, as it does in the reference).  Each row is one
    ``occ_confusion_counts`` launch over that label's scores (float64 compare on the device)."""
    s = scores if isinstance(scores, torch.Tensor) else torch.as_tensor(np.asarray(scores, dtype=np.float64))
    s = s.to("cuda", torch.float64).reshape(-1)
    lab = np.asarray(list(labels), dtype=object)
    if lab.shape[0] != s.shape[0]:
        raise ValueError("confusion_matrix: %d labels for %d scores" % (lab.shape[0], s.shape[0]))
    rows = sorted(set(lab.tolist()))
    above = below = 0
    per_row = []
    for name in rows:
        idx = torch.from_numpy(np.nonzero(lab == name)[0]).to(s.device)
        gt, _, _, le = calculate_confusion_matrix(s[idx], s[:0], threshold)      # (#>thr, -, -, #<=thr)
        per_row.append((gt, le))
        above += gt
        below += le
    names = sorted(set(rows) | ({"spoof"} if above else set()) | ({"bonafide"} if below else set()))
    cm = np.zeros((len(names), len(names)), dtype=np.int64)
    for name, (gt, le) in zip(rows, per_row):
        r = names.index(name)
        if gt:
            cm[r, names.index("spoof")] = gt
        if le:
            cm[r, names.index("bonafide")] = le
    return cm


def main(argv: Optional[List[str]] = None) -> None:
    parser = argparse.ArgumentParser()
    parser.add_argument("--score_file", type=str, default="score.txt")
    parser.add_argument("--protocol_file", type=str, default="protocol.txt")
    parser.add_argument("--metadata_file", type=str, default="metadata.txt")
    parser.add_argument("--threshold", type=float, default=0.1)
    args = parser.parse_args(argv)
    proto = load_proto_file(args.protocol_file)
    scores = load_score_file(args.score_file)
    labels = load_metadata_from_proto(args.metadata_file, args.protocol_file)
    if len(scores) < len(proto):
        raise IndexError("list index out of range")          # the reference's scores[i] for i over the protocol
    print(f"bona fide = {labels.count('bonafide')}")
    print(f"spoof = {labels.count('spoof')}")
    cm = confusion_matrix(labels, scores[:len(proto)], args.threshold)
    print(cm)
    print(f"TP = {cm[0][0]}")
    print(f"TN = {cm[1][1]}")
    print(f"FP = {cm[0][1]}")
    print(f"FN = {cm[1][0]}")
    # evaluate.py:199 zips ALL scores with the labels (zip stops at the shorter list)
    n = min(len(scores), len(labels))
    calculate_EER(scores[:n], labels[:n], verbose=True)


if __name__ == "__main__":
    main()
