"""GPU ``compute_eer`` -- drop-in for ``/root/reference/evaluate_metrics.py:34-40``.

``compute_eer(target_scores, nontarget_scores) -> (eer, threshold)`` accepts numpy arrays, Python
sequences or torch tensors (CPU or CUDA).  Scores are taken to float64 exactly like
``np.array(list_of_python_floats)`` does at ``evaluate.py:138-142``; the sort + scan + argmin run on
the GPU (``occ_eer``: stable LSD radix sort over order-preserving 64-bit keys) and the result is
bit-identical to the numpy reference on the same float64 inputs, ties included.
"""
from __future__ import annotations

import ctypes as C
from typing import Tuple

import numpy as np
import torch

from . import _lib


def _to_dev_f64(x, device) -> torch.Tensor:
    if isinstance(x, torch.Tensor):
        return x.detach().to(device=device, dtype=torch.float64).reshape(-1).contiguous()
    a = np.asarray(x, dtype=np.float64).reshape(-1)
    return torch.from_numpy(np.ascontiguousarray(a)).to(device)


def compute_eer_device(target_scores, nontarget_scores, device=None) -> torch.Tensor:
    """Returns a CUDA float64 tensor ``[eer, threshold]`` without synchronising."""
    if device is None:
        for x in (target_scores, nontarget_scores):
            if isinstance(x, torch.Tensor) and x.is_cuda:
                device = x.device
                break
        else:
            device = torch.device("cuda", torch.cuda.current_device())
    device = torch.device(device)
    if device.type != "cuda":
        raise ValueError("occm_b200.compute_eer runs on a CUDA device (no CPU path)")
    lib = _lib.load()
    t = _to_dev_f64(target_scores, device)
    n = _to_dev_f64(nontarget_scores, device)
    if t.numel() < 1 or n.numel() < 1:
        raise ValueError("compute_eer: need at least one target and one nontarget score")
    res = torch.empty(2, device=device, dtype=torch.float64)
    work = torch.empty(int(lib.occ_eer_workspace_bytes(t.numel() + n.numel())), device=device, dtype=torch.uint8)
    with torch.cuda.device(device):
        stream = torch.cuda.current_stream(device).cuda_stream
        _lib.check(lib.occ_eer(t.data_ptr(), t.numel(), n.data_ptr(), n.numel(), res.data_ptr(), work.data_ptr(), stream),
                   None, "occ_eer")
    return res


def compute_det_curve_device(target_scores, nontarget_scores, device=None):
    """(frr, far, thresholds) as CUDA float64 tensors of n + 1 points -- evaluate_metrics.py:3-21 (``occ_det_curve``)."""
    if device is None:
        for x in (target_scores, nontarget_scores):
            if isinstance(x, torch.Tensor) and x.is_cuda:
                device = x.device
                break
        else:
            device = torch.device("cuda", torch.cuda.current_device())
    device = torch.device(device)
    if device.type != "cuda":
        raise ValueError("occm_b200.compute_det_curve runs on a CUDA device (no CPU path)")
    lib = _lib.load()
    t = _to_dev_f64(target_scores, device)
    n = _to_dev_f64(nontarget_scores, device)
    if t.numel() < 1 or n.numel() < 1:
        raise ValueError("compute_det_curve: need at least one target and one nontarget score")
    m = t.numel() + n.numel() + 1
    frr, far, thr = (torch.empty(m, device=device, dtype=torch.float64) for _ in range(3))
    work = torch.empty(int(lib.occ_eer_workspace_bytes(m - 1)), device=device, dtype=torch.uint8)
    with torch.cuda.device(device):
        stream = torch.cuda.current_stream(device).cuda_stream
        _lib.check(lib.occ_det_curve(t.data_ptr(), t.numel(), n.data_ptr(), n.numel(), frr.data_ptr(), far.data_ptr(), thr.data_ptr(),
                                     work.data_ptr(), stream), None, "occ_det_curve")
    return frr, far, thr


def compute_det_curve(target_scores, nontarget_scores):
    """(frr, far, thresholds) as numpy float64 arrays, like the reference's ``compute_det_curve``."""
    frr, far, thr = compute_det_curve_device(target_scores, nontarget_scores)
    return frr.cpu().numpy(), far.cpu().numpy(), thr.cpu().numpy()


def compute_eer(target_scores, nontarget_scores) -> Tuple[float, float]:
    """(eer, threshold) as Python floats -- evaluate_metrics.py:34-40."""
    r = compute_eer_device(target_scores, nontarget_scores).cpu()
    return float(r[0]), float(r[1])


def calculate_confusion_matrix(target_scores, nontarget_scores, threshold) -> Tuple[int, int, int, int]:
    """(tp, tn, fp, fn) -- ``/root/reference/evaluate_metrics.py:23-32``: tp = #(target > thr), tn = #(nontarget <= thr),
    fn = #(target <= thr), fp = #(nontarget > thr), compared in float64 on the GPU (``occ_confusion_counts``)."""
    device = None
    for x in (target_scores, nontarget_scores):
        if isinstance(x, torch.Tensor) and x.is_cuda:
            device = x.device
            break
    if device is None:
        device = torch.device("cuda", torch.cuda.current_device())
    lib = _lib.load()
    t = _to_dev_f64(target_scores, device)
    n = _to_dev_f64(nontarget_scores, device)
    counts = torch.zeros(4, device=device, dtype=torch.int64)
    with torch.cuda.device(device):
        stream = torch.cuda.current_stream(device).cuda_stream
        _lib.check(lib.occ_confusion_counts(t.data_ptr() if t.numel() else None, t.numel(), n.data_ptr() if n.numel() else None,
                                            n.numel(), float(threshold), counts.data_ptr(), stream), None, "occ_confusion_counts")
    tp, tn, fp, fn = (int(v) for v in counts.cpu())
    return tp, tn, fp, fn
