"""Waveform ingest on the GPU -- the loader side of the scoring path.

The reference decodes one FLAC / WAV file per ``__getitem__`` with librosa on the host
(``/root/reference/oc_classifier.py:85-105``: ``librosa.load(path, sr=None)`` -> fp32 in [-1, 1), i.e. PCM16 / 32768) and the
training datasets bring every utterance to 64 600 samples with ``pad`` (``/root/reference/data_utils_SSL.py:47-54``:
truncate, or ``np.tile`` the utterance).  Here the 16-bit samples travel to the device as they are (half the bytes of fp32)
and one kernel (``occ_ingest_pcm16``) produces the fp32 batch: scaled, and either zero-padded (variable-length batch; pass
``lengths`` to ``AModel.forward``) or repeat-padded exactly like ``pad``.  File decoding itself (FLAC) stays on the host.
"""
from __future__ import annotations

from typing import List, Optional, Sequence, Tuple

import numpy as np
import torch

from . import _lib


def batch_from_pcm16(utterances: Sequence, max_len: int = 64600, mode: str = "repeat", device=None
                     ) -> Tuple[torch.Tensor, List[int]]:
    """utterances: sequence of 1-D int16 arrays / tensors (ragged).  Returns (wav [B, max_len] fp32 CUDA, lengths) where
    lengths[b] = min(len_b, max_len) (``mode="zero"``: the valid length to hand to ``forward(x, lengths)``) or max_len
    (``mode="repeat"``: every row is full after ``pad``)."""
    if mode not in ("repeat", "zero"):
        raise ValueError("mode must be 'repeat' (data_utils_SSL.pad) or 'zero'")
    dev = torch.device(device) if device is not None else torch.device("cuda", torch.cuda.current_device())
    if dev.type != "cuda":
        raise ValueError("occm_b200.ingest runs on a CUDA device (no CPU path)")
    arrs = []
    for u in utterances:
        a = u.detach().cpu().numpy() if isinstance(u, torch.Tensor) else np.asarray(u)
        if a.dtype != np.int16 or a.ndim != 1:
            raise ValueError("utterances must be 1-D int16 PCM")
        arrs.append(np.ascontiguousarray(a))
    B = len(arrs)
    lens = np.array([len(a) for a in arrs], dtype=np.int32)
    if B == 0:
        return torch.empty(0, max_len, device=dev), []
    if (lens < 1).any():
        raise ValueError("empty utterance")      # the reference divides by x_len (data_utils_SSL.py:52)
    offs = np.zeros(B, dtype=np.int64)
    offs[1:] = np.cumsum(lens[:-1], dtype=np.int64)
    flat = torch.from_numpy(np.concatenate(arrs)).pin_memory()
    pcm = flat.to(dev, non_blocking=True)
    offs_d = torch.from_numpy(offs).to(dev)
    lens_d = torch.from_numpy(lens).to(dev)
    wav = torch.empty(B, max_len, device=dev, dtype=torch.float32)
    lib = _lib.load()
    with torch.cuda.device(dev):
        stream = torch.cuda.current_stream(dev).cuda_stream
        _lib.check(lib.occ_ingest_pcm16(pcm.data_ptr(), offs_d.data_ptr(), lens_d.data_ptr(), B, max_len, 1 if mode == "repeat" else 0,
                                        wav.data_ptr(), stream), None, "occ_ingest_pcm16")
    out_lens = [max_len] * B if mode == "repeat" else [int(min(n, max_len)) for n in lens]
    return wav, out_lens


def pad(x, max_len: int = 64600, device=None) -> torch.Tensor:
    """Drop-in for ``data_utils_SSL.pad`` on one int16 utterance: returns the fp32 [max_len] CUDA tensor."""
    wav, _ = batch_from_pcm16([x], max_len, "repeat", device)
    return wav[0]
