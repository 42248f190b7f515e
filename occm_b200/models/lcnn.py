"""Drop-in surface of the reference's ``models/lcnn.py`` on B200 (other backends on the shared frontend).

``ssl_lcnn(device)`` (``/root/reference/models/lcnn.py:236-260``): XLS-R frontend + ``LCNN(asoftmax=False)``; ``forward(x)``
returns the two logits ``[B, 2]``.  The module holds only parameters, under the reference's key layout
(``frontend.model.*`` + ``lcnn.*``); the arithmetic runs in ``libocc_b200.so`` (``occ_forward_lcnn`` / ``occ_lcnn_backend``:
the frontend's tcgen05 kernels + the fused fp32 LCNN kernels of ``csrc/lcnn.cu``).  Eval mode only; no CPU path.
"""
from __future__ import annotations

from typing import Optional

import torch

from .. import _lib, synth
from .sslassist import SSLModel, _EngineOwner, _Tree, _as_wave


class ssl_lcnn(_EngineOwner):
    """``ssl_lcnn(device)`` of ``models/lcnn.py:236-260``.  Random-init from the seeded synthetic checkpoints
    (``synth.make_state_dict`` / ``synth.make_lcnn_state_dict``); load real weights with ``load_state_dict``."""

    def __init__(self, device, seed: int = 0, dtype=None):
        super().__init__()
        self.frontend = SSLModel(device, seed=seed, dtype=dtype, _owner=self)
        self.lcnn = _Tree.build({k[len(synth.LCNN_PREFIX):]: v for k, v in synth.make_lcnn_state_dict(seed).items()})
        self.device = device
        self._init_engine_state(dtype)
        for m in self.modules():
            m.training = False
        if device is not None and torch.device(device).type == "cuda":
            self.to(device)

    def _engine_state_dict(self):
        return self.state_dict()          # frontend.model.* (accepted by the library as ssl_model.model.*) + lcnn.*

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        """[B,N] / [B,N,1] CUDA fp32 waveforms -> logits [B,2] (``ssl_lcnn.forward``, lcnn.py:242-260)."""
        x = _as_wave(x)
        B, N = x.shape
        eng = self._get_engine(B, N, x.device)
        out = torch.empty(B, 2, device=x.device, dtype=torch.float32)
        stream = torch.cuda.current_stream(x.device).cuda_stream
        with torch.cuda.device(x.device):
            _lib.check(eng.lib.occ_forward_lcnn(eng.handle, x.data_ptr(), B, N, out.data_ptr(), stream), eng.handle, "occ_forward_lcnn")
        return out

    def lcnn_forward(self, feat: torch.Tensor) -> torch.Tensor:
        """``LCNN.forward`` alone (lcnn.py:189-213): XLS-R features ``[B,T,1024]`` or ``[B,1,T,1024]`` (CUDA fp32) -> logits ``[B,2]``."""
        if feat.ndim == 4 and feat.shape[1] == 1:
            feat = feat[:, 0]
        if feat.ndim != 3 or feat.shape[-1] != 1024 or not feat.is_cuda:
            raise ValueError("occm_b200: lcnn_forward expects CUDA features [B,T,1024] (or [B,1,T,1024])")
        f = feat.to(torch.float32).contiguous()
        B, T = f.shape[0], f.shape[1]
        eng = self._get_engine(B, max(400 * T, 64600), f.device)     # workspace for at least T frames (320 samples per frame)
        out = torch.empty(B, 2, device=f.device, dtype=torch.float32)
        stream = torch.cuda.current_stream(f.device).cuda_stream
        with torch.cuda.device(f.device):
            _lib.check(eng.lib.occ_lcnn_backend(eng.handle, f.data_ptr(), B, T, out.data_ptr(), stream), eng.handle, "occ_lcnn_backend")
        return out
