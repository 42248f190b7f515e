"""Drop-in surface of the reference's ``models/senet.py`` on B200 (other backends on the shared frontend).

``ssl_resnet34(device)`` (``/root/reference/models/senet.py:162-185``): XLS-R frontend + ``se_resnet34()``; ``forward(x)``
returns ``(com [B,128], des [B,2])`` like ``ResNet.forward`` (``:121-143``).  Parameters only, under the reference's key
layout (``frontend.model.*`` + ``resnet34.*``); the arithmetic runs in ``libocc_b200.so`` (``occ_forward_resnet34`` /
``occ_resnet34_backend``: the frontend's tcgen05 kernels + the fp32 kernels of ``csrc/senet.cu``).  Eval mode only.
"""
from __future__ import annotations

from typing import Tuple

import torch

from .. import _lib, synth
from .sslassist import SSLModel, _EngineOwner, _Tree, _as_wave


def _features(feat: torch.Tensor) -> torch.Tensor:
    if feat.ndim == 4 and feat.shape[1] == 1:
        feat = feat[:, 0]
    if feat.ndim != 3 or feat.shape[-1] != 1024 or not feat.is_cuda:
        raise ValueError("occm_b200: expected CUDA features [B,T,1024] (or [B,1,T,1024])")
    return feat.to(torch.float32).contiguous()


class ssl_resnet34(_EngineOwner):
    """``ssl_resnet34(device)`` of ``models/senet.py:162-185`` (random-init from the seeded synthetic checkpoints)."""

    def __init__(self, device, seed: int = 0, dtype=None):
        super().__init__()
        self.frontend = SSLModel(device, seed=seed, dtype=dtype, _owner=self)
        self.resnet34 = _Tree.build({k[len(synth.RESNET_PREFIX):]: v for k, v in synth.make_resnet34_state_dict(seed).items()})
        self.device = device
        self._init_engine_state(dtype)
        for m in self.modules():
            m.training = False
        if device is not None and torch.device(device).type == "cuda":
            self.to(device)

    def _engine_state_dict(self):
        return self.state_dict()

    def forward(self, x: torch.Tensor) -> Tuple[torch.Tensor, torch.Tensor]:
        x = _as_wave(x)
        B, N = x.shape
        eng = self._get_engine(B, N, x.device)
        com = torch.empty(B, 128, device=x.device, dtype=torch.float32)
        des = torch.empty(B, 2, device=x.device, dtype=torch.float32)
        stream = torch.cuda.current_stream(x.device).cuda_stream
        with torch.cuda.device(x.device):
            _lib.check(eng.lib.occ_forward_resnet34(eng.handle, x.data_ptr(), B, N, com.data_ptr(), des.data_ptr(), stream),
                       eng.handle, "occ_forward_resnet34")
        return com, des

    def resnet34_forward(self, feat: torch.Tensor) -> Tuple[torch.Tensor, torch.Tensor]:
        """``ResNet.forward`` alone (senet.py:121-143) on XLS-R features ``[B,T,1024]`` / ``[B,1,T,1024]``."""
        return _resnet34_on_features(self, feat)


def _resnet34_on_features(owner: _EngineOwner, feat: torch.Tensor):
    f = _features(feat)
    B, T = f.shape[0], f.shape[1]
    eng = owner._get_engine(B, max(400 * T, 64600), f.device)
    com = torch.empty(B, 128, device=f.device, dtype=torch.float32)
    des = torch.empty(B, 2, device=f.device, dtype=torch.float32)
    stream = torch.cuda.current_stream(f.device).cuda_stream
    with torch.cuda.device(f.device):
        _lib.check(eng.lib.occ_resnet34_backend(eng.handle, f.data_ptr(), B, T, com.data_ptr(), des.data_ptr(), stream),
                   eng.handle, "occ_resnet34_backend")
    return com, des
