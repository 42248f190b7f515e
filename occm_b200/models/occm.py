"""Drop-in surface of the reference's ``models/occm.py`` on B200: the two-branch ``OCCM`` model (``:48-67``).

``OCCM(device)``: one XLS-R frontend pass, then the SE-ResNet34 branch and the LCNN branch on the same features;
``forward(x)`` returns ``((com [B,128], des [B,2]), lcnn_logits [B,2])`` like the reference (whose two debugging
``print(x.shape)`` calls at ``:62,64`` are not reproduced).  State-dict layout as in the reference: ``frontend.model.*``,
``senet34_branch.*``, ``lcnn_branch.*``.  Eval mode only; all arithmetic in ``libocc_b200.so``.
"""
from __future__ import annotations

import torch

from .. import _lib, synth
from .senet import _resnet34_on_features
from .sslassist import SSLModel, _EngineOwner, _Tree


class OCCM(_EngineOwner):
    def __init__(self, device, seed: int = 0, dtype=None):
        super().__init__()
        self.frontend = SSLModel(device, seed=seed, dtype=dtype, _owner=self)
        self.senet34_branch = _Tree.build({k[len(synth.RESNET_PREFIX):]: v for k, v in synth.make_resnet34_state_dict(seed).items()})
        self.lcnn_branch = _Tree.build({k[len(synth.LCNN_PREFIX):]: v for k, v in synth.make_lcnn_state_dict(seed).items()})
        self.device = device
        self._init_engine_state(dtype)
        for m in self.modules():
            m.training = False
        if device is not None and torch.device(device).type == "cuda":
            self.to(device)

    def _engine_state_dict(self):
        return self.state_dict()

    def forward(self, x: torch.Tensor):
        feat = self.frontend.extract_feat(x)                       # one frontend pass (occ_extract_feat) feeds both branches
        senet34_output = _resnet34_on_features(self, feat)
        B, T = feat.shape[0], feat.shape[1]
        eng = self._engine
        lcnn_output = torch.empty(B, 2, device=feat.device, dtype=torch.float32)
        stream = torch.cuda.current_stream(feat.device).cuda_stream
        with torch.cuda.device(feat.device):
            _lib.check(eng.lib.occ_lcnn_backend(eng.handle, feat.data_ptr(), B, T, lcnn_output.data_ptr(), stream), eng.handle,
                       "occ_lcnn_backend")
        return senet34_output, lcnn_output
