"""ORACLE (test infrastructure only) -- CPU restatement of the reference's LCNN backend on XLS-R features
(`/root/reference/models/lcnn.py`: `mfm` :123-137, `group` :140-150, `LCNN.forward` :189-213, `ssl_lcnn.forward` :242-260),
eval mode (Dropout = identity, BatchNorm on running statistics).  `group.bn` is constructed but never called in the reference.

Pinned against the UNMODIFIED `models/lcnn.py` (imported in the build container by `oracle/gen_golden_r3.py`); the
outputs are committed in `tests/golden/lcnn_golden.pt`.  Only tests / smoke may import this.
"""
from __future__ import annotations

from typing import Dict

import torch
import torch.nn.functional as F

P = "lcnn."


def _mfm_conv(sd, name, x, padding):
    y = F.conv2d(x, sd[P + name + ".filter.weight"], sd[P + name + ".filter.bias"], padding=padding)     # lcnn.py:128, 134
    c = y.shape[1] // 2
    return torch.max(y[:, :c], y[:, c:])                                                                # :135-137


def _mfm_fc(sd, name, x):
    y = F.linear(x, sd[P + name + ".filter.0.weight"], sd[P + name + ".filter.0.bias"])                 # :130-132 (Dropout: eval)
    c = y.shape[1] // 2
    return torch.max(y[:, :c], y[:, c:])


def _bn(sd, name, x):
    return F.batch_norm(x, sd[P + name + ".running_mean"], sd[P + name + ".running_var"], sd[P + name + ".weight"],
                        sd[P + name + ".bias"], training=False, eps=1e-5)


def lcnn_backend(sd: Dict[str, torch.Tensor], feat: torch.Tensor) -> torch.Tensor:
    """feat [B,T,1024] (XLS-R output) -> logits [B,2]   (ssl_lcnn.forward :257-259, LCNN.forward :189-213)."""
    with torch.no_grad():
        x = feat.unsqueeze(1)                                              # [B,1,T,1024]
        x = F.max_pool2d(_mfm_conv(sd, "layer1.0", x, 2), 2, 2)            # layer1: mfm(1,4,5,1,2) + MaxPool2d(2,2)
        for n in (2, 3):
            x = _mfm_conv(sd, f"layer{n}.0.conv_a", x, 0)                  # group: 1x1 mfm, (bn unused), 3x3 mfm
            x = _mfm_conv(sd, f"layer{n}.0.conv", x, 1)
            x = _bn(sd, f"layer{n}.2", F.max_pool2d(x, 2, 2))
        x = F.adaptive_avg_pool2d(x, (1, 64)).reshape(x.shape[0], -1)       # [B, 16*64]
        x = _mfm_fc(sd, "fc0.0", x)
        x = _mfm_fc(sd, "fc1.0", x)
        x = _mfm_fc(sd, "fc2.0", x)
        return F.linear(x, sd[P + "fc3.weight"], sd[P + "fc3.bias"])
