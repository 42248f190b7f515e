"""ORACLE (test infrastructure only) -- CPU restatement of the reference's SE-ResNet34 backend on XLS-R features
(`/root/reference/models/senet.py`: `SELayer` :13-28, `SEBasicBlock` :31-61, `ResNet.forward` :121-143, `ssl_resnet34` :162-185),
eval mode (BatchNorm on running statistics).  Pinned against the UNMODIFIED file by `oracle/gen_golden_r3.py`
(`tests/golden/senet_golden.pt`).  Only tests / smoke may import this.
"""
from __future__ import annotations

from typing import Dict, Tuple

import torch
import torch.nn.functional as F

P = "resnet34."
LAYERS = ((16, 3, 1), (32, 4, 2), (64, 6, 2), (128, 3, 2))


def _bn(sd, name, x):
    return F.batch_norm(x, sd[P + name + "running_mean"], sd[P + name + "running_var"], sd[P + name + "weight"],
                        sd[P + name + "bias"], training=False, eps=1e-5)


def resnet34_backend(sd: Dict[str, torch.Tensor], feat: torch.Tensor) -> Tuple[torch.Tensor, torch.Tensor]:
    """feat [B,T,1024] -> (com [B,128], des [B,2])   (ssl_resnet34.forward :180-184, ResNet.forward :121-143)."""
    with torch.no_grad():
        x = feat.unsqueeze(1)
        x = F.relu(_bn(sd, "bn1.", F.conv2d(x, sd[P + "conv1.weight"], None, stride=2, padding=3)))
        x = F.max_pool2d(x, kernel_size=3, stride=2, padding=1)
        for li, (planes, blocks, stride) in enumerate(LAYERS, start=1):
            for b in range(blocks):
                p = f"layer{li}.{b}."
                st = stride if b == 0 else 1
                out = F.relu(_bn(sd, p + "bn1.", F.conv2d(x, sd[P + p + "conv1.weight"], None, stride=st, padding=1)))
                out = _bn(sd, p + "bn2.", F.conv2d(out, sd[P + p + "conv2.weight"], None, stride=1, padding=1))
                y = out.mean(dim=(2, 3))                                                   # SELayer: squeeze
                y = torch.sigmoid(F.linear(F.relu(F.linear(y, sd[P + p + "se.fc.0.weight"])), sd[P + p + "se.fc.2.weight"]))
                out = out * y[:, :, None, None]
                if (P + p + "downsample.0.weight") in sd:
                    x = _bn(sd, p + "downsample.1.", F.conv2d(x, sd[P + p + "downsample.0.weight"], None, stride=st))
                x = F.relu(out + x)
        v = x.mean(dim=(2, 3))
        return F.linear(v, sd[P + "embedding.weight"], sd[P + "embedding.bias"]), F.linear(v, sd[P + "classifier.weight"], sd[P + "classifier.bias"])
