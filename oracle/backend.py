"""ORACLE (test infrastructure only) -- CPU fp32 restatement of the AASIST
backend half of `AModel.forward` (`/root/reference/models/sslassist.py:509-597`),
written as pure functions over a reference-layout state dict, eval mode only
(dropout = identity, BatchNorm = running statistics).

Pinned by `oracle/gen_golden.py`, which imports the *unmodified* reference file
in this container (fairseq shimmed by torchaudio's XLS-R) and checks this
restatement against it on the same weights/inputs; the resulting vectors are
committed under `tests/golden/` and re-checked by `tests/test_oracle.py`.
The reference's own tests pin nothing on this path.

Only tests / smoke / bench's cpu_baseline leg may import this.
"""
from __future__ import annotations

from typing import Dict, Optional, Tuple

import torch
import torch.nn.functional as F

SD = Dict[str, torch.Tensor]


def _bn(sd: SD, p: str, x: torch.Tensor, ch_dim: int = 1) -> torch.Tensor:
    """eval-mode BatchNorm: (x - mean) / sqrt(var + 1e-5) * w + b, via F.batch_norm."""
    if ch_dim != 1:
        x = x.transpose(1, ch_dim)
    y = F.batch_norm(x, sd[p + "running_mean"], sd[p + "running_var"], sd[p + "weight"], sd[p + "bias"],
                     False, 0.1, 1e-5)
    if ch_dim != 1:
        y = y.transpose(1, ch_dim)
    return y


def _lin(sd: SD, p: str, x: torch.Tensor) -> torch.Tensor:
    return F.linear(x, sd[p + ".weight"], sd[p + ".bias"])


def front_pool(sd: SD, feat: torch.Tensor) -> torch.Tensor:
    """sslassist.py:509-516: LL 1024->128, [B,T,128] -> [B,1,128,T], max_pool2d(3,3),
    first_bn (eval), SELU -> [B,1,42,T//3]."""
    x = _lin(sd, "LL", feat)
    x = x.transpose(1, 2).unsqueeze(1)
    x = F.max_pool2d(x, (3, 3))
    x = _bn(sd, "first_bn.", x)
    return F.selu(x)


def residual_block(sd: SD, p: str, x: torch.Tensor) -> torch.Tensor:
    """sslassist.py:406-429.  NOTE the observable quirk at :408-415: bn1+selu are
    computed and then discarded -- conv1 is applied to the raw block input."""
    out = F.conv2d(x, sd[p + "conv1.weight"], sd[p + "conv1.bias"], padding=(1, 1))
    out = F.selu(_bn(sd, p + "bn2.", out))
    out = F.conv2d(out, sd[p + "conv2.weight"], sd[p + "conv2.bias"], padding=(0, 1))
    if (p + "conv_downsample.weight") in sd:
        x = F.conv2d(x, sd[p + "conv_downsample.weight"], sd[p + "conv_downsample.bias"], padding=(0, 1))
    return out + x


def res_encoder(sd: SD, x: torch.Tensor) -> torch.Tensor:
    """sslassist.py:519-521: six residual blocks then first_bn1 + SELU."""
    for n in range(6):
        x = residual_block(sd, f"encoder.{n}.0.", x)
    return F.selu(_bn(sd, "first_bn1.", x))


def spectro_temporal_pool(sd: SD, x: torch.Tensor) -> Tuple[torch.Tensor, torch.Tensor]:
    """sslassist.py:523-538: 1x1-conv attention map w; softmax over time -> e_S (+pos_S);
    softmax over frequency -> e_T."""
    w = F.conv2d(x, sd["attention.0.weight"], sd["attention.0.bias"])
    w = F.selu(w)
    w = _bn(sd, "attention.2.", w)
    w = F.conv2d(w, sd["attention.3.weight"], sd["attention.3.bias"])
    w1 = F.softmax(w, dim=-1)
    e_S = torch.sum(x * w1, dim=-1).transpose(1, 2) + sd["pos_S"]
    w2 = F.softmax(w, dim=-2)
    e_T = torch.sum(x * w2, dim=-2).transpose(1, 2)
    return e_S, e_T


def gat_layer(sd: SD, p: str, x: torch.Tensor, temp: float) -> torch.Tensor:
    """GraphAttentionLayer, sslassist.py:84-146.  x [B,n,D]."""
    pair = x.unsqueeze(2) * x.unsqueeze(1)                      # :97-108  [B,n,n,D]
    att = torch.tanh(_lin(sd, p + "att_proj", pair))            # :117
    att = torch.matmul(att, sd[p + "att_weight"]) / temp        # :119-122  [B,n,n,1]
    att = F.softmax(att, dim=-2)                                # :124
    y = _lin(sd, p + "proj_with_att", torch.matmul(att.squeeze(-1), x)) + _lin(sd, p + "proj_without_att", x)
    y = _bn(sd, p + "bn.", y, ch_dim=2)                         # :135-141 (BatchNorm1d over the feature dim)
    return F.selu(y)


def graph_pool(sd: SD, p: str, h: torch.Tensor, ratio: float = 0.5) -> torch.Tensor:
    """GraphPool, sslassist.py:341-368: sigmoid(Linear(D->1)); keep top int(n*ratio) nodes in
    descending-score order (torch.topk sorted=True); output rows are h*score."""
    s = torch.sigmoid(_lin(sd, p + ".proj", h))
    k = max(int(h.shape[1] * ratio), 1)
    _, idx = torch.topk(s, k, dim=1)
    return torch.gather(h * s, 1, idx.expand(-1, -1, h.shape[2]))


def htrg_gat_layer(sd: SD, p: str, x1: torch.Tensor, x2: torch.Tensor, master: Optional[torch.Tensor],
                   temp: float) -> Tuple[torch.Tensor, torch.Tensor, torch.Tensor]:
    """HtrgGraphAttentionLayer, sslassist.py:191-329."""
    n1, n2 = x1.shape[1], x2.shape[1]
    x = torch.cat([_lin(sd, p + "proj_type1", x1), _lin(sd, p + "proj_type2", x2)], dim=1)   # :201-206
    if master is None:
        master = x.mean(dim=1, keepdim=True)                                                   # :209-210
    pair = x.unsqueeze(2) * x.unsqueeze(1)
    att = torch.tanh(_lin(sd, p + "att_proj", pair))                                           # :278
    board = torch.zeros_like(att[..., :1])
    board[:, :n1, :n1] = torch.matmul(att[:, :n1, :n1], sd[p + "att_weight11"])                # :284-285
    board[:, n1:, n1:] = torch.matmul(att[:, n1:, n1:], sd[p + "att_weight22"])                # :286-287
    board[:, :n1, n1:] = torch.matmul(att[:, :n1, n1:], sd[p + "att_weight12"])                # :288-289
    board[:, n1:, :n1] = torch.matmul(att[:, n1:, :n1], sd[p + "att_weight12"])                # :290-291 (12 reused)
    att_map = F.softmax(board / temp, dim=-2)                                                  # :298-300
    # master update, :234-239 / :255-270 / :310-316
    am = torch.tanh(_lin(sd, p + "att_projM", x * master))
    am = F.softmax(torch.matmul(am, sd[p + "att_weightM"]) / temp, dim=-2)
    new_master = _lin(sd, p + "proj_with_attM", torch.matmul(am.squeeze(-1).unsqueeze(1), x)) \
        + _lin(sd, p + "proj_without_attM", master)
    y = _lin(sd, p + "proj_with_att", torch.matmul(att_map.squeeze(-1), x)) + _lin(sd, p + "proj_without_att", x)
    y = F.selu(_bn(sd, p + "bn.", y, ch_dim=2))
    return y[:, :n1], y[:, n1:n1 + n2], new_master


def graph_head(sd: SD, e_S: torch.Tensor, e_T: torch.Tensor) -> Tuple[torch.Tensor, torch.Tensor]:
    """sslassist.py:541-597: GAT_S/GAT_T + pools, two HS-GAL branches, element-wise
    branch max, readout -> emb [B,160], out [B,2]."""
    out_S = graph_pool(sd, "pool_S", gat_layer(sd, "GAT_layer_S.", e_S, 2.0))
    out_T = graph_pool(sd, "pool_T", gat_layer(sd, "GAT_layer_T.", e_T, 2.0))
    branches = []
    for b in ("1", "2"):
        T1, S1, m = htrg_gat_layer(sd, f"HtrgGAT_layer_ST{b}1.", out_T, out_S, sd["master" + b], 100.0)
        S1 = graph_pool(sd, "pool_hS" + b, S1)
        T1 = graph_pool(sd, "pool_hT" + b, T1)
        Ta, Sa, ma = htrg_gat_layer(sd, f"HtrgGAT_layer_ST{b}2.", T1, S1, m, 100.0)
        branches.append((T1 + Ta, S1 + Sa, m + ma))
    oT = torch.max(branches[0][0], branches[1][0])       # :580 element-wise, order-sensitive
    oS = torch.max(branches[0][1], branches[1][1])
    mm = torch.max(branches[0][2], branches[1][2])
    T_max, _ = torch.max(torch.abs(oT), dim=1)
    S_max, _ = torch.max(torch.abs(oS), dim=1)
    emb = torch.cat([T_max, oT.mean(dim=1), S_max, oS.mean(dim=1), mm.squeeze(1)], dim=1)
    out = _lin(sd, "out_layer", emb)
    return emb, out


def aasist_backend(sd: SD, feat: torch.Tensor, stages: Optional[dict] = None):
    """feat [B,T,1024] (frontend output) -> (emb [B,160], out [B,2])."""
    x0 = front_pool(sd, feat)
    x = res_encoder(sd, x0)
    e_S, e_T = spectro_temporal_pool(sd, x)
    emb, out = graph_head(sd, e_S, e_T)
    if stages is not None:
        stages.update(pool=x0, enc=x, e_S=e_S, e_T=e_T)
    return emb, out
