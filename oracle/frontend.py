"""ORACLE (test infrastructure only) -- CPU fp32 restatement of the XLS-R /
wav2vec 2.0 frontend that `/root/reference/models/sslassist.py:48` calls as
``fairseq_model(x, mask=False, features_only=True)['x']``.

The arithmetic lives in an un-vendored dependency: fairseq, pinned at
``git+https://github.com/facebookresearch/fairseq.git@a54021305d6b3c4c5959ac9395135f63202db8f1``
(`/root/reference/requirement.txt:33`).  fairseq is not installable here (no
network), so this file restates the published wav2vec 2.0 "large / layer_norm"
forward from the only in-container statement of it, torchaudio 2.11
``torchaudio/models/wav2vec2/components.py`` (line numbers cited per function),
driven directly by a *fairseq-named* state dict.  Pinning: `oracle/gen_golden.py`
checks this file bit-for-bit against ``torchaudio.models.wav2vec2_xlsr_300m()``
loaded with the same weights, and `tests/test_oracle.py` repeats that check on
every run (torchaudio is on both machines).  The reference's own tests pin
nothing at this boundary: **parity unpinned by the reference**.

Only `tests/`, `__graft_entry__.smoke()` and `bench.py`'s cpu_baseline /
`--impl reference` legs may import this package.  Nothing under `occm_b200/`
does.
"""
from __future__ import annotations

from typing import Dict, List, Optional

import torch
import torch.nn.functional as F

CONV_SPECS = [(10, 5)] + [(3, 2)] * 4 + [(2, 2)] * 2
P = "ssl_model.model."
HEADS = 16


def frame_lengths(lengths: torch.Tensor) -> torch.Tensor:
    """components.py:95-98 -- valid frames after each conv block."""
    L = lengths.clone()
    for k, s in CONV_SPECS:
        L = torch.div(L - k, s, rounding_mode="floor") + 1
        L = torch.clamp(L, min=0)
    return L


def conv_feature_encoder(sd: Dict[str, torch.Tensor], wav: torch.Tensor, p: str = P) -> torch.Tensor:
    """components.py:53-143 (ConvLayerBlock x7, layer_norm mode, conv bias):
    Conv1d -> LayerNorm over the 512 channels of each frame -> erf-GELU.
    wav [B, N] -> [B, T, 512]."""
    h = wav.unsqueeze(1)
    for i, (_, s) in enumerate(CONV_SPECS):
        q = f"{p}feature_extractor.conv_layers.{i}."
        h = F.conv1d(h, sd[q + "0.weight"], sd[q + "0.bias"], stride=s)
        h = h.transpose(-2, -1)
        h = F.layer_norm(h, (h.shape[-1],), sd[q + "2.1.weight"], sd[q + "2.1.bias"], 1e-5)
        h = h.transpose(-2, -1)
        h = F.gelu(h)
    return h.transpose(1, 2)


def feature_projection(sd, feats: torch.Tensor, p: str = P) -> torch.Tensor:
    """components.py:146-182: LayerNorm(512) -> Linear 512->1024 (dropout = identity in eval)."""
    x = F.layer_norm(feats, (feats.shape[-1],), sd[p + "layer_norm.weight"], sd[p + "layer_norm.bias"], 1e-5)
    return F.linear(x, sd[p + "post_extract_proj.weight"], sd[p + "post_extract_proj.bias"])


def pos_conv_weight(sd, p: str = P) -> torch.Tensor:
    """weight_norm(dim=2) (components.py:211): W[:,:,k] = g[k] * v[:,:,k] / ||v[:,:,k]||_F.
    Uses torch._weight_norm, the op the parametrization itself calls."""
    return torch._weight_norm(sd[p + "encoder.pos_conv.0.weight_v"], sd[p + "encoder.pos_conv.0.weight_g"], 2)


def pos_conv(sd, x: torch.Tensor, p: str = P) -> torch.Tensor:
    """components.py:185-234: grouped Conv1d(1024,1024,k=128,pad=64,groups=16),
    drop the last frame (even kernel), erf-GELU.  x [B,T,1024] -> same."""
    w = pos_conv_weight(sd, p)
    y = F.conv1d(x.transpose(-2, -1), w, sd[p + "encoder.pos_conv.0.bias"], padding=64, groups=16)
    y = y[..., :-1]
    y = F.gelu(y)
    return y.transpose(-2, -1)


def encoder_layer(sd, x: torch.Tensor, i: int, mask: Optional[torch.Tensor], p: str = P) -> torch.Tensor:
    """components.py:345-401 with layer_norm_first=True; attention :237-310; FFN :313-342."""
    q = f"{p}encoder.layers.{i}."
    B, T, E = x.shape
    res = x
    h = F.layer_norm(x, (E,), sd[q + "self_attn_layer_norm.weight"], sd[q + "self_attn_layer_norm.bias"], 1e-5)
    shape = (B, T, HEADS, E // HEADS)
    qq = F.linear(h, sd[q + "self_attn.q_proj.weight"], sd[q + "self_attn.q_proj.bias"]).view(*shape).transpose(2, 1)
    kk = F.linear(h, sd[q + "self_attn.k_proj.weight"], sd[q + "self_attn.k_proj.bias"]).view(*shape).transpose(2, 1)
    vv = F.linear(h, sd[q + "self_attn.v_proj.weight"], sd[q + "self_attn.v_proj.bias"]).view(*shape).transpose(2, 1)
    a = F.scaled_dot_product_attention(qq, kk, vv, attn_mask=mask, dropout_p=0.0, is_causal=False)
    a = a.transpose(1, 2).reshape(B, -1, E)
    a = F.linear(a, sd[q + "self_attn.out_proj.weight"], sd[q + "self_attn.out_proj.bias"])
    x = res + a
    h = F.layer_norm(x, (E,), sd[q + "final_layer_norm.weight"], sd[q + "final_layer_norm.bias"], 1e-5)
    h = F.linear(h, sd[q + "fc1.weight"], sd[q + "fc1.bias"])
    h = F.gelu(h)
    h = F.linear(h, sd[q + "fc2.weight"], sd[q + "fc2.bias"])
    return x + h


def xlsr_forward(sd: Dict[str, torch.Tensor], wav: torch.Tensor, lengths: Optional[torch.Tensor] = None,
                 all_layers: bool = False, p: str = P, n_layers: int = 24):
    """Wav2Vec2Model.forward / extract_features (torchaudio model.py:73-130,
    components.py:476-510, Transformer :404-463 with the inverted flag at :756:
    no LN before the layers, final LN after layer 24).

    Returns the final-LN'd [B,T,1024] (what fairseq returns as ['x']); with
    ``all_layers`` also the 24 pre-final-LN layer outputs (fairseq layer_results /
    torchaudio extract_features)."""
    feats = conv_feature_encoder(sd, wav, p)
    x = feature_projection(sd, feats, p)
    mask = None
    if lengths is not None:
        fl = frame_lengths(lengths)
        B, T, _ = x.shape
        pad = torch.arange(T).expand(B, T) >= fl[:, None]
        x = x.clone()
        x[pad] = 0.0
        mask = (-10000.0 * pad[:, None, None, :].to(x.dtype)).expand(B, 1, T, T)
    x = x + pos_conv(sd, x, p)
    outs: List[torch.Tensor] = []
    for i in range(n_layers):
        x = encoder_layer(sd, x, i, mask, p)
        if all_layers:
            outs.append(x)
    y = F.layer_norm(x, (x.shape[-1],), sd[p + "encoder.layer_norm.weight"], sd[p + "encoder.layer_norm.bias"], 1e-5)
    return (y, outs) if all_layers else y


def to_torchaudio_state_dict(sd: Dict[str, torch.Tensor], p: str = P) -> Dict[str, torch.Tensor]:
    """fairseq names -> torchaudio names, following
    torchaudio/models/wav2vec2/utils/import_fairseq.py:51-115; pos-conv g/v go to the
    torch>=2.1 parametrization spelling."""
    from torchaudio.models.wav2vec2.utils.import_fairseq import _map_key

    out = {}
    for k, v in sd.items():
        if not k.startswith(p):
            continue
        m = _map_key(k[len(p):])
        if m is None:
            continue
        if m.endswith("pos_conv_embed.conv.weight_g"):
            m = m[: -len("weight_g")] + "parametrizations.weight.original0"
        elif m.endswith("pos_conv_embed.conv.weight_v"):
            m = m[: -len("weight_v")] + "parametrizations.weight.original1"
        out[m] = v
    return out
