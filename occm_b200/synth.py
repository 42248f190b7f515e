"""Synthetic weights and waveforms for the SSL-AASIST scoring path.

BASELINE.json's configs are all "random-init" on synthetic waveforms, so both
the product (bench.py, smoke) and the tests need the *same* reproducible
random state dict in the reference's checkpoint layout
(`/root/reference/oc_classifier.py:340` loads exactly this layout: fairseq
XLS-R keys under ``ssl_model.model.*`` plus the AASIST backend keys of
`/root/reference/models/sslassist.py:433-504`).

Every tensor is drawn from its own ``torch.Generator`` seeded by
``(seed, crc32(key))`` so the result does not depend on module construction
order, on the torch version's ``nn.Module`` init code, or on which subset of
keys is requested.  BatchNorm running statistics, affine parameters, ``pos_S``
and the master nodes get non-trivial values on purpose (default init would
hide BN-folding bugs).
"""
from __future__ import annotations

import math
import zlib
from collections import OrderedDict

import torch

# (out_channels, kernel, stride) of the wav2vec 2.0 feature encoder
# (fairseq `conv_feature_layers`, torchaudio/models/wav2vec2/model.py:1432-1479).
CONV_SPECS = [(512, 10, 5)] + [(512, 3, 2)] * 4 + [(512, 2, 2)] * 2
EMBED = 1024
FFN = 4096
HEADS = 16
LAYERS = 24
POS_K = 128
POS_G = 16
SSL_PREFIX = "ssl_model.model."


def num_frames(n_samples: int) -> int:
    """T = f(N): seven applications of L <- floor((L-k)/s)+1
    (torchaudio/models/wav2vec2/components.py:96)."""
    L = int(n_samples)
    for _, k, s in CONV_SPECS:
        L = (L - k) // s + 1
        if L < 0:
            L = 0
    return L


def _gen(seed: int, key: str) -> torch.Generator:
    g = torch.Generator(device="cpu")
    g.manual_seed((int(seed) * 1000003 + zlib.crc32(key.encode())) % (2**63 - 1))
    return g


def _normal(seed, key, shape, std=1.0, mean=0.0):
    return torch.randn(*shape, generator=_gen(seed, key), dtype=torch.float32) * std + mean


def _uniform(seed, key, shape, lo, hi):
    return torch.rand(*shape, generator=_gen(seed, key), dtype=torch.float32) * (hi - lo) + lo


def frontend_shapes() -> "OrderedDict[str, tuple]":
    """fairseq Wav2Vec2Model (XLS-R 300M) state-dict keys and shapes, un-prefixed.

    Key spelling: fairseq@a540213 `fairseq/models/wav2vec/wav2vec2.py`; the
    fairseq->torchaudio map in torchaudio/models/wav2vec2/utils/import_fairseq.py:51-115
    is the in-container statement of the same names."""
    s: "OrderedDict[str, tuple]" = OrderedDict()
    s["mask_emb"] = (EMBED,)
    cin = 1
    for i, (co, k, _) in enumerate(CONV_SPECS):
        s[f"feature_extractor.conv_layers.{i}.0.weight"] = (co, cin, k)
        s[f"feature_extractor.conv_layers.{i}.0.bias"] = (co,)
        s[f"feature_extractor.conv_layers.{i}.2.1.weight"] = (co,)
        s[f"feature_extractor.conv_layers.{i}.2.1.bias"] = (co,)
        cin = co
    s["post_extract_proj.weight"] = (EMBED, 512)
    s["post_extract_proj.bias"] = (EMBED,)
    # present in a real XLS-R checkpoint, unused with features_only=True
    s["quantizer.vars"] = (1, 640, 384)
    s["quantizer.weight_proj.weight"] = (640, 512)
    s["quantizer.weight_proj.bias"] = (640,)
    s["project_q.weight"] = (768, 768)
    s["project_q.bias"] = (768,)
    s["encoder.pos_conv.0.bias"] = (EMBED,)
    s["encoder.pos_conv.0.weight_g"] = (1, 1, POS_K)
    s["encoder.pos_conv.0.weight_v"] = (EMBED, EMBED // POS_G, POS_K)
    for i in range(LAYERS):
        p = f"encoder.layers.{i}."
        for n in ("k_proj", "v_proj", "q_proj", "out_proj"):
            s[p + f"self_attn.{n}.weight"] = (EMBED, EMBED)
            s[p + f"self_attn.{n}.bias"] = (EMBED,)
        s[p + "self_attn_layer_norm.weight"] = (EMBED,)
        s[p + "self_attn_layer_norm.bias"] = (EMBED,)
        s[p + "fc1.weight"] = (FFN, EMBED)
        s[p + "fc1.bias"] = (FFN,)
        s[p + "fc2.weight"] = (EMBED, FFN)
        s[p + "fc2.bias"] = (EMBED,)
        s[p + "final_layer_norm.weight"] = (EMBED,)
        s[p + "final_layer_norm.bias"] = (EMBED,)
    s["encoder.layer_norm.weight"] = (EMBED,)
    s["encoder.layer_norm.bias"] = (EMBED,)
    s["layer_norm.weight"] = (512,)
    s["layer_norm.bias"] = (512,)
    s["final_proj.weight"] = (768, EMBED)
    s["final_proj.bias"] = (768,)
    return s


FRONTEND_UNUSED = ("mask_emb", "quantizer.", "project_q.", "final_proj.")

# AASIST hyper-parameters, hard-coded at /root/reference/models/sslassist.py:438-441
RES_FILTS = [(1, 32), (32, 32), (32, 64), (64, 64), (64, 64), (64, 64)]
GAT_DIMS = (64, 32)
N_SPEC = 42  # pos_S rows, sslassist.py:473
TEMPERATURES = (2.0, 2.0, 100.0, 100.0)


def _bn(s, p, c):
    s[p + "weight"] = (c,)
    s[p + "bias"] = (c,)
    s[p + "running_mean"] = (c,)
    s[p + "running_var"] = (c,)
    s[p + "num_batches_tracked"] = ()


def backend_shapes() -> "OrderedDict[str, tuple]":
    """AASIST backend keys in `AModel.state_dict()` order
    (/root/reference/models/sslassist.py:446-504)."""
    s: "OrderedDict[str, tuple]" = OrderedDict()
    s["pos_S"] = (1, N_SPEC, 64)
    s["master1"] = (1, 1, GAT_DIMS[0])
    s["master2"] = (1, 1, GAT_DIMS[0])
    s["LL.weight"] = (128, EMBED)
    s["LL.bias"] = (128,)
    _bn(s, "first_bn.", 1)
    _bn(s, "first_bn1.", 64)
    for n, (ci, co) in enumerate(RES_FILTS):
        p = f"encoder.{n}.0."
        if n > 0:
            _bn(s, p + "bn1.", ci)
        s[p + "conv1.weight"] = (co, ci, 2, 3)
        s[p + "conv1.bias"] = (co,)
        _bn(s, p + "bn2.", co)
        s[p + "conv2.weight"] = (co, co, 2, 3)
        s[p + "conv2.bias"] = (co,)
        if ci != co:
            s[p + "conv_downsample.weight"] = (co, ci, 1, 3)
            s[p + "conv_downsample.bias"] = (co,)
    s["attention.0.weight"] = (128, 64, 1, 1)
    s["attention.0.bias"] = (128,)
    _bn(s, "attention.2.", 128)
    s["attention.3.weight"] = (64, 128, 1, 1)
    s["attention.3.bias"] = (64,)
    for name in ("GAT_layer_S", "GAT_layer_T"):
        p = name + "."
        s[p + "att_weight"] = (64, 1)
        for lin in ("att_proj", "proj_with_att", "proj_without_att"):
            s[p + lin + ".weight"] = (64, 64)
            s[p + lin + ".bias"] = (64,)
        _bn(s, p + "bn.", 64)
    for name, din in (("HtrgGAT_layer_ST11", 64), ("HtrgGAT_layer_ST12", 32),
                      ("HtrgGAT_layer_ST21", 64), ("HtrgGAT_layer_ST22", 32)):
        p = name + "."
        dout = GAT_DIMS[1]
        for w in ("att_weight11", "att_weight22", "att_weight12", "att_weightM"):
            s[p + w] = (dout, 1)
        for lin in ("proj_type1", "proj_type2"):
            s[p + lin + ".weight"] = (din, din)
            s[p + lin + ".bias"] = (din,)
        for lin in ("att_proj", "att_projM", "proj_with_att", "proj_without_att",
                    "proj_with_attM", "proj_without_attM"):
            s[p + lin + ".weight"] = (dout, din)
            s[p + lin + ".bias"] = (dout,)
        _bn(s, p + "bn.", dout)
    for name, d in (("pool_S", 64), ("pool_T", 64), ("pool_hS1", 32), ("pool_hT1", 32),
                    ("pool_hS2", 32), ("pool_hT2", 32)):
        s[name + ".proj.weight"] = (1, d)
        s[name + ".proj.bias"] = (1,)
    s["out_layer.weight"] = (2, 5 * GAT_DIMS[1])
    s["out_layer.bias"] = (2,)
    return s


LCNN_PREFIX = "lcnn."


def lcnn_shapes() -> "OrderedDict[str, tuple]":
    """`LCNN(asoftmax=False)` keys in `state_dict()` order (/root/reference/models/lcnn.py:151-187; c_s = [128,64,32,16,8,4,2]),
    as they sit under ``lcnn.`` in ``ssl_lcnn`` (`:236-240`).  ``layerN.0.bn`` exists in the reference but is never called."""
    s: "OrderedDict[str, tuple]" = OrderedDict()
    s["layer1.0.filter.weight"] = (8, 1, 5, 5)
    s["layer1.0.filter.bias"] = (8,)
    for n, (ci, co) in ((2, (4, 8)), (3, (8, 16))):
        p = f"layer{n}.0."
        s[p + "conv_a.filter.weight"] = (2 * ci, ci, 1, 1)
        s[p + "conv_a.filter.bias"] = (2 * ci,)
        _bn(s, p + "bn.", ci)
        s[p + "conv.filter.weight"] = (2 * co, ci, 3, 3)
        s[p + "conv.filter.bias"] = (2 * co,)
        _bn(s, f"layer{n}.2.", co)
    s["fc0.0.filter.0.weight"] = (64, 1024)
    s["fc0.0.filter.0.bias"] = (64,)
    s["fc1.0.filter.0.weight"] = (64, 32)
    s["fc1.0.filter.0.bias"] = (64,)
    s["fc2.0.filter.0.weight"] = (16, 32)
    s["fc2.0.filter.0.bias"] = (16,)
    s["fc3.weight"] = (2, 8)
    s["fc3.bias"] = (2,)
    return s


def make_lcnn_state_dict(seed: int = 0) -> "OrderedDict[str, torch.Tensor]":
    """Seeded LCNN backend checkpoint under ``lcnn.*`` (non-trivial BatchNorm statistics and affine)."""
    sd: "OrderedDict[str, torch.Tensor]" = OrderedDict()
    for k, shp in lcnn_shapes().items():
        full = LCNN_PREFIX + k
        leaf = k.rsplit(".", 1)[-1]
        if (k.startswith(("layer2.2.", "layer3.2.")) or ".bn." in k) and leaf == "weight":
            sd[full] = _uniform(seed, full, shp, 0.5, 1.5)
        elif (k.startswith(("layer2.2.", "layer3.2.")) or ".bn." in k) and leaf == "bias":
            sd[full] = _normal(seed, full, shp, 0.1)
        else:
            sd[full] = _draw(seed, full, shp)
    return sd


RESNET_PREFIX = "resnet34."
RESNET_LAYERS = ((16, 3, 1), (32, 4, 2), (64, 6, 2), (128, 3, 2))          # (planes, blocks, stride of the first block)


def resnet34_shapes() -> "OrderedDict[str, tuple]":
    """`se_resnet34()` keys in `state_dict()` order (/root/reference/models/senet.py:64-156: channels [16,16,32,64,128],
    layers [3,4,6,3], SE reduction 16), as they sit under ``resnet34.`` in ``ssl_resnet34`` (`:162-166`)."""
    s: "OrderedDict[str, tuple]" = OrderedDict()
    s["conv1.weight"] = (16, 1, 7, 7)
    _bn(s, "bn1.", 16)
    inpl = 16
    for li, (planes, blocks, stride) in enumerate(RESNET_LAYERS, start=1):
        for b in range(blocks):
            p = f"layer{li}.{b}."
            cin = inpl if b == 0 else planes
            s[p + "conv1.weight"] = (planes, cin, 3, 3)
            _bn(s, p + "bn1.", planes)
            s[p + "conv2.weight"] = (planes, planes, 3, 3)
            _bn(s, p + "bn2.", planes)
            s[p + "se.fc.0.weight"] = (planes // 16, planes)
            s[p + "se.fc.2.weight"] = (planes, planes // 16)
            if b == 0 and (stride != 1 or inpl != planes):
                s[p + "downsample.0.weight"] = (planes, inpl, 1, 1)
                _bn(s, p + "downsample.1.", planes)
        inpl = planes
    s["embedding.weight"] = (128, 128)
    s["embedding.bias"] = (128,)
    s["classifier.weight"] = (2, 128)
    s["classifier.bias"] = (2,)
    return s


def make_resnet34_state_dict(seed: int = 0) -> "OrderedDict[str, torch.Tensor]":
    """Seeded SE-ResNet34 backend checkpoint under ``resnet34.*`` (kaiming-scale conv weights, non-trivial BatchNorm)."""
    sd: "OrderedDict[str, torch.Tensor]" = OrderedDict()
    for k, shp in resnet34_shapes().items():
        full = RESNET_PREFIX + k
        leaf = k.rsplit(".", 1)[-1]
        is_bn = ".bn" in "." + k or "downsample.1." in k
        if leaf == "num_batches_tracked":
            sd[full] = torch.tensor(100, dtype=torch.int64)
        elif is_bn and leaf == "weight":
            sd[full] = _uniform(seed, full, shp, 0.5, 1.5)
        elif is_bn and leaf == "bias":
            sd[full] = _normal(seed, full, shp, 0.1)
        elif is_bn and leaf == "running_mean":
            sd[full] = _normal(seed, full, shp, 0.1)
        elif is_bn and leaf == "running_var":
            sd[full] = _uniform(seed, full, shp, 0.5, 1.5)
        elif len(shp) == 4:                                   # conv: kaiming_normal_(fan_out) as senet.py:88-90
            sd[full] = _normal(seed, full, shp, math.sqrt(2.0 / (shp[0] * shp[2] * shp[3])))
        else:
            sd[full] = _draw(seed, full, shp)
    return sd


def _draw(seed: int, key: str, shape: tuple) -> torch.Tensor:
    """One tensor of the synthetic checkpoint; distribution chosen by key suffix."""
    leaf = key.rsplit(".", 1)[-1]
    if leaf == "num_batches_tracked":
        return torch.tensor(100, dtype=torch.int64)
    if leaf == "running_mean":
        return _normal(seed, key, shape, 0.1)
    if leaf == "running_var":
        return _uniform(seed, key, shape, 0.5, 1.5)
    if key.endswith(("bn.weight", "bn1.weight", "bn2.weight", "first_bn.weight",
                     "first_bn1.weight", "attention.2.weight")):
        return _uniform(seed, key, shape, 0.5, 1.5)
    if key.endswith(("bn.bias", "bn1.bias", "bn2.bias", "first_bn.bias",
                     "first_bn1.bias", "attention.2.bias")):
        return _normal(seed, key, shape, 0.1)
    if "layer_norm" in key or ".2.1." in key:  # LayerNorm affine
        return _normal(seed, key, shape, 0.1, 1.0) if leaf == "weight" else _normal(seed, key, shape, 0.05)
    if key.endswith(("pos_S", "master1", "master2")):
        return _normal(seed, key, shape, 1.0)
    if "att_weight" in leaf:  # xavier_normal_ on (out_dim, 1)
        return _normal(seed, key, shape, math.sqrt(2.0 / (shape[0] + 1)))
    if leaf == "weight_v":
        return _normal(seed, key, shape, math.sqrt(4.0 / (POS_K * EMBED)))
    if leaf == "weight_g":
        raise RuntimeError("weight_g is derived from weight_v")
    if key.startswith(SSL_PREFIX):
        if "feature_extractor" in key and leaf == "weight":
            fan_in = shape[1] * shape[2]
            return _normal(seed, key, shape, math.sqrt(2.0 / fan_in))
        if leaf == "bias":
            return _normal(seed, key, shape, 0.02)
        if len(shape) >= 2 and leaf == "weight":
            return _normal(seed, key, shape, 0.02)
        return _uniform(seed, key, shape, 0.0, 1.0)  # mask_emb, quantizer.vars
    # backend Linear / Conv2d: torch default init bound 1/sqrt(fan_in)
    if leaf == "weight":
        fan_in = 1
        for d in shape[1:]:
            fan_in *= d
        b = 1.0 / math.sqrt(fan_in)
        return _uniform(seed, key, shape, -b, b)
    if leaf == "bias":
        return _uniform(seed, key, shape, -0.1, 0.1)
    raise KeyError(key)


def make_state_dict(seed: int = 0, frontend: bool = True, backend: bool = True,
                    include_unused: bool = True) -> "OrderedDict[str, torch.Tensor]":
    """Reference-layout random checkpoint (fp32, CPU)."""
    sd: "OrderedDict[str, torch.Tensor]" = OrderedDict()
    if frontend:
        for k, shp in frontend_shapes().items():
            if not include_unused and k.startswith(FRONTEND_UNUSED):
                continue
            full = SSL_PREFIX + k
            if k.endswith("weight_g"):
                continue
            sd[full] = _draw(seed, full, shp)
            if k.endswith("weight_v"):
                v = sd[full]
                # weight_norm(dim=2): g[k] scales v[:, :, k] / ||v[:, :, k]||_F
                # (torchaudio/models/wav2vec2/components.py:211)
                g = v.pow(2).sum(dim=(0, 1), keepdim=True).sqrt()
                g = g * _uniform(seed, full + ".gscale", (1, 1, POS_K), 0.8, 1.2)
                sd[full[: -len("weight_v")] + "weight_g"] = g
    if backend:
        for k, shp in backend_shapes().items():
            sd[k] = _draw(seed, k, shp)
    return sd


def make_waveforms(seed: int, batch: int, n_samples: int, amplitude: float = 0.1) -> torch.Tensor:
    """`amplitude * randn(batch, n_samples)` fp32 on CPU."""
    g = torch.Generator(device="cpu")
    g.manual_seed(int(seed))
    return torch.randn(batch, n_samples, generator=g, dtype=torch.float32) * amplitude


# ------------------------------------------------------------------------------------------------ BASELINE.json configs[2]
CONFIG3_UTTERANCES = 71237          # ASVspoof2019-LA-eval size
CONFIG3_BONAFIDE = 7355             # bona fide utterances come first (protocol order), the other 63 882 are spoof


def config3_n_bonafide(n_utterances: int) -> int:
    return max(1, round(n_utterances * CONFIG3_BONAFIDE / CONFIG3_UTTERANCES))


def config3_batch(lo: int, hi: int, n_bonafide: int, device, n_samples: int = 64600) -> torch.Tensor:
    """Utterances ``lo .. hi-1`` of the config-3 synthetic evaluation set, generated on ``device``.

    Utterance ``i`` depends only on the counter-based seed ``1000 + i``, so any sharding / batching sees the same
    data and the 18.4 GB set never has to exist at once.  Bona fide (``i < n_bonafide``): white noise, sigma 0.1.  Spoof: the
    same noise plus a tone ``a sin(2 pi f t)`` with ``a ~ U(0.03, 0.40)``, ``f ~ U(300, 3000) Hz`` drawn from the utterance's own
    generator.  An amplitude change (0.10 vs 0.12, round 1) vanishes in the per-frame LayerNorm of the conv encoder and gave a
    chance-level EER; a spectral line survives it, and the weak-tone utterances overlap the bona fide scores, so the set's
    fp32-mode EER lands between 5 and 20 % -- informative about whether reduced precision preserves the ranking."""
    dev = torch.device(device)
    gen = torch.Generator(device=dev)
    n = hi - lo
    w = torch.empty(n, n_samples, device=dev, dtype=torch.float32)
    u = torch.empty(n, 2, device=dev, dtype=torch.float32)
    for j, i in enumerate(range(lo, hi)):
        gen.manual_seed(1000 + i)
        w[j].normal_(0.0, 0.10, generator=gen)
        u[j].uniform_(0.0, 1.0, generator=gen)
    spoof = (torch.arange(lo, hi, device=dev) >= n_bonafide).to(torch.float32)
    amp = (0.03 + 0.37 * u[:, 0]) * spoof
    freq = 300.0 + 2700.0 * u[:, 1]
    t = torch.arange(n_samples, device=dev, dtype=torch.float32) / 16000.0
    w += amp[:, None] * torch.sin((2.0 * math.pi) * freq[:, None] * t[None, :])
    return w


def config3_train_batch(device, n: int = 32, n_samples: int = 64600) -> torch.Tensor:
    """The separate bona fide "train" batch whose mean embedding is the one-class reference (seed 2)."""
    dev = torch.device(device)
    gen = torch.Generator(device=dev)
    gen.manual_seed(2)
    return torch.empty(n, n_samples, device=dev, dtype=torch.float32).normal_(0.0, 0.10, generator=gen)


# ------------------------------------------------------------------------------------------------ BASELINE.json configs[3]
def config4_lengths(n_utterances: int, seed: int = 3, lo: int = 16000, hi: int = 160000) -> list:
    """``N_i ~ U{16 000 .. 160 000}`` samples (1-10 s), seed 3."""
    g = torch.Generator(device="cpu")
    g.manual_seed(int(seed))
    return torch.randint(lo, hi + 1, (n_utterances,), generator=g).tolist()
