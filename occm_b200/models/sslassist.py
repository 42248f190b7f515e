"""Drop-in surface of the reference's ``models/sslassist.py`` on B200.

``AModel(args, device)`` (``/root/reference/models/sslassist.py:432-597``; ``Model`` is the alias
BASELINE.json names) and ``SSLModel(device)`` (``:20-49``) keep the reference's constructor,
``forward`` / ``extract_feat`` signatures, return values and ``state_dict()`` key layout
(fairseq XLS-R keys under ``ssl_model.model.*`` + the AASIST backend keys).  The modules hold only
parameters; all arithmetic runs in ``libocc_b200.so`` (hand-written sm_100a kernels) through the
C ABI of ``include/occ_b200.h``.  There is no PyTorch / CPU compute path: a non-CUDA input raises
``ValueError`` and a missing library raises ``OccLibraryMissing``.

Eval-mode only (the scoring path of ``oc_classifier.py:243-312``): dropout is identity, BatchNorm
uses running statistics, exactly what ``model.eval()`` gives in the reference.
"""
from __future__ import annotations

import ctypes as C
import os
from typing import Dict, Optional, Tuple

import torch
import torch.nn as nn

from .. import _lib, synth

_BUFFER_LEAVES = ("running_mean", "running_var", "num_batches_tracked")


class _Tree(nn.Module):
    """Parameter container whose ``state_dict()`` keys are the dotted names it was built from."""

    def __init__(self):
        super().__init__()

    def _insert(self, key: str, value: torch.Tensor) -> None:
        head, _, rest = key.partition(".")
        if rest:
            if head not in self._modules:
                self.add_module(head, _Tree())
            self._modules[head]._insert(rest, value)
        elif head in _BUFFER_LEAVES:
            self.register_buffer(head, value)
        else:
            self.register_parameter(head, nn.Parameter(value, requires_grad=False))

    @staticmethod
    def build(sd: Dict[str, torch.Tensor]) -> "_Tree":
        t = _Tree()
        for k, v in sd.items():
            t._insert(k, v)
        return t


def _dtype_code(dtype) -> int:
    if dtype in (None, "bf16", "bfloat16", torch.bfloat16, _lib.OCC_DTYPE_BF16):
        return _lib.OCC_DTYPE_BF16
    if dtype in ("fp32", "float32", torch.float32):
        return _lib.OCC_DTYPE_FP32
    raise ValueError(f"occm_b200: unknown compute dtype {dtype!r} (use 'bf16' or 'fp32')")


class _Engine:
    """One ``occ_handle``: packed weights + activation workspace on one GPU."""

    def __init__(self, state_dict: Dict[str, torch.Tensor], device: torch.device, dtype: int,
                 max_batch: int, max_samples: int):
        self.lib = _lib.load()
        self.device = device
        self.dtype = dtype
        self.max_batch = max_batch
        self.max_samples = max_samples
        self.handle = C.c_void_p()
        idx = device.index if device.index is not None else torch.cuda.current_device()
        _lib.check(self.lib.occ_create(C.byref(self.handle), idx, dtype, max_batch, max_samples), None, "occ_create")
        # occ_load_weights copies with a blocking cudaMemcpy on the legacy stream and occ_finalize_weights packs on stream 0: the
        # tensors staged by the caller (``v.to(device)`` on torch's current, possibly non-blocking, stream) must be complete first.
        torch.cuda.current_stream(device).synchronize()
        try:
            for k, v in state_dict.items():
                if v.dtype == torch.int64:
                    continue   # num_batches_tracked
                t = v.detach().to(torch.float32).contiguous()
                shape = (C.c_int64 * max(t.dim(), 1))(*t.shape)
                _lib.check(self.lib.occ_load_weights(self.handle, k.encode(), t.data_ptr(), shape, t.dim(), 0),
                           self.handle, f"occ_load_weights({k})")
            _lib.check(self.lib.occ_finalize_weights(self.handle), self.handle, "occ_finalize_weights")
        except Exception:
            self.close()
            raise

    def close(self):
        if self.handle:
            self.lib.occ_destroy(self.handle)
            self.handle = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    @property
    def launches(self) -> int:
        return int(self.lib.occ_kernel_launches(self.handle))

    PROF_TAGS = ("xf_gemm", "conv_gemm", "pos_gemm", "attn", "norm", "conv0", "backend", "proj_gemm",
                 "xf_qkv", "xf_out", "xf_fc1", "xf_fc2")          # xf_gemm = the sum of the last four

    def profile(self, enable: bool) -> None:
        _lib.check(self.lib.occ_profile(self.handle, 1 if enable else 0), self.handle, "occ_profile")

    def profile_read(self) -> Dict[str, dict]:
        """{family: {ms, launches, flops}} summed over the ranges recorded since profile(True)."""
        out = {}
        for tag, name in enumerate(self.PROF_TAGS):
            ms, n, fl = C.c_double(), C.c_int64(), C.c_double()
            _lib.check(self.lib.occ_profile_read(self.handle, tag, C.byref(ms), C.byref(n), C.byref(fl)), self.handle,
                       "occ_profile_read")
            out[name] = dict(ms=ms.value, launches=n.value, flops=fl.value)
        return out


def _as_wave(x: torch.Tensor) -> torch.Tensor:
    if not isinstance(x, torch.Tensor):
        raise ValueError("occm_b200: input must be a torch.Tensor")
    if x.ndim == 3:
        x = x[:, :, 0]           # sslassist.py:42-43
    if x.ndim != 2:
        raise ValueError(f"occm_b200: expected [batch, samples] (or [batch, samples, 1]), got {tuple(x.shape)}")
    if not x.is_cuda:
        raise ValueError("occm_b200: input must be a CUDA tensor (this build has no CPU path)")
    return x.to(torch.float32).contiguous()


def _lengths_arg(lengths, B: int, N: int):
    if lengths is None:
        return None, None
    ls = [int(v) for v in (lengths.tolist() if isinstance(lengths, torch.Tensor) else lengths)]
    if len(ls) != B:
        raise ValueError("occm_b200: lengths must have one entry per utterance")
    arr = (C.c_int32 * B)(*ls)
    return arr, ls


class _EngineOwner(nn.Module):
    """Shared lazy-engine logic of ``AModel`` and the stand-alone ``SSLModel``."""

    def _init_engine_state(self, dtype):
        self._occ_dtype = _dtype_code(dtype if dtype is not None else os.environ.get("OCC_DTYPE"))
        self._engine: Optional[_Engine] = None
        self._weights_epoch = 0          # bumped whenever the parameters may have changed
        self._engine_epoch = -1
        # The packed copy inside the engine is rebuilt after ``load_state_dict`` (any sub-module: the frontend shares its owner's
        # engine) and after ``.to()`` / ``.cuda()`` / ``.float()`` -- the ways the reference's callers change weights
        # (oc_classifier.py:335-340).  A hook instead of walking ~500 parameter versions on every forward.  In-place edits of
        # ``param.data`` are not seen: call ``invalidate_engine()`` after them.
        self.register_load_state_dict_post_hook(lambda module, incompatible: module.invalidate_engine())

    def invalidate_engine(self) -> None:
        self._weights_epoch += 1
        owner = self.__dict__.get("_owner")
        if owner is not None:
            owner.invalidate_engine()

    def _apply(self, fn, *args, **kwargs):
        out = super()._apply(fn, *args, **kwargs)
        if "_weights_epoch" in self.__dict__:
            self.invalidate_engine()
        return out

    def _engine_state_dict(self) -> Dict[str, torch.Tensor]:
        raise NotImplementedError

    def _get_engine(self, B: int, N: int, device: torch.device) -> _Engine:
        e = self._engine
        if e is not None and (e.device != device or e.max_batch < B or e.max_samples < N or self._engine_epoch != self._weights_epoch):
            e.close()
            e = self._engine = None
        if e is None:
            mb = max(B, int(os.environ.get("OCC_MAX_BATCH", "8")))
            ms = max(N, int(os.environ.get("OCC_MAX_SAMPLES", "64600")))
            sd = {k: v.to(device) for k, v in self._engine_state_dict().items()}
            e = self._engine = _Engine(sd, device, self._occ_dtype, mb, ms)
            self._engine_epoch = self._weights_epoch
        return e

    def reserve(self, max_batch: int, max_samples: int, device=None) -> None:
        """Pre-build the engine (weight packing + workspace) for the given maximum shapes."""
        dev = torch.device(device) if device is not None else next(self.parameters()).device
        if dev.type != "cuda":
            raise ValueError("occm_b200: reserve() needs a CUDA device")
        if dev.index is None:
            dev = torch.device("cuda", torch.cuda.current_device())
        self._get_engine(max_batch, max_samples, dev)

    @property
    def kernel_launches(self) -> int:
        return self._engine.launches if self._engine is not None else 0

    def train(self, mode: bool = True):
        if mode:
            raise RuntimeError("occm_b200 accelerates the eval-mode scoring path only "
                               "(oc_classifier.py:243-312); training is out of scope")
        return super().train(False)


class SSLModel(_EngineOwner):
    """``SSLModel(device)`` of ``models/sslassist.py:20-49`` / ``models/xlsr.py:25-52``.

    ``self.model`` holds the fairseq XLS-R-300M parameters (random-init from
    ``synth.make_state_dict(seed)`` -- BASELINE.json's configs are random-init; the reference's
    hard-coded checkpoint path is never touched); load real weights with ``load_state_dict``.
    """

    def __init__(self, device, seed: int = 0, dtype=None, _owner: Optional[_EngineOwner] = None):
        super().__init__()
        sd = synth.make_state_dict(seed, frontend=True, backend=False)
        self.model = _Tree.build({k[len(synth.SSL_PREFIX):]: v for k, v in sd.items()})
        self.device = device
        self.out_dim = 1024
        self._init_engine_state(dtype)
        self.__dict__["_owner"] = _owner          # AModel shares its engine with its frontend
        self.training = False
        for m in self.modules():
            m.training = False
        if _owner is None and device is not None and torch.device(device).type == "cuda":
            self.to(device)

    def _engine_state_dict(self):
        return {"ssl_model." + k: v for k, v in self.state_dict().items()}

    def extract_feat(self, input_data: torch.Tensor, lengths=None, all_layers: bool = False):
        """[B,N] / [B,N,1] CUDA fp32 -> [B,T,1024] (fairseq ``['x']``: final LayerNorm applied).
        Extensions: ``lengths`` (valid samples per utterance) and ``all_layers`` (also return the
        24 per-layer outputs [24,B,T,1024], pre final LN = fairseq ``layer_results``)."""
        x = _as_wave(input_data)
        B, N = x.shape
        owner = self.__dict__.get("_owner") or self
        eng = owner._get_engine(B, N, x.device)
        T = eng.lib.occ_num_frames(N)
        feat = torch.empty(B, T, 1024, device=x.device, dtype=torch.float32)
        hidden = torch.empty(24, B, T, 1024, device=x.device, dtype=torch.float32) if all_layers else None
        larr, _ = _lengths_arg(lengths, B, N)
        stream = torch.cuda.current_stream(x.device).cuda_stream
        with torch.cuda.device(x.device):
            _lib.check(eng.lib.occ_extract_feat(eng.handle, x.data_ptr(), larr, B, N, feat.data_ptr(),
                                                hidden.data_ptr() if hidden is not None else None, stream),
                       eng.handle, "occ_extract_feat")
        return (feat, hidden) if all_layers else feat

    def forward(self, input_data, lengths=None):
        return self.extract_feat(input_data, lengths)


class AModel(_EngineOwner):
    """``AModel(args, device)`` of ``models/sslassist.py:432-597``: XLS-R frontend + AASIST backend.

    ``forward(x)`` returns ``(emb [B,160], output [B,2])`` like the reference.  ``args`` may be
    ``None`` (as at ``oc_classifier.py:335``) or carry ``seed`` / ``occ_dtype`` attributes.
    Compute dtype: ``'bf16'`` (default; bf16 tensor-core operands, fp32 accumulate/LN/softmax/backend)
    or ``'fp32'`` (split-bf16 3-MMA GEMMs, fp32-accurate); also selectable with ``OCC_DTYPE``.
    """

    def __init__(self, args, device, dtype=None, seed: Optional[int] = None):
        super().__init__()
        if seed is None:
            seed = int(getattr(args, "seed", 0) or 0) if args is not None else 0
        if dtype is None and args is not None:
            dtype = getattr(args, "occ_dtype", None)
        self.device = device
        self.ssl_model = SSLModel(device, seed=seed, dtype=dtype, _owner=self)
        be = synth.make_state_dict(seed, frontend=False, backend=True)
        for k, v in be.items():
            head, _, rest = k.partition(".")
            if not rest:
                self.register_parameter(head, nn.Parameter(v, requires_grad=False))
            else:
                if head not in self._modules:
                    self.add_module(head, _Tree())
                self._modules[head]._insert(rest, v)
        self._init_engine_state(dtype)
        for m in self.modules():
            m.training = False

    def _engine_state_dict(self):
        return self.state_dict()

    def forward(self, x: torch.Tensor, lengths=None) -> Tuple[torch.Tensor, torch.Tensor]:
        if x.ndim == 3 and x.shape[-1] == 1:
            x = x.squeeze(-1)                      # sslassist.py:508
        x = _as_wave(x)
        B, N = x.shape
        eng = self._get_engine(B, N, x.device)
        emb = torch.empty(B, 160, device=x.device, dtype=torch.float32)
        out = torch.empty(B, 2, device=x.device, dtype=torch.float32)
        larr, _ = _lengths_arg(lengths, B, N)
        stream = torch.cuda.current_stream(x.device).cuda_stream
        with torch.cuda.device(x.device):
            _lib.check(eng.lib.occ_forward(eng.handle, x.data_ptr(), larr, B, N, emb.data_ptr(), out.data_ptr(),
                                           None, stream), eng.handle, "occ_forward")
        return emb, out

    def forward_host(self, wav_host: torch.Tensor, ref_host: Optional[torch.Tensor] = None, eps: float = 1e-6,
                     lengths=None, device=None):
        """End-to-end with HOST buffers (the loop body of ``score_eval_set_1c2``, oc_classifier.py:256-261):
        H2D of the waveforms, forward, one-class distance to ``ref_host`` and D2H of (emb, out, score)."""
        if wav_host.is_cuda or wav_host.dtype != torch.float32 or wav_host.ndim != 2 or not wav_host.is_contiguous():
            raise ValueError("occm_b200: forward_host expects a contiguous CPU fp32 [B,N] tensor")
        B, N = wav_host.shape
        dev = torch.device(device) if device is not None else next(self.parameters()).device
        if dev.index is None:
            dev = torch.device("cuda", torch.cuda.current_device())
        eng = self._get_engine(B, N, dev)
        emb = torch.empty(B, 160, dtype=torch.float32)
        out = torch.empty(B, 2, dtype=torch.float32)
        score = torch.empty(B, dtype=torch.float32) if ref_host is not None else None
        refp = None
        if ref_host is not None:
            ref_c = ref_host.detach().to("cpu", torch.float32).reshape(-1).contiguous()
            if ref_c.numel() != 160:
                raise ValueError("occm_b200: reference embedding must have 160 elements")
            refp = ref_c.data_ptr()
        larr, _ = _lengths_arg(lengths, B, N)
        stream = torch.cuda.current_stream(dev).cuda_stream
        with torch.cuda.device(dev):
            _lib.check(eng.lib.occ_forward_host(eng.handle, wav_host.data_ptr(), larr, B, N, refp, eps, emb.data_ptr(),
                                                out.data_ptr(), score.data_ptr() if score is not None else None, stream),
                       eng.handle, "occ_forward_host")
        return emb, out, score


Model = AModel   # the name BASELINE.json's north_star uses
