"""ctypes binding of ``libocc_b200.so`` (C ABI: ``include/occ_b200.h``).

The shared library is the product; there is no Python/CPU fallback.  If it has not been
built (``make`` at the repo root or ``__graft_entry__.build()``), importing a symbol from
here raises ``OccLibraryMissing`` loudly.
"""
from __future__ import annotations

import ctypes as C
import os
from typing import Optional

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libocc_b200.so")

OCC_DTYPE_FP32 = 0
OCC_DTYPE_BF16 = 1
EMB_DIM = 160
FEAT_DIM = 1024


class OccLibraryMissing(RuntimeError):
    pass


class OccError(RuntimeError):
    pass


# name -> (restype, argtypes); every symbol declared in include/occ_b200.h
_vp, _i, _i64, _f = C.c_void_p, C.c_int, C.c_int64, C.c_float
PROTOTYPES = {
    "occ_version": (_i, []),
    "occ_create": (_i, [C.POINTER(_vp), _i, _i, _i, _i]),
    "occ_destroy": (None, [_vp]),
    "occ_last_error": (C.c_char_p, [_vp]),
    "occ_load_weights": (_i, [_vp, C.c_char_p, _vp, C.POINTER(_i64), _i, _i]),
    "occ_finalize_weights": (_i, [_vp]),
    "occ_forward": (_i, [_vp, _vp, _vp, _i, _i, _vp, _vp, _vp, _vp]),
    "occ_extract_feat": (_i, [_vp, _vp, _vp, _i, _i, _vp, _vp, _vp]),
    "occ_forward_resnet34": (_i, [_vp, _vp, _i, _i, _vp, _vp, _vp]),
    "occ_resnet34_backend": (_i, [_vp, _vp, _i, _i, _vp, _vp, _vp]),
    "occ_forward_lcnn": (_i, [_vp, _vp, _i, _i, _vp, _vp]),
    "occ_lcnn_backend": (_i, [_vp, _vp, _i, _i, _vp, _vp]),
    "occ_num_frames": (_i, [_i]),
    "occ_conv_encoder": (_i, [_vp, _vp, _i, _i, _vp, _vp]),
    "occ_pos_conv": (_i, [_vp, _vp, _i, _i, _vp, _vp]),
    "occ_transformer": (_i, [_vp, _vp, _i, _i, _i, _vp, _vp, _vp]),
    "occ_aasist_backend": (_i, [_vp, _vp, _i, _i, _vp, _vp, _vp]),
    "occ_score_l2": (_i, [_vp, _vp, _i, _i, _f, _vp, _vp]),
    "occ_score_cos": (_i, [_vp, _vp, _i, _i, _vp, _vp]),
    "occ_ref_embedding": (_i, [_vp, _i, _i, _f, _vp, _vp, _vp]),
    "occ_eer_workspace_bytes": (_i64, [_i64]),
    "occ_eer": (_i, [_vp, _i64, _vp, _i64, _vp, _vp, _vp]),
    "occ_eer_noalloc": (_i, [_vp, _i64, _vp, _i64, _vp, _vp]),
    "occ_det_curve": (_i, [_vp, _i64, _vp, _i64, _vp, _vp, _vp, _vp, _vp]),
    "occ_allgather_scores": (_i, [_vp, _i64, _vp, _vp, _vp]),
    "occ_ingest_pcm16": (_i, [_vp, _vp, _vp, _i, _i, _i, _vp, _vp]),
    "occ_confusion_counts": (_i, [_vp, _i64, _vp, _i64, C.c_double, _vp, _vp]),
    "occ_eer_host": (_i, [_vp, _i64, _vp, _i64, C.POINTER(C.c_double), C.POINTER(C.c_double), _i]),
    "occ_forward_host": (_i, [_vp, _vp, _vp, _i, _i, _vp, _f, _vp, _vp, _vp, _vp]),
    "occ_debug_gemm": (_i, [_i, _i, _i, _i, _i, _vp, _vp, _vp, _vp, _i, _vp, _vp]),
    "occ_debug_gemm_bench": (_i, [_i, _i, _i, _i, _i, _i, _i, _i, C.POINTER(C.c_double)]),
    "occ_debug_attention_trace": (_i, [_vp, _i]),
    "occ_debug_attention": (_i, [_i, _vp, _i, _i, _vp, _vp, _vp]),
    "occ_kernel_launches": (_i64, [_vp]),
    "occ_profile": (_i, [_vp, _i]),
    "occ_profile_read": (_i, [_vp, _i, C.POINTER(C.c_double), C.POINTER(_i64), C.POINTER(C.c_double)]),
}

_lib: Optional[C.CDLL] = None


def load() -> C.CDLL:
    """Load the library once and attach prototypes."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise OccLibraryMissing(
            f"{LIB_PATH} not found: build it with `make` (or `python -c 'import __graft_entry__ as g; g.build()'`). "
            "occm_b200 has no CPU / PyTorch fallback.")
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in PROTOTYPES.items():
        fn = getattr(lib, name)       # AttributeError here = header / library mismatch
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


def last_error(handle=None) -> str:
    msg = load().occ_last_error(handle)
    return msg.decode("utf-8", "replace") if msg else ""


def check(rc: int, handle=None, what: str = "") -> None:
    if rc == 0:
        return
    msg = last_error(handle)
    if rc == -1:
        raise ValueError(f"{what}: {msg}")
    raise OccError(f"{what} failed ({rc}): {msg}")
