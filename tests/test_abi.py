"""CPU: libocc_b200.so loads and exports every symbol include/occ_b200.h declares; error behaviour without a GPU."""
import ctypes as C
import os
import re

import pytest

from occm_b200 import _lib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def header_symbols():
    src = open(os.path.join(ROOT, "include", "occ_b200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(occ_[a-z0-9_]+)\s*\(", src)))


def test_every_declared_symbol_is_exported_and_bound():
    lib = _lib.load()
    syms = header_symbols()
    assert len(syms) >= 20
    for s in syms:
        assert hasattr(lib, s), f"{s} declared in include/occ_b200.h but not exported"
        assert s in _lib.PROTOTYPES, f"{s} has no ctypes prototype in occm_b200/_lib.py"
    for s in _lib.PROTOTYPES:
        assert s in syms, f"{s} bound in _lib.py but not declared in the header"


def test_version_and_frames():
    lib = _lib.load()
    assert lib.occ_version() >= 100
    for n, t in ((64600, 201), (16000, 49), (160000, 499), (480000, 1499), (399, 0)):
        assert lib.occ_num_frames(n) == t
    assert lib.occ_eer_workspace_bytes(71237) > 71237 * 24


def test_create_fails_loudly_without_gpu_or_with_bad_args():
    import torch
    lib = _lib.load()
    h = C.c_void_p()
    assert lib.occ_create(C.byref(h), 0, 7, 1, 64600) == -1          # bad dtype
    assert b"dtype" in lib.occ_last_error(None)
    if not torch.cuda.is_available():
        rc = lib.occ_create(C.byref(h), 0, 1, 1, 64600)
        assert rc == -2 and not h.value
        assert b"no CPU path" in lib.occ_last_error(None)


def test_null_arguments_are_errors_not_crashes():
    lib = _lib.load()
    assert lib.occ_forward(None, None, None, 1, 64600, None, None, None, None) == -1
    assert lib.occ_eer(None, 0, None, 0, None, None, None) == -1
    assert lib.occ_score_l2(None, None, 1, 160, 1e-6, None, None) == -1
    with pytest.raises(ValueError):
        _lib.check(-1, None, "x")
    with pytest.raises(_lib.OccError):
        _lib.check(-2, None, "x")
