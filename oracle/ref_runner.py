"""Runs the UNMODIFIED reference model from `oracle/_ref` (see `oracle/build_ref.py`).  Test / benchmark infrastructure only.

`load_reference_amodel(device)` does what `oc_classifier.py:335-340` does -- `AModel(None, device).to(device)`,
`load_state_dict`, `.eval()` -- with the seeded synthetic checkpoint of `occm_b200.synth` (BASELINE.json: random-init).
"""
from __future__ import annotations

import importlib
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF_DIR = os.path.join(ROOT, "oracle", "_ref")
ARCHIVE = os.path.join(REF_DIR, "reference_py.zip")      # the unmodified reference sources, packed by oracle/build_ref.py


def available() -> bool:
    return os.path.exists(ARCHIVE) and os.path.exists(os.path.join(REF_DIR, "fairseq", "__init__.py"))


def _paths():
    if not available():
        raise RuntimeError("oracle/_ref is missing: run `python -m oracle.build_ref` where /root/reference exists")
    for p in (REF_DIR, ARCHIVE):                             # REF_DIR: the fairseq stand-in; ARCHIVE: zipimport of the reference
        if p not in sys.path:
            sys.path.insert(0, p)
    for name in ("fairseq", "fairseq.checkpoint_utils", "models", "models.sslassist", "evaluate_metrics"):
        mod = sys.modules.get(name)
        if mod is not None and not str(getattr(mod, "__file__", "")).startswith(REF_DIR):
            del sys.modules[name]


def _import_reference():
    _paths()
    return importlib.import_module("models.sslassist")


def load_reference_amodel(device="cpu", seed: int = 0):
    from occm_b200 import synth
    from oracle.frontend import to_torchaudio_state_dict
    ref = _import_reference()
    model = ref.AModel(None, device)
    sd = synth.make_state_dict(seed)
    model.ssl_model.model.m.load_state_dict(to_torchaudio_state_dict(sd), strict=True)
    be = {k: v for k, v in sd.items() if not k.startswith("ssl_model.")}
    res = model.load_state_dict(be, strict=False)
    assert not res.unexpected_keys and all(k.startswith("ssl_model.") for k in res.missing_keys)
    return model.to(device).eval()


def reference_compute_eer():
    _paths()
    return importlib.import_module("evaluate_metrics").compute_eer
