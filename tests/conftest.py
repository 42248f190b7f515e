import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (B200); run with -m gpu")


def pytest_collection_modifyitems(config, items):
    try:
        import torch
        has_gpu = torch.cuda.is_available()
    except Exception:
        has_gpu = False
    if has_gpu:
        return
    skip = pytest.mark.skip(reason="no CUDA device")
    for it in items:
        if "gpu" in it.keywords:
            it.add_marker(skip)


@pytest.fixture(scope="session")
def golden():
    import torch
    return torch.load(os.path.join(GOLDEN, "model_golden.pt"), weights_only=False)


@pytest.fixture(scope="session")
def eer_kats():
    import json
    import numpy as np
    out = []
    for k in json.load(open(os.path.join(GOLDEN, "eer_kat.json"))):
        if "target" in k:
            t, n = np.array(k["target"], dtype=np.float64), np.array(k["nontarget"], dtype=np.float64)
        elif k["name"] == "KAT3":
            rng = np.random.default_rng(1234)
            t = rng.normal(1, 1, 7355).astype(np.float32).astype(np.float64)
            n = rng.normal(-1, 1, 63882).astype(np.float32).astype(np.float64)
        else:
            r = np.random.default_rng(k["seed"])
            q = k["q"]
            t = np.round(r.normal(0.3, 1, k["nt"]) * q) / q
            n = np.round(r.normal(-0.3, 1, k["nn"]) * q) / q
        out.append((k, t, n))
    return out


@pytest.fixture(scope="session")
def synth_sd():
    from occm_b200 import synth
    return synth.make_state_dict(0)
