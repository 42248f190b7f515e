"""GPU (-m gpu), collected last: ``occ_eer_noalloc`` -- the EER entry point that allocates nothing (its sort scratch is a
file-scope ``static __device__`` array inside libocc_b200.so, limit OCC_EER_NOALLOC_MAX_SCORES) -- against ``occ_eer``
(caller-provided workspace) through the C ABI.  Integer / float64 path: the bar is bit-exact, as in test_gpu_eer_kernels.py.
No counterpart in the reference."""
import numpy as np
import pytest
import torch

from occm_b200 import _lib

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
MAX_SCORES = 131072      # OCC_EER_NOALLOC_MAX_SCORES (include/occ_b200.h)


def _scores(seed, nt, nn, q=0):
    r = np.random.default_rng(seed)
    t, n = r.normal(0.4, 1, nt), r.normal(-0.4, 1, nn)
    if q:
        t, n = np.round(t * q) / q, np.round(n * q) / q          # heavy ties, -0.0 values
    return torch.from_numpy(t).to(DEV), torch.from_numpy(n).to(DEV)


def _eer_with_workspace(lib, t, n):
    work = torch.empty(int(lib.occ_eer_workspace_bytes(t.numel() + n.numel())), device=DEV, dtype=torch.uint8)
    res = torch.empty(2, device=DEV, dtype=torch.float64)
    stream = torch.cuda.current_stream(t.device).cuda_stream
    _lib.check(lib.occ_eer(t.data_ptr(), t.numel(), n.data_ptr(), n.numel(), res.data_ptr(), work.data_ptr(), stream), None, "occ_eer")
    return tuple(res.cpu().tolist())


def test_noalloc_eer_matches_occ_eer_on_consecutive_calls():
    """Two calls in a row (the second, smaller one runs over whatever the first left in the shared scratch), then the first
    input again; every result bit-identical to occ_eer's.  The device's free memory does not move across the calls."""
    lib = _lib.load()
    torch.cuda.set_device(0)
    cases = [_scores(1, 7355, 63882), _scores(2, 311, 4097, q=8), _scores(1, 7355, 63882), _scores(3, 1, 1), _scores(4, 65536, 65536, q=64)]
    want = [_eer_with_workspace(lib, t, n) for t, n in cases]
    res = torch.full((len(cases), 2), float("nan"), device=DEV, dtype=torch.float64)
    stream = torch.cuda.current_stream(torch.device(DEV)).cuda_stream
    # set-up call (first use on this device resolves the scratch address), then measure free memory around the rest
    _lib.check(lib.occ_eer_noalloc(cases[0][0].data_ptr(), cases[0][0].numel(), cases[0][1].data_ptr(), cases[0][1].numel(),
                                   res[0].data_ptr(), stream), None, "occ_eer_noalloc")
    torch.cuda.synchronize()
    free_before = torch.cuda.mem_get_info(0)[0]
    for i, (t, n) in enumerate(cases):
        _lib.check(lib.occ_eer_noalloc(t.data_ptr(), t.numel(), n.data_ptr(), n.numel(), res[i].data_ptr(), stream), None, "occ_eer_noalloc")
    torch.cuda.synchronize()
    free_after = torch.cuda.mem_get_info(0)[0]
    got = [tuple(r) for r in res.cpu().tolist()]
    assert got == want
    assert got[0] == got[2]
    assert free_after == free_before, (free_before, free_after)


def test_noalloc_eer_rejects_sizes_beyond_the_static_scratch():
    lib = _lib.load()
    t, n = _scores(5, 1000, MAX_SCORES - 999)                    # one score too many
    res = torch.zeros(2, device=DEV, dtype=torch.float64)
    rc = lib.occ_eer_noalloc(t.data_ptr(), t.numel(), n.data_ptr(), n.numel(), res.data_ptr(), None)
    assert rc == -1
    msg = _lib.last_error()
    assert "131072" in msg and "occ_eer_noalloc" in msg, msg
    with pytest.raises(ValueError):
        _lib.check(rc, None, "occ_eer_noalloc")
    assert lib.occ_eer_noalloc(None, 0, None, 0, None, None) == -1
    # exactly at the limit is accepted and still bit-identical
    t, n = _scores(6, 1000, MAX_SCORES - 1000, q=16)
    torch.cuda.synchronize()
    _lib.check(lib.occ_eer_noalloc(t.data_ptr(), t.numel(), n.data_ptr(), n.numel(), res.data_ptr(),
                                   torch.cuda.current_stream(torch.device(DEV)).cuda_stream), None, "occ_eer_noalloc")
    assert tuple(res.cpu().tolist()) == _eer_with_workspace(lib, t, n)
