"""GPU (-m gpu): EER bit-exactness, one-class / cosine scores, GEMM and attention kernels through the C ABI."""
import ctypes as C

import numpy as np
import pytest
import torch

from occm_b200 import _lib
from oracle import scoring as osc

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def test_eer_kats_bit_exact(eer_kats):
    from occm_b200.evaluate_metrics import compute_eer
    lib = _lib.load()
    for k, t, n in eer_kats:
        assert compute_eer(t, n) == (k["eer"], k["thr"]), k["name"]
        e, th = C.c_double(), C.c_double()
        assert lib.occ_eer_host(t.ctypes.data, t.size, n.ctypes.data, n.size, C.byref(e), C.byref(th), 0) == 0
        assert (e.value, th.value) == (k["eer"], k["thr"])
        if "eer_swapped" in k:
            assert compute_eer(n, t) == (k["eer_swapped"], k["thr_swapped"])


@pytest.mark.parametrize("seed", range(6))
def test_eer_random_vs_oracle_bit_exact(seed):
    """Random sizes, heavy ties (quantised), negative zero, duplicates across classes, tensors on the GPU."""
    from occm_b200.evaluate_metrics import compute_eer
    r = np.random.default_rng(100 + seed)
    nt, nn = int(r.integers(1, 9000)), int(r.integers(1, 70000))
    q = [0, 4, 64, 0, 1, 1024][seed]
    t, n = r.normal(0.2, 1, nt), r.normal(-0.2, 1, nn)
    if q:
        t, n = np.round(t * q) / q, np.round(n * q) / q       # produces -0.0 values too
    want = osc.compute_eer(t, n)
    assert compute_eer(t, n) == want
    assert compute_eer(torch.from_numpy(t).to(DEV), torch.from_numpy(n).to(DEV)) == want
    assert compute_eer(list(t), list(n)) == want


def test_eer_edge_cases():
    from occm_b200.evaluate_metrics import compute_eer
    for t, n in (([0.0], [0.0]), ([1.0], [-1.0]), ([-1.0], [1.0]), ([0.0, -0.0], [-0.0, 0.0]), ([1e300, -1e300], [5e-324, -5e-324])):
        assert compute_eer(t, n) == osc.compute_eer(np.array(t), np.array(n)), (t, n)
    with pytest.raises(ValueError):
        compute_eer([], [1.0])
    # fp32 scores are promoted exactly like np.array(list of python floats) (evaluate.py:138-142)
    r = np.random.default_rng(3)
    t32, n32 = r.normal(1, 1, 500).astype(np.float32), r.normal(-1, 1, 900).astype(np.float32)
    assert compute_eer(torch.from_numpy(t32).to(DEV), torch.from_numpy(n32).to(DEV)) == \
        osc.compute_eer(t32.astype(np.float64), n32.astype(np.float64))


def test_eer_full_size_properties():
    """ASVspoof2019-LA-eval size: argument swap mirrors the curve; shifting all scores shifts only the threshold."""
    from occm_b200.evaluate_metrics import compute_eer
    r = np.random.default_rng(5)
    t = r.normal(1, 1, 63882).astype(np.float32).astype(np.float64)
    n = r.normal(-1, 1, 7355).astype(np.float32).astype(np.float64)
    e, th = compute_eer(t, n)
    assert (e, th) == osc.compute_eer(t, n)
    e2, th2 = compute_eer(t + 8.0, n + 8.0)
    assert e2 == e and abs((th2 - 8.0) - th) < 1e-12


def test_one_class_and_cosine_scores():
    from occm_b200.scoring import one_class_scores, reference_from_embeddings
    from occm_b200.losses.custom_loss import OCSoftmax, cosine_scores
    g = torch.Generator().manual_seed(0)
    emb = torch.randn(37, 160, generator=g)
    ref = emb.mean(0, keepdim=True) + 0.01 * torch.randn(1, 160, generator=g)
    got = one_class_scores(ref.to(DEV), emb.to(DEV)).cpu()
    want = osc.one_class_scores(ref, emb)
    assert torch.allclose(got, want, rtol=2e-6, atol=0)
    z = torch.zeros(1, 160)
    assert abs(float(one_class_scores(z.to(DEV), z.to(DEV))) - float(osc.one_class_scores(z, z))) < 1e-11
    r2, thr = reference_from_embeddings(emb.to(DEV))
    o_ref, o_thr = osc.reference_embedding([e[None] for e in emb])
    assert torch.allclose(r2.cpu(), o_ref, rtol=1e-6, atol=1e-7) and abs(float(thr) - float(o_thr)) < 1e-5
    w = torch.randn(160, generator=g)
    assert torch.allclose(cosine_scores(emb.to(DEV), w.to(DEV)).cpu(), osc.ocsoftmax_scores(emb, w), atol=2e-6)
    oc = OCSoftmax(160).to(DEV)
    labels = torch.tensor([0, 1] * 18 + [0])
    loss, s = oc(emb.to(DEV), labels.to(DEV))
    want_loss = osc.ocsoftmax_loss(osc.ocsoftmax_scores(emb, oc.center.detach().cpu()), labels)
    assert abs(float(loss) - float(want_loss)) < 1e-4


GEMM_CASES = [(128, 64, 64, 0, 0, 0), (300, 256, 128, 0, 0, 0), (402, 1024, 1024, 0, 0, 1), (402, 3072, 1024, 0, 0, 0),
              (402, 4096, 1024, 0, 1, 0), (402, 1024, 4096, 0, 0, 1), (402, 128, 1024, 1, 0, 0), (1000, 1024, 512, 1, 1, 1),
              (12864, 1024, 1024, 0, 0, 1), (1, 64, 64, 0, 0, 0), (129, 192, 64, 1, 0, 0)]


@pytest.mark.parametrize("impl", [0, 1])
@pytest.mark.parametrize("case", GEMM_CASES)
def test_gemm_tcgen05_vs_torch(impl, case):
    """impl 1 = tcgen05/TMEM/TMA kernel, impl 0 = SIMT cross-check.  Plain mode: exact bf16 products, fp32 accumulate
    -> 1e-5; split mode (hi/lo bf16, 3 MMAs) is fp32-accurate -> 2e-5 vs float64."""
    M, N, K, split, gelu, use_res = case
    lib = _lib.load()
    g = torch.Generator(device=DEV).manual_seed(M * 7 + N)
    A = torch.randn(M, K, device=DEV, generator=g)
    W = torch.randn(N, K, device=DEV, generator=g) * 0.05
    bias = torch.randn(N, device=DEV, generator=g)
    res = torch.randn(M, N, device=DEV, generator=g) if use_res else None
    out = torch.full((M, N), float("nan"), device=DEV)
    rc = lib.occ_debug_gemm(impl, M, N, K, split, A.data_ptr(), W.data_ptr(), bias.data_ptr(),
                            res.data_ptr() if use_res else None, gelu, out.data_ptr(), None)
    assert rc == 0, _lib.last_error()
    ref = (A.double() @ W.double().t()) if split else (A.bfloat16().double() @ W.bfloat16().double().t())
    ref = ref + bias.double()
    if gelu:
        ref = torch.nn.functional.gelu(ref)
    if use_res:
        ref = ref + res.double()
    err = float((out.double() - ref).abs().max() / ref.abs().max())
    assert err < 2e-5, err


@pytest.mark.parametrize("M,N,K,split", [(12864, 1024, 1024, 0), (12864, 1024, 4096, 0), (1000, 1024, 1024, 1), (77, 256, 64, 0)])
def test_gemm_residual_in_place(M, N, K, split):
    """The out-proj / FC2 form of the transformer: x <- x + (A W^T + b) with the residual read (prefetched one chunk ahead)
    and written through the same buffer, several tiles per CTA pair."""
    lib = _lib.load()
    g = torch.Generator(device=DEV).manual_seed(M + N + K)
    A = torch.randn(M, K, device=DEV, generator=g)
    W = torch.randn(N, K, device=DEV, generator=g) * 0.05
    bias = torch.randn(N, device=DEV, generator=g)
    x = torch.randn(M, N, device=DEV, generator=g)
    ref = (A.double() @ W.double().t()) if split else (A.bfloat16().double() @ W.bfloat16().double().t())
    ref = ref + bias.double() + x.double()
    rc = lib.occ_debug_gemm(1, M, N, K, split, A.data_ptr(), W.data_ptr(), bias.data_ptr(), x.data_ptr(), 0, x.data_ptr(), None)
    assert rc == 0, _lib.last_error()
    err = float((x.double() - ref).abs().max() / ref.abs().max())
    assert err < 2e-5, err


@pytest.mark.parametrize("B,T,masked", [(2, 49, False), (2, 201, False), (3, 201, True), (1, 499, False), (2, 64, True),
                                        (2, 256, True), (2, 130, False), (2, 17, True)])
def test_attention_kernels_vs_torch(B, T, masked):
    """impl 0 = fp32 SIMT (fp32 mode, tol 2e-5), impl 1 = bf16 attention (tcgen05 kernel for T <= 256, mma.sync flash kernel beyond; bf16 mode, tol 1e-2:
    bf16 rounding of q,k,v and of the probabilities)."""
    lib = _lib.load()
    g = torch.Generator(device=DEV).manual_seed(T)
    qkv = torch.randn(B * T, 3072, device=DEV, generator=g)
    qkv[:, :1024] *= 0.125
    klen = torch.tensor([max(T - 7 * i, 1) for i in range(B)], dtype=torch.int32, device=DEV) if masked else None
    q, k, v = (qkv[:, i * 1024:(i + 1) * 1024].reshape(B, T, 16, 64).transpose(1, 2).double() for i in range(3))
    s = q @ k.transpose(-1, -2)
    valid = torch.ones(B * T, dtype=torch.bool, device=DEV)
    if masked:
        m = torch.arange(T, device=DEV)[None, :] >= klen[:, None]
        s = s.masked_fill(m[:, None, None, :], float("-inf"))
        valid = (~m).reshape(-1)
    ref = (torch.softmax(s, -1) @ v).transpose(1, 2).reshape(B * T, 1024)
    for impl, tol in ((0, 2e-5), (1, 1e-2)):
        ctx = torch.full((B * T, 1024), float("nan"), device=DEV)
        rc = lib.occ_debug_attention(impl, qkv.data_ptr(), B, T, klen.data_ptr() if masked else None, ctx.data_ptr(), None)
        assert rc == 0, _lib.last_error()
        err = float((ctx[valid].double() - ref[valid]).abs().max() / ref[valid].abs().max())
        assert err < tol, (impl, err)


@pytest.mark.gpu
@pytest.mark.parametrize("case", ["ragged", "clean"])
def test_evaluate_cli_stdout_equals_reference(case, tmp_path, capsys):
    """`python -m occm_b200.evaluate` prints exactly what the unmodified /root/reference/evaluate.py printed on the same files
    (tests/golden/evalproto_kat.json, made by oracle/gen_ingest_golden.py): label counts, the sklearn-ordered confusion
    matrix (3x3 when protocol entries have no metadata), TP/TN/FP/FN as the reference indexes them, EER and threshold to
    the last digit."""
    import json
    import os
    from occm_b200 import evaluate
    kat = json.load(open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "evalproto_kat.json")))["cli"][case]
    pf, mf, sf = tmp_path / "p.txt", tmp_path / "m.txt", tmp_path / "s.txt"
    pf.write_text("\n".join(kat["proto"]) + "\n")
    mf.write_text("\n".join(kat["meta"]) + "\n")
    sf.write_text("\n".join(kat["score_lines"]) + "\n")
    evaluate.main(["--score_file", str(sf), "--protocol_file", str(pf), "--metadata_file", str(mf),
                   "--threshold", repr(kat["threshold"])])
    assert capsys.readouterr().out == kat["stdout"]
