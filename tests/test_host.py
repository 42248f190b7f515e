"""CPU: host-side logic -- module surface / state_dict layout, score-file formats, sharding and the
world_size-2 gather over gloo."""
import json
import os

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from occm_b200 import parallel, scoring, synth
GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


@pytest.fixture(scope="module")
def cpu_model():
    from occm_b200.models.sslassist import AModel
    return AModel(None, "cpu").eval()


def test_state_dict_layout_matches_reference(cpu_model):
    want = json.load(open(os.path.join(GOLDEN, "state_dict_keys.json")))
    got = {k: list(v.shape) for k, v in cpu_model.state_dict().items()}
    exp = dict(want["backend"])
    exp.update(want["frontend"])
    assert set(got) == set(exp), (sorted(set(got) ^ set(exp))[:10])
    for k, s in exp.items():
        assert got[k] == s, (k, got[k], s)


def test_model_surface_and_no_cpu_path(cpu_model):
    from occm_b200.models import AModel, Model, SSLModel
    from occm_b200.models import xlsr
    assert Model is AModel and xlsr.SSLModel is SSLModel
    assert cpu_model.ssl_model.out_dim == 1024
    with pytest.raises(ValueError, match="CUDA"):
        cpu_model(torch.zeros(1, 64600))
    with pytest.raises(ValueError):
        cpu_model.ssl_model.extract_feat(torch.zeros(1, 64600))
    with pytest.raises(RuntimeError, match="eval"):
        cpu_model.train()
    assert cpu_model.eval() is cpu_model and not cpu_model.training


def test_load_state_dict_roundtrip(cpu_model):
    sd = synth.make_state_dict(3, frontend=False)
    res = cpu_model.load_state_dict(sd, strict=False)
    assert not res.unexpected_keys
    assert torch.equal(cpu_model.state_dict()["LL.weight"], sd["LL.weight"])
    assert torch.equal(cpu_model.state_dict()["encoder.3.0.bn2.running_var"], sd["encoder.3.0.bn2.running_var"])


def test_score_file_formats(tmp_path, golden):
    lines = scoring.format_score_lines_1c(golden["c1_dist"].tolist(), float(golden["c1_thr"]))
    assert lines == golden["c1_lines"]                       # produced by the oracle of oc_classifier.py:262-265
    p = tmp_path / "scores.txt"
    p.write_text("".join(lines) + "0.5\n")
    got = scoring.load_score_file(str(p))
    assert got[:-1] == [float(np.float32(v)) for v in golden["c1_dist"].tolist()] and got[-1] == 0.5
    assert scoring.format_score_lines_2c([float(np.float32(-1.25))]) == ["-1.25\n"]
    q = tmp_path / "utt.txt"
    scoring.write_scores_utt(str(q), ["LA_E_1", "LA_E_2"], [0.25, float(np.float32(0.1))])
    assert q.read_text() == "LA_E_1 0.25\nLA_E_2 0.10000000149011612\n"
    assert scoring.load_scores_utt(str(q), parse=float) == (["LA_E_1", "LA_E_2"], [0.25, float(np.float32(0.1))])
    # the default reader reproduces pandas' C parser (calculate_eer.py:19), which is one ulp off on this very value
    assert scoring.load_scores_utt(str(q)) == (["LA_E_1", "LA_E_2"], [0.25, 0.1000000014901161])


def test_shard_bounds_cover_and_balance():
    for n in (0, 1, 7, 64, 71237):
        for w in (1, 2, 3, 4, 8):
            b = [parallel.shard_bounds(n, w, r) for r in range(w)]
            assert b[0][0] == 0 and b[-1][1] == n
            assert all(b[i][1] == b[i + 1][0] for i in range(w - 1))
            sizes = [hi - lo for lo, hi in b]
            assert max(sizes) - min(sizes) <= 1 and sizes == parallel.shard_sizes(n, w)
    with pytest.raises(ValueError):
        parallel.shard_bounds(10, 2, 2)


def _gather_worker(rank, world, port, n, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        allv = torch.arange(n, dtype=torch.float32) * 0.5 - 3.0
        lo, hi = parallel.shard_bounds(n, world, rank)
        got = parallel.all_gather_scores(allv[lo:hi].clone(), n)
        q.put((rank, bool(torch.equal(got, allv))))
        dist.barrier()
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("n", [7, 10])
def test_all_gather_scores_gloo_world2(n):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29500 + (os.getpid() + n) % 2000
    procs = [ctx.Process(target=_gather_worker, args=(r, 2, port, n, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = sorted(q.get(timeout=120) for _ in range(2))
    for p in procs:
        p.join(60)
    assert res == [(0, True), (1, True)]


def test_all_gather_single_process_passthrough():
    x = torch.arange(5.0)
    assert parallel.all_gather_scores(x, 5) is x
    with pytest.raises(ValueError):
        parallel.all_gather_scores(x, 6)


def test_metadata_loaders_match_reference_golden(tmp_path):
    """evaluate.py:7-84 loaders against outputs of the unmodified reference functions (oracle/gen_ingest_golden.py):
    duplicate protocol entries fill the first position only, absent files stay "", the last metadata line wins."""
    kat = json.load(open(os.path.join(GOLDEN, "evalproto_kat.json")))
    pp, mp_ = tmp_path / "proto.txt", tmp_path / "meta.txt"
    pp.write_text("\n".join(kat["proto"]) + "\n")
    mp_.write_text("\n".join(kat["meta"]) + "\n")
    assert scoring.load_proto_file(str(pp)) == kat["load_proto_file"]
    assert scoring.load_metadata(str(mp_)) == kat["load_metadata"]
    assert scoring.load_metadata_from_proto(str(mp_), str(pp)) == kat["load_metadata_from_proto"]
    assert scoring.eval_dict(str(mp_)) == kat["eval_dict"]


def test_metadata_from_proto_large_is_linear(tmp_path):
    """200k-entry protocol (ASVspoof2021 DF is 611 829): the dict-based loader finishes in well under a second and agrees
    with a direct restatement of the list.index semantics on a sample."""
    import time
    n = 200_000
    rng = np.random.default_rng(0)
    perm = rng.permutation(n)
    pp, mp_ = tmp_path / "proto.txt", tmp_path / "meta.txt"
    pp.write_text("".join("DF_E_%07d\n" % i for i in perm))
    mp_.write_text("".join("LA_%04d DF_E_%07d c d e %s g\n" % (i % 97, i, "bonafide" if i % 5 == 0 else "spoof") for i in range(0, n, 2)))
    t0 = time.time()
    labels = scoring.load_metadata_from_proto(str(mp_), str(pp))
    assert time.time() - t0 < 5.0
    want = ["" if i % 2 else ("bonafide" if i % 5 == 0 else "spoof") for i in perm.tolist()]
    assert labels == want


def test_create_reference_embedding2_cache_hit_needs_no_model(tmp_path):
    """oc_classifier.py:165-169: when both .pt files exist they are returned as stored, the model is never touched."""
    ref, thr = torch.arange(160, dtype=torch.float32).reshape(1, 160), torch.tensor(3.25)
    torch.save(ref, tmp_path / "reference_embedding.pt")
    torch.save(thr, tmp_path / "threshold.pt")
    r, t = scoring.create_reference_embedding2(None, None, "cpu", cache_dir=str(tmp_path))
    assert torch.equal(r, ref) and torch.equal(t, thr) and not (tmp_path / "distances.txt").exists()


# ------------------------------------------------------------------------------------ round 2: S2 join, DET curve, buckets
def test_s2_join_matches_reference_pandas_merge(tmp_path):
    """scoring.load_protocol_5col / load_scores_utt / join_scores_with_protocol against the stdout of the reference's own
    calculate_eer.py (pandas merge + compute_eer) on the same files: tests/golden/s2_join_kat.json (oracle/gen_golden_r2.py)."""
    from oracle import scoring as osc
    kat = json.load(open(os.path.join(GOLDEN, "s2_join_kat.json")))
    pp, sp = tmp_path / "proto.txt", tmp_path / "scores.txt"
    pp.write_text("\n".join(kat["protocol"]) + "\n")
    sp.write_text("\n".join(kat["scores"]) + "\n")
    rows = scoring.load_protocol_5col(str(pp))
    utts, sc = scoring.load_scores_utt(str(sp))
    assert len(rows) == len(kat["protocol"]) and len(utts) == len(kat["scores"])
    bona, spoof = scoring.join_scores_with_protocol(rows, utts, sc)
    eer, thr = osc.compute_eer(np.asarray(bona, dtype=np.float64), np.asarray(spoof, dtype=np.float64))
    assert f"EER = {eer * 100.0}, threshold = {thr}" == kat["stdout_last_line"]
    # a correctly rounded reader gives a different last digit: the reference's pandas parse is not round-trip exact
    _, sc_rt = scoring.load_scores_utt(str(sp), parse=float)
    assert sc_rt != sc
    # the writer produces the format the readers parse back
    p2 = tmp_path / "rt.txt"
    scoring.write_scores_utt(str(p2), utts, sc_rt)
    assert scoring.load_scores_utt(str(p2), parse=float) == (utts, sc_rt)


def test_pandas_float_equals_pandas_read_csv():
    """scoring.pandas_float == the double pandas' default C parser yields (calculate_eer.py:19), checked against pandas itself."""
    pd = pytest.importorskip("pandas")
    import io
    rng = np.random.default_rng(0)
    vals = np.concatenate([rng.normal(0, 1, 20000).astype(np.float32).astype(np.float64), rng.normal(0, 1e-3, 3000),
                           rng.normal(0, 1e4, 3000), 10.0 ** rng.uniform(-300, 300, 3000), [0.0, 1.0, -1.0, 1e22, 1e23, 5e-324]])
    strs = [repr(float(v)) for v in vals] + ["3", "-7", "1e5", "0.1e-2", "123456789012345678901234.5", "+2.5"]
    df = pd.read_csv(io.StringIO("\n".join(f"u{i} {s}" for i, s in enumerate(strs))), sep=" ", header=None)
    want = np.asarray(df[1].values, dtype=np.float64)
    got = np.array([scoring.pandas_float(s) for s in strs])
    assert np.array_equal(got, want)
    assert int((want[: len(vals)] != vals).sum()) > 1000          # ... and it really is not float(): thousands differ by an ulp


def test_det_curve_oracle_vs_reference_golden():
    from oracle import scoring as osc
    for k in json.load(open(os.path.join(GOLDEN, "det_kat.json"))):
        if "target" in k:
            t, n = np.asarray(k["target"], dtype=np.float64), np.asarray(k["nontarget"], dtype=np.float64)
        else:
            r = np.random.default_rng(7)
            t = np.round(r.normal(0.3, 1, 300) * 16) / 16
            n = np.round(r.normal(-0.3, 1, 500) * 16) / 16
        frr, far, thr = osc.compute_det_curve(t, n)
        assert frr.tolist() == k["frr"] and far.tolist() == k["far"] and thr.tolist() == k["thr"], k["name"]


def test_bucket_by_length_properties():
    lens = synth.config4_lengths(300)
    assert min(lens) >= 16000 and max(lens) <= 160000 and lens[:3] == [26558, 78939, 42616]        # seed 3, pinned
    buckets = scoring.bucket_by_length(lens, max_batch=64, max_padded_samples=64 * 64600)
    flat = [i for b in buckets for i in b]
    assert sorted(flat) == list(range(300))                                   # every utterance exactly once
    for b in buckets:
        longest = max(lens[i] for i in b)
        assert len(b) <= 64 and len(b) * longest <= 64 * 64600
        assert lens[b[0]] == longest                                          # sorted by length inside / across buckets
    heads = [lens[b[0]] for b in buckets]
    assert heads == sorted(heads, reverse=True)
    padded = sum(len(b) * max(lens[i] for i in b) for b in buckets)
    assert 1.0 - sum(lens) / padded < 0.15                                    # padding waste stays small
    # degenerate inputs
    assert scoring.bucket_by_length([]) == []
    assert scoring.bucket_by_length([5, 5, 5], max_batch=2) == [[0, 1], [2]]
    assert scoring.bucket_by_length([64600] * 100, max_batch=64) == [list(range(64)), list(range(64, 100))]
    with pytest.raises(ValueError):
        scoring.bucket_by_length([10, 0])


def test_config3_generator_is_index_addressed():
    """Utterance i depends only on i: any batching of the config-3 set yields the same waveforms (CPU generator)."""
    a = synth.config3_batch(5, 9, 7, "cpu", n_samples=2000)
    b = torch.cat([synth.config3_batch(5, 6, 7, "cpu", n_samples=2000), synth.config3_batch(6, 9, 7, "cpu", n_samples=2000)])
    assert torch.equal(a, b)
    assert abs(float(a[0].std()) - 0.1) < 0.01 and float(a[3].std()) > float(a[0].std())      # index 8 >= n_bona: tone added
    assert synth.config3_n_bonafide(71237) == 7355


def test_reference_archive_runs_unmodified_reference(golden):
    """oracle/_ref (oracle/build_ref.py): the UNMODIFIED reference AModel, imported from the packed archive with the fairseq
    stand-in, reproduces the committed config-1 golden -- this is what bench.py's reference arm / cpu_baseline time."""
    from oracle import ref_runner
    if not ref_runner.available():
        pytest.skip("oracle/_ref not built (python -m oracle.build_ref needs /root/reference)")
    model = ref_runner.load_reference_amodel("cpu")
    wav = synth.make_waveforms(seed=1, batch=8, n_samples=64600)[:1]
    with torch.no_grad():
        emb, out = model(wav)
    assert float((emb - golden["c1_emb"][:1]).abs().max()) < 5e-6
    t, n = np.array([0.9, 0.8, 0.35, 0.7]), np.array([0.1, 0.4, 0.35, 0.2, 0.05])
    assert tuple(float(v) for v in ref_runner.reference_compute_eer()(t, n)) == (0.225, 0.35)


# ------------------------------------------------------------------------------------ LCNN backend
def _lcnn_feats():
    g = torch.Generator().manual_seed(7)
    return {tag: torch.randn(B, T, 1024, generator=g) for tag, (B, T) in (("f49", (3, 49)), ("f201", (2, 201)), ("f499", (1, 499)))}


def test_lcnn_oracle_and_state_dict_vs_reference_golden():
    """oracle/lcnn.py against outputs of the UNMODIFIED models/lcnn.py (tests/golden/lcnn_golden.pt, oracle/gen_golden_r3.py), and
    the ssl_lcnn module's lcnn.* keys against the reference class's state_dict."""
    from oracle import lcnn as olc
    from occm_b200.models.lcnn import ssl_lcnn
    gold = torch.load(os.path.join(GOLDEN, "lcnn_golden.pt"), weights_only=False)
    sd = synth.make_lcnn_state_dict(0)
    for tag, feat in _lcnn_feats().items():
        assert torch.equal(olc.lcnn_backend(sd, feat), gold[f"{tag}_out"]), tag
        assert gold[f"{tag}_oracle_maxabs"] == 0.0
    m = ssl_lcnn("cpu")
    got = {k[len("lcnn."):]: list(v.shape) for k, v in m.state_dict().items() if k.startswith("lcnn.")}
    assert got == gold["keys"]
    assert any(k.startswith("frontend.model.encoder.layers.23.") for k in m.state_dict())
    with pytest.raises(ValueError, match="CUDA"):
        m(torch.zeros(1, 16000))


def test_senet_oracle_and_state_dict_vs_reference_golden():
    """oracle/senet.py against outputs of the UNMODIFIED models/senet.py (tests/golden/senet_golden.pt), and the key layouts of
    ssl_resnet34 / OCCM against the reference classes' state_dicts."""
    from oracle import senet as ose
    from occm_b200.models.occm import OCCM
    from occm_b200.models.senet import ssl_resnet34
    gold = torch.load(os.path.join(GOLDEN, "senet_golden.pt"), weights_only=False)
    sd = synth.make_resnet34_state_dict(0)
    g = torch.Generator().manual_seed(8)
    for tag, (B, T) in (("f49", (3, 49)), ("f201", (2, 201)), ("f312", (1, 312))):
        feat = torch.randn(B, T, 1024, generator=g)
        if tag == "f49":
            com, des = ose.resnet34_backend(sd, feat)
            assert torch.equal(com, gold["f49_com"]) and torch.equal(des, gold["f49_des"])
        assert gold[f"{tag}_oracle_maxabs"] == 0.0
    m = ssl_resnet34("cpu")
    got = {k[len("resnet34."):]: list(v.shape) for k, v in m.state_dict().items() if k.startswith("resnet34.")}
    assert got == gold["keys"]
    o = OCCM("cpu")
    ks = list(o.state_dict())
    assert {k[len("senet34_branch."):] for k in ks if k.startswith("senet34_branch.")} == set(gold["keys"])
    assert any(k.startswith("lcnn_branch.fc3.") for k in ks) and any(k.startswith("frontend.model.") for k in ks)


def test_gelu_polynomial_of_the_kernels_matches_erf():
    """The CUDA kernels' erf-GELU (csrc/common.cuh: gelu_erf_pair / gelu_erf_pairs) is q = 0.5 erfc(|x|/sqrt 2) = 2^(-P(u)),
    u = |x| sqrt(log2(e)/2), gelu = 0.5 x + |x| (0.5 - q).  The coefficients are read from the source and the same fp32 Horner steps
    are run in numpy: the result must stay within 2e-7 * max(|x|, 1) of torch's erf form (the form it replaced sat at 4e-7), tend to
    0 / x for large |x| and never blow up (negative leading coefficient)."""
    import math
    import re
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    src = open(os.path.join(root, "occm_b200", "csrc", "common.cuh")).read()
    body = src[src.index("__device__ __forceinline__ uint64_t gelu_erf_pair(uint64_t x)"):]
    body = body[:body.index("return f2_fma(u, h")]
    consts = [float(c) for c in re.findall(r"f2_splat\((-?[0-9.]+(?:e-?[0-9]+)?)f\)", body)]
    scale = float(re.search(r"fabsf\(x0\) \* ([0-9.]+)f", body).group(1))
    assert len(consts) == 11 and abs(scale - math.sqrt(math.log2(math.e) / 2)) < 1e-7
    lead, rest, (hm, hc) = consts[0], consts[1:9], consts[9:]
    assert lead < 0                                            # q -> 0 beyond the fitted interval
    f = np.float32
    x = np.concatenate([np.linspace(-12, 12, 400001), [-50.0, 50.0, -1e4, 1e4, -1e30, 1e30, 0.0]]).astype(f)
    u = np.abs(x) * f(scale)
    p = u * f(lead) + f(rest[0])
    with np.errstate(over="ignore", invalid="ignore"):
        for c in rest[1:]:
            p = (p * u + f(c)).astype(f)
        q = np.exp2(p.astype(np.float64)).astype(f)
        y = (u * (q * f(hm) + f(hc)).astype(f) + f(0.5) * x).astype(f)
    xd = torch.from_numpy(x.astype(np.float64))
    ref = (0.5 * xd * (1 + torch.erf(xd / math.sqrt(2)))).numpy()
    assert np.isfinite(y).all()
    err = np.abs(y - ref) / np.maximum(np.abs(x.astype(np.float64)), 1.0)
    assert err.max() < 2e-7, err.max()
