"""One-class scoring, score files and EER drivers -- the functions of the reference's
``oc_classifier.py`` / ``evaluate.py`` / ``calculate_eer.py`` that lie on the hot path, batched.

Reference behaviour kept (file:line):
  * ``create_reference_embedding``  <- ``create_reference_embedding2`` oc_classifier.py:159-202
        ref = mean of the bona-fide embeddings (index order), threshold = max ``||ref - emb_i + 1e-6||``.
  * ``one_class_scores``            <- ``F.pairwise_distance(ref, emb, p=2)`` oc_classifier.py:261
  * ``score_eval_set_1c`` / ``score_eval_set_2c`` <- ``score_eval_set_1c2`` :243-265 / ``score_eval_set_2c2`` :293-312
        score-file lines ``"<repr(float)>, <0|1> \\n"`` (note the trailing space) / ``"<repr(float)>\\n"``.
  * ``write_scores_utt`` / ``load_scores_utt``    <- the ``utt score`` format ``calculate_eer.py:13-25`` reads.
  * ``load_score_file``             <- evaluate.py:85-101 (``float(line.split(",")[0])``)
  * ``calculate_EER``               <- evaluate.py:127-146 (spoof = target class) and
    ``calculate_EER_utt``           <- calculate_eer.py:6-26 (bona fide = target class; argument order differs!).

The reference runs batch 1 with a device->host sync per utterance; here utterances are batched and
scores stay on the device until the file is written.
"""
from __future__ import annotations

import ctypes as C
from typing import Iterable, List, Optional, Sequence, Tuple

import torch

from . import _lib
from .evaluate_metrics import compute_eer

EPS = 1e-6   # torch.nn.functional.pairwise_distance default, added to the difference


def _stream(dev):
    return torch.cuda.current_stream(dev).cuda_stream


def one_class_scores(reference_embedding: torch.Tensor, emb: torch.Tensor, eps: float = EPS) -> torch.Tensor:
    """``||ref - emb[b] + eps||_2`` per row, on the GPU (``occ_score_l2``).  emb [B,D] CUDA -> [B]."""
    if not emb.is_cuda:
        raise ValueError("occm_b200: one_class_scores needs CUDA tensors (no CPU path)")
    lib = _lib.load()
    e = emb.detach().to(torch.float32).contiguous()
    r = reference_embedding.detach().to(device=e.device, dtype=torch.float32).reshape(-1).contiguous()
    B, D = e.shape
    if r.numel() != D:
        raise ValueError("occm_b200: reference embedding and embeddings differ in dimension")
    out = torch.empty(B, device=e.device, dtype=torch.float32)
    with torch.cuda.device(e.device):
        _lib.check(lib.occ_score_l2(e.data_ptr(), r.data_ptr(), B, D, eps, out.data_ptr(), _stream(e.device)), None, "occ_score_l2")
    return out


def reference_from_embeddings(embs: torch.Tensor, eps: float = EPS) -> Tuple[torch.Tensor, torch.Tensor]:
    """embs [n,D] CUDA -> (reference_embedding [1,D], threshold scalar tensor), oc_classifier.py:189-197."""
    if not embs.is_cuda:
        raise ValueError("occm_b200: reference_from_embeddings needs CUDA tensors (no CPU path)")
    lib = _lib.load()
    e = embs.detach().to(torch.float32).contiguous()
    n, D = e.shape
    ref = torch.empty(1, D, device=e.device, dtype=torch.float32)
    thr = torch.empty(1, device=e.device, dtype=torch.float32)
    with torch.cuda.device(e.device):
        _lib.check(lib.occ_ref_embedding(e.data_ptr(), n, D, eps, ref.data_ptr(), thr.data_ptr(), _stream(e.device)),
                   None, "occ_ref_embedding")
    return ref, thr[0]


def _batches(dataloader_or_tensor, batch_size: int, device) -> Iterable[torch.Tensor]:
    """Yields CUDA [b,N] batches from a [n,N] tensor or from a dataloader of (data, target) pairs."""
    if isinstance(dataloader_or_tensor, torch.Tensor):
        x = dataloader_or_tensor
        for i in range(0, x.shape[0], batch_size):
            yield x[i:i + batch_size].to(device, non_blocking=True)
    else:
        for item in dataloader_or_tensor:
            data = item[0] if isinstance(item, (tuple, list)) else item
            yield data.to(device, non_blocking=True)


def embed_all(model, data, device, batch_size: int = 64) -> Tuple[torch.Tensor, torch.Tensor]:
    embs, outs = [], []
    with torch.no_grad():
        for x in _batches(data, batch_size, device):
            e, o = model(x)
            embs.append(e)
            outs.append(o)
    return torch.cat(embs), torch.cat(outs)


def create_reference_embedding(model, data, device, batch_size: int = 64):
    """Batched ``create_reference_embedding2`` (no .pt caching side effects: the caller saves if wanted)."""
    model.eval()
    embs, _ = embed_all(model, data, device, batch_size)
    return reference_from_embeddings(embs)


def score_host_batches(model, host_batches, reference_embedding, device=None, eps: float = EPS):
    """Pipelined end-to-end scoring of HOST batches (the loop of ``score_eval_set_1c2``, oc_classifier.py:243-265, at batch
    size > 1): yields ``(emb, out, distance)`` CPU tensors per batch, in order.

    ``host_batches``: iterable of pinned (or pageable) CPU fp32 ``[b, N]`` tensors.  The H2D copy of batch i+1 runs on a
    copy stream while batch i is in the kernels, and the D2H of batch i's results is read one batch later, so PCIe time is
    hidden behind compute instead of being added to it (the reference pays a blocking ``.to(device)`` and a
    ``float(distance)`` sync per utterance).  Two device input buffers, two pinned result buffers."""
    dev = torch.device(device) if device is not None else next(model.parameters()).device
    if dev.index is None:
        dev = torch.device("cuda", torch.cuda.current_device())
    ref = reference_embedding.detach().to(dev, torch.float32).reshape(1, -1)
    copy_stream = torch.cuda.Stream(dev)
    main = torch.cuda.current_stream(dev)
    dbuf, ready, consumed = [None, None], [None, None], [None, None]
    res = [None, None]
    pending = None                                     # (slot, n, done_event) of the batch whose results are in flight

    def flush(p):
        slot, n, ev = p
        ev.synchronize()
        r = res[slot]
        return r[:n, :160].clone(), r[:n, 160:162].clone(), r[:n, 162].clone()

    with torch.no_grad():
        for i, hb in enumerate(host_batches):
            if hb.is_cuda or hb.dtype != torch.float32 or hb.ndim != 2:
                raise ValueError("occm_b200: score_host_batches expects CPU fp32 [b,N] batches")
            slot = i & 1
            n = hb.shape[0]
            with torch.cuda.stream(copy_stream):
                if consumed[slot] is not None:
                    copy_stream.wait_event(consumed[slot])         # the forward that read this buffer has finished with it
                if dbuf[slot] is None or dbuf[slot].shape != hb.shape:
                    dbuf[slot] = torch.empty(hb.shape, device=dev, dtype=torch.float32)
                dbuf[slot].copy_(hb, non_blocking=True)
                ready[slot] = copy_stream.record_event()
            main.wait_event(ready[slot])
            emb, out = model(dbuf[slot])
            consumed[slot] = main.record_event()
            d = one_class_scores(ref, emb)
            if res[slot] is None or res[slot].shape[0] < n:
                res[slot] = torch.empty(n, 163, dtype=torch.float32).pin_memory()
            packed = torch.cat([emb, out, d[:, None]], dim=1)
            res[slot][:n].copy_(packed, non_blocking=True)
            done = main.record_event()
            if pending is not None:
                yield flush(pending)
            pending = (slot, n, done)
        if pending is not None:
            yield flush(pending)


def format_score_lines_1c(distances: Sequence[float], threshold: float) -> List[str]:
    return [f"{float(d)}, {1 if float(d) > threshold else 0} \n" for d in distances]


def format_score_lines_2c(logits0: Sequence[float]) -> List[str]:
    return [f"{float(v)}\n" for v in logits0]


def score_eval_set_1c(model, data, device, reference_embedding, threshold, path: Optional[str] = "scores.txt",
                      batch_size: int = 64) -> torch.Tensor:
    """ONE-CLASS scoring (``score_eval_set_1c2``).  Returns the fp32 distances (CUDA) and writes ``path``."""
    model.eval()
    embs, _ = embed_all(model, data, device, batch_size)
    dist = one_class_scores(reference_embedding, embs)
    if path is not None:
        with open(path, "w") as f:
            f.writelines(format_score_lines_1c(dist.cpu().tolist(), float(threshold)))
    return dist


def score_eval_set_2c(model, data, device, path: Optional[str] = "scores.txt", batch_size: int = 64) -> torch.Tensor:
    """TWO-CLASS scoring (``score_eval_set_2c2``): the bona-fide logit ``out[:,0]``."""
    model.eval()
    _, outs = embed_all(model, data, device, batch_size)
    s = outs[:, 0].contiguous()
    if path is not None:
        with open(path, "w") as f:
            f.writelines(format_score_lines_2c(s.cpu().tolist()))
    return s


# ---------------------------------------------------------------------------------- score files
def load_score_file(file_path: str) -> List[float]:
    """evaluate.py:85-101."""
    with open(file_path, "r") as f:
        return [float(line.strip().split(",")[0]) for line in f if line.strip()]


def write_scores_utt(path: str, utts: Sequence[str], scores: Sequence[float]) -> None:
    """``utt score`` lines, the format calculate_eer.py:19-20 parses with ``sep=" "``."""
    with open(path, "w") as f:
        for u, s in zip(utts, scores):
            f.write(f"{u} {float(s)}\n")


_POW10 = [float("1e%d" % i) for i in range(309)]


def pandas_float(text: str) -> float:
    """The double ``pd.read_csv`` produces for a score field -- NOT always ``float(text)``.

    calculate_eer.py:19 reads the score file with ``pd.read_csv(..., sep=" ")``; since pandas 1.2 (the reference pins 2.0.2,
    requirement.txt:78) the C parser's default ``float_precision`` is "high", i.e. ``precise_xstrtod``
    (pandas/_libs/src/parser/tokenizer.c): at most 17 significant digits are accumulated in a double
    (``number = number * 10 + digit``, which rounds once the value passes 2^53), further digits only move the exponent, and
    the result is scaled by ONE multiplication / division with an exact-table power of ten.  That is double rounding: for
    about a third of the 17-digit ``repr`` strings of fp32 scores it is one ulp away from the correctly rounded
    ``float(text)``, and the EER threshold the reference prints carries that ulp.  Restated here so the S2 reader is
    bit-identical to the reference (checked against pandas itself in tests/test_host.py)."""
    s = text.strip()
    n = len(s)
    p = 0
    neg = False
    if p < n and s[p] in "+-":
        neg = s[p] == "-"
        p += 1
    low = s[p:].lower()
    if low in ("nan", "inf", "infinity", "-nan"):
        return float(s)
    number, exponent, digits, decimals = 0.0, 0, 0, 0
    while p < n and "0" <= s[p] <= "9":
        if digits < 17:
            number = number * 10.0 + (ord(s[p]) - 48)
            digits += 1
        else:
            exponent += 1
        p += 1
    if p < n and s[p] == ".":
        p += 1
        while p < n and "0" <= s[p] <= "9":
            if digits < 17:
                number = number * 10.0 + (ord(s[p]) - 48)
                digits += 1
                decimals += 1
            p += 1
        exponent -= decimals
    if digits == 0:
        raise ValueError(f"could not convert string to float: {text!r}")
    if neg:
        number = -number
    if p < n and s[p] in "eE":
        p += 1
        eneg = False
        if p < n and s[p] in "+-":
            eneg = s[p] == "-"
            p += 1
        e = 0
        while p < n and "0" <= s[p] <= "9":
            e = e * 10 + (ord(s[p]) - 48)
            p += 1
        exponent += -e if eneg else e
    if p != n:
        raise ValueError(f"could not convert string to float: {text!r}")
    if exponent > 308:
        return float("-inf") if neg else float("inf")
    if exponent > 0:
        return number * _POW10[exponent]
    if exponent < -308:
        if exponent < -616:
            return number * 0.0
        return number / _POW10[-308 - exponent] / _POW10[308]
    return number / _POW10[-exponent]


def load_scores_utt(path: str, parse=pandas_float) -> Tuple[List[str], List[float]]:
    """``utt score`` lines as calculate_eer.py:19-20 reads them (``pd.read_csv(sep=" ")``): scores converted like pandas'
    C parser does (``pandas_float``; pass ``parse=float`` for correctly rounded values)."""
    utts, scores = [], []
    with open(path, "r") as f:
        for line in f:
            p = line.split()
            if len(p) >= 2:
                utts.append(p[0])
                scores.append(parse(p[1]))
    return utts, scores


def load_protocol_5col(path: str):
    """``sid utt phy attack label`` rows (calculate_eer.py:13-17)."""
    rows = []
    with open(path, "r") as f:
        for line in f:
            p = line.split()
            if len(p) >= 5:
                rows.append(tuple(p[:5]))
    return rows


# ---------------------------------------------------------------------------------- EER drivers
def calculate_EER(scores: Sequence[float], labels: Sequence[str], verbose: bool = False) -> Tuple[float, float]:
    """evaluate.py:127-146: label == "spoof" -> target class; everything else non-target."""
    s = torch.as_tensor(list(scores) if not isinstance(scores, torch.Tensor) else scores, dtype=torch.float64)
    is_spoof = torch.tensor([l == "spoof" for l in labels], dtype=torch.bool, device=s.device)
    eer, thr = compute_eer(s[is_spoof], s[~is_spoof])
    if verbose:
        print(f"EER = {eer * 100.0}, threshold = {thr}")
    return eer, thr


def join_scores_with_protocol(protocol_rows, utts: Sequence[str], scores: Sequence[float]) -> Tuple[List[float], List[float]]:
    """The pandas inner join of calculate_eer.py:13-24 as a dict join: ``pd.merge(protocol, scores, on="utt")`` keeps the
    protocol's row order, drops utterances present on one side only, and repeats a protocol row once per score line of its
    utterance (in score-file order).  Returns ``(bonafide_scores, spoof_scores)`` in that order."""
    by_utt = {}
    for u, s in zip(utts, scores):
        by_utt.setdefault(u, []).append(float(s))
    bona, spoof = [], []
    for _, utt, _, _, label in protocol_rows:
        for s in by_utt.get(utt, ()):
            if label == "spoof":
                spoof.append(s)
            elif label == "bonafide":
                bona.append(s)
    return bona, spoof


def calculate_EER_utt(protocol_rows, utts: Sequence[str], scores: Sequence[float], verbose: bool = False):
    """calculate_eer.py:6-26: inner join on ``utt`` (protocol order), ``compute_eer(bonafide, spoof)``."""
    bona, spoof = join_scores_with_protocol(protocol_rows, utts, scores)
    eer, thr = compute_eer(bona, spoof)
    if verbose:
        print(f"EER = {eer * 100.0}, threshold = {thr}")
    return eer, thr


# ---------------------------------------------------------------------------------- evaluate.py metadata loaders
def load_metadata(file_path: str) -> List[str]:
    """``evaluate.py:7-22``: the label (6th space-separated column) of every line of a trial-metadata file."""
    with open(file_path, "r") as f:
        return [line.strip().split(" ")[5] for line in f.readlines()]


def load_proto_file(file_path: str) -> List[str]:
    """``evaluate.py:71-83``: one (stripped) file name per line."""
    with open(file_path, "r") as f:
        return [line.strip() for line in f.readlines()]


def eval_dict(file_path: str) -> dict:
    """``evaluate.py:50-69``: {file name (2nd column): label (6th column)}."""
    d = {}
    with open(file_path, "r") as f:
        for line in f:
            parts = line.strip().split(" ")
            d[parts[1]] = parts[5]
    return d


def load_metadata_from_proto(meta_file_path: str, proto_file_path: str) -> List[str]:
    """``evaluate.py:24-48`` with the same result in O(N) instead of O(N^2): labels in protocol order; a file absent from the
    metadata keeps ``""``; a file listed twice in the protocol is filled at its FIRST position only (``list.index``); the
    last metadata line of a file wins."""
    protos = load_proto_file(proto_file_path)
    first = {}
    for i, name in enumerate(protos):
        first.setdefault(name, i)
    labels = [""] * len(protos)
    with open(meta_file_path, "r") as f:
        for line in f.readlines():
            parts = line.strip().split(" ")
            i = first.get(parts[1])
            if i is not None:
                labels[i] = parts[5]
    return labels


def create_reference_embedding2(model, dataloader, device, cache_dir: str = "."):
    """``oc_classifier.py:159-202`` including its side effects: loads ``reference_embedding.pt`` / ``threshold.pt`` from
    ``cache_dir`` when both exist; otherwise embeds the bona fide set (batched), takes the mean embedding and the maximum
    distance to it (``occ_ref_embedding``), appends every distance to ``distances.txt`` and saves the two ``.pt`` files
    (``torch.load``-compatible: reference [1,160] and a 0-d threshold tensor, as the reference stores them)."""
    import os
    rp, tp = os.path.join(cache_dir, "reference_embedding.pt"), os.path.join(cache_dir, "threshold.pt")
    if os.path.exists(rp) and os.path.exists(tp):
        return torch.load(rp), torch.load(tp)
    model.eval()
    embs, _ = embed_all(model, dataloader, device, batch_size=64)
    ref, thr = reference_from_embeddings(embs)
    dist = one_class_scores(ref, embs)
    with open(os.path.join(cache_dir, "distances.txt"), "a") as f:
        f.writelines(f"{float(d)}\n" for d in dist.cpu().tolist())
    torch.save(ref, rp)
    torch.save(thr, tp)
    return ref, thr


# ---------------------------------------------------------------------------------- variable length (BASELINE.json configs[3])
def bucket_by_length(lengths: Sequence[int], max_batch: int = 64, max_padded_samples: int = 64 * 64600,
                     max_waste: float = 0.15, min_padded_samples: int = 16 * 64600) -> List[List[int]]:
    """Length-bucketed batching of variable-length utterances.

    The reference scores one un-padded utterance per forward (``oc_classifier.py:350,356``: ``batch_size=1``); batching them
    needs padding, and padded frames cost real GEMM rows.  Indices are sorted by length (descending, stable), then cut into
    consecutive buckets: a bucket closes when it holds ``max_batch`` utterances, when one more utterance would push
    ``count * longest`` past ``max_padded_samples`` (the workspace / step size), or -- once it carries at least
    ``min_padded_samples`` of work, enough to fill the GPU -- when the next utterance is more than ``max_waste`` shorter than
    the bucket's longest.  Returns the buckets as lists of ORIGINAL indices; every index appears exactly once, so results
    scattered back through them are in protocol order."""
    n = len(lengths)
    order = sorted(range(n), key=lambda i: -int(lengths[i]))
    buckets: List[List[int]] = []
    cur: List[int] = []
    longest = 0
    for i in order:
        L = int(lengths[i])
        if L <= 0:
            raise ValueError("bucket_by_length: lengths must be positive")
        if cur:
            full = len(cur) >= max_batch or (len(cur) + 1) * longest > max_padded_samples
            wasteful = L < (1.0 - max_waste) * longest and len(cur) * longest >= min_padded_samples
            if full or wasteful:
                buckets.append(cur)
                cur = []
        if not cur:
            longest = L
        cur.append(i)
    if cur:
        buckets.append(cur)
    return buckets


def score_varlen(model, waves: Sequence[torch.Tensor], device, reference_embedding: Optional[torch.Tensor] = None,
                 max_batch: int = 64, max_padded_samples: int = 64 * 64600, max_waste: float = 0.15):
    """Scores variable-length utterances (1-D fp32 tensors, CPU or CUDA) through length buckets with padding masks: each
    bucket is one zero-padded ``[b, N_max]`` batch with per-utterance lengths (masks in the conv encoder, the attention and
    every backend kernel), which reproduces the reference's batch-1 un-padded result per utterance.  Returns
    ``(emb [n,160], out [n,2], distance [n] or None)`` on ``device`` in the ORIGINAL order of ``waves``."""
    dev = torch.device(device)
    lengths = [int(w.numel()) for w in waves]
    n = len(waves)
    emb = torch.empty(n, 160, device=dev, dtype=torch.float32)
    out = torch.empty(n, 2, device=dev, dtype=torch.float32)
    with torch.no_grad():
        for bucket in bucket_by_length(lengths, max_batch, max_padded_samples, max_waste):
            nmax = max(lengths[i] for i in bucket)
            x = torch.zeros(len(bucket), nmax, device=dev, dtype=torch.float32)
            for r, i in enumerate(bucket):
                x[r, : lengths[i]].copy_(waves[i].reshape(-1), non_blocking=True)
            same = all(lengths[i] == nmax for i in bucket)
            e, o = model(x) if same else model(x, lengths=[lengths[i] for i in bucket])
            idx = torch.tensor(bucket, device=dev, dtype=torch.long)
            emb.index_copy_(0, idx, e)
            out.index_copy_(0, idx, o)
    dist = one_class_scores(reference_embedding, emb) if reference_embedding is not None else None
    return emb, out, dist


# ---------------------------------------------------------------------------------- indexed synthetic sets (configs[2])
def score_indexed_set(model, make_batch, lo: int, hi: int, reference_embedding: torch.Tensor, batch_size: int = 64,
                      pregenerated: Optional[torch.Tensor] = None) -> torch.Tensor:
    """One-class distances of utterances ``lo .. hi-1`` of an indexed set (the loop of ``score_eval_set_1c2``,
    oc_classifier.py:243-265, batched): ``make_batch(a, b)`` returns the CUDA ``[b-a, N]`` waveforms of utterances a..b-1, or
    ``pregenerated`` already holds ``[hi-lo, N]``.  Scores stay on the device; one tensor of ``hi - lo`` fp32 values."""
    dev = reference_embedding.device
    scores = torch.empty(hi - lo, device=dev, dtype=torch.float32)
    with torch.no_grad():
        for s in range(lo, hi, batch_size):
            t = min(s + batch_size, hi)
            x = pregenerated[s - lo:t - lo] if pregenerated is not None else make_batch(s, t)
            emb, _ = model(x)
            scores[s - lo:t - lo] = one_class_scores(reference_embedding, emb)
    return scores
