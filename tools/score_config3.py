#!/usr/bin/env python
"""BASELINE.json configs[2]: an ASVspoof2019-LA-eval-sized synthetic set (71 237 x 64 600 samples, 7 355 bona fide first)
sharded over 1/2/4/8 B200, scores all-gathered, global EER.

    python tools/score_config3.py [--utterances 71237] [--batch 64] [--dtype bf16]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P tools/score_config3.py ...

Utterance i is generated ON DEVICE from the counter-based seed 1000 + i (occm_b200.synth.config3_batch), so any
sharding sees identical data.  Bona fide utterances are white noise, spoof utterances carry an added tone of random amplitude
and frequency -- a perturbation that survives the conv encoder's per-frame LayerNorm, so the EER is informative (~10 %).  Prints one JSON line: utt/s (device-timed, max over ranks), EER / threshold
from the GPU kernel (bit-exact against the reference's numpy compute_eer: tests/test_gpu_eer_kernels.py), and a
SHA-256 of the gathered fp32 scores -- equal across world sizes because per-utterance results do not depend on the batch.
Mirrors the loop of /root/reference/oc_classifier.py:243-265 (score_eval_set_1c2) + evaluate.py:127-146.
"""
from __future__ import annotations

import argparse
import hashlib
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
N_SAMPLES = 64600


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--utterances", type=int, default=71237)
    ap.add_argument("--batch", type=int, default=64)
    ap.add_argument("--dtype", default="bf16", choices=["bf16", "fp32"])
    a = ap.parse_args()
    import torch
    import torch.distributed as dist
    from occm_b200 import parallel, synth
    from occm_b200.evaluate_metrics import compute_eer
    from occm_b200.models.sslassist import AModel
    from occm_b200.scoring import score_indexed_set

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    dev = torch.device("cuda", local)
    torch.cuda.set_device(dev)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)
    n, B = a.utterances, a.batch
    n_bona = synth.config3_n_bonafide(n)
    model = AModel(None, dev, dtype=a.dtype).to(dev).eval()
    model.reserve(B, N_SAMPLES, dev)

    with torch.no_grad():
        # reference embedding: mean embedding of 32 separate "train" bona fide utterances (seed 2), every rank alike
        ref = model(synth.config3_train_batch(dev))[0].mean(dim=0, keepdim=True)
        lo, hi = parallel.shard_bounds(n, world, rank)
        # this rank's shard is generated BEFORE the timed region (one RNG launch per utterance is synthetic-data cost, not the
        # path's): 64 600 x 4 B per utterance, 18.4 GB for the whole set on one GPU
        shard = torch.empty(hi - lo, N_SAMPLES, device=dev)
        for s in range(lo, hi, 512):
            t = min(s + 512, hi)
            shard[s - lo:t - lo] = synth.config3_batch(s, t, n_bona, dev)
        torch.cuda.synchronize(dev)
        if world > 1:
            dist.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        scores = score_indexed_set(model, None, lo, hi, ref, batch_size=B, pregenerated=shard)
        allsc = parallel.all_gather_scores(scores, n)            # ONE gather per shard (occ_allgather_scores over NCCL)
        is_bona = torch.arange(n, device=dev) < n_bona
        # evaluate.py:145 calls compute_eer(spoof, bonafide): spoof scores are the "target" set
        eer, thr = compute_eer(allsc[~is_bona], allsc[is_bona])
        e1.record()
        torch.cuda.synchronize(dev)
        ms = torch.tensor([e0.elapsed_time(e1)], device=dev, dtype=torch.float64)
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    if rank == 0:
        sc = allsc.cpu().numpy()
        line = {"config": "BASELINE.json configs[2]", "n_utterances": n, "n_bonafide": n_bona, "n_gpus": world, "dtype": a.dtype,
                "batch_per_gpu": B, "seconds": float(ms) / 1e3, "utterances_per_s": n / (float(ms) / 1e3),
                "note": "timed region = forward + one-class score over the rank's shard (waveforms pre-generated in HBM), ONE all-gather, global EER; max over ranks",
                "eer": eer, "threshold": thr, "scores_sha256": hashlib.sha256(sc.tobytes()).hexdigest()}
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
