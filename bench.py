#!/usr/bin/env python
"""bench.py -- utterances/s of the SSL-AASIST + one-class scoring hot path (BASELINE.json's metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--batch 64] [--dtype bf16|fp32]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P bench.py --gpus N ...

A "step" = one pass of the hot path over one batch: `batch` x 64 600-sample synthetic waveforms -> XLS-R frontend ->
AASIST backend -> (emb, out) -> one-class L2 distance + OC-Softmax cosine score.  Workload = BASELINE.json configs[1]
(batch 64, bf16, random-init XLS-R-300M arch).  N > 1: one process per GPU, each rank scores its own batch (weak
scaling), the step's scores are all-gathered over NCCL (the path's only exchange), global EER on the gathered scores.

One JSON line on stdout (rank 0): value (device-timed, inputs resident in HBM), e2e (host buffers through
scoring.score_host_batches: H2D + forward + score + D2H inside the timed region, pipelined), roofline of the dominant kernel family
(transformer GEMMs, tcgen05), cpu_baseline (the oracle on the host cores, bounded sample), clocks, gpu_launches.
`--impl reference` times the UNMODIFIED reference (models/sslassist.py AModel.forward + F.pairwise_distance) on the host cores,
imported from oracle/_ref -- the archive oracle/build_ref.py packs from /root/reference in the build container (fairseq stood
in by torchaudio's XLS-R-300M); if the archive did not travel it falls back to the oracle port and says so.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

N_SAMPLES = 64600
METRIC = "utterances/s SSL-AASIST+OC-Softmax scoring @64,600 samples"
GFLOP_PER_UTT = 150.18          # algorithmic work per 64 600-sample utterance
XF_GFLOP_PER_UTT = 121.40       # transformer linear GEMMs only
# dram__bytes_read.sum + dram__bytes_write.sum of the four transformer GEMMs of one layer (QKV 64.9 MB, out-proj + in-place residual
# + bf16 image 108.4 MB, FC1 94.8 MB, FC2 + residual + bf16 image 218.8 MB) from one `ncu --set full` capture at B=64 bf16 (ncu
# flushes the caches before each kernel, so reads are cold and writes still sit in L2): profiles/r4_layer_ncu_full.txt; averaged
# per launch.  Algorithmic bytes for comparison (operands once + outputs once): 112 / 134 / 140 / 219 MB -- no wasted re-reads.
XF_GEMM_DRAM_BYTES_PER_LAUNCH = (64.91e6 + 108.47e6 + 94.83e6 + 218.83e6) / 4


def load_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return dict(bf16_sustained=d.get("bf16_tflops_sustained", 1400.0), bf16_burst=d.get("bf16_tflops", 1590.0),
                    hbm=d.get("hbm_gbs", 6650.0), source="measured (MEASURED_PEAKS.json)")
    return dict(bf16_sustained=1400.0, bf16_burst=1590.0, hbm=6650.0, source="fallback (B200_PROFILING.md)")


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md recipe)."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.rows, self.proc = [], None
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(index), f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "100"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append((time.time(), [x.strip() for x in line.split(",")]))

    def stop(self, t0: float, t1: float) -> dict:
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        deadline = time.time() + 2.0            # nvidia-smi needs a moment to enumerate an 8-GPU box: wait for its first rows
        while time.time() < deadline and not any(len(r) >= 7 for _, r in self.rows):
            time.sleep(0.05)
        time.sleep(0.15)
        self.proc.terminate()
        rows = [r for ts, r in self.rows if t0 <= ts <= t1 + 0.2 and len(r) >= 7] or [r for _, r in self.rows if len(r) >= 7]
        if not rows:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["no samples"]}
        mhz = sorted(float(r[0]) for r in rows)
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for i, n in enumerate(names) if any(r[3 + i].lower().startswith("active") for r in rows)]
        pw = [float(r[2]) for r in rows if r[2].replace(".", "", 1).isdigit()]
        return {"sm_mhz": mhz[len(mhz) // 2], "sm_max_mhz": float(rows[0][1]), "power_w_max": max(pw) if pw else None,
                "samples": len(rows), "reasons": reasons}


# ------------------------------------------------------------------------------------------------ CPU legs (reference / oracle)
REF_BATCH = 8          # BASELINE.json configs[0]: the reference's own CPU-runnable case is batch 8 x 64 600, fp32


def workload_string(batch: int, dtype: str) -> str:
    return (f"SSL-AASIST (XLSR-300M arch, random-init) + one-class scoring, batch {batch} x {N_SAMPLES}-sample "
            f"16 kHz synthetic waveforms per GPU, {dtype} (BASELINE.json configs[1])")


def reference_step_fn(n_utt: int):
    """One bounded-sample step of the reference's CPU path on all host threads.  kind "reference": the UNMODIFIED
    models/sslassist.py from oracle/_ref (packed by oracle/build_ref.py where /root/reference exists; fairseq stood in by the
    torchaudio shim of SURVEY Appendix A) + F.pairwise_distance, exactly the loop body of oc_classifier.py:256-261 at batch
    n_utt.  kind "port": the oracle restatement, only when oracle/_ref did not travel."""
    import torch
    from occm_b200 import synth
    torch.set_num_threads(os.cpu_count() or 1)
    wav = synth.make_waveforms(1, n_utt, N_SAMPLES)
    ref = torch.zeros(1, 160)
    from oracle import ref_runner
    if ref_runner.available():
        import torch.nn.functional as F
        model = ref_runner.load_reference_amodel("cpu")

        def step():
            with torch.no_grad():
                emb, out = model(wav)
                return F.pairwise_distance(ref, emb, p=2)
        return step, "reference"
    from oracle import scoring as osc
    sd = synth.make_state_dict(0)

    def step_port():
        emb, out = osc.amodel_forward(sd, wav)
        return osc.one_class_scores(ref, emb)
    return step_port, "port"


def cpu_baseline(budget_s: float = 20.0) -> dict:
    """The reference timed on this box's host cores on a bounded sample of the same workload (about 10-30 s)."""
    import torch
    step, kind = reference_step_fn(REF_BATCH)
    step()                                                           # warm-up pass (thread pool, allocator)
    t0 = time.perf_counter(); step(); one = time.perf_counter() - t0
    reps = max(1, min(64, int(budget_s / max(one, 1e-3))))           # about budget_s seconds of CPU work
    t0 = time.perf_counter()
    for _ in range(reps):
        step()
    dt = time.perf_counter() - t0
    return {"value": REF_BATCH * reps / dt, "unit": "utterances/s", "cores": torch.get_num_threads(), "kind": kind,
            "sample": f"{reps} x {REF_BATCH} utterances x {N_SAMPLES} samples fp32 ({dt:.1f} s) after 2 warm-up passes; "
                      + ("the unmodified reference models/sslassist.py AModel.forward (oracle/_ref, fairseq stood in by torchaudio's XLS-R-300M) "
                         if kind == "reference" else "oracle port of AModel.forward ")
                      + f"+ pairwise_distance, batch {REF_BATCH} = BASELINE configs[0]; os.cpu_count()={os.cpu_count()}"}


def run_reference(args) -> None:
    """--impl reference: the reference's own CPU implementation of the path on the host cores (rank 0 only; it does not shard,
    so the other ranks exit without work and the line says so)."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    import torch
    n_utt = REF_BATCH
    step, kind = reference_step_fn(n_utt)
    t0 = time.perf_counter(); step(); first = time.perf_counter() - t0
    if first * (args.steps + args.warmup) > 240.0:                  # keep the whole run within a few minutes on a slow host
        n_utt = 2
        step, kind = reference_step_fn(n_utt)
    for _ in range(max(args.warmup - 1, 0)):
        step()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        step()
    dt = time.perf_counter() - t0
    v = n_utt * args.steps / dt
    sample = (f"{n_utt} utterances x {N_SAMPLES} samples per step (BASELINE configs[0] batch), fp32, {torch.get_num_threads()} host threads, "
              f"rank 0 only: the CPU arm is one process and does not scale with --gpus")
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": v, "unit": "utterances/s", "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1e3 * dt / args.steps, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "fp32", "data": "synthetic",
        "config": {"workload": workload_string(args.batch, args.dtype), "batch_per_gpu": args.batch, "n_samples": N_SAMPLES,
                   "sample": sample},
        "cpu_baseline": {"value": v, "unit": "utterances/s", "cores": torch.get_num_threads(), "kind": kind, "sample": sample},
        "e2e": {"value": v, "unit": "utterances/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0}), flush=True)


# ------------------------------------------------------------------------------------------------ GPU arm
def run_ours(args) -> None:
    import torch
    import torch.distributed as dist
    from occm_b200 import parallel, synth
    from occm_b200.evaluate_metrics import compute_eer
    from occm_b200.losses.custom_loss import cosine_scores
    from occm_b200.models.sslassist import AModel
    from occm_b200.scoring import one_class_scores

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus:
        if world == 1 and args.gpus > 1:
            raise SystemExit("bench.py --gpus N>1 must be launched with torch.distributed.run --nproc-per-node N")
    dev = torch.device("cuda", local)
    torch.cuda.set_device(dev)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)
    B, K, W = args.batch, args.steps, max(args.warmup, 3)
    peaks = load_peaks()

    model = AModel(None, dev, dtype=args.dtype).to(dev).eval()
    model.reserve(B, N_SAMPLES, dev)
    eng = model._engine
    # synthetic inputs: NB distinct batches per rank (rotated each step), seeded by global utterance index
    NB = 4
    host = [synth.make_waveforms(1000 + 17 * (rank * NB + j), B, N_SAMPLES).pin_memory() for j in range(NB)]
    devb = [h.to(dev) for h in host]
    g = torch.Generator().manual_seed(2)
    ref = (0.1 * torch.randn(1, 160, generator=g)).to(dev)
    center = torch.randn(160, generator=g).to(dev)
    n_total = world * B

    def step(i: int):
        emb, out = model(devb[i % NB])
        d = one_class_scores(ref, emb)
        c = cosine_scores(emb, center)
        if world > 1:
            d = parallel.all_gather_scores(d, n_total)
        return d, c

    def sync():
        torch.cuda.synchronize(dev)
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize(dev)

    with torch.no_grad():
        sampler = ClockSampler(local) if rank == 0 else None      # started before the warm-up: running by the time the timed region starts
        for i in range(W):
            step(i)
        sync()
        l0 = eng.launches
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t_wall0 = time.time()
        e0.record()
        for i in range(K):
            d, c = step(i)
        e1.record()
        sync()
        t_wall1 = time.time()
        ms = torch.tensor([e0.elapsed_time(e1)], device=dev, dtype=torch.float64)
        launches = eng.launches - l0 + 2 * K            # + the two score kernels per step
        clocks = sampler.stop(t_wall0, t_wall1) if sampler else None
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        ms_total = float(ms)

        # ---- per-family device timing (CUDA events on the launch stream) over a second timed pass -> roofline
        eng.profile(True)
        for i in range(min(K, 5)):
            model(devb[i % NB])
        torch.cuda.synchronize(dev)
        eng.profile(False)
        prof = eng.profile_read()

        # ---- e2e: host (pinned) buffers through the public scoring call (scoring.score_host_batches): per step the H2D of that
        # step's waveforms, forward, one-class score and the D2H of (emb, out, score) all fall inside the timed region; the
        # call pipelines them (copy stream + double buffering), as a user scoring an evaluation set would run it.
        from occm_b200.scoring import score_host_batches
        ref_h = ref.cpu()
        for _ in score_host_batches(model, (host[i % NB] for i in range(3)), ref_h, dev):
            pass
        # three passes of K steps each, the MEDIAN pass is reported (every pass listed in e2e.passes_ms): the region is 0.25 s of
        # wall clock on a shared host, and one pass of one record run read half the usual figure with the device-timed value unchanged
        e2e_passes = []
        for _ in range(3):
            sync()
            t0 = time.perf_counter()
            n_done = 0
            for emb_h, out_h, score_h in score_host_batches(model, (host[i % NB] for i in range(K)), ref_h, dev):
                n_done += emb_h.shape[0]
            torch.cuda.synchronize(dev)
            e2e_passes.append((time.perf_counter() - t0) * 1e3)
            assert n_done == B * K
        e2e_ms = torch.tensor([sorted(e2e_passes)[1]], device=dev, dtype=torch.float64)
        if world > 1:
            dist.all_reduce(e2e_ms, op=dist.ReduceOp.MAX)

        # ---- global EER on the last step's gathered scores (sanity; first half labelled bona fide)
        labels = torch.arange(d.numel(), device=dev) % 2 == 0
        eer, thr = compute_eer(d[~labels], d[labels])

    if rank == 0:
        xf = prof["xf_gemm"]
        ach = xf["flops"] / (xf["ms"] * 1e-3) / 1e12 if xf["ms"] > 0 else 0.0
        fam_ms = {k: round(v["ms"] / max(min(K, 5), 1), 3) for k, v in prof.items()}
        # per-family rooflines: tensor families against the sustained bf16 figure, memory-bound families (LayerNorm kernels,
        # conv0, backend) as achieved HBM GB/s on their ALGORITHMIC bytes (inputs read once + outputs written once; for the
        # backend: the LL operand planes + the residual-encoder activations of its 11 conv launches)
        Tfr = synth.num_frames(N_SAMPLES)
        Wp = Tfr // 3 + 2
        backend_bytes = min(K, 5) * B * (Tfr * 1024 * 4 + Tfr * 128 * 8 + 11 * 45 * Wp * 64 * 2 * 2 * 2 + 45 * Wp * 64 * 4 * 2)
        fams = {}
        for name, v in prof.items():
            if v["ms"] <= 0:
                continue
            if name in ("norm", "conv0", "backend"):
                byts = backend_bytes if name == "backend" else v["flops"]
                a = byts / (v["ms"] * 1e-3) / 1e9
                fams[name] = {"bound": "hbm", "achieved": round(a, 1), "peak": peaks["hbm"], "unit": "GB/s", "frac": round(a / peaks["hbm"], 4),
                              "ms_per_step": fam_ms[name], "launches_per_step": v["launches"] // max(min(K, 5), 1)}
            else:
                a = v["flops"] / (v["ms"] * 1e-3) / 1e12
                fams[name] = {"bound": "tensor", "achieved": round(a, 1), "peak": peaks["bf16_sustained"], "unit": "TFLOP/s",
                              "frac": round(a / peaks["bf16_sustained"], 4), "ms_per_step": fam_ms[name],
                              "launches_per_step": v["launches"] // max(min(K, 5), 1)}
        value = n_total * K / (ms_total * 1e-3)
        line = {
            "metric": METRIC, "value": value, "unit": "utterances/s", "n_gpus": world, "steps": K, "warmup": W,
            "ms_per_step": ms_total / K, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": args.dtype, "data": "synthetic",
            "config": {"workload": workload_string(B, args.dtype),
                       "batch_per_gpu": B, "n_samples": N_SAMPLES, "parallelism": f"shard{world}",
                       "l2": f"no explicit flush: {NB} rotating input batches; per-step working set (631 MB bf16 weights + "
                             f">2 GB activations) exceeds the 126 MB L2",
                       "eer_sanity": [eer, thr], "score_gather": parallel.LAST_GATHER_PATH if world > 1 else "single process: none"},
            "e2e": {"value": n_total * K / (float(e2e_ms) * 1e-3), "unit": "utterances/s",
                    "h2d_bytes_per_step": B * N_SAMPLES * 4 + 160 * 4, "d2h_bytes_per_step": B * (160 + 2 + 1) * 4,
                    "passes_ms": [round(x, 3) for x in e2e_passes],
                    "note": "median of three passes of K steps (rank 0's passes listed); "
                            "scoring.score_host_batches per rank (pinned host buffers -> H2D -> forward -> score -> D2H, pipelined); "
                            "at N > 1 each rank scores its own shard and the score all-gather (once per shard in the real driver, "
                            "tools/score_config3.py) is OUTSIDE this region"},
            "gpu_launches": int(launches),
            "frac_of_model_roofline": value / world * GFLOP_PER_UTT * 1e9 / (peaks["bf16_sustained"] * 1e12),
            "roofline": {"bound": "tensor", "kernel": "gemm_tc_kernel<256> (transformer QKV/out-proj/FC1/FC2, tcgen05+TMA)",
                         "achieved": ach, "peak": peaks["bf16_sustained"], "unit": "TFLOP/s",
                         "frac": ach / peaks["bf16_sustained"],
                         "traffic": XF_GEMM_DRAM_BYTES_PER_LAUNCH if (B == 64 and args.dtype == "bf16") else None,
                         "traffic_unit": "bytes/launch (ncu dram read+write, profiles/r4_layer_ncu_full.txt)",
                         "launches_timed": xf["launches"], "avg_launch_ms": xf["ms"] / max(xf["launches"], 1),
                         "peak_source": peaks["source"] + ", sustained figure (kernel timed inside a long step)",
                         "family_ms_per_step": fam_ms},
            "rooflines": fams,
            "clocks": clocks,
        }
        if not args.no_cpu_baseline and world == 1:
            line["cpu_baseline"] = cpu_baseline()
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--batch", type=int, default=64)
    ap.add_argument("--dtype", default="bf16", choices=["bf16", "fp32"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
