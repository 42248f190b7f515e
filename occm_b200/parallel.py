"""Multi-GPU scoring: one process per GPU, utterances sharded, ONE exchange step.

Replaces the reference's ``nn.DataParallel`` wrap (``/root/reference/oc_classifier.py:343``), which at
batch 1 only ever uses GPU 0.  Utterances are independent in eval mode, so each rank scores a
contiguous block of the protocol order with no data-path collective; the only exchange is an
``all_gather`` of the fp32 scores (71 237 x 4 B for ASVspoof2019-LA-eval) over NCCL/NVLink, after
which every rank holds the scores in protocol order and rank 0 runs the global EER / writes the
score file.  Per-utterance results do not depend on batch composition (no split-K across
utterances, fixed reduction order per row), so scores and EER are bit-identical for any world size.
"""
from __future__ import annotations

from typing import List, Tuple

import torch
import torch.distributed as dist


LAST_GATHER_PATH = "none"      # which exchange the last all_gather_scores used (bench.py reports it)


def shard_bounds(n_items: int, world_size: int, rank: int) -> Tuple[int, int]:
    """Contiguous block [lo, hi) of rank ``rank``; the first ``n % world`` ranks get one extra item."""
    if world_size < 1 or not (0 <= rank < world_size):
        raise ValueError("bad rank / world_size")
    base, rem = divmod(int(n_items), world_size)
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def shard_sizes(n_items: int, world_size: int) -> List[int]:
    return [shard_bounds(n_items, world_size, r)[1] - shard_bounds(n_items, world_size, r)[0] for r in range(world_size)]


def all_gather_scores(local_scores: torch.Tensor, n_items: int, group=None) -> torch.Tensor:
    """local_scores: this rank's block (1-D, any float dtype) -> all ``n_items`` scores in global order on
    every rank.  Ragged blocks are padded to the largest block for a single fixed-size all_gather."""
    if not (dist.is_available() and dist.is_initialized()):
        if local_scores.numel() != n_items:
            raise ValueError("single process: local block must hold every item")
        return local_scores
    world = dist.get_world_size(group)
    rank = dist.get_rank(group)
    sizes = shard_sizes(n_items, world)
    if local_scores.numel() != sizes[rank]:
        raise ValueError(f"rank {rank}: expected {sizes[rank]} local scores, got {local_scores.numel()}")
    m = max(sizes)
    buf = torch.zeros(m, dtype=local_scores.dtype, device=local_scores.device)
    buf[: sizes[rank]] = local_scores
    out = torch.empty(world * m, dtype=local_scores.dtype, device=local_scores.device)
    global LAST_GATHER_PATH
    comm = _nccl_comm_ptr(group, local_scores) if local_scores.is_cuda and local_scores.dtype == torch.float32 else 0
    LAST_GATHER_PATH = "occ_allgather_scores (C ABI, ncclAllGather on torch's communicator)" if comm else "torch.distributed all_gather"
    if comm:
        # the C-ABI exchange (include/occ_b200.h: occ_allgather_scores) on torch's communicator and current stream
        import ctypes as C
        from . import _lib
        lib = _lib.load()
        with torch.cuda.device(local_scores.device):
            _lib.check(lib.occ_allgather_scores(buf.data_ptr(), m, out.data_ptr(), C.c_void_p(comm),
                                                torch.cuda.current_stream(local_scores.device).cuda_stream), None, "occ_allgather_scores")
    elif hasattr(dist, "all_gather_into_tensor") and local_scores.is_cuda:
        dist.all_gather_into_tensor(out, buf, group=group)
    else:
        dist.all_gather(list(out.view(world, m).unbind(0)), buf, group=group)      # gloo (CPU tests)
    out = out.view(world, m)
    return torch.cat([out[r, : sizes[r]] for r in range(world)])


def _nccl_comm_ptr(group, tensor: torch.Tensor) -> int:
    """ncclComm_t of the (default) process group for ``tensor``'s device, or 0 when the backend is not NCCL / torch does not
    expose it (then torch.distributed performs the gather)."""
    try:
        pg = group if group is not None else dist.distributed_c10d._get_default_group()
        if dist.get_backend(pg) != "nccl":
            return 0
        be = pg._get_backend(tensor.device)
        return int(be._comm_ptr())
    except Exception:
        return 0


def split_by_label(scores: torch.Tensor, is_target: torch.Tensor) -> Tuple[torch.Tensor, torch.Tensor]:
    """(target_scores, nontarget_scores) in protocol order (what evaluate.py:138-142 builds)."""
    return scores[is_target], scores[~is_target]
