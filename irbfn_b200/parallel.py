"""Data-parallel plumbing for the IRBFN path: one process per GPU, queries sharded by rank,
parameters replicated, one all-reduce of one flat gradient buffer per training step
(SURVEY.md 8e).  Forward and input-gradient sweeps need no collective at all.

The reference is single-device (scripts/train_nmpc.py:43 pins CUDA_VISIBLE_DEVICES); the only
behaviour to preserve is "same gradients as one device on the concatenated batch".
"""
from __future__ import annotations

import os

import torch
import torch.distributed as dist


def env_world():
    """(rank, local_rank, world_size) from the torchrun environment (1 process -> 0, 0, 1)."""
    return (int(os.environ.get("RANK", "0")), int(os.environ.get("LOCAL_RANK", "0")),
            int(os.environ.get("WORLD_SIZE", "1")))


def init_distributed(backend: str = None):
    """Initialise torch.distributed from the environment if WORLD_SIZE > 1.  Returns
    (rank, local_rank, world_size).  NCCL on CUDA, gloo otherwise."""
    rank, local_rank, world = env_world()
    if world > 1 and not dist.is_initialized():
        if backend is None:
            backend = "nccl" if torch.cuda.is_available() else "gloo"
        if backend == "nccl":
            torch.cuda.set_device(local_rank)
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        kw = {"device_id": torch.device("cuda", local_rank)} if backend == "nccl" else {}
        dist.init_process_group(backend=backend, rank=rank, world_size=world, **kw)
    elif torch.cuda.is_available():
        torch.cuda.set_device(local_rank)
    return rank, local_rank, world


def shard_range(n: int, rank: int, world: int):
    """Contiguous, balanced [lo, hi) of n items for ``rank`` (the first n % world ranks get one
    more).  The union over ranks is [0, n) with no overlap, in rank order."""
    if world <= 0 or not (0 <= rank < world):
        raise ValueError("bad rank %d / world %d" % (rank, world))
    base, extra = divmod(n, world)
    lo = rank * base + min(rank, extra)
    return lo, lo + base + (1 if rank < extra else 0)


def allreduce_sum_(flat: torch.Tensor, group=None) -> torch.Tensor:
    """In-place sum of one flat buffer over all ranks (no-op for a single process)."""
    if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.all_reduce(flat, op=dist.ReduceOp.SUM, group=group)
    return flat


def allreduce_max_scalar(value: float, device=None) -> float:
    """Max of a python float over ranks (used for max-over-ranks timing)."""
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
        return float(value)
    t = torch.tensor([float(value)], dtype=torch.float64,
                     device=device if device is not None else ("cuda" if dist.get_backend() == "nccl" else "cpu"))
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def barrier():
    if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
        dist.barrier()
