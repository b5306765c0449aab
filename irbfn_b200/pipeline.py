"""On-device input pipeline for training.

The reference's ``train_epoch`` (scripts/train_nmpc_frenet.py:487-540; scripts/train_nmpc.py:455-495)
permutes on the device, gathers ``train_x[perm]`` in NumPy on the HOST, copies the batch to the
device every step and calls ``jax.device_get`` on three scalars per step for wandb.  At ~0.8 ms of
compute per 262,144-sample step that host work dominates.  Here the whole table lives in HBM (a
180 GB device holds ~2.5 G rows of an 8-in / 10-out table), the permutation is drawn on the device,
each step's batch is one gather launch (irbfn_gather_rows), and the per-step losses stay on the
device until the epoch ends.

Data parallel: every rank draws the SAME permutation (same seed) and gathers only its
``parallel.shard_range`` slice of each batch, so the union over ranks is exactly the one-device batch.
The permutation itself comes from torch's device generator; the reference's comes from JAX's threefry
PRNG, which cannot be reproduced here (no JAX) -- same distribution, different stream, documented as
parity-unpinned.  The gather is index-exact.
"""
from __future__ import annotations

from typing import Iterator, Optional, Tuple

import numpy as np
import torch

from . import ops, parallel


class DeviceTable:
    """The training table ``(train_x (N,D), train_y (N,O))`` resident on one device."""

    def __init__(self, train_x, train_y, device=None, infeasible_value: Optional[float] = None):
        device = torch.device(device) if device is not None else torch.device("cuda", torch.cuda.current_device())
        x = np.ascontiguousarray(np.asarray(train_x), dtype=np.float32) if not isinstance(train_x, torch.Tensor) else train_x
        y = np.ascontiguousarray(np.asarray(train_y), dtype=np.float32) if not isinstance(train_y, torch.Tensor) else train_y
        x = torch.as_tensor(x).to(device=device, dtype=torch.float32).contiguous()
        y = torch.as_tensor(y).to(device=device, dtype=torch.float32).contiguous()
        if x.dim() != 2 or y.dim() != 2 or x.shape[0] != y.shape[0]:
            raise ValueError("train_x (N,D) and train_y (N,O) must have the same number of rows")
        if infeasible_value is not None:
            # valid_ind = np.unique(np.where(outputs != -999)[0]) (scripts/train_nmpc_frenet.py:49-52):
            # a row survives if ANY of its outputs differs from the infeasible marker; once, on the device
            keep = (y != float(infeasible_value)).any(dim=1)
            x, y = x[keep].contiguous(), y[keep].contiguous()
        self.x, self.y, self.device = x, y, device
        self.num_rows = int(x.shape[0])

    def permutation(self, seed: int) -> torch.Tensor:
        """A device-resident permutation of the rows; identical on every rank for the same seed."""
        g = torch.Generator(device=self.device)
        g.manual_seed(int(seed))
        return torch.randperm(self.num_rows, generator=g, device=self.device)

    def batches(self, batch_size: int, seed: int, rank: int = 0, world: int = 1
                ) -> Iterator[Tuple[torch.Tensor, torch.Tensor]]:
        """The epoch's batches as train_epoch cuts them: ``perms[: steps * bs].reshape(steps, bs)``
        (train_nmpc_frenet.py:491-497).  With ``world > 1`` each rank gets its contiguous slice of
        every batch.  Two staging buffers alternate, so a yielded batch stays valid for one more step."""
        bs = int(batch_size)
        steps = self.num_rows // bs
        if steps == 0:
            return
        perm = self.permutation(seed)[: steps * bs].view(steps, bs)
        lo, hi = parallel.shard_range(bs, rank, world)
        n = hi - lo
        D, O = self.x.shape[1], self.y.shape[1]
        bufs = [(torch.empty((n, D), device=self.device), torch.empty((n, O), device=self.device)) for _ in range(2)]
        for s in range(steps):
            xb, yb = bufs[s & 1]
            ops.gather_rows(self.x, self.y, perm[s, lo:hi], x_out=xb, y_out=yb)
            yield xb, yb


def train_epoch(state, table: DeviceTable, batch_size: int, epoch_seed: int, loss: str = "frenet_fullint",
                rank: Optional[int] = None, world: Optional[int] = None):
    """One epoch of the reference's train_epoch with the gather on the device and no per-step host
    synchronisation.  Returns (mean batch loss, per-batch losses), both device tensors."""
    from .train import train_step
    r, _, w = parallel.env_world()
    rank = r if rank is None else rank
    world = w if world is None else world
    losses = []
    for xb, yb in table.batches(batch_size, epoch_seed, rank, world):
        losses.append(train_step(state, xb, yb, global_batch=batch_size, loss=loss).reshape(()))
    if not losses:
        z = torch.zeros((), device=table.device)
        return z, z.reshape(1)
    stacked = torch.stack(losses)
    return stacked.mean(), stacked
