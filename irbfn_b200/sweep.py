"""The explicit-vs-IRBFN lattice sweep on the device.

Mirrors scripts/irbfn_vs_explicit_frenet.ipynb cells 8-13 (and irbfn_vs_explicit.ipynb cells 7-12):

    test_inputs      = lattice over two state dimensions, the others held          (cell 8)
    pred_u_irbfn     = pred_step(state, test_inputs)                               (cell 10)
    inds_explicit    = vmap(argmin(norm(inputs - q, axis=1)))(test_inputs)         (cells 4, 11)
    pred_u_explicit  = outputs[inds_explicit].reshape((-1, O), order="F")          (cell 12)
    errors           = pred_u_irbfn - pred_u_explicit                              (cell 13)

The reference evaluates 400 lattice points against a table that lives on the host; here the table is
resident in HBM, the lattice can have 2^20 points (BASELINE config C2), the network evaluation is the fused
forward kernel, the lookup is irbfn_nearest_row and the row fetch irbfn_gather_rows.  Queries shard across
ranks with no collective (every rank holds the table).  Plotting (seaborn heat maps) stays with the caller.
"""
from __future__ import annotations

from typing import Dict, Optional, Sequence

import numpy as np
import torch

from . import ops, parallel


def lattice_queries(axes: Dict[int, Sequence[float]], fixed: Sequence[float]) -> np.ndarray:
    """(prod n_i, D) float32 queries: a meshgrid (indexing="ij", like cell 8) over the columns named in
    ``axes`` (column -> axis values, in insertion order), every other column held at ``fixed[col]``."""
    cols = list(axes.keys())
    grids = np.meshgrid(*[np.asarray(axes[c], np.float64) for c in cols], indexing="ij")
    n = grids[0].size if grids else 1
    x = np.tile(np.asarray(fixed, np.float32)[None, :], (n, 1))
    for c, g in zip(cols, grids):
        x[:, c] = g.reshape(-1)
    return x


class ExplicitTable:
    """The explicit NMPC table on one device: ``inputs (N, D)`` and ``outputs`` flattened per row in Fortran
    order, as ``outputs[inds].reshape((-1, O), order="F")`` does for (N, H, 2) control sequences."""

    def __init__(self, inputs, outputs, device=None):
        device = torch.device(device) if device is not None else torch.device("cuda", torch.cuda.current_device())
        inp = np.ascontiguousarray(np.asarray(inputs), dtype=np.float32)
        out = np.asarray(outputs)
        if out.shape[0] != inp.shape[0]:
            raise ValueError("inputs and outputs must have the same number of rows")
        # a = outputs[inds] is (B, H, 2); a.reshape((-1, O), order="F")[b, h + H*c] == a[b, h, c]: rows stay rows,
        # each flattened in Fortran order = trailing axes reversed, then C order.  Done once for the whole table.
        if out.ndim == 1:
            out = out[:, None]
        flat = np.transpose(out, (0,) + tuple(range(out.ndim - 1, 0, -1))).reshape(out.shape[0], -1)
        self.inputs = torch.from_numpy(inp).to(device)
        self.outputs = torch.from_numpy(np.ascontiguousarray(flat, dtype=np.float32)).to(device)
        self.device = device
        self.num_rows, self.in_features = inp.shape
        self.out_features = self.outputs.shape[1]


def irbfn_vs_explicit(net, params, table: ExplicitTable, queries, rank: Optional[int] = None,
                      world: Optional[int] = None) -> Dict[str, torch.Tensor]:
    """Evaluate this rank's shard of ``queries`` (NumPy or CUDA tensor, (B, D)) both ways.  Returns device
    tensors: ``pred_irbfn (b, O)``, ``index (b,)`` int64 nearest table row, ``distance (b,)``,
    ``pred_explicit (b, O_table)``, and -- when the widths agree -- ``errors = pred_irbfn - pred_explicit``
    with ``mean_abs_error (O,)`` of the shard."""
    r, _, w = parallel.env_world()
    rank = r if rank is None else rank
    world = w if world is None else world
    dev = table.device
    q = queries if isinstance(queries, torch.Tensor) else torch.from_numpy(np.ascontiguousarray(queries, dtype=np.float32))
    lo, hi = parallel.shard_range(q.shape[0], rank, world)
    q = q[lo:hi].to(device=dev, dtype=torch.float32).contiguous()
    out = net.apply(params, q)
    pred = out[0] if isinstance(out, tuple) else out          # ClusterWCRBFNet returns (y, logits)
    idx, dist = ops.nearest_row(q, table.inputs)
    explicit = ops.gather_rows(table.outputs, None, idx)
    res = {"queries": q, "pred_irbfn": pred, "index": idx, "distance": dist, "pred_explicit": explicit}
    if explicit.shape[1] == pred.shape[1]:
        err = pred - explicit
        res["errors"] = err
        res["mean_abs_error"] = err.abs().mean(dim=0)
    return res
