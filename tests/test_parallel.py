"""Data-parallel host logic on the CPU: world_size-2 gloo.  What is checked is the contract of
SURVEY.md 8e: queries shard by rank with no overlap, the step's only collective is one all-reduce
of one flat gradient buffer, and the reduced gradient equals the one-device gradient on the
concatenated batch (the oracle stands in for the device kernels, which these tests cannot run)."""
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from conftest import GOLDEN_DIR, load_golden, rel_l2

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


def test_shard_range_partitions_exactly():
    from irbfn_b200.parallel import shard_range
    for n in (0, 1, 7, 8, 262144, 1000003):
        for world in (1, 2, 3, 8):
            spans = [shard_range(n, r, world) for r in range(world)]
            assert spans[0][0] == 0 and spans[-1][1] == n
            assert all(a[1] == b[0] for a, b in zip(spans, spans[1:]))          # contiguous, no overlap
            sizes = [hi - lo for lo, hi in spans]
            assert max(sizes) - min(sizes) <= 1                                  # balanced
    with pytest.raises(ValueError):
        shard_range(10, 2, 2)


def test_env_world_defaults_to_single_process(monkeypatch):
    from irbfn_b200 import parallel
    for k in ("RANK", "LOCAL_RANK", "WORLD_SIZE"):
        monkeypatch.delenv(k, raising=False)
    assert parallel.env_world() == (0, 0, 1)
    t = torch.arange(4.0)
    assert parallel.allreduce_sum_(t) is t and parallel.allreduce_max_scalar(3.5) == 3.5   # no-ops
    parallel.barrier()


def _worker(rank, world, port, out_dir):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    os.environ.update(RANK=str(rank), LOCAL_RANK=str(rank), WORLD_SIZE=str(world),
                      MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    from irbfn_b200 import parallel
    from irbfn_b200.train import loss_and_cotangent
    from oracle import irbfn_oracle as orc
    r, lr, w = parallel.init_distributed("gloo")
    assert (r, w) == (rank, world) and dist.get_backend() == "gloo"

    cfg, params, arr = load_golden(os.path.join(GOLDEN_DIR, "c2_frenet_12x100.npz"))
    x = arr["x"][:512].astype(np.float64)
    tgt = np.random.default_rng(7).normal(size=(512, cfg["out_features"]))
    lo, hi = parallel.shard_range(512, rank, world)

    # this rank's shard: forward, cotangent of the GLOBAL-mean loss, VJP
    y = orc.wcrbf_forward(x[lo:hi], params, cfg)
    lval, dy = loss_and_cotangent(torch.from_numpy(y), torch.from_numpy(tgt[lo:hi]), 512, "l2")
    g = orc.wcrbf_vjp(x[lo:hi], params, cfg, dy.numpy())
    flat = torch.from_numpy(np.concatenate([g[k].ravel() for k in ("centers", "log_sigs", "kernel", "bias")]))
    buf = torch.cat([flat, lval.reshape(1)])            # the loss rides along: still ONE collective
    parallel.allreduce_sum_(buf)
    tmax = parallel.allreduce_max_scalar(float(rank + 1))
    parallel.barrier()
    if rank == 0:
        np.savez(os.path.join(out_dir, "dp.npz"), flat=buf[:-1].numpy(), loss=buf[-1].numpy(), tmax=tmax)
    dist.destroy_process_group()


def test_gloo_world2_gradients_equal_single_device(tmp_path):
    from irbfn_b200.train import loss_and_cotangent
    from oracle import irbfn_oracle as orc
    port = _free_port()
    mp.spawn(_worker, args=(2, port, str(tmp_path)), nprocs=2, join=True)
    z = np.load(tmp_path / "dp.npz")

    cfg, params, arr = load_golden(os.path.join(GOLDEN_DIR, "c2_frenet_12x100.npz"))
    x = arr["x"][:512].astype(np.float64)
    tgt = np.random.default_rng(7).normal(size=(512, cfg["out_features"]))
    y = orc.wcrbf_forward(x, params, cfg)
    lval, dy = loss_and_cotangent(torch.from_numpy(y), torch.from_numpy(tgt), 512, "l2")
    g = orc.wcrbf_vjp(x, params, cfg, dy.numpy())
    ref = np.concatenate([g[k].ravel() for k in ("centers", "log_sigs", "kernel", "bias")])
    assert z["flat"].shape == ref.shape
    assert rel_l2(z["flat"], ref) < 1e-12            # same gradients as one device on the concatenated batch
    assert abs(float(z["loss"]) - float(lval)) < 1e-12 * max(1.0, abs(float(lval)))
    assert float(z["tmax"]) == 2.0                   # max-over-ranks timing helper
