"""The explicit-vs-IRBFN lattice sweep (irbfn_b200/sweep.py) against the NumPy restatement of the
notebook cells (oracle/table_oracle.py lattice / explicit_lookup, oracle/irbfn_oracle.py forward).

Bars: lattice construction and the Fortran-order row flatten are exact; the nearest-row index follows
tests/test_table_gpu.py's rule (float64 argmin unless the runner-up ties within float32 rounding); the
network column is within 1e-5 norm-wise of the float64 oracle (BASELINE.json tolerance)."""
import os

import numpy as np
import pytest
import torch

from conftest import GOLDEN_DIR, load_golden, rel_max
from oracle import irbfn_oracle as orc
from oracle import table_oracle as tor


def _frenet_lattice(n=20):
    # scripts/irbfn_vs_explicit_frenet.ipynb cell 8: ey x epsi lattice, vx_car = 1.076, vx_goal = 4.0
    axes = {0: np.linspace(0.0, 2.0, n), 6: np.linspace(0.0, 1.0, n)}
    fixed = [0.0, 0.0, 1.076, 0.0, 4.0, 0.0, 0.0, 0.0]
    return axes, fixed


def test_lattice_matches_the_notebook_construction():
    from irbfn_b200.sweep import lattice_queries
    axes, fixed = _frenet_lattice()
    got = lattice_queries(axes, fixed)
    # the notebook's own spelling: meshgrid ij, flatten, column_stack with zero columns
    ey, epsi = np.meshgrid(axes[0], axes[6], indexing="ij")
    ey, epsi = ey.flatten(), epsi.flatten()
    zs = np.zeros_like(ey)
    want = np.column_stack((ey, zs, 1.076 * np.ones_like(ey), zs, 4.0 * np.ones_like(ey), zs, epsi, zs))
    assert got.shape == (400, 8) and got.dtype == np.float32
    assert np.array_equal(got, want.astype(np.float32))
    assert np.array_equal(got, tor.lattice(axes, fixed))
    one = lattice_queries({2: [1.0, 2.0, 3.0]}, [9.0, 8.0, 7.0])
    assert np.array_equal(one, np.array([[9, 8, 1], [9, 8, 2], [9, 8, 3]], np.float32))


def test_fortran_row_flatten_is_what_the_notebook_reshape_does():
    """outputs[inds].reshape((-1, O), order="F") keeps rows and flattens each (H, 2) block column-major."""
    rng = np.random.default_rng(0)
    out = rng.normal(size=(50, 5, 2)).astype(np.float32)
    inds = rng.integers(0, 50, 17)
    want = out[inds].reshape((-1, 10), order="F")
    flat = np.transpose(out, (0, 2, 1)).reshape(50, -1)          # what ExplicitTable stores
    assert np.array_equal(flat[inds], want)
    assert np.array_equal(want[:, :5], out[inds][:, :, 0]) and np.array_equal(want[:, 5:], out[inds][:, :, 1])


@pytest.mark.gpu
@pytest.mark.parametrize("fixture,n", [("c2b_frenet_1x1000_o10.npz", 20), ("c2b_frenet_1x1000_o10.npz", 64)])
def test_sweep_matches_oracle(fixture, n, cuda_device):
    from irbfn_b200.basis import gaussian, inverse_quadratic
    from irbfn_b200.model import WCRBFNet, device_params
    from irbfn_b200.sweep import ExplicitTable, irbfn_vs_explicit, lattice_queries
    cfg, params, arr = load_golden(os.path.join(GOLDEN_DIR, fixture))
    bf = gaussian if cfg["basis_func"] == "gaussian" else inverse_quadratic
    net = WCRBFNet(cfg["in_features"], cfg["out_features"], cfg["num_kernels"], bf, cfg["num_regions"],
                   cfg["lower_bounds"], cfg["upper_bounds"], cfg["dimension_ranges"], cfg["activation_idx"],
                   cfg["delta"])
    dp = device_params({"params": params}, cuda_device)
    rng = np.random.default_rng(n)
    N, H = 30000, 5
    lo, hi = np.array([0, -0.4, 0.5, -1, 0.5, -2, -0.5, -0.2]), np.array([2, 0.4, 8, 1, 8, 2, 1, 0.2])
    tin = (lo + (hi - lo) * rng.random((N, 8))).astype(np.float32)
    tout = rng.normal(size=(N, H, 2)).astype(np.float32)
    axes, fixed = _frenet_lattice(n)
    q = lattice_queries(axes, fixed)
    table = ExplicitTable(tin, tout, cuda_device)
    assert table.out_features == 10 and table.num_rows == N
    res = irbfn_vs_explicit(net, dp, table, q, rank=0, world=1)
    torch.cuda.synchronize()

    y64 = orc.wcrbf_forward(q.astype(np.float64), params, cfg)
    assert rel_max(res["pred_irbfn"].cpu().numpy(), y64) < 1e-5
    inds, explicit, d1, d2 = tor.explicit_lookup(q, tin, tout, 10)
    got_i = res["index"].cpu().numpy()
    clear = (d2 - d1) > 1e-5 * np.maximum(d1, 1e-30)
    assert clear.mean() > 0.95
    assert np.array_equal(got_i[clear], inds[clear])
    dsel = np.linalg.norm(tin[got_i].astype(np.float64) - q.astype(np.float64), axis=1)
    assert np.allclose(dsel, d1, rtol=1e-5, atol=1e-12)
    assert np.allclose(res["distance"].cpu().numpy(), dsel, rtol=1e-5, atol=1e-7)
    # the row fetch is exact for whichever row was picked
    assert np.array_equal(res["pred_explicit"].cpu().numpy(), tout[got_i].reshape((-1, 10), order="F"))
    err = res["pred_irbfn"].cpu().numpy() - res["pred_explicit"].cpu().numpy()
    assert np.array_equal(res["errors"].cpu().numpy(), err)
    assert np.allclose(res["mean_abs_error"].cpu().numpy(), np.abs(err).mean(axis=0), rtol=1e-5)


@pytest.mark.gpu
def test_sweep_shards_tile_the_lattice(cuda_device):
    """Two ranks' shards, concatenated in rank order, are the one-rank sweep (no collective on the path)."""
    from irbfn_b200.basis import gaussian
    from irbfn_b200.model import WCRBFNet, device_params
    from irbfn_b200.sweep import ExplicitTable, irbfn_vs_explicit, lattice_queries
    cfg, params, arr = load_golden(os.path.join(GOLDEN_DIR, "c2_frenet_12x100.npz"))
    net = WCRBFNet(cfg["in_features"], cfg["out_features"], cfg["num_kernels"], gaussian, cfg["num_regions"],
                   cfg["lower_bounds"], cfg["upper_bounds"], cfg["dimension_ranges"], cfg["activation_idx"],
                   cfg["delta"])
    dp = device_params({"params": params}, cuda_device)
    rng = np.random.default_rng(1)
    tin = rng.random((5000, 8)).astype(np.float32)
    tout = rng.normal(size=(5000, 2)).astype(np.float32)
    table = ExplicitTable(tin, tout, cuda_device)
    axes, fixed = _frenet_lattice(21)                    # 441 points: an odd split
    q = lattice_queries(axes, fixed)
    whole = irbfn_vs_explicit(net, dp, table, q, rank=0, world=1)
    parts = [irbfn_vs_explicit(net, dp, table, q, rank=r, world=2) for r in range(2)]
    assert parts[0]["index"].shape[0] == 221 and parts[1]["index"].shape[0] == 220
    for k in ("index", "pred_explicit"):
        assert torch.equal(torch.cat([p[k] for p in parts]), whole[k]), k
    # small batches split the kernel range by batch size, so the summation order differs between 441 and 221 rows
    for k in ("pred_irbfn", "errors"):
        assert torch.allclose(torch.cat([p[k] for p in parts]), whole[k], rtol=0, atol=2e-6 * float(whole["pred_irbfn"].abs().max())), k
