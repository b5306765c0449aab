"""Table-side kernels on the GPU against oracle/table_oracle.py.

gather_rows / grid_index: bit-exact.  nearest_row: the returned index must be the float64 argmin
unless the float64 runner-up lies within float32 rounding of the winner (then either is accepted
and the returned row's float64 distance must match the minimum to 1e-6 relative); the returned
distance is within 1e-6 relative of the float64 distance."""
import numpy as np
import pytest
import torch

from irbfn_b200 import _lib, ops
from irbfn_b200.pipeline import DeviceTable
from oracle import table_oracle as tor

pytestmark = pytest.mark.gpu


def _dev(a, dev, dtype=np.float32):
    return torch.tensor(np.ascontiguousarray(a, dtype=dtype), device=dev)


@pytest.mark.parametrize("N,D,O,n", [(1000, 8, 10, 4096), (50000, 7, 2, 1), (17, 5, 3, 100), (300000, 8, 10, 262144)])
def test_gather_rows_is_exact(N, D, O, n, cuda_device):
    rng = np.random.default_rng(N + n)
    X, Y = rng.normal(size=(N, D)).astype(np.float32), rng.normal(size=(N, O)).astype(np.float32)
    perm = rng.integers(0, N, n)
    xr, yr = tor.gather(X, Y, perm)
    xb, yb = ops.gather_rows(_dev(X, cuda_device), _dev(Y, cuda_device), _dev(perm, cuda_device, np.int64), check=True)
    assert np.array_equal(xb.cpu().numpy(), xr) and np.array_equal(yb.cpu().numpy(), yr)
    only_x = ops.gather_rows(_dev(X, cuda_device), None, _dev(perm, cuda_device, np.int64))
    assert np.array_equal(only_x.cpu().numpy(), xr)


def test_gather_rows_flags_bad_indices_without_reading_them(cuda_device):
    X = torch.ones((10, 4), device=cuda_device)
    Y = torch.ones((10, 2), device=cuda_device)
    idx = torch.tensor([0, 10, -1, 9], device=cuda_device)
    with pytest.raises(IndexError):
        ops.gather_rows(X, Y, idx, check=True)
    xb, yb = ops.gather_rows(X, Y, idx)
    assert xb.cpu().tolist() == [[1.0] * 4, [0.0] * 4, [0.0] * 4, [1.0] * 4]
    assert ops.gather_rows(X, Y, idx[:0])[0].shape == (0, 4)


def test_device_table_epoch_matches_reference_batching(cuda_device):
    """Same permutation cut the reference's way; rank slices tile each batch; -999 rows dropped."""
    rng = np.random.default_rng(3)
    X, Y = rng.normal(size=(1000, 8)).astype(np.float32), rng.normal(size=(1000, 10)).astype(np.float32)
    Y[::7] = -999.0
    Y[1, 3] = -999.0                       # one marker only: the reference keeps this row
    Xv, Yv = tor.remove_infeasible(X, Y)
    tab = DeviceTable(X, Y, cuda_device, infeasible_value=-999)
    assert tab.num_rows == Xv.shape[0]
    assert np.array_equal(tab.x.cpu().numpy(), Xv) and np.array_equal(tab.y.cpu().numpy(), Yv)
    perm = tab.permutation(11).cpu().numpy()
    assert np.array_equal(np.sort(perm), np.arange(tab.num_rows))
    ref = tor.epoch_batches(perm, tab.num_rows, 96)
    got = [(x.clone(), y.clone()) for x, y in tab.batches(96, 11)]
    assert len(got) == ref.shape[0] == tab.num_rows // 96
    for s, (xb, yb) in enumerate(got):
        assert np.array_equal(xb.cpu().numpy(), Xv[ref[s]]) and np.array_equal(yb.cpu().numpy(), Yv[ref[s]])
    for s, parts in enumerate(zip(*[[(x.clone(), y.clone()) for x, y in tab.batches(96, 11, r, 3)] for r in range(3)])):
        assert torch.equal(torch.cat([p[0] for p in parts]), got[s][0])
        assert torch.equal(torch.cat([p[1] for p in parts]), got[s][1])


def test_train_epoch_on_device_table_runs_and_lowers_the_loss(cuda_device):
    import os
    from conftest import GOLDEN_DIR, load_golden
    from irbfn_b200.basis import gaussian
    from irbfn_b200.model import WCRBFNet, device_params
    from irbfn_b200.pipeline import train_epoch
    from irbfn_b200.train import TrainState
    from oracle import irbfn_oracle as orc
    cfg, params, arr = load_golden(os.path.join(GOLDEN_DIR, "c2b_frenet_1x1000_o10.npz"))
    net = WCRBFNet(cfg["in_features"], cfg["out_features"], cfg["num_kernels"], gaussian, cfg["num_regions"],
                   cfg["lower_bounds"], cfg["upper_bounds"], cfg["dimension_ranges"], cfg["activation_idx"],
                   cfg["delta"])
    x = arr["x"].astype(np.float32)
    y = (arr["y64"] + 0.05).astype(np.float32)          # targets a constant offset away from the network
    tab = DeviceTable(x, y, cuda_device)
    state = TrainState(net, device_params({"params": params}, cuda_device), lr=1e-3, device=cuda_device)
    first, _ = train_epoch(state, tab, 256, 0, loss="frenet_fullint", rank=0, world=1)
    for e in range(1, 6):
        last, per_batch = train_epoch(state, tab, 256, e, loss="frenet_fullint", rank=0, world=1)
    assert per_batch.shape == (4,)
    assert float(last) < float(first)


@pytest.mark.parametrize("B,N,D", [(1, 5000, 8), (400, 100000, 8), (5000, 3001, 7), (3, 2, 4), (70000, 513, 5)])
def test_nearest_row_matches_float64_argmin(B, N, D, cuda_device):
    rng = np.random.default_rng(B + N)
    table = rng.uniform(-2, 2, (N, D)).astype(np.float32)
    q = rng.uniform(-2, 2, (B, D)).astype(np.float32)
    Bc = min(B, 3000)                                    # the oracle is O(B N D) on the host
    idx_ref, d1, d2 = tor.nearest_row(q[:Bc], table)
    idx, dist = ops.nearest_row(_dev(q, cuda_device), _dev(table, cuda_device))
    idx, dist = idx.cpu().numpy(), dist.cpu().numpy()
    assert idx.min() >= 0 and idx.max() < N
    clear = (d2 - d1) > 1e-5 * np.maximum(d1, 1e-3)
    assert np.array_equal(idx[:Bc][clear], idx_ref[clear])
    got64 = np.linalg.norm(table[idx[:Bc]].astype(np.float64) - q[:Bc].astype(np.float64), axis=1)
    assert np.all(np.abs(got64 - d1) <= 1e-5 * np.maximum(d1, 1e-3))
    assert np.all(np.abs(dist[:Bc] - d1) <= 1e-5 * np.maximum(d1, 1e-3))


def test_nearest_row_ties_take_the_first_row_and_exact_hits_are_found(cuda_device):
    rng = np.random.default_rng(8)
    base = rng.integers(-5, 5, (600, 8)).astype(np.float32)        # integer lattice: fp32 distances are exact
    table = np.concatenate([base, base])                             # every row appears twice
    q = base[rng.integers(0, 600, 2000)] + 0.25
    idx_ref, _, _ = tor.nearest_row(q, table)
    idx, dist = ops.nearest_row(_dev(q, cuda_device), _dev(table, cuda_device))
    assert np.array_equal(idx.cpu().numpy(), idx_ref)               # argmin's first-minimum rule
    assert idx.max().item() < 600
    idx2, dist2 = ops.nearest_row(_dev(base, cuda_device), _dev(table, cuda_device))
    assert float(dist2.max()) == 0.0


def test_grid_index_matches_searchsorted(cuda_device):
    rng = np.random.default_rng(4)
    keys = [np.sort(rng.uniform(-3, 3, n)).astype(np.float32) for n in (20, 1, 7, 33, 5, 2, 11)]
    q = rng.uniform(-4, 4, (20000, 7)).astype(np.float32)
    q[:200, 0] = keys[0][rng.integers(0, 20, 200)]                   # exact key hits: side="right"
    q[5, 3] = np.nan
    ref = tor.grid_index(q, keys)
    gk = ops.GridKeys(keys, cuda_device)
    out = ops.grid_index(_dev(q, cuda_device), gk)
    assert out.dtype == torch.int32 and np.array_equal(out.cpu().numpy(), ref)


def test_grid_index_f64_is_index_exact_where_float32_keys_are_not(cuda_device):
    """irbfn_grid_index_f64: float64 keys and values, as explicit_planner.py:165-175 bisects them.  Keys that are not
    representable in float32 and values one float64 ulp either side of a key: the float64 entry point gives numpy's
    cell for every value; the float32 one (same inputs narrowed) provably does not on these."""
    rng = np.random.default_rng(9)
    keys = [np.sort(rng.uniform(-3, 3, n)) + 1e-9 for n in (20, 1, 7, 33)]        # float64, off the float32 lattice
    assert all((k.astype(np.float32).astype(np.float64) != k).all() for k in keys)
    q = rng.uniform(-4, 4, (20000, 4))
    for d, k in enumerate(keys):                                                   # at, just below, just above keys
        pick = k[rng.integers(0, len(k), 3000)]
        q[:3000, d] = pick
        q[3000:6000, d] = np.nextafter(pick, -np.inf)
        q[6000:9000, d] = np.nextafter(pick, np.inf)
    q[7, 2] = np.nan
    ref = tor.grid_index(q, keys)
    gk = ops.GridKeys(keys, cuda_device, dtype=torch.float64)
    out = ops.grid_index(torch.tensor(q, device=cuda_device), gk)
    assert out.dtype == torch.int32 and np.array_equal(out.cpu().numpy(), ref)
    out32 = ops.grid_index(_dev(q.astype(np.float32), cuda_device), ops.GridKeys(keys, cuda_device))
    assert not np.array_equal(out32.cpu().numpy(), ref)       # why the float64 entry point exists
    with pytest.raises(ValueError):                            # dtypes of keys and values must agree
        ops.grid_index(_dev(q.astype(np.float32), cuda_device), gk)
    assert ops.grid_index(torch.empty((0, 4), device=cuda_device, dtype=torch.float64), gk).shape == (0, 4)


def test_table_errors(cuda_device):
    q = torch.zeros((4, 8), device=cuda_device)
    with pytest.raises(ValueError):
        ops.nearest_row(q, torch.zeros((10, 7), device=cuda_device))
    with pytest.raises(ValueError):
        ops.nearest_row(q.cpu(), torch.zeros((10, 8), device=cuda_device))
    with pytest.raises(TypeError):
        ops.gather_rows(q, None, torch.zeros(3, device=cuda_device, dtype=torch.int32))
    with pytest.raises(ValueError):
        ops.grid_index(q, ops.GridKeys([[0.0, 1.0]], cuda_device))


def test_lattice_queries_match_the_numpy_meshgrid(cuda_device):
    """irbfn_lattice_queries against irbfn_b200.sweep.lattice_queries (NumPy, the notebook's spelling): bit-exact."""
    from irbfn_b200 import ops
    from irbfn_b200.sweep import lattice_queries
    rng = np.random.default_rng(0)
    for D, axes in ((8, {0: np.linspace(-0.2, 2.2, 7), 6: np.linspace(-1, 1, 5), 2: np.linspace(1, 7, 3)}),
                    (5, {4: rng.normal(size=11)}), (3, {}), (8, {c: rng.normal(size=n) for c, n in zip(range(8), (2, 3, 2, 5, 2, 3, 2, 2))})):
        axes = {c: np.asarray(v, np.float32) for c, v in axes.items()}
        fixed = rng.normal(size=D).astype(np.float32)
        ref = lattice_queries(axes, fixed)
        got = ops.lattice_queries(axes, fixed, cuda_device).cpu().numpy()
        assert got.shape == ref.shape and np.array_equal(got, ref)


def test_apply_lattice_into_a_pinned_host_buffer(cuda_device):
    """apply_lattice(out=pinned): the kernel stores y straight into host memory; same values as the device-resident
    evaluation of the same lattice (unsplit kernel: bit-identical), ragged last tile included."""
    import os
    from conftest import GOLDEN_DIR, load_golden
    from irbfn_b200 import ops
    from irbfn_b200.basis import gaussian
    from irbfn_b200.model import WCRBFNet, device_params
    cfg, params, _ = load_golden(os.path.join(GOLDEN_DIR, "c2_frenet_12x100.npz"))
    net = WCRBFNet(cfg["in_features"], cfg["out_features"], cfg["num_kernels"], gaussian, cfg["num_regions"],
                   cfg["lower_bounds"], cfg["upper_bounds"], cfg["dimension_ranges"], cfg["activation_idx"], cfg["delta"])
    dp = device_params({"params": params}, cuda_device)
    lo = [min(b) for b in cfg["lower_bounds"]]
    hi = [max(b) for b in cfg["upper_bounds"]]
    axes = {0: np.linspace(lo[0], hi[0], 37).astype(np.float32), 6: np.linspace(lo[6], hi[6], 31).astype(np.float32),
            2: np.linspace(lo[2], hi[2], 5).astype(np.float32)}
    fixed = np.zeros(8, np.float32)
    fixed[4] = 4.0
    B = 37 * 31 * 5
    yh = torch.full((B, 2), float("nan")).pin_memory()
    for _ in range(2):                                   # the flag / counter words are re-armed on every call
        x, out = net.apply_lattice(dp, axes, fixed, out=yh)
        torch.cuda.synchronize()
        rb, lin = dp["params"]["rbf_list"], dp["params"]["linear"]
        ref = ops.forward(x, rb["centers"], rb["log_sigs"], lin["kernel"], lin["bias"], gd=net._gate_pack(cuda_device),
                          basis=0, split=False)
        assert out is yh and torch.equal(yh, ref.cpu())
        yh.fill_(float("nan"))
