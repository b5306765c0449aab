"""GPU parity of the forward (irbfn_forward / irbfn_region_gate / irbfn_softmax_gate, called
through the C ABI via ctypes) against the oracle and the committed golden fixtures.

Tolerance (BASELINE.json north_star): outputs norm-wise
max|y - y_ref| / max|y_ref| <= 1e-5 against the float64 reference values.
"""
import os

import numpy as np
import pytest
import torch

from conftest import golden_paths, load_golden, rel_l2, rel_max
from oracle import irbfn_oracle as orc

pytestmark = pytest.mark.gpu

OUT_TOL = 1e-5


def _ops():
    from irbfn_b200 import ops
    return ops


def _t(a, dev):
    return torch.tensor(np.ascontiguousarray(a, dtype=np.float32), device=dev)


def _gd(cfg, dev):
    return _ops().GatePack(cfg["lower_bounds"], cfg["upper_bounds"], cfg["delta"],
                           cfg["dimension_ranges"], dev)


@pytest.mark.parametrize("path", golden_paths(), ids=[os.path.basename(p)[:-4] for p in golden_paths()])
def test_forward_matches_golden(path, cuda_device):
    ops = _ops()
    cfg, params, arr = load_golden(path)
    dev = cuda_device
    x = _t(arr["x"], dev)
    rb = params["rbf_list"]
    basis = orc.BASIS_IDS[cfg["basis_func"]]
    if cfg["kind"] == "wcrbf":
        y = ops.forward(x, _t(rb["centers"], dev), _t(rb["log_sigs"], dev), _t(params["linear"]["kernel"], dev),
                        _t(params["linear"]["bias"], dev), gd=_gd(cfg, dev), basis=basis)
        # the fixtures' few thousand queries take the latency kernel (irbfn_forward_path); asking for the hidden
        # activations as well keeps the throughput kernel on every fixture too
        y_t, _ = ops.forward(x, _t(rb["centers"], dev), _t(rb["log_sigs"], dev), _t(params["linear"]["kernel"], dev),
                             _t(params["linear"]["bias"], dev), gd=_gd(cfg, dev), basis=basis, want_hidden=True)
        assert rel_max(y_t.cpu().numpy(), arr["y64"]) < OUT_TOL
    elif cfg["kind"] == "cluster":
        gate, logits = ops.softmax_gate(x, _t(params["cluster"]["kernel"], dev), _t(params["cluster"]["bias"], dev))
        assert rel_max(logits.cpu().numpy(), arr["logits64"]) < OUT_TOL
        y = ops.forward(x, _t(rb["centers"], dev), _t(rb["log_sigs"], dev), _t(params["linear"]["kernel"], dev),
                        _t(params["linear"]["bias"], dev), gate=gate, basis=basis)
    else:  # deeper: hidden activations from the kernel, 64-wide MLP tail in plain fp32 matmuls
        h = ops.forward(x, _t(rb["centers"], dev), _t(rb["log_sigs"], dev), gd=_gd(cfg, dev), basis=basis,
                        want_y=False, want_hidden=True)
        # the kernel's own output: hidden activations against the float64 oracle
        gamma = orc.region_activation(arr["x"].astype(np.float64), cfg["num_regions"], cfg["lower_bounds"],
                                      cfg["upper_bounds"], cfg["delta"], cfg["dimension_ranges"])
        K = rb["centers"].shape[1]
        _, h_ref = orc.forward(arr["x"].astype(np.float64), rb["centers"], rb["log_sigs"], np.eye(K), np.zeros(K),
                               gamma, basis, return_hidden=True)
        assert rel_max(h.cpu().numpy(), h_ref) < OUT_TOL
        p = {k: {kk: _t(vv, dev) for kk, vv in v.items()} for k, v in params.items() if k != "rbf_list"}
        h = torch.relu(h @ p["linear_pre1"]["kernel"] + p["linear_pre1"]["bias"])
        h = torch.relu(h @ p["linear_pre2"]["kernel"] + p["linear_pre2"]["bias"])
        y = h @ p["linear"]["kernel"] + p["linear"]["bias"]
        # The shipped 64-wide tail amplifies a 1e-6 perturbation of h to ~2e-3 of y (weights up to
        # 437), and float32 NumPy evaluation of the whole net is itself 1.5e-5 from float64.  The
        # tail is plain fp32 matmuls outside the kernel, so the composite gets the adjudication
        # bound of SURVEY.md 8c (2x the float32 reference's own error, rounded up).
        assert rel_max(y.cpu().numpy(), arr["y64"]) < 1e-4
        return
    assert rel_max(y.cpu().numpy(), arr["y64"]) < OUT_TOL


@pytest.mark.parametrize("path", [p for p in golden_paths() if "deeper" not in p and "cluster" not in p][:4])
def test_region_gate_matches_oracle(path, cuda_device):
    ops = _ops()
    cfg, _, arr = load_golden(path)
    gam = ops.region_gate(_t(arr["x"], cuda_device), _gd(cfg, cuda_device), cfg["num_regions"]).cpu().numpy()
    ref = orc.region_activation(arr["x"].astype(np.float64), cfg["num_regions"], cfg["lower_bounds"],
                                cfg["upper_bounds"], cfg["delta"], cfg["dimension_ranges"])
    assert np.abs(gam - ref).max() < 1e-5


def _synthetic(seed, B, R, K, D, O, S=2, nb=(3, 2), num_ranges=None):
    rng = np.random.default_rng(seed)
    S = min(S, D)
    lbs, ubs = [], []
    for d in range(S):
        edges = np.linspace(-1.2, 1.2, nb[d % len(nb)] + 1)
        lbs.append([float(v - 0.05) for v in edges[:-1]])
        ubs.append([float(v + 0.05) for v in edges[1:]])
    import itertools
    combos = list(itertools.product(*[range(len(l)) for l in lbs]))
    ranges = [list(combos[i % len(combos)]) for i in range(R if num_ranges is None else num_ranges)]
    cfg = dict(num_regions=R, lower_bounds=lbs, upper_bounds=ubs, delta=[5.0 + 3 * d for d in range(S)],
               dimension_ranges=ranges, in_features=D, out_features=O, num_kernels=K)
    p = {"centers": rng.uniform(-1.2, 1.2, (R, K, D)).astype(np.float32),
         "log_sigs": rng.uniform(-0.7, 0.7, (R, K)).astype(np.float32),
         "kernel": rng.normal(size=(K, O)).astype(np.float32),
         "bias": rng.normal(size=O).astype(np.float32)}
    x = rng.uniform(-1.3, 1.3, (B, D)).astype(np.float32)
    return cfg, p, x


SHAPES = [
    # B, R, K, D, O
    (1, 1, 1, 1, 1), (3, 2, 5, 3, 3), (1023, 3, 7, 4, 5), (1025, 6, 10, 5, 7), (5000, 4, 33, 6, 9),
    (257, 2, 100, 7, 10), (2048, 12, 100, 8, 2), (300, 1, 1000, 8, 16), (200, 3, 9, 8, 18),
    (129, 2, 6, 2, 40),
]


@pytest.mark.parametrize("basis", [0, 1])
@pytest.mark.parametrize("shape", SHAPES, ids=["B%d_R%d_K%d_D%d_O%d" % s for s in SHAPES])
def test_forward_ragged_shapes_match_oracle(shape, basis, cuda_device):
    ops = _ops()
    B, R, K, D, O = shape
    cfg, p, x = _synthetic(11, B, R, K, D, O)
    dev = cuda_device
    gamma = orc.region_activation(x.astype(np.float64), R, cfg["lower_bounds"], cfg["upper_bounds"],
                                  cfg["delta"], cfg["dimension_ranges"])
    y_ref, h_ref = orc.forward(x, p["centers"], p["log_sigs"], p["kernel"], p["bias"], gamma, basis,
                               return_hidden=True)
    y, h = ops.forward(_t(x, dev), _t(p["centers"], dev), _t(p["log_sigs"], dev), _t(p["kernel"], dev),
                       _t(p["bias"], dev), gd=_gd(cfg, dev), basis=basis, want_hidden=True)
    assert rel_max(y.cpu().numpy(), y_ref) < OUT_TOL
    assert rel_max(h.cpu().numpy(), h_ref) < OUT_TOL
    # y alone: these batches take the latency kernel where the output window allows (O <= 16)
    assert ops.forward_path(B, R, K, D, O) == ("small" if O <= 16 else ops.forward_path(B, R, K, D, O, False))
    y1 = ops.forward(_t(x, dev), _t(p["centers"], dev), _t(p["log_sigs"], dev), _t(p["kernel"], dev),
                     _t(p["bias"], dev), gd=_gd(cfg, dev), basis=basis)
    assert rel_max(y1.cpu().numpy(), y_ref) < OUT_TOL
    # the same through an externally supplied gate
    y2 = ops.forward(_t(x, dev), _t(p["centers"], dev), _t(p["log_sigs"], dev), _t(p["kernel"], dev),
                     _t(p["bias"], dev), gate=_t(gamma, dev), basis=basis)
    assert rel_max(y2.cpu().numpy(), y_ref) < OUT_TOL


def test_forward_parameters_larger_than_shared_memory(cuda_device):
    """64 x 4096 kernels (the C5 stress shape, 7.6 MB of parameters) are streamed through shared
    memory in (region, kernel) chunks."""
    ops = _ops()
    cfg, p, x = _synthetic(5, 96, 64, 4096, 6, 16, S=3, nb=(4, 4, 4))
    dev = cuda_device
    gamma = orc.region_activation(x.astype(np.float64), 64, cfg["lower_bounds"], cfg["upper_bounds"],
                                  cfg["delta"], cfg["dimension_ranges"])
    y_ref = orc.forward(x, p["centers"], p["log_sigs"], p["kernel"], p["bias"], gamma, 0, chunk=8)
    y = ops.forward(_t(x, dev), _t(p["centers"], dev), _t(p["log_sigs"], dev), _t(p["kernel"], dev),
                    _t(p["bias"], dev), gd=_gd(cfg, dev), basis=0)
    assert rel_max(y.cpu().numpy(), y_ref) < OUT_TOL


def test_fewer_ranges_than_regions(cuda_device):
    ops = _ops()
    cfg, p, x = _synthetic(2, 100, 8, 6, 4, 3, num_ranges=5)
    dev = cuda_device
    gamma = orc.region_activation(x.astype(np.float64), 8, cfg["lower_bounds"], cfg["upper_bounds"],
                                  cfg["delta"], cfg["dimension_ranges"])
    assert np.all(gamma[:, 5:] == 0)
    y_ref = orc.forward(x, p["centers"], p["log_sigs"], p["kernel"], p["bias"], gamma, 0)
    y = ops.forward(_t(x, dev), _t(p["centers"], dev), _t(p["log_sigs"], dev), _t(p["kernel"], dev),
                    _t(p["bias"], dev), gd=_gd(cfg, dev), basis=0)
    assert rel_max(y.cpu().numpy(), y_ref) < OUT_TOL


def test_empty_batch_and_errors(cuda_device):
    ops = _ops()
    from irbfn_b200._lib import IrbfnError
    cfg, p, x = _synthetic(2, 4, 2, 3, 4, 2)
    dev = cuda_device
    args = (_t(p["centers"], dev), _t(p["log_sigs"], dev), _t(p["kernel"], dev), _t(p["bias"], dev))
    y = ops.forward(_t(x[:0], dev), *args, gd=_gd(cfg, dev))
    assert tuple(y.shape) == (0, 2)
    with pytest.raises(IrbfnError) as e:
        ops.forward(_t(x, dev), *args, gd=_gd(cfg, dev), basis=5)
    assert e.value.code == -3
    with pytest.raises(IrbfnError):
        ops.forward(_t(x, dev), *args)  # no gate at all
    with pytest.raises(ValueError):
        ops.forward(torch.tensor(x), *args, gd=_gd(cfg, dev))  # CPU tensor: no fallback
    big = torch.zeros((4, 9), device=dev)
    with pytest.raises(IrbfnError) as e:
        ops.forward(big, torch.zeros((2, 3, 9), device=dev), args[1], args[2], args[3], gate=torch.ones((4, 2), device=dev))
    assert e.value.code == -2


def test_forward_is_linear_in_the_readout(cuda_device):
    """Size-independent property at the full C2 size (2^20 queries): y(W1 + W2, b1 + b2) =
    y(W1, b1) + y(W2, b2), and doubling the gate doubles h."""
    ops = _ops()
    cfg, params, arr = load_golden([p for p in golden_paths() if "c2_frenet_12x100" in p][0])
    dev = cuda_device
    rb = params["rbf_list"]
    B = 1 << 20
    x = _t(orc.uniform_queries(cfg, B, seed=1), dev)
    c, ls = _t(rb["centers"], dev), _t(rb["log_sigs"], dev)
    W1, b1 = _t(params["linear"]["kernel"], dev), _t(params["linear"]["bias"], dev)
    g = torch.Generator(device="cpu").manual_seed(0)
    W2 = torch.randn(W1.shape, generator=g).to(dev)
    b2 = torch.randn(b1.shape, generator=g).to(dev)
    gd = _gd(cfg, dev)
    y1 = ops.forward(x, c, ls, W1, b1, gd=gd)
    y2 = ops.forward(x, c, ls, W2, b2, gd=gd)
    y12 = ops.forward(x, c, ls, W1 + W2, b1 + b2, gd=gd)
    scale = y12.abs().max().item()
    assert (y12 - (y1 + y2)).abs().max().item() / scale < 1e-5
    # a subsample against the oracle
    idx = np.arange(0, B, 4099)
    y_ref = orc.wcrbf_forward(x[idx].cpu().numpy().astype(np.float64), params, cfg)
    assert rel_max(y1[idx].cpu().numpy(), y_ref) < OUT_TOL
    # determinism: bit-identical across launches
    assert torch.equal(y1, ops.forward(x, c, ls, W1, b1, gd=gd))


@pytest.mark.parametrize("B", [1, 400, 4096, 40000])
def test_small_batch_split_over_kernels_matches_unsplit_and_oracle(B, cuda_device):
    """Batches that cannot give every SM a query tile split the kernels over gridDim.y and reduce
    the partial readouts in a fixed order (workspace-backed).  Same numbers as the unsplit launch
    (which runs when no workspace is passed), both within tolerance of the oracle; hidden
    activations come out right as well."""
    import ctypes as C
    from irbfn_b200 import _lib
    ops = _ops()
    cfg, p, x = _synthetic(41, B, 3, 250, 7, 10)
    dev = cuda_device
    gamma = orc.region_activation(x.astype(np.float64), 3, cfg["lower_bounds"], cfg["upper_bounds"],
                                  cfg["delta"], cfg["dimension_ranges"])
    y_ref, h_ref = orc.forward(x, p["centers"], p["log_sigs"], p["kernel"], p["bias"], gamma, 0, return_hidden=True)
    xs, c, ls, W, b = (_t(a, dev) for a in (x, p["centers"], p["log_sigs"], p["kernel"], p["bias"]))
    gd = _gd(cfg, dev)
    lib = _lib.load()
    need = lib.irbfn_forward_workspace_bytes(B, 3, 250, 7, 10, 1)
    assert (need > 0) == (B <= 40000 and B * 2 <= 148 * 1024)
    # (the hidden activations are asked for as well: y alone would take the latency kernel at the smaller sizes)
    assert ops.forward_path(B, 3, 250, 7, 10, y_only=False) == ("split" if need > 0 else "tiled")
    y, h = ops.forward(xs, c, ls, W, b, gd=gd, basis=0, want_hidden=True)
    y0, h0 = torch.empty_like(y), torch.empty_like(h)
    rc = lib.irbfn_forward(C.c_void_p(xs.data_ptr()), C.c_void_p(c.data_ptr()), C.c_void_p(ls.data_ptr()),
                           C.c_void_p(W.data_ptr()), C.c_void_p(b.data_ptr()), None, gd.ref(),
                           C.c_void_p(y0.data_ptr()), C.c_void_p(h0.data_ptr()), B, 3, 250, 7, 10, 0, None, 0,
                           C.c_void_p(torch.cuda.current_stream().cuda_stream))
    assert rc == 0
    assert rel_max(y.cpu().numpy(), y_ref) < OUT_TOL
    assert rel_max(y0.cpu().numpy(), y_ref) < OUT_TOL
    assert rel_max(h.cpu().numpy(), h_ref) < OUT_TOL
    assert rel_max(h0.cpu().numpy(), h_ref) < OUT_TOL
    y2, _ = ops.forward(xs, c, ls, W, b, gd=gd, basis=0, want_hidden=True)
    assert torch.equal(y, y2)          # fixed-order reduction: bit-identical run to run


SMALL_SHAPES = [
    # B, R, K, D, O: one query (pred_step), CTA-count edges (148 / 149: one vs four queries per CTA), K below and
    # above the CTA's 256 threads, K = 1, many regions, every output window width
    (1, 12, 100, 8, 2), (1, 1, 1000, 7, 2), (2, 4, 100, 5, 10), (5, 3, 1, 4, 1), (147, 2, 257, 6, 3),
    (148, 12, 100, 8, 2), (149, 12, 100, 8, 2), (400, 1, 1000, 8, 10), (593, 5, 64, 3, 16), (1001, 7, 33, 8, 7),
    (3, 500, 10, 8, 10), (4099, 12, 100, 8, 2), (7, 16, 1024, 6, 16),
]


@pytest.mark.parametrize("basis", [0, 1])
@pytest.mark.parametrize("shape", SMALL_SHAPES, ids=["B%d_R%d_K%d_D%d_O%d" % s for s in SMALL_SHAPES])
def test_latency_kernel_matches_oracle_and_the_throughput_kernel(shape, basis, cuda_device):
    """The pair-parallel small-batch kernel (irbfn_fwd_small.cu; pred_step, irbfn_planner.py:29-32): against the
    float64 oracle with the tanh gate and with an external gate, against the throughput kernel on the same inputs
    (taken when the hidden activations are requested), bit-identical run to run, and it needs no workspace."""
    import ctypes as C
    from irbfn_b200 import _lib
    ops = _ops()
    B, R, K, D, O = shape
    nb = (3, 2, 2) if R <= 12 else (8, 8, 8)
    cfg, p, x = _synthetic(23, B, R, K, D, O, S=3, nb=nb)
    dev = cuda_device
    assert ops.forward_path(B, R, K, D, O) == "small"
    gamma = orc.region_activation(x.astype(np.float64), R, cfg["lower_bounds"], cfg["upper_bounds"],
                                  cfg["delta"], cfg["dimension_ranges"])
    y_ref = orc.forward(x, p["centers"], p["log_sigs"], p["kernel"], p["bias"], gamma, basis)
    xs, c, ls, W, b = (_t(a, dev) for a in (x, p["centers"], p["log_sigs"], p["kernel"], p["bias"]))
    gd = _gd(cfg, dev)
    y = ops.forward(xs, c, ls, W, b, gd=gd, basis=basis)
    assert rel_max(y.cpu().numpy(), y_ref) < OUT_TOL
    assert torch.equal(y, ops.forward(xs, c, ls, W, b, gd=gd, basis=basis))
    y_t, _ = ops.forward(xs, c, ls, W, b, gd=gd, basis=basis, want_hidden=True)
    assert rel_max(y.cpu().numpy(), y_t.cpu().numpy().astype(np.float64)) < OUT_TOL
    y_g = ops.forward(xs, c, ls, W, b, gate=_t(gamma, dev), basis=basis)
    assert rel_max(y_g.cpu().numpy(), y_ref) < OUT_TOL
    # through the C ABI with no workspace at all, output bracketed by canaries
    buf = torch.full((B + 2, O), 7.0, device=dev)
    rc = _lib.load().irbfn_forward(C.c_void_p(xs.data_ptr()), C.c_void_p(c.data_ptr()), C.c_void_p(ls.data_ptr()),
                                   C.c_void_p(W.data_ptr()), C.c_void_p(b.data_ptr()), None, gd.ref(),
                                   C.c_void_p(buf[1:].data_ptr()), None, B, R, K, D, O, basis, None, 0,
                                   C.c_void_p(torch.cuda.current_stream().cuda_stream))
    assert rc == 0
    torch.cuda.synchronize()
    assert torch.equal(buf[1:B + 1], y) and bool((buf[0] == 7.0).all()) and bool((buf[B + 1] == 7.0).all())


def test_host_pipeline_variants_agree(cuda_device):
    """apply_host: default pipeline, two upload streams, and the CUDA-graph replay of the whole chunk
    pipeline (twice: capture, then replay with new input VALUES in the same pinned buffers) all give the
    device-resident result."""
    import os
    from conftest import GOLDEN_DIR
    from irbfn_b200.basis import gaussian
    from irbfn_b200.model import WCRBFNet, device_params
    cfg, params, arr = load_golden(os.path.join(GOLDEN_DIR, "c2_frenet_12x100.npz"))
    net = WCRBFNet(cfg["in_features"], cfg["out_features"], cfg["num_kernels"], gaussian, cfg["num_regions"],
                   cfg["lower_bounds"], cfg["upper_bounds"], cfg["dimension_ranges"], cfg["activation_idx"],
                   cfg["delta"])
    dp = device_params({"params": params}, cuda_device)
    x = np.tile(arr["x"], (5, 1))[:5000].astype(np.float32)
    xh = torch.from_numpy(x.copy()).pin_memory()
    yh = torch.empty((5000, 2)).pin_memory()
    ref = net.apply(dp, torch.from_numpy(x).to(cuda_device)).cpu()
    first = None
    for kw in (dict(streamed=False), dict(), dict(streamed=True), dict(copy_streams=2), dict(graph=True), dict(graph=True),
               dict(run_streams=2), dict(run_streams=3, graph=True), dict(run_streams=3, graph=True)):
        yh.zero_()
        net.apply_host(dp, xh, out=yh, chunk=2048, **kw)
        torch.cuda.synchronize()
        # 2,048-query chunks take the latency kernel, the streamed launch and several chunk kernels in flight the
        # tiled one, the 5,000-query call behind `ref` the split one: three summation orders over the kernels
        assert torch.allclose(yh, ref, rtol=0, atol=2e-6 * float(ref.abs().max())), kw
        if first is None:
            first = yh.clone()
        elif kw.get("run_streams", 1) == 1 and not kw.get("streamed"):
            assert torch.equal(yh, first), kw         # the chunk pipeline's variants agree to the bit
    xh.mul_(0.5)                                      # new values, same buffers: the graph must pick them up
    ref2 = net.apply(dp, xh.to(cuda_device)).cpu()
    net.apply_host(dp, xh, out=yh, chunk=2048, graph=True)
    torch.cuda.synchronize()
    assert torch.allclose(yh, ref2, rtol=0, atol=2e-6 * float(ref2.abs().max()))
    assert np.array_equal(net.apply_host(dp, x * 0.5, chunk=2048), yh.numpy())


def test_pred_step_sized_host_calls_replay_one_graph(cuda_device):
    """module.apply on a small NumPy batch (the planner's pred_step, irbfn_planner.py:29-32; float64 states): upload,
    latency kernel and download captured once per batch size (hostpath.PredStepPipeline).  First calls (eager, capture),
    replays with new values, parameters updated in place, and a batch beyond the staging buffers (chunk pipeline)."""
    import os
    from conftest import GOLDEN_DIR
    from irbfn_b200.basis import gaussian
    from irbfn_b200.hostpath import PredStepPipeline
    from irbfn_b200.model import WCRBFNet, device_params
    cfg, params, arr = load_golden(os.path.join(GOLDEN_DIR, "c2_frenet_12x100.npz"))
    net = WCRBFNet(cfg["in_features"], cfg["out_features"], cfg["num_kernels"], gaussian, cfg["num_regions"],
                   cfg["lower_bounds"], cfg["upper_bounds"], cfg["dimension_ranges"], cfg["activation_idx"],
                   cfg["delta"])
    dp = device_params({"params": params}, cuda_device)
    x_all = np.tile(arr["x"].astype(np.float64), (5, 1))
    ref = lambda xx, pp: orc.wcrbf_forward(xx.astype(np.float32).astype(np.float64), pp, cfg)
    for B in (1, 400, 4096, 1):
        for rep in range(4):                                   # eager, eager + capture, replay, replay -- new values every time
            x = x_all[rep * 7: rep * 7 + B] * (1.0 - 0.05 * rep)
            y = net.apply(dp, x)
            assert isinstance(y, np.ndarray) and y.shape == (B, cfg["out_features"]) and y.dtype == np.float32
            assert rel_max(y, ref(x, params)) < OUT_TOL, (B, rep)
    pipe = net._const[("pred", str(torch.device(cuda_device)))]
    assert isinstance(pipe, PredStepPipeline) and len(pipe._graphs) == 3
    # parameters updated in place (same buffers): the captured launch reads the new values
    dp["params"]["linear"]["bias"].add_(0.25)
    y = net.apply(dp, x_all[:400])
    p2 = {k: dict(v) for k, v in params.items()}
    p2["linear"]["bias"] = np.asarray(params["linear"]["bias"]) + 0.25
    assert rel_max(y, ref(x_all[:400], p2)) < OUT_TOL
    dp["params"]["linear"]["bias"].sub_(0.25)
    # one row more than the staging buffers hold: the chunk pipeline
    y = net.apply(dp, x_all[:4097])
    assert y.shape == (4097, cfg["out_features"]) and rel_max(y, ref(x_all[:4097], params)) < OUT_TOL
    assert len(pipe._graphs) == 3


def test_streamed_host_path_matches_the_device_resident_forward(cuda_device):
    """irbfn_forward_streamed through apply_host: ONE launch whose query tiles wait for their chunk of the upload and
    which stores y straight into pinned host memory.  Bit-identical to the unsplit device-resident forward; calling it
    again with new values in the same buffers (flags cleared and re-armed) and with a pageable NumPy input."""
    import os
    from conftest import GOLDEN_DIR
    from irbfn_b200.basis import gaussian, inverse_quadratic
    from irbfn_b200.model import WCRBFNet, device_params
    ops = _ops()
    for name, B, chunk in (("c2_frenet_12x100.npz", 300 * 1024 + 333, 65536), ("dnmpc_4regions_iq.npz", 70001, 4096),
                           ("c2b_frenet_1x1000_o10.npz", 33000, 1024)):
        cfg, params, arr = load_golden(os.path.join(GOLDEN_DIR, name))
        bf = gaussian if cfg["basis_func"] == "gaussian" else inverse_quadratic
        net = WCRBFNet(cfg["in_features"], cfg["out_features"], cfg["num_kernels"], bf, cfg["num_regions"],
                       cfg["lower_bounds"], cfg["upper_bounds"], cfg["dimension_ranges"], cfg["activation_idx"],
                       cfg["delta"])
        dp = device_params({"params": params}, cuda_device)
        x = _big_queries(arr, B, cuda_device, 3).cpu()
        xh = x.clone().pin_memory()
        yh = torch.full((B, cfg["out_features"]), float("nan")).pin_memory()
        rb, lin = dp["params"]["rbf_list"], dp["params"]["linear"]
        # (hidden activations requested: the tiled kernel also for the 5,001-query call below)
        dense = lambda xx: ops.forward(xx.to(cuda_device), rb["centers"], rb["log_sigs"], lin["kernel"], lin["bias"],
                                       gd=_gd(cfg, cuda_device), basis=orc.BASIS_IDS[cfg["basis_func"]], split=False,
                                       want_hidden=True)[0].cpu()
        net.apply_host(dp, xh, out=yh, chunk=chunk, streamed=True)
        torch.cuda.synchronize()
        assert torch.equal(yh, dense(x)), name
        xh.mul_(0.7)
        yh.fill_(float("nan"))
        net.apply_host(dp, xh, out=yh, chunk=chunk, streamed=True)
        torch.cuda.synchronize()
        assert torch.equal(yh, dense(xh)), name
        y_np = net.apply_host(dp, x.numpy()[:5001], chunk=chunk, streamed=True)
        assert np.array_equal(y_np, dense(x[:5001]).numpy()), name
        if B >= (1 << 18):      # "auto" times both pipelines once on these buffers and keeps the faster; same values
            for _ in range(2):
                yh.fill_(float("nan"))
                net.apply_host(dp, xh, out=yh, streamed="auto")
                torch.cuda.synchronize()
                # (the chunk pipeline's ragged last chunk takes the split-over-kernels path: rounding, not bit equality)
                assert rel_max(yh.numpy(), dense(xh).numpy().astype(np.float64)) < 2e-6, name
            assert ("auto", str(torch.device("cuda", torch.cuda.current_device())), B) in net._const


def test_streamed_forward_gives_up_when_an_upload_never_arrives(cuda_device):
    """irbfn_forward_streamed whose second piece is never flagged: the launch ends within about one wait budget (not
    one per tile) with the status word raised, a launch that finds the word still raised gives up at once, and after
    clearing it the same buffers work again (include/irbfn_b200.h, `status`)."""
    import os
    import time
    from conftest import GOLDEN_DIR
    ops = _ops()
    dev = cuda_device
    cfg, params, arr = load_golden(os.path.join(GOLDEN_DIR, "c2_frenet_12x100.npz"))
    rb, lin = params["rbf_list"], params["linear"]
    c, ls = _t(rb["centers"], dev), _t(rb["log_sigs"], dev)
    W, b = _t(lin["kernel"], dev), _t(lin["bias"], dev)
    gd = _gd(cfg, dev)
    basis = orc.BASIS_IDS[cfg["basis_func"]]
    B, piece = 600 * 1024, 300 * 1024               # 600 tiles on 148 CTAs: several waits per CTA
    x = _big_queries(arr, B, dev, 11)
    y = torch.zeros((B, cfg["out_features"]), device=dev)
    status = torch.zeros(1, dtype=torch.int32, device=dev)
    ready = torch.zeros(3, dtype=torch.int32, device=dev)
    ready[0] = 1                                     # piece 0 "arrived", piece 1 never does
    for _ in range(2):                               # the second launch finds the status word still raised
        ready[2] = 0
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        ops.forward_streamed(x, c, ls, W, b, y, ready, piece, gd=gd, basis=basis, status=status)
        torch.cuda.synchronize()
        assert time.perf_counter() - t0 < 5.0
        assert int(status.item()) == 1
    status.zero_()
    ready[:2] = 1
    ready[2] = 0
    ops.forward_streamed(x, c, ls, W, b, y, ready, piece, gd=gd, basis=basis, status=status)
    torch.cuda.synchronize()
    assert int(status.item()) == 0
    assert torch.equal(y, ops.forward(x, c, ls, W, b, gd=gd, basis=basis, split=False))


def test_forward_also_writes_the_gate_for_the_backward(cuda_device):
    """irbfn_forward_gate: same y as irbfn_forward, and the gate it writes is the tanh gate (model.py:70-95);
    the backward fed with it as an external gate gives the parameter gradients of the descriptor form."""
    ops = _ops()
    dev = cuda_device
    for path in [p for p in golden_paths() if "c2_frenet" in p or "warmstart_27" in p or "c1b" in p]:
        cfg, params, arr = load_golden(path)
        rb, lin = params["rbf_list"], params["linear"]
        basis = orc.BASIS_IDS[cfg["basis_func"]]
        x = _t(arr["x"], dev)
        args = (x, _t(rb["centers"], dev), _t(rb["log_sigs"], dev), _t(lin["kernel"], dev), _t(lin["bias"], dev))
        gd = _gd(cfg, dev)
        y0, _ = ops.forward(*args, gd=gd, basis=basis, want_hidden=True)     # the same (tiled / split) kernel family
        gate = torch.full((x.shape[0], cfg["num_regions"]), float("nan"), device=dev)
        y1 = ops.forward(*args, gd=gd, basis=basis, gate_out=gate)
        assert torch.equal(y0, y1)
        assert rel_max(ops.forward(*args, gd=gd, basis=basis).cpu().numpy(), y1.cpu().numpy().astype(np.float64)) < OUT_TOL
        ref = orc.region_activation(arr["x"].astype(np.float64), cfg["num_regions"], cfg["lower_bounds"],
                                    cfg["upper_bounds"], cfg["delta"], cfg["dimension_ranges"])
        assert np.abs(gate.cpu().numpy() - ref).max() <= 5e-6
        dy = _t(np.random.default_rng(1).normal(size=arr["y64"].shape), dev)
        ga = ops.backward(*args[:4], dy, gd=gd, basis=basis, grad_mask=ops.GRAD_PARAMS)
        gb = ops.backward(*args[:4], dy, gate=gate, basis=basis, grad_mask=ops.GRAD_PARAMS)
        for k in ("centers", "log_sigs", "kernel", "bias"):
            assert rel_l2(gb[k].cpu().numpy(), ga[k].cpu().numpy()) < 1e-5, k


def test_gaussian_wide_and_wider_run_on_the_gaussian_kernel(cuda_device):
    """flax_rbf.py:40-47: exp(-0.1 a^2), exp(-0.01 a^2) -- the gaussian kernel with shifted log-widths; checked
    against the formula itself (float64), forward and log_sig gradient."""
    from irbfn_b200 import basis as pbasis
    from irbfn_b200.model import WCRBFNet, device_params
    cfg, params, arr = load_golden([p for p in golden_paths() if "c2_frenet" in p][0])
    x = arr["x"][:512].astype(np.float64)
    rb, lin = params["rbf_list"], params["linear"]
    gamma = orc.region_activation(x, cfg["num_regions"], cfg["lower_bounds"], cfg["upper_bounds"], cfg["delta"],
                                  cfg["dimension_ranges"])
    c, ls = rb["centers"].astype(np.float64), rb["log_sigs"].astype(np.float64)
    alpha2 = ((x[:, None, None, :] - c[None]) ** 2).sum(-1) / np.exp(ls)[None] ** 2
    for fn, coef in ((pbasis.gaussian_wide, 0.1), (pbasis.gaussian_wider, 0.01)):
        phi = np.exp(-coef * alpha2)
        ref = (gamma[:, :, None] * phi).sum(1) @ lin["kernel"].astype(np.float64) + lin["bias"].astype(np.float64)
        net = WCRBFNet(cfg["in_features"], cfg["out_features"], cfg["num_kernels"], fn, cfg["num_regions"],
                       cfg["lower_bounds"], cfg["upper_bounds"], cfg["dimension_ranges"], cfg["activation_idx"], cfg["delta"])
        p = device_params({"params": params}, cuda_device)
        p["params"]["rbf_list"]["log_sigs"].requires_grad_(True)
        y = net.apply(p, torch.tensor(arr["x"][:512], device=cuda_device))
        assert rel_max(y.detach().cpu().numpy(), ref) < 1e-5
        (g,) = torch.autograd.grad((y ** 2).mean(), [p["params"]["rbf_list"]["log_sigs"]])
        # d/d log_sig of exp(-coef a^2) = 2 coef a^2 phi
        dh = (2 * ref / ref.size) @ lin["kernel"].astype(np.float64).T
        gref = (dh[:, None, :] * gamma[:, :, None] * (2 * coef * alpha2 * phi)).sum(0)
        assert rel_l2(g.cpu().numpy(), gref) < 1e-4


# ---------------------------------------------------------------------------------------------------
# Large batches: the persistent kernel walks several 1,024-query tiles per CTA.  A query's result must not depend on
# which tile / CTA / call it was evaluated in: one call over the whole batch must agree BIT FOR BIT with the same batch
# evaluated in pieces of 64 tiles (split=False: the same unsplit kernel, so the same order of additions), within
# rounding with the pieces' default split-over-kernels path, and with the float64 oracle on a sample that includes the
# ragged tail.
# ---------------------------------------------------------------------------------------------------
def _big_queries(arr, n, dev, seed):
    """n queries spread over (and a little beyond) the fixture's own input range."""
    x0 = arr["x"].astype(np.float32)
    lo, hi = x0.min(0), x0.max(0)
    g = torch.Generator(device="cpu").manual_seed(seed)
    u = torch.rand((n, x0.shape[1]), generator=g, dtype=torch.float32)
    x = torch.tensor(lo) + u * torch.tensor((hi - lo) * 1.1 + 1e-3) - torch.tensor((hi - lo) * 0.05)
    return x.to(dev).contiguous()


@pytest.mark.parametrize("name", ["c2_frenet_12x100", "c1_cartesian_1x1000", "dnmpc_4regions_iq",
                                  "c2b_frenet_1x1000_o10", "warmstart_1x500_o10", "cluster_500x10",
                                  "dnmpc_9regions_multiv"])
def test_large_batch_equals_piecewise_evaluation_and_oracle(name, cuda_device):
    ops = _ops()
    path = [p for p in golden_paths() if os.path.basename(p)[:-4] == name][0]
    cfg, params, arr = load_golden(path)
    dev = cuda_device
    sms = torch.cuda.get_device_properties(dev).multi_processor_count
    n = 2 * sms * 1024 + 777                       # ragged last tile, two tiles and a bit per SM
    x = _big_queries(arr, n, dev, 11)
    rb = params["rbf_list"]
    basis = orc.BASIS_IDS[cfg["basis_func"]]
    args = (_t(rb["centers"], dev), _t(rb["log_sigs"], dev), _t(params["linear"]["kernel"], dev),
            _t(params["linear"]["bias"], dev))
    if cfg["kind"] == "cluster":
        gate, _ = ops.softmax_gate(x, _t(params["cluster"]["kernel"], dev), _t(params["cluster"]["bias"], dev))
        run = lambda lo, hi, **kw: ops.forward(x[lo:hi], *args, gate=gate[lo:hi].contiguous(), basis=basis, **kw)
    else:
        gd = _gd(cfg, dev)
        run = lambda lo, hi, **kw: ops.forward(x[lo:hi], *args, gd=gd, basis=basis, **kw)
    y = run(0, n, split=False)
    assert rel_max(run(0, n).cpu().numpy(), y.cpu().numpy().astype(np.float64)) < OUT_TOL   # the default call
    step = 65536                                   # 64 tiles per piece
    y_plain = torch.cat([run(lo, min(n, lo + step), split=False) for lo in range(0, n, step)])
    assert torch.equal(y, y_plain)
    y_split = torch.cat([run(lo, min(n, lo + step)) for lo in range(0, n, step)])
    assert rel_max(y_split.cpu().numpy(), y.cpu().numpy().astype(np.float64)) < OUT_TOL
    # against the float64 oracle on a strided sample (including the ragged tail)
    idx = np.unique(np.concatenate([np.arange(0, n, 97), np.arange(n - 800, n)]))
    xs = x[torch.tensor(idx, device=dev)].cpu().numpy().astype(np.float64)
    if cfg["kind"] == "cluster":
        y_ref, _ = orc.cluster_forward(xs, params, cfg)
    else:
        gamma = orc.region_activation(xs, cfg["num_regions"], cfg["lower_bounds"], cfg["upper_bounds"], cfg["delta"],
                                      cfg["dimension_ranges"])
        y_ref = orc.forward(xs, rb["centers"], rb["log_sigs"], params["linear"]["kernel"], params["linear"]["bias"],
                            gamma, basis)
    assert rel_max(y[torch.tensor(idx, device=dev)].cpu().numpy(), y_ref) < OUT_TOL
