"""Gate-sparse parameter gradients (irbfn_backward_sparse) against the dense backward at the C3 shape: time and
per-leaf deviation for a few thresholds; uniform queries inside the gate's bounds (what bench.py's C3 draws)."""
import sys, os, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from irbfn_b200 import ops

dev = torch.device("cuda:0")
z = np.load(os.path.join(os.path.dirname(__file__), "..", "tests", "golden", "c2_frenet_12x100.npz"))
cfg = json.loads(str(z["cfg_json"]))
to = lambda a: torch.from_numpy(np.ascontiguousarray(a, dtype=np.float32)).to(dev)
c, ls, W, b = to(z["p/rbf_list/centers"]), to(z["p/rbf_list/log_sigs"]), to(z["p/linear/kernel"]), to(z["p/linear/bias"])
gd = ops.GatePack(cfg["lower_bounds"], cfg["upper_bounds"], cfg["delta"], cfg["dimension_ranges"], dev)
lo = np.array([min(v) for v in cfg["lower_bounds"]], np.float32)
hi = np.array([max(v) for v in cfg["upper_bounds"]], np.float32)


def timed(fn, n=50, w=5):
    for _ in range(w):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n * 1e3


for B in (262144, 32768):
    x = (torch.rand((B, 8)) * torch.from_numpy(hi - lo) + torch.from_numpy(lo)).to(dev)
    dy = torch.randn((B, 2), device=dev) / B
    gate = ops.region_gate(x, gd, 12)
    print("B=%d  active (gamma > 2^-24): %.2f of 12 regions per query" % (B, float((gate > 2.0 ** -24).sum(1).float().mean())))
    dense = ops.backward(x, c, ls, W, dy, gate=gate, basis=0)
    t0 = timed(lambda: ops.backward(x, c, ls, W, dy, gate=gate, basis=0))
    print("  dense            %7.1f us" % t0)
    for tol in (0.0, 1e-12, 2.0 ** -24, 1e-6, 1e-4):
        g = ops.backward(x, c, ls, W, dy, gate=gate, basis=0, gate_tol=tol)
        t = timed(lambda: ops.backward(x, c, ls, W, dy, gate=gate, basis=0, gate_tol=tol))
        err = {k: float((g[k] - dense[k]).norm() / dense[k].norm()) for k in ("centers", "log_sigs", "kernel", "bias")}
        print("  gate_tol %-8.2g %7.1f us  (%.2fx)  rel-L2 vs dense: %s  skipped pairs %.3f"
              % (tol, t, t0 / t, " ".join("%s %.1e" % kv for kv in err.items()), float((gate <= tol).float().mean())))


def run_shape(name, x, c, ls, W, gate, tols, n=10):
    B, O = x.shape[0], W.shape[1]
    dy = torch.randn((B, O), device=dev) / B
    R = c.shape[0]
    print("%s B=%d  active (gamma > 2^-24): %.2f of %d regions per query"
          % (name, B, float((gate > 2.0 ** -24).sum(1).float().mean()), R))
    dense = ops.backward(x, c, ls, W, dy, gate=gate, basis=0)
    t0 = timed(lambda: ops.backward(x, c, ls, W, dy, gate=gate, basis=0), n, 2)
    print("  dense            %9.1f us" % t0)
    for tol in tols:
        g = ops.backward(x, c, ls, W, dy, gate=gate, basis=0, gate_tol=tol)
        t = timed(lambda: ops.backward(x, c, ls, W, dy, gate=gate, basis=0, gate_tol=tol), n, 2)
        err = {k: float((g[k] - dense[k]).norm() / dense[k].norm()) for k in ("centers", "log_sigs", "kernel")}
        print("  gate_tol %-8.2g %9.1f us  (%.2fx)  rel-L2 vs dense: %s" % (tol, t, t0 / t, " ".join("%s %.1e" % kv for kv in err.items())))


# the shipped 27-region, 10-output shape (dense: region-lane kernel irbfn_bwd_pw)
z = np.load(os.path.join(os.path.dirname(__file__), "..", "tests", "golden", "warmstart_27x500_o10.npz"))
cfg = json.loads(str(z["cfg_json"]))
c, ls, W = to(z["p/rbf_list/centers"]), to(z["p/rbf_list/log_sigs"]), to(z["p/linear/kernel"])
gd = ops.GatePack(cfg["lower_bounds"], cfg["upper_bounds"], cfg["delta"], cfg["dimension_ranges"], dev)
lo = np.array([min(v) for v in cfg["lower_bounds"]], np.float32)
hi = np.array([max(v) for v in cfg["upper_bounds"]], np.float32)
x = (torch.rand((262144, 8)) * torch.from_numpy(hi - lo) + torch.from_numpy(lo)).to(dev)
run_shape("27x500x8->10", x, c, ls, W, ops.region_gate(x, gd, 27), (0.0, 2.0 ** -24, 1e-6))

# C5 stress shape (bench.py)
R, K, D, O = 64, 4096, 6, 16
rng = np.random.default_rng(1234)
edges = np.linspace(-1.0, 1.0, 5)
gd5 = ops.GatePack([list(edges[:-1])] * 3, [list(edges[1:])] * 3, [100.0] * 3,
                   [[i, j, k] for i in range(4) for j in range(4) for k in range(4)], dev)
c5, ls5, W5 = to(rng.uniform(-1, 1, (R, K, D))), to(rng.uniform(0, 2, (R, K))), to(rng.standard_normal((K, O)))
x5 = torch.rand((1 << 18, D), device=dev) * 2 - 1
run_shape("C5 64x4096x6->16", x5, c5, ls5, W5, ops.region_gate(x5, gd5, 64), (0.0, 2.0 ** -24, 1e-6), n=3)
