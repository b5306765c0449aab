"""Gate-sparse query-side backward (irbfn_backward_sparse, GRAD_X) against the dense one at the C4 shape."""
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


def timed(fn, n=20, w=3):
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


B = 1000000
x = (torch.rand((B, 8)) * torch.from_numpy(hi - lo) + torch.from_numpy(lo)).to(dev)
dy = torch.ones((B, 2), device=dev)
dense = ops.backward(x, c, ls, W, dy, gd=gd, basis=0, grad_mask=ops.GRAD_X)["x"]
t0 = timed(lambda: ops.backward(x, c, ls, W, dy, gd=gd, basis=0, grad_mask=ops.GRAD_X))
print("C4 dx dense %8.1f us" % t0)
for tol in (0.0, 2.0 ** -24, 1e-6):
    g = ops.backward(x, c, ls, W, dy, gd=gd, basis=0, grad_mask=ops.GRAD_X, gate_tol=tol)["x"]
    t = timed(lambda: ops.backward(x, c, ls, W, dy, gd=gd, basis=0, grad_mask=ops.GRAD_X, gate_tol=tol))
    print("  gate_tol %-8.2g %8.1f us (%.2fx)  rel-L2 vs dense %.2e  max|d|/max|dx| %.2e"
          % (tol, t, t0 / t, float((g - dense).norm() / dense.norm()), float((g - dense).abs().max() / dense.abs().max())))

# one-pass Jacobian on the same queries
jd = ops.jacobian(x, c, ls, W, gd=gd, basis=0)
tj = timed(lambda: ops.jacobian(x, c, ls, W, gd=gd, basis=0), n=10)
print("C4 Jacobian dense %8.1f us" % tj)
for tol in (0.0, 2.0 ** -24, 1e-6):
    js = ops.jacobian(x, c, ls, W, gd=gd, basis=0, gate_tol=tol)
    t = timed(lambda: ops.jacobian(x, c, ls, W, gd=gd, basis=0, gate_tol=tol), n=10)
    print("  gate_tol %-8.2g %8.1f us (%.2fx)  rel-L2 vs dense %.2e  max|d|/max|J| %.2e"
          % (tol, t, tj / t, float((js - jd).norm() / jd.norm()), float((js - jd).abs().max() / jd.abs().max())))
