"""Dev: cost of writing the gate from the forward (irbfn_forward_gate) against the plain forward + gate_fill."""
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
lo = np.array([min(v) for v in cfg["lower_bounds"]], np.float32); hi = np.array([max(v) for v in cfg["upper_bounds"]], np.float32)
def timed(fn, n=100, w=10):
    for _ in range(w): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n * 1e3
for B in (32768, 262144, 1 << 20):
    x = (torch.rand((B, 8)) * torch.from_numpy(hi - lo) + torch.from_numpy(lo)).to(dev)
    y = torch.empty((B, 2), device=dev); g = torch.empty((B, 12), device=dev)
    t0 = timed(lambda: ops.forward(x, c, ls, W, b, gd=gd, basis=0, out=y))
    t1 = timed(lambda: ops.forward(x, c, ls, W, b, gd=gd, basis=0, out=y, gate_out=g))
    t2 = timed(lambda: ops.region_gate(x, gd, 12))
    print("B %8d  forward %8.1f us   forward+gate_out %8.1f us   region_gate alone %6.1f us" % (B, t0, t1, t2))
