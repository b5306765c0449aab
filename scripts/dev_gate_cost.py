"""Dev: what the fused gate costs per tile -- forward with the fused tanh gate vs the same gate passed in."""
import sys, os, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from irbfn_b200 import ops
dev = torch.device("cuda:0")
def timed(fn, n=100, w=10):
    for _ in range(w): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n * 1e3
z = np.load(os.path.join(os.path.dirname(__file__), "..", "tests", "golden", "c2_frenet_12x100.npz"))
cfg = json.loads(str(z["cfg_json"]))
to = lambda a: torch.from_numpy(np.ascontiguousarray(a, dtype=np.float32)).to(dev)
c, ls, W, b = to(z["p/rbf_list/centers"]), to(z["p/rbf_list/log_sigs"]), to(z["p/linear/kernel"]), to(z["p/linear/bias"])
gd = ops.GatePack(cfg["lower_bounds"], cfg["upper_bounds"], cfg["delta"], cfg["dimension_ranges"], dev)
for B in (131072, 1 << 20):
    x = torch.rand((B, 8), device=dev)
    g = ops.region_gate(x, gd, 12)
    y = torch.empty((B, 2), device=dev)
    t_f = timed(lambda: ops.forward(x, c, ls, W, b, gd=gd, basis=0, out=y))
    t_e = timed(lambda: ops.forward(x, c, ls, W, b, gate=g, basis=0, out=y))
    t_g = timed(lambda: ops.region_gate(x, gd, 12))
    print("B=%d fused gate %.1f us   external gate %.1f us   region_gate kernel alone %.1f us" % (B, t_f, t_e, t_g), flush=True)
