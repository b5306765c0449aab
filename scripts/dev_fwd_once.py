"""Dev: a few forward launches of one golden config at one batch size (for ncu)."""
import sys, os, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from irbfn_b200 import ops
dev = torch.device("cuda:0")
fx = sys.argv[1] if len(sys.argv) > 1 else "c2_frenet_12x100.npz"
B = int(sys.argv[2]) if len(sys.argv) > 2 else 32768
z = np.load(os.path.join(os.path.dirname(__file__), "..", "tests", "golden", fx))
cfg = json.loads(str(z["cfg_json"]))
to = lambda a: torch.from_numpy(np.ascontiguousarray(a, dtype=np.float32)).to(dev)
c, ls, W, b = to(z["p/rbf_list/centers"]), to(z["p/rbf_list/log_sigs"]), to(z["p/linear/kernel"]), to(z["p/linear/bias"])
gd = ops.GatePack(cfg["lower_bounds"], cfg["upper_bounds"], cfg["delta"], cfg["dimension_ranges"], dev)
x = torch.rand((B, c.shape[2]), device=dev)
for _ in range(4):
    y = ops.forward(x, c, ls, W, b, gd=gd, basis=0)
torch.cuda.synchronize()
print(float(y.sum()))
