"""Dev: a few C3 grad steps at one shard size, for an ncu launch list."""
import sys, os, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from irbfn_b200.basis import gaussian
from irbfn_b200.model import WCRBFNet, device_params
from irbfn_b200.train import TrainState, grad_step
dev = torch.device("cuda:0")
B = int(sys.argv[1]) if len(sys.argv) > 1 else 32768
z = np.load(os.path.join(os.path.dirname(__file__), "..", "tests", "golden", "c2_frenet_12x100.npz"))
cfg = json.loads(str(z["cfg_json"]))
params = {"params": {"rbf_list": {"centers": z["p/rbf_list/centers"], "log_sigs": z["p/rbf_list/log_sigs"]},
                     "linear": {"kernel": z["p/linear/kernel"], "bias": z["p/linear/bias"]}}}
net = WCRBFNet(cfg["in_features"], cfg["out_features"], cfg["num_kernels"], gaussian, cfg["num_regions"],
               cfg["lower_bounds"], cfg["upper_bounds"], cfg["dimension_ranges"], cfg["activation_idx"], cfg["delta"])
state = TrainState(net, device_params(params, dev), lr=1e-3, device=dev)
x, y = torch.rand((B, 8), device=dev), torch.randn((B, 2), device=dev)
for _ in range(3):
    grad_step(state, x, y, B)
torch.cuda.synchronize()
