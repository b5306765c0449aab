"""C3 grad_step + update at ONE per-GPU shard size (default 32,768 = the 8-GPU share), a few eager steps: run under
`ncu --metrics gpu__time_duration.sum` to list what a rank launches per step and what each launch costs."""
import sys, os, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from irbfn_b200.basis import gaussian
from irbfn_b200.model import WCRBFNet, device_params
from irbfn_b200.train import TrainState, train_step

B = int(sys.argv[1]) if len(sys.argv) > 1 else 32768
dev = torch.device("cuda:0")
z = np.load(os.path.join(os.path.dirname(__file__), "..", "tests", "golden", "c2_frenet_12x100.npz"))
cfg = json.loads(str(z["cfg_json"]))
params = {"params": {"rbf_list": {"centers": z["p/rbf_list/centers"], "log_sigs": z["p/rbf_list/log_sigs"]},
                     "linear": {"kernel": z["p/linear/kernel"], "bias": z["p/linear/bias"]}}}
net = WCRBFNet(cfg["in_features"], cfg["out_features"], cfg["num_kernels"], gaussian, cfg["num_regions"],
               cfg["lower_bounds"], cfg["upper_bounds"], cfg["dimension_ranges"], cfg["activation_idx"], cfg["delta"])
state = TrainState(net, device_params(params, dev), lr=1e-3, device=dev)
lo = np.array([min(b) for b in cfg["lower_bounds"]], np.float32)
hi = np.array([max(b) for b in cfg["upper_bounds"]], np.float32)
x = (torch.rand((B, 8)) * torch.from_numpy(hi - lo) + torch.from_numpy(lo)).to(dev)
y = torch.randn((B, 2), device=dev)
for _ in range(4):
    train_step(state, x, y)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(50):
    train_step(state, x, y)
e1.record()
torch.cuda.synchronize()
print("B=%d eager step %.1f us" % (B, e0.elapsed_time(e1) / 50 * 1e3))
