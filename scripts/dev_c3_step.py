"""Dev: whole C3 optimiser steps (forward, loss, VJP, clip + Adam) at one shard size on one GPU.
Prints the CUDA-graph replay time; run under `ncu --metrics gpu__time_duration.sum` for the launch list."""
import sys, os, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from irbfn_b200.basis import gaussian
from irbfn_b200.model import WCRBFNet, device_params
from irbfn_b200.train import GraphedTrainStep, TrainState, train_step
dev = torch.device("cuda:0")
B = int(sys.argv[1]) if len(sys.argv) > 1 else 32768
loss = sys.argv[2] if len(sys.argv) > 2 else "l2"
z = np.load(os.path.join(os.path.dirname(__file__), "..", "tests", "golden", "c2_frenet_12x100.npz"))
cfg = json.loads(str(z["cfg_json"]))
params = {"params": {"rbf_list": {"centers": z["p/rbf_list/centers"], "log_sigs": z["p/rbf_list/log_sigs"]},
                     "linear": {"kernel": z["p/linear/kernel"], "bias": z["p/linear/bias"]}}}
net = WCRBFNet(cfg["in_features"], cfg["out_features"], cfg["num_kernels"], gaussian, cfg["num_regions"],
               cfg["lower_bounds"], cfg["upper_bounds"], cfg["dimension_ranges"], cfg["activation_idx"], cfg["delta"])
state = TrainState(net, device_params(params, dev), lr=1e-3, device=dev)
x, y = torch.rand((B, 8), device=dev), torch.randn((B, 2), device=dev)
for _ in range(3):
    train_step(state, x, y, B, loss)
torch.cuda.synchronize()
if os.environ.get("IRBFN_DEV_GRAPH", "1") == "1":
    step = GraphedTrainStep(state, B, loss=loss).load(x, y)
    for _ in range(20):
        step()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(200):
        step()
    e1.record()
    torch.cuda.synchronize()
    print("B", B, loss, "graph-replayed step us", e0.elapsed_time(e1) / 200 * 1e3)
