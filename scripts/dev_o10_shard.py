"""The 27 x 500 x 8 -> 10 full-integration step at per-GPU shard sizes (one GPU, CUDA graph): dense, gate-sparse VJP,
gate-sparse forward + VJP -- where does each variant pay?"""
import sys, os, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from irbfn_b200 import basis
from irbfn_b200.model import WCRBFNet, device_params
from irbfn_b200.train import GraphedTrainStep, TrainState

dev = torch.device("cuda:0")
z = np.load(os.path.join(os.path.dirname(__file__), "..", "tests", "golden", "warmstart_27x500_o10.npz"))
cfg = json.loads(str(z["cfg_json"]))
R, K, D, O = cfg["num_regions"], cfg["num_kernels"], cfg["in_features"], cfg["out_features"]
net = WCRBFNet(D, O, K, getattr(basis, cfg["basis_func"]), R, cfg["lower_bounds"], cfg["upper_bounds"],
               cfg["dimension_ranges"], cfg["activation_idx"], cfg["delta"])
params = device_params({"params": {"rbf_list": {"centers": z["p/rbf_list/centers"], "log_sigs": z["p/rbf_list/log_sigs"]},
                                   "linear": {"kernel": z["p/linear/kernel"], "bias": z["p/linear/bias"]}}}, dev)
lo = np.array([min(b) for b in cfg["lower_bounds"]] + [-1.0] * D, np.float32)[:D]
hi = np.array([max(b) for b in cfg["upper_bounds"]] + [1.0] * D, np.float32)[:D]


def timed(fn, n=10, w=3):
    for _ in range(w):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n


for B in (262144, 131072, 65536, 32768):
    x = (torch.rand((B, D)) * torch.from_numpy(hi - lo) + torch.from_numpy(lo)).to(dev)
    y = torch.randn((B, O), device=dev)
    ys = float(net.apply(params, x).abs().max())
    out = []
    for kw in (dict(), dict(gate_tol=2.0 ** -24), dict(gate_tol=2.0 ** -24, fwd_tol_abs=1e-6 * ys)):
        st = TrainState(net, params, lr=1e-3, device=dev, **kw)
        g = GraphedTrainStep(st, B, 262144, loss="frenet_fullint").load(x, y)
        out.append(timed(lambda: g()))
    print("B=%6d  dense %.3f ms   sparse VJP %.3f ms   sparse forward + VJP %.3f ms" % (B, *out))
