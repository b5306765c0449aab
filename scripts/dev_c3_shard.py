"""C3 step at the per-GPU shard sizes of the strong-scaling run (262144 / N samples), one GPU, no collective:
per-launch times of forward, loss, VJP, and the graph-replayed whole, to see what bounds the step at N = 8."""
import sys, os, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from irbfn_b200 import ops
from irbfn_b200.basis import gaussian
from irbfn_b200.model import WCRBFNet, device_params
from irbfn_b200.train import TrainState, grad_step, loss_and_cotangent

dev = torch.device("cuda:0")
z = np.load(os.path.join(os.path.dirname(__file__), "..", "tests", "golden", "c2_frenet_12x100.npz"))
cfg = json.loads(str(z["cfg_json"]))
params = {"params": {"rbf_list": {"centers": z["p/rbf_list/centers"], "log_sigs": z["p/rbf_list/log_sigs"]},
                     "linear": {"kernel": z["p/linear/kernel"], "bias": z["p/linear/bias"]}}}
net = WCRBFNet(cfg["in_features"], cfg["out_features"], cfg["num_kernels"], gaussian, cfg["num_regions"],
               cfg["lower_bounds"], cfg["upper_bounds"], cfg["dimension_ranges"], cfg["activation_idx"], cfg["delta"])
dp = device_params(params, dev)
state = TrainState(net, dp, lr=1e-3, device=dev)
lo = np.array([min(b) for b in cfg["lower_bounds"]], np.float32)
hi = np.array([max(b) for b in cfg["upper_bounds"]], np.float32)


def timed(fn, n=200, w=20):
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


for N in (1, 2, 4, 8, 16):
    B = 262144 // N
    x = (torch.rand((B, 8)) * torch.from_numpy(hi - lo) + torch.from_numpy(lo)).to(dev)
    y = torch.randn((B, 2), device=dev)
    v = state.views
    c, ls = v["centers"], v["log_sigs"]
    gd = net._gate_pack(dev)
    dy = torch.randn((B, 2), device=dev)
    yp = torch.empty((B, 2), device=dev)
    t_f = timed(lambda: ops.forward(x, c, ls, v["kernel"], v["bias"], gd=gd, basis=0, out=yp))
    t_l = timed(lambda: loss_and_cotangent(yp, y, 262144))
    t_b = timed(lambda: ops.backward(x, c, ls, v["kernel"], dy, gd=gd, basis=0, grad_mask=state.mask, flat_out=state.grad))
    t_s = timed(lambda: grad_step(state, x, y, B))
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        grad_step(state, x, y, B)
    t_g = timed(lambda: g.replay())
    print("N=%2d B=%6d  fwd %7.1f us  loss(torch) %6.1f us  vjp %7.1f us  grad_step eager %7.1f us  graph %7.1f us"
          % (N, B, t_f, t_l, t_b, t_s, t_g), flush=True)
