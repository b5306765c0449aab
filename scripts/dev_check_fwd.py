"""Dev check: forward parity on the golden fixtures + a first timing (run under gpurun)."""
import glob, json, os, sys, time
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from irbfn_b200 import ops

dev = torch.device("cuda:0")
for path in sorted(glob.glob("tests/golden/*.npz")):
    z = np.load(path)
    cfg = json.loads(str(z["cfg_json"]))
    if cfg["kind"] != "wcrbf":
        continue
    t = lambda a: torch.tensor(a, device=dev)
    gd = ops.GatePack(cfg["lower_bounds"], cfg["upper_bounds"], cfg["delta"], cfg["dimension_ranges"], dev)
    basis = {"gaussian": 0, "inverse_quadratic": 1}[cfg["basis_func"]]
    x = t(z["x"])
    y = ops.forward(x, t(z["p/rbf_list/centers"]), t(z["p/rbf_list/log_sigs"]), t(z["p/linear/kernel"]),
                    t(z["p/linear/bias"]), gd=gd, basis=basis)
    torch.cuda.synchronize()
    y64 = z["y64"]
    err = np.abs(y.cpu().numpy().astype(np.float64) - y64).max() / np.abs(y64).max()
    gam = ops.region_gate(x, gd, cfg["num_regions"]).cpu().numpy()
    print("%-30s R=%d K=%d D=%d O=%d  y err %.2e  gate sum range [%.3f, %.3f]" % (
        os.path.basename(path), cfg["num_regions"], cfg["num_kernels"], cfg["in_features"], cfg["out_features"],
        err, gam.sum(1).min(), gam.sum(1).max()))

# timing at C2 shape
z = np.load("tests/golden/c2_frenet_12x100.npz")
cfg = json.loads(str(z["cfg_json"]))
t = lambda a: torch.tensor(a, device=dev)
gd = ops.GatePack(cfg["lower_bounds"], cfg["upper_bounds"], cfg["delta"], cfg["dimension_ranges"], dev)
B = 1 << 20
lo = np.array([min(b) for b in cfg["lower_bounds"]]); hi = np.array([max(b) for b in cfg["upper_bounds"]])
rng = np.random.default_rng(0)
x = t((lo + rng.random((B, 8)) * (hi - lo)).astype(np.float32))
args = (x, t(z["p/rbf_list/centers"]), t(z["p/rbf_list/log_sigs"]), t(z["p/linear/kernel"]), t(z["p/linear/bias"]))
y = torch.empty((B, 2), device=dev)
for _ in range(3):
    ops.forward(*args, gd=gd, basis=0, out=y)
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(10):
    ops.forward(*args, gd=gd, basis=0, out=y)
e1.record(); torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / 10
lane_ops = B * 1200 * 18
print("C2 fwd B=2^20: %.3f ms  %.3f Gq/s  fp32 lane-ops %.1f T/s (peak ~36.3)" % (ms, B / ms / 1e6, lane_ops / ms / 1e9))
