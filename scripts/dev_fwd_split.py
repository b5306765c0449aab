"""Dev: forward time against the kernels-per-split choice (needs the IRBFN_DEV_KPS hook in irbfn_api.cu)."""
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
for fx in ("c2_frenet_12x100.npz", "c1_cartesian_1x1000.npz", "c2b_frenet_1x1000_o10.npz"):
    z = np.load(os.path.join(os.path.dirname(__file__), "..", "tests", "golden", fx))
    cfg = json.loads(str(z["cfg_json"]))
    to = lambda a: torch.from_numpy(np.ascontiguousarray(a, dtype=np.float32)).to(dev)
    c, ls, W, b = to(z["p/rbf_list/centers"]), to(z["p/rbf_list/log_sigs"]), to(z["p/linear/kernel"]), to(z["p/linear/bias"])
    gd = ops.GatePack(cfg["lower_bounds"], cfg["upper_bounds"], cfg["delta"], cfg["dimension_ranges"], dev)
    R, K, D = c.shape
    print(fx, "R K D O", R, K, D, W.shape[1], flush=True)
    for B in (8192, 16384, 32768, 65536, 131072, 262144):
        x = torch.rand((B, D), device=dev)
        row = []
        os.environ.pop("IRBFN_DEV_KPS", None)
        row.append(("auto", timed(lambda: ops.forward(x, c, ls, W, b, gd=gd, basis=0))))
        Kt = (K + 3) // 4
        seen = set()
        for ns in (1, 2, 3, 4, 5, 6, 7, 8, 9, 10, 12, 13, 16, 20, 25):
            kps = 4 * ((Kt + ns - 1) // ns)
            if kps in seen or kps < 4: continue
            seen.add(kps)
            os.environ["IRBFN_DEV_KPS"] = str(kps)
            row.append(("%d/%d" % (kps, (K + kps - 1) // kps), timed(lambda: ops.forward(x, c, ls, W, b, gd=gd, basis=0))))
        print("B=%6d " % B + "  ".join("%s:%.0f" % r for r in row), flush=True)
