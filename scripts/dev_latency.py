"""Small-batch latency of the forward (pred_step, irbfn_planner.py:29-32: B = 1 per control step, 400 in the notebooks,
4,096 at BASELINE config 1) on the C1 and C2 checkpoint shapes: per call eagerly and replayed from a CUDA graph."""
import sys, os, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from irbfn_b200 import ops

dev = torch.device("cuda:0")
G = os.path.join(os.path.dirname(__file__), "..", "tests", "golden")


PROFILE = "--profile" in sys.argv      # under ncu: two launches per case, no graphs


def timed(fn, n=200, w=20):
    if PROFILE:
        n, w = 2, 0
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


for name in ("c1_cartesian_1x1000", "c2_frenet_12x100", "c2b_frenet_1x1000_o10"):
    z = np.load(os.path.join(G, name + ".npz"))
    cfg = json.loads(str(z["cfg_json"]))
    to = lambda a: torch.from_numpy(np.ascontiguousarray(a, dtype=np.float32)).to(dev)
    c, ls, W, b = (to(z["p/rbf_list/centers"]), to(z["p/rbf_list/log_sigs"]), to(z["p/linear/kernel"]),
                   to(z["p/linear/bias"]))
    gd = ops.GatePack(cfg["lower_bounds"], cfg["upper_bounds"], cfg["delta"], cfg["dimension_ranges"], dev)
    basis = 0 if cfg["basis_func"] == "gaussian" else 1
    D, O = c.shape[2], W.shape[1]
    xs = np.asarray(z["x"], np.float32)                     # the fixture's 4,096 queries inside the config's box
    for B in [int(v) for v in os.environ.get("LAT_B", "1,32,400,1024,2048,4096,8192,32768").split(",")]:
        x = torch.from_numpy(np.ascontiguousarray(np.tile(xs, ((B + len(xs) - 1) // len(xs), 1))[:B])).to(dev)
        y = torch.empty((B, O), device=dev)
        fn = lambda: ops.forward(x, c, ls, W, b, gd=gd, basis=basis, out=y)
        t_eager = timed(fn)
        y_ref = y.clone()
        if PROFILE:
            continue
        g = torch.cuda.CUDAGraph()
        s = torch.cuda.Stream(dev)
        with torch.cuda.stream(s):
            fn()
            s.synchronize()
            with torch.cuda.graph(g, stream=s):
                for _ in range(10):
                    fn()
        t_graph = timed(g.replay, n=50, w=5) / 10
        y0 = ops.forward(x, c, ls, W, b, gd=gd, basis=basis, split=False)
        err = float((y_ref - y0).abs().max() / y0.abs().max().clamp_min(1e-30))
        print("%-24s B=%6d  eager %6.1f us  graph %6.2f us/call   split-vs-unsplit %.1e" % (name, B, t_eager, t_graph, err))
    if name == "c2_frenet_12x100" and not PROFILE:
        # the planner's call: module.apply on a NumPy batch, wall clock (the call waits for its result)
        import time
        from irbfn_b200.basis import gaussian
        from irbfn_b200.model import WCRBFNet, device_params
        net = WCRBFNet(cfg["in_features"], cfg["out_features"], cfg["num_kernels"], gaussian, cfg["num_regions"],
                       cfg["lower_bounds"], cfg["upper_bounds"], cfg["dimension_ranges"], cfg["activation_idx"], cfg["delta"])
        dp = device_params({"params": {"rbf_list": {"centers": z["p/rbf_list/centers"], "log_sigs": z["p/rbf_list/log_sigs"]},
                                       "linear": {"kernel": z["p/linear/kernel"], "bias": z["p/linear/bias"]}}}, dev)
        for B in (1, 400, 4096):
            xn = np.tile(xs, (4, 1))[:B].astype(np.float64)
            for _ in range(20):
                net.apply(dp, xn)
            t0 = time.perf_counter()
            for _ in range(300):
                net.apply(dp, xn)
            t_new = (time.perf_counter() - t0) / 300 * 1e6
            old = lambda: net.apply_host(dp, xn, chunk=148 * 1024)      # the chunk pipeline (what apply did before)
            for _ in range(20):
                old()
            t0 = time.perf_counter()
            for _ in range(300):
                old()
            t_old = (time.perf_counter() - t0) / 300 * 1e6
            print("module.apply(NumPy float64 x), B=%5d: %6.1f us per call, chunk pipeline %6.1f us (wall clock, result "
                  "on the host)" % (B, t_new, t_old))
