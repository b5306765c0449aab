"""Dev: dense vs gate-sparse forward on the bench's C2 lattice and on the C5 stress shape."""
import sys, os, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import bench
from irbfn_b200 import ops
dev = torch.device("cuda:0")
to = lambda a: torch.from_numpy(np.ascontiguousarray(a, dtype=np.float32)).to(dev)
def timed(fn, n=10, w=3):
    for _ in range(w): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n
cfg, p = bench.load_c2()
gd = ops.GatePack(cfg["lower_bounds"], cfg["upper_bounds"], cfg["delta"], cfg["dimension_ranges"], dev)
c, ls, W, b = to(p["rbf_list"]["centers"]), to(p["rbf_list"]["log_sigs"]), to(p["linear"]["kernel"]), to(p["linear"]["bias"])
lo = np.array([min(v) for v in cfg["lower_bounds"]], np.float32); hi = np.array([max(v) for v in cfg["upper_bounds"]], np.float32)
for name, x in (("C2 lattice", to(bench.lattice(cfg))),
                ("C2 uniform", (torch.rand((1 << 20, 8)) * torch.from_numpy(hi - lo) + torch.from_numpy(lo)).to(dev))):
    yd = ops.forward(x, c, ls, W, b, gd=gd, basis=0)
    scale = float(yd.abs().max())
    cnt = torch.zeros(2, dtype=torch.int64, device=dev)
    for tol in (1e-6 * scale, 1e-7 * scale):
        ys = ops.forward_sparse(x, c, ls, W, b, gd=gd, basis=0, tol_abs=tol, counters=cnt)
        err = float((ys - yd).abs().max()) / scale
        td = timed(lambda: ops.forward(x, c, ls, W, b, gd=gd, basis=0, out=yd))
        ts = timed(lambda: ops.forward_sparse(x, c, ls, W, b, gd=gd, basis=0, tol_abs=tol, out=ys))
        sk, dn = (int(v) for v in cnt.cpu())
        print("%s tol %.1e*max|y|: dense %.3f ms  sparse %.3f ms  (%.2fx)  rel err %.1e  skipped %.1f%%" %
              (name, tol / scale, td, ts, td / ts, err, 100.0 * sk / (sk + dn)), flush=True)
R, K, D, O = 64, 4096, 6, 16
rng = np.random.default_rng(1234)
edges = np.linspace(-1.0, 1.0, 5)
gd5 = ops.GatePack([list(edges[:-1])] * 3, [list(edges[1:])] * 3, [100.0] * 3,
                   [[i, j, k] for i in range(4) for j in range(4) for k in range(4)], dev)
c5, ls5, W5, b5 = to(rng.uniform(-1, 1, (R, K, D))), to(rng.uniform(0, 2, (R, K))), to(rng.standard_normal((K, O))), torch.zeros(O, device=dev)
B5 = 1 << 22
x5 = torch.rand((B5, D), device=dev) * 2 - 1
yd = ops.forward(x5, c5, ls5, W5, b5, gd=gd5, basis=0)
scale = float(yd.abs().max())
cnt = torch.zeros(2, dtype=torch.int64, device=dev)
ys = ops.forward_sparse(x5, c5, ls5, W5, b5, gd=gd5, basis=0, tol_abs=1e-6 * scale, counters=cnt)
err = float((ys - yd).abs().max()) / scale
td = timed(lambda: ops.forward(x5, c5, ls5, W5, b5, gd=gd5, basis=0, out=yd), 2, 1)
ts = timed(lambda: ops.forward_sparse(x5, c5, ls5, W5, b5, gd=gd5, basis=0, tol_abs=1e-6 * scale, out=ys), 3, 1)
sk, dn = (int(v) for v in cnt.cpu())
print("C5 2^22 queries: dense %.1f ms  sparse %.1f ms  (%.2fx)  rel err %.1e  skipped %.1f%%" %
      (td, ts, td / ts, err, 100.0 * sk / (sk + dn)), flush=True)
