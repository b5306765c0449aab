"""Streamed host path (irbfn_forward_streamed) variants at C2, 2^20 queries: y stored by the kernel into pinned host
memory vs y on the device + one download; upload chunk size."""
import sys, os, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from irbfn_b200 import ops
from irbfn_b200.hostpath import StreamedPipeline

dev = torch.device("cuda:0")
z = np.load(os.path.join(os.path.dirname(__file__), "..", "tests", "golden", "c2_frenet_12x100.npz"))
cfg = json.loads(str(z["cfg_json"]))
to = lambda a: torch.from_numpy(np.ascontiguousarray(a, dtype=np.float32)).to(dev)
c, ls, W, b = to(z["p/rbf_list/centers"]), to(z["p/rbf_list/log_sigs"]), to(z["p/linear/kernel"]), to(z["p/linear/bias"])
gd = ops.GatePack(cfg["lower_bounds"], cfg["upper_bounds"], cfg["delta"], cfg["dimension_ranges"], dev)
B = 1 << 20
xh = torch.rand((B, 8)).pin_memory()
yh = torch.empty((B, 2)).pin_memory()
yd = torch.empty((B, 2), device=dev)


def timed(fn, n=20, w=5):
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


xd = xh.to(dev)
print("device-resident kernel        %7.1f us" % timed(lambda: ops.forward(xd, c, ls, W, b, gd=gd, basis=0, out=yd)))
print("device x, y -> pinned host    %7.1f us" % timed(lambda: ops.forward_streamed(
    xd, c, ls, W, b, yh, torch.cat([torch.ones(16, dtype=torch.int32, device=dev), torch.zeros(1, dtype=torch.int32, device=dev)]),
    65536, gd=gd, basis=0)))
print("H2D 33.5 MB alone             %7.1f us" % timed(lambda: xd.copy_(xh, non_blocking=True)))
for chunk, ns in ((1 << 14, 1), (1 << 14, 2), (1 << 14, 3), (1 << 15, 1), (1 << 15, 2), (1 << 15, 3), (1 << 15, 4), (1 << 16, 2),
                  (1 << 16, 3), (1 << 17, 2)):
    pipe = StreamedPipeline(dev, 8, 2, chunk, ns)
    t_host = timed(lambda: pipe.run(xh, yh, lambda x_, y_, r_, q_, s_: ops.forward_streamed(x_, c, ls, W, b, y_, r_, q_, gd=gd, basis=0, status=s_)))
    print("pieces of %6d rows on %d copy streams: %7.1f us   status %d" % (chunk, ns, t_host, int(pipe.status.item())))
