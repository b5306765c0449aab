"""Host-path chunk sweep (run on the GPU box): e2e ms per 2^20-query forward for several chunk sizes,
plus raw pinned H2D / D2H copy times of the same payloads."""
import sys, os, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import bench
from irbfn_b200.basis import gaussian
from irbfn_b200.model import WCRBFNet, device_params

dev = torch.device("cuda:0")
cfg, p_np = bench.load_c2()
net = WCRBFNet(cfg["in_features"], cfg["out_features"], cfg["num_kernels"], gaussian, cfg["num_regions"],
               cfg["lower_bounds"], cfg["upper_bounds"], cfg["dimension_ranges"], cfg["activation_idx"], cfg["delta"])
params = device_params({"params": p_np}, dev)
B = 1 << 20
xh = torch.from_numpy(bench.lattice(cfg)).pin_memory()
yh = torch.empty((B, 2), dtype=torch.float32).pin_memory()
xd = torch.empty((B, 8), device=dev); yd = torch.empty((B, 2), device=dev)

def timed(fn, n=20, w=5):
    for _ in range(w): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n

print("H2D 33.5MB ms", timed(lambda: xd.copy_(xh, non_blocking=True)))
print("D2H 8.4MB ms", timed(lambda: yh.copy_(yd, non_blocking=True)))
sms = torch.cuda.get_device_properties(dev).multi_processor_count
for graph in (False, True):
    for ncopy in (1, 2):
        for chunk in (sms * 1024, sms * 2048):
            print("graph", graph, "copy streams", ncopy, "chunk", chunk, "e2e ms",
                  timed(lambda: net.apply_host(params, xh, out=yh, chunk=chunk, copy_streams=ncopy, graph=graph)))
ref = net.apply_host(params, xh, out=torch.empty((B, 2)).pin_memory(), graph=False).clone()
torch.cuda.synchronize()
got = net.apply_host(params, xh, out=yh, graph=True); torch.cuda.synchronize()
print("graph == eager:", torch.equal(ref, got))
