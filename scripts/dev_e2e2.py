"""Host-path sweep 2: chunk size x concurrent kernel streams x graph replay (e2e ms per 2^20-query forward)."""
import sys, os
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
def timed(fn, n=20, w=5):
    for _ in range(w): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n
ref = net.apply_host(params, xh, out=torch.empty((B, 2)).pin_memory()).clone()
torch.cuda.synchronize()
for graph in (False, True):
    for tiles, nrun in ((148, 1), (148, 2), (74, 2), (74, 3), (37, 4), (37, 6), (50, 3), (50, 4), (25, 6), (296, 1)):
        ms = timed(lambda: net.apply_host(params, xh, out=yh, chunk=tiles * 1024, graph=graph, run_streams=nrun))
        torch.cuda.synchronize()
        ok = torch.equal(yh, ref)
        print("graph %-5s chunk %3d tiles  run streams %d  e2e %.3f ms  equal %s" % (graph, tiles, nrun, ms, ok), flush=True)
