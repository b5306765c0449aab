"""Zero-copy experiment: the forward kernel reads its queries straight from pinned host memory."""
import sys, os, ctypes as C
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import bench
from irbfn_b200 import _lib, ops
from irbfn_b200.basis import gaussian
from irbfn_b200.model import WCRBFNet, device_params

dev = torch.device("cuda:0")
cfg, p_np = bench.load_c2()
net = WCRBFNet(cfg["in_features"], cfg["out_features"], cfg["num_kernels"], gaussian, cfg["num_regions"],
               cfg["lower_bounds"], cfg["upper_bounds"], cfg["dimension_ranges"], cfg["activation_idx"], cfg["delta"])
params = device_params({"params": p_np}, dev)["params"]
gd = net._gate_pack(dev)
B = 1 << 20
xh = torch.from_numpy(bench.lattice(cfg)).pin_memory()
yh = torch.empty((B, 2), dtype=torch.float32).pin_memory()
xd = xh.to(dev); yd = torch.empty((B, 2), device=dev)
lib = _lib.load()
c, ls, W, b = params["rbf_list"]["centers"], params["rbf_list"]["log_sigs"], params["linear"]["kernel"], params["linear"]["bias"]
P = lambda t: C.c_void_p(t.data_ptr())
def fwd(xp, yp):
    rc = lib.irbfn_forward(xp, P(c), P(ls), P(W), P(b), None, gd.ref(), yp, None, B, 12, 100, 8, 2, 0, None, 0,
                           C.c_void_p(torch.cuda.current_stream().cuda_stream))
    assert rc == 0, rc
def timed(fn, n=10, w=3):
    for _ in range(w): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n
print("device x, device y   ms", timed(lambda: fwd(P(xd), P(yd))))
print("HOST x,   device y   ms", timed(lambda: fwd(P(xh), P(yd))))
print("HOST x,   HOST y     ms", timed(lambda: fwd(P(xh), P(yh))))
print("device x, HOST y     ms", timed(lambda: fwd(P(xd), P(yh))))
ref = yd.clone(); fwd(P(xh), P(yh)); torch.cuda.synchronize()
fwd(P(xd), P(yd)); torch.cuda.synchronize()
print("equal", torch.equal(yd.cpu(), yh))
