"""C5 stress shape, backward only, small batch (run under ncu launch list on the GPU box)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from irbfn_b200 import ops
dev = torch.device("cuda:0")
R, K, D, O = 64, 4096, 6, 16
B = int(sys.argv[1]) if len(sys.argv) > 1 else 1 << 18
rng = np.random.default_rng(1234)
to = lambda a: torch.from_numpy(np.ascontiguousarray(a, dtype=np.float32)).to(dev)
edges = np.linspace(-1.0, 1.0, 5)
gd = ops.GatePack([list(edges[:-1])] * 3, [list(edges[1:])] * 3, [100.0] * 3,
                  [[i, j, k] for i in range(4) for j in range(4) for k in range(4)], dev)
c, ls, W = to(rng.uniform(-1, 1, (R, K, D))), to(rng.uniform(0, 2, (R, K))), to(rng.standard_normal((K, O)))
x = torch.rand((B, D), device=dev) * 2 - 1
dy = torch.randn((B, O), device=dev)
flat = torch.empty(R * K * (D + 1) + K * O + O, device=dev)
for _ in range(2):
    ops.backward(x, c, ls, W, dy, gd=gd, basis=0, grad_mask=ops.GRAD_PARAMS, flat_out=flat)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
ops.backward(x, c, ls, W, dy, gd=gd, basis=0, grad_mask=ops.GRAD_PARAMS, flat_out=flat)
e1.record(); torch.cuda.synchronize()
print("B", B, "backward ms", e0.elapsed_time(e1))
y = torch.empty((B, O), device=dev)
b0 = torch.zeros(O, device=dev)
for _ in range(2):
    ops.forward(x, c, ls, W, b0, gd=gd, basis=0, out=y)
torch.cuda.synchronize()
e0.record()
ops.forward(x, c, ls, W, b0, gd=gd, basis=0, out=y)
e1.record(); torch.cuda.synchronize()
print("B", B, "forward ms", e0.elapsed_time(e1))
