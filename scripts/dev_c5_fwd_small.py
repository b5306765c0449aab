import sys, os
root = os.path.abspath(sys.argv[1])
sys.path.insert(0, root)
import numpy as np, torch
from irbfn_b200 import ops
dev = torch.device("cuda:0")
R, K, D, O = 64, 4096, 6, 16
B = 1 << 19
rng = np.random.default_rng(1234)
to = lambda a: torch.from_numpy(np.ascontiguousarray(a, dtype=np.float32)).to(dev)
edges = np.linspace(-1.0, 1.0, 5)
gd = ops.GatePack([list(edges[:-1])] * 3, [list(edges[1:])] * 3, [100.0] * 3,
                  [[i, j, k] for i in range(4) for j in range(4) for k in range(4)], dev)
c, ls, W = to(rng.uniform(-1, 1, (R, K, D))), to(rng.uniform(0, 2, (R, K))), to(rng.standard_normal((K, O)))
b0 = torch.zeros(O, device=dev)
x = torch.rand((B, D), device=dev) * 2 - 1
y = torch.empty((B, O), device=dev)
for rep in range(2):
    ops.forward(x, c, ls, W, b0, gd=gd, basis=0, out=y)
torch.cuda.synchronize()
