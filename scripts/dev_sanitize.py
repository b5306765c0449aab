"""Small invocations of every kernel, meant to run under compute-sanitizer (memcheck / racecheck)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import numpy as np, torch
from irbfn_b200 import ops, _lib
from test_forward_gpu import _synthetic, _gd, _t

dev = torch.device("cuda:0")
P = ops.default_dyn_params()
for (B, R, K, D, O) in [(700, 3, 37, 5, 3), (2100, 2, 300, 8, 10), (300, 12, 100, 8, 2), (130, 16, 2048, 8, 2), (50, 2, 9, 3, 18),
                        (260, 4, 512, 6, 16), (40000, 12, 100, 8, 2)]:
    cfg, p, x = _synthetic(1, B, R, K, D, O)
    xs, c, ls, W, b = (_t(a, dev) for a in (x, p["centers"], p["log_sigs"], p["kernel"], p["bias"]))
    gd = _gd(cfg, dev)
    for basis in (0, 1):
        y, h = ops.forward(xs, c, ls, W, b, gd=gd, basis=basis, want_hidden=True)
        dy = torch.randn_like(y)
        ops.backward(xs, c, ls, W, dy, gd=gd, basis=basis, grad_mask=ops.GRAD_PARAMS | ops.GRAD_X)
        ops.backward(xs, c, ls, None, None, dhidden=torch.randn_like(h), gd=gd, basis=basis,
                     grad_mask=ops.GRAD_CENTERS | ops.GRAD_LOG_SIGS | ops.GRAD_X)
        gate = ops.region_gate(xs, gd, R)
        ops.backward(xs, c, ls, W, dy, gate=gate, basis=basis, grad_mask=ops.GRAD_PARAMS | ops.GRAD_X | ops.GRAD_GATE)
        ops.jacobian(xs, c, ls, W, b, gd=gd, basis=basis, want_y=True)
        ops.jacobian(xs, c, ls, W, b, gate=gate, basis=basis, want_y=True, want_region_y=True)
    print("ok", B, R, K, D, O)
x8 = torch.rand((999, 8), device=dev)
for mode, O in [(0, 2), (0, 10), (1, 10), (2, 2), (3, 10)]:
    ops.dynamics_loss(mode, x8, torch.randn((999, O), device=dev), torch.randn((999, O), device=dev), P)
for model, W_ in [(0, 18), (1, 9), (2, 17), (3, 9)]:
    ops.integrate(model, torch.rand((777, W_), device=dev) + 0.5, P)
X, Y = torch.rand((5000, 8), device=dev), torch.rand((5000, 10), device=dev)
ops.gather_rows(X, Y, torch.randint(0, 5000, (3333,), device=dev))
ops.nearest_row(torch.rand((333, 8), device=dev), X)
ops.nearest_row(torch.rand((70000, 8), device=dev)[:5000].contiguous(), X[:1001].contiguous())
gk = ops.GridKeys([np.linspace(0, 1, n) for n in (5, 1, 9)], dev)
ops.grid_index(torch.rand((1000, 3), device=dev), gk)
ops.softmax_gate(torch.rand((100, 8), device=dev), torch.rand((8, 50), device=dev), torch.rand(50, device=dev))
for kind in ("l2", "l1"):
    ops.prediction_loss(torch.randn((1234, 10), device=dev), torch.randn((1234, 10), device=dev), kind=kind)
n = 11003
prm, g, mu, nu = (torch.randn(n, device=dev) for _ in range(4))
nu.abs_()
cnt = torch.zeros(1, device=dev, dtype=torch.float64)
for max_norm in (0.0, 1.0):
    ops.clip_adam_(prm, g, mu, nu, cnt, 1e-3, max_norm=max_norm)
torch.cuda.synchronize()
print("all kernels ran")
