"""Dev: sustained timing of each kernel with SM clock / power sampled through NVML (run under gpurun)."""
import json, os, sys, threading, time
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
from irbfn_b200 import ops
from irbfn_b200.basis import gaussian
from irbfn_b200.model import WCRBFNet, device_params
import pynvml
pynvml.nvmlInit(); h = pynvml.nvmlDeviceGetHandleByIndex(0)

dev = torch.device("cuda:0")
cfg, p_np = bench.load_c2()
net = WCRBFNet(cfg["in_features"], cfg["out_features"], cfg["num_kernels"], gaussian, cfg["num_regions"],
               cfg["lower_bounds"], cfg["upper_bounds"], cfg["dimension_ranges"], cfg["activation_idx"], cfg["delta"])
p = device_params({"params": p_np}, dev)["params"]
gd = net._gate_pack(dev)
B = 1 << 20
x = torch.from_numpy(bench.lattice(cfg)).to(dev)
dy = torch.ones((B, 2), device=dev)
c, ls, W, b = p["rbf_list"]["centers"], p["rbf_list"]["log_sigs"], p["linear"]["kernel"], p["linear"]["bias"]
y = torch.empty((B, 2), device=dev)
runs = {
    "fwd": lambda: ops.forward(x, c, ls, W, b, gd=gd, basis=0, out=y),
    "bwd_x": lambda: ops.backward(x, c, ls, W, dy, gd=gd, basis=0, grad_mask=ops.GRAD_X),
    "bwd_p": lambda: ops.backward(x, c, ls, W, dy, gd=gd, basis=0, grad_mask=ops.GRAD_PARAMS),
}
for name, fn in runs.items():
    for n in (3, 50, 400):
        for _ in range(3): fn()
        torch.cuda.synchronize()
        samples = []; stop = [False]
        def samp():
            while not stop[0]:
                samples.append((pynvml.nvmlDeviceGetClockInfo(h, pynvml.NVML_CLOCK_SM), pynvml.nvmlDeviceGetPowerUsage(h) / 1000.0,
                                pynvml.nvmlDeviceGetCurrentClocksThrottleReasons(h)))
                time.sleep(0.002)
        t = threading.Thread(target=samp); t.start()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(n): fn()
        e1.record(); torch.cuda.synchronize(); stop[0] = True; t.join()
        ms = e0.elapsed_time(e1) / n
        clk = np.median([s[0] for s in samples]); pw = max(s[1] for s in samples); rs = 0
        for s in samples: rs |= s[2]
        print("%-6s n=%4d  %.4f ms/iter  sm_mhz median %.0f  power max %.0f W  reasons 0x%x  samples %d" % (name, n, ms, clk, pw, rs, len(samples)))
