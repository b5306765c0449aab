"""Dev: FP32 / XU issue-rate micro-benchmark table (run under gpurun)."""
import os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from irbfn_b200 import peak

names = ["FFMA", "FFMA2", "FADD2+FFMA2", "MUFU.EX2", "packed mix D=8", "scalar mix D=8", "FFMA2 + 1/9 MUFU"]
p = torch.cuda.get_device_properties(0)
print(p.name, "SMs", p.multi_processor_count)
for rep in range(2):
    for mode, nm in enumerate(names):
        f, m, ms = peak.fp32_microbench(mode, iters=20000)
        sm = p.multi_processor_count
        print("mode %d %-18s %.2f ms  fp32 lane-ops %.2f T/s (%.1f /SM/clk@1.965)  mufu %.2f T/s (%.2f /SM/clk)" % (
            mode, nm, ms, f / 1e12, f / sm / 1.965e9, m / 1e12, m / sm / 1.965e9))
