#!/usr/bin/env python
"""Turn gpurun_out/ ncu artefacts into the tracked summaries under profiles/.

usage: python scripts/summarize_profiles.py <tag>
  gpurun_out/launches_<tag>.csv          -> profiles/launches_<tag>.md   (per-kernel totals and shares)
  gpurun_out/prof_{fwd,bwdp,bwdx}_<tag>.ncu-rep -> profiles/ncu_<kernel>_<tag>.md (key metrics + stalls)
"""
import collections
import csv
import io
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GO = os.path.join(ROOT, "gpurun_out")
PR = os.path.join(ROOT, "profiles")

KEYS = [
    "gpu__time_duration.sum", "launch__grid_size", "launch__block_size", "launch__registers_per_thread",
    "launch__shared_mem_per_block_dynamic", "sm__cycles_elapsed.max", "sm__cycles_active.avg",
    "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__warps_active.avg.per_cycle_active",
    "smsp__warps_eligible.avg.per_cycle_active", "smsp__issue_active.avg.pct_of_peak_sustained_active",
    "sm__inst_executed.avg.per_cycle_active", "smsp__inst_executed.sum",
    "sm__throughput.avg.pct_of_peak_sustained_elapsed",
    "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active",
    "sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
    "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
    "dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
]


def launches(tag):
    path = os.path.join(GO, "launches_%s.csv" % tag)
    if not os.path.exists(path):
        return
    rows = list(csv.reader(open(path)))
    for i, r in enumerate(rows):
        if "Kernel Name" in r:
            hdr, start = r, i + 1
            break
    ki, vi = hdr.index("Kernel Name"), hdr.index("Metric Value")
    agg = collections.OrderedDict()
    for r in rows[start:]:
        if len(r) > vi:
            agg.setdefault(r[ki].split("(")[0][:80], []).append(float(r[vi].replace(",", "")))
    tot = sum(sum(v) for v in agg.values())
    with open(os.path.join(PR, "launches_%s.md" % tag), "w") as f:
        f.write("# ncu launch list, tag %s\n\n" % tag)
        f.write("Command: `python bench.py --steps 2 --warmup 3 --no-cpu` under `ncu --metrics "
                "gpu__time_duration.sum --clock-control none` (cold-cache, serialised: compare shares).\n"
                "The run contains the headline forward steps, the host-pipeline chunks, the C3 train-step\n"
                "and C4 input-gradient side measurements and the FP32/XU peak micro-benchmark.\n\n")
        f.write("| kernel | launches | total us | mean us | share |\n|---|---:|---:|---:|---:|\n")
        for k, v in sorted(agg.items(), key=lambda kv: -sum(kv[1])):
            f.write("| `%s` | %d | %.1f | %.1f | %.1f%% |\n" % (k, len(v), sum(v) / 1e3, sum(v) / len(v) / 1e3,
                                                              100 * sum(v) / tot))


def report(kind, tag):
    path = os.path.join(GO, "prof_%s_%s.ncu-rep" % (kind, tag))
    if not os.path.exists(path):
        return
    out = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    hdr, units, vals = rows[0], rows[1], rows[2]
    kn = vals[hdr.index("Kernel Name")] if "Kernel Name" in hdr else kind
    with open(os.path.join(PR, "ncu_%s_%s.md" % (kind, tag)), "w") as f:
        f.write("# ncu --set full, %s, tag %s\n\nKernel: `%s`\n\n| metric | value | unit |\n|---|---:|---|\n" % (kind, tag, kn))
        for k in KEYS:
            if k in hdr:
                i = hdr.index(k)
                f.write("| %s | %s | %s |\n" % (k, vals[i], units[i]))
        st = []
        for h, v in zip(hdr, vals):
            if "issue_stalled" in h and h.endswith("per_issue_active.ratio"):
                try:
                    st.append((float(v), h))
                except ValueError:
                    pass
        f.write("\nWarp stall reasons (warps stalled per issue-active cycle, top 10):\n\n| ratio | reason |\n|---:|---|\n")
        for v, h in sorted(st, reverse=True)[:10]:
            f.write("| %.3f | %s |\n" % (v, h.replace("smsp__average_warps_issue_stalled_", "").replace("_per_issue_active.ratio", "")))


if __name__ == "__main__":
    tag = sys.argv[1]
    os.makedirs(PR, exist_ok=True)
    launches(tag)
    for kind in ("fwd", "bwdp", "bwdx", "jac", "nearest", "bwdps", "bwdxs", "fwdsmall"):
        report(kind, tag)
    print(sorted(os.listdir(PR)))
