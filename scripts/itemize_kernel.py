#!/usr/bin/env python
"""Itemise one ncu --set full capture (--import-source on, kernels built with -lineinfo) by SOURCE SECTION:
where the executed warp-instructions and the stall samples of a kernel go.

usage: python scripts/itemize_kernel.py gpurun_out/prof_x.ncu-rep profiles/itemize_x.md

Sections are line ranges of the kernel's source found by marker comments (so the table survives edits)."""
import collections
import csv
import io
import re
import subprocess
import sys

# (section name, file suffix, start marker regex, end marker regex) -- first match wins, order matters
SECTIONS = [
    # ncu attributes an inlined instruction to its INNERMOST frame: the packed-arithmetic and MUFU wrappers of
    # irbfn_common.cuh collect the arithmetic of the hot loop's lambdas (distances, blend, scale, readout)
    ("packed fp32 arithmetic of the steps (FFMA2 / FADD2 / FMUL2: distances, scale, blend, readout)", "irbfn_common.cuh",
     r"^// ---- packed fp32", r"^// ---- XU \(MUFU\) wrappers"),
    ("basis function + gate sigmoids (MUFU wrappers)", "irbfn_common.cuh", r"^// ---- XU \(MUFU\) wrappers", r"^// Device-side copy of IrbfnGateDesc"),
    ("gate table (tanh factors, region products)", "irbfn_common.cuh", r"gate_columns_impl", r"^// consecutive rows"),
    ("query rows -> registers", "irbfn_common.cuh", r"void load_queries4", r"^// Region gate of a thread"),
    ("gate-sparse masks / scheduling", "irbfn_fwd.cuh", r"SPARSE: which regions of the chunk matter", r"const bool compact ="),
    ("gate written out (training)", "irbfn_fwd.cuh", r"training: the gate goes out", r"for \(int kch = 0"),
    ("parameter staging (transpose, exp)", "irbfn_fwd.cuh", r"---- stage parameters of", r"---- hot loop:"),
    ("readout y += h.W", "irbfn_fwd.cuh", r"auto readout = ", r"auto store_hidden = "),
    ("hidden store", "irbfn_fwd.cuh", r"auto store_hidden = ", r"// basis function of the PREVIOUS step"),
    ("basis function (MUFU)", "irbfn_fwd.cuh", r"auto mufu = ", r"// distances of one step"),
    ("distances (FADD2 / FFMA2)", "irbfn_fwd.cuh", r"auto distances = ", r"// blend of the PREVIOUS step"),
    ("blend h += gamma*phi", "irbfn_fwd.cuh", r"auto blend = ", r"// arguments of this step's basis function"),
    ("scale (basis argument)", "irbfn_fwd.cuh", r"auto scale = ", r"const size_t rstride ="),
    ("step loops, record / gate loads, drain", "irbfn_fwd.cuh", r"const size_t rstride =", r"// ---- store y ----"),
    ("store y", "irbfn_fwd.cuh", r"// ---- store y ----", r"^// Host-side: pick the instantiation"),
    ("tile loop, query load, bias init", "irbfn_fwd.cuh", r"const bool warp_sched", r"for \(int rch = 0"),
]


def section_ranges():
    import os
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    out = []
    for name, suffix, a, b in SECTIONS:
        path = os.path.join(root, "irbfn_b200", "csrc", suffix)
        lines = open(path).read().splitlines()
        lo = next((i + 1 for i, l in enumerate(lines) if re.search(a, l)), None)
        hi = next((i + 1 for i, l in enumerate(lines) if lo and i + 1 > lo and re.search(b, l)), None)
        if lo and hi:
            out.append((name, suffix, lo, hi))
    return out


def main():
    rep, dst = sys.argv[1], sys.argv[2]
    txt = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"],
                         capture_output=True, text=True).stdout
    ranges = section_ranges()
    cur_file, hdr, kernel = None, None, ""
    # an inlined SASS instruction is listed once per inline frame (wrapper line, lambda line, call-site line):
    # attribute every ADDRESS once, to the most specific section among its frames
    prio = {name: i for i, (name, _, _, _) in enumerate(ranges)}
    generic = {"step loops, record / gate loads, drain": 100, "tile loop, query load, bias init": 101}
    per_addr = {}
    line_no = None
    for row in csv.reader(io.StringIO(txt)):
        if not row:
            continue
        if row[0] in ("File Path", "File Name"):
            cur_file = row[1]
            continue
        if row[0] == "Function Name":
            kernel = row[1]
            continue
        if row[0] == "Line No":
            hdr = row
            continue
        if hdr is None or len(row) < len(hdr):
            continue
        if row[0].strip():
            try:
                line_no = int(row[0])
            except ValueError:
                continue
        try:
            n = int(row[hdr.index("Instructions Executed")] or 0)
            sm = int(row[hdr.index("Warp Stall Sampling (All Samples)")] or 0)
        except ValueError:
            continue
        addr, sass = row[2], row[3]
        sec, rank = "other (address arithmetic, setup)", 1000
        if cur_file and cur_file.endswith("sm_100_rt.hpp"):
            # __ffma2_rn / __fadd2_rn / __fmul2_rn: the packed arithmetic of the hot loop's lambdas
            sec, rank = "packed fp32 arithmetic of the steps (FFMA2 / FADD2 / FMUL2: distances, scale, blend, readout)", 50
        elif cur_file and cur_file.endswith("sm_32_intrinsics.hpp"):
            sec, rank = "read-only loads (__ldg: query rows, parameters, gate tables)", 51
        for name, suffix, lo, hi in ranges:
            if cur_file and cur_file.endswith(suffix) and lo <= (line_no or -1) < hi:
                sec, rank = name, generic.get(name, prio[name])
                break
        old = per_addr.get(addr)
        if old is None or rank < old[0]:
            per_addr[addr] = (rank, sec, n, sm, sass)
    inst = collections.Counter()
    samp = collections.Counter()
    fma = collections.Counter()
    for rank, sec, n, sm, sass in per_addr.values():
        inst[sec] += n
        samp[sec] += sm
        if re.search(r"\b(FFMA2|FADD2|FMUL2|FFMA|FADD|FMUL)\b", sass):
            fma[sec] += n
    ti, ts = sum(inst.values()) or 1, sum(samp.values()) or 1
    with open(dst, "w") as f:
        f.write("# Source-section itemisation\n\nKernel: `%s`\n\nFrom `%s` (ncu --set full, --import-source on; warp-level "
                "instructions executed and warp-stall samples per source section; FP32 = FFMA/FADD/FMUL, packed or "
                "scalar).\n\n| section | warp instructions | share | FP32-pipe instr. | stall samples | share |\n"
                "|---|---:|---:|---:|---:|---:|\n" % (kernel, rep))
        for sec, n in sorted(inst.items(), key=lambda kv: -kv[1]):
            f.write("| %s | %d | %.1f%% | %d | %d | %.1f%% |\n" % (sec, n, 100.0 * n / ti, fma[sec], samp[sec],
                                                                100.0 * samp[sec] / ts))
        f.write("| **total** | %d | 100%% | %d | %d | 100%% |\n" % (ti, sum(fma.values()), ts))
    print(open(dst).read())


if __name__ == "__main__":
    main()
