#!/usr/bin/env python
"""bench.py -- IRBFN queries/s on B200 (BASELINE.json metric), one JSON line on stdout.

Workload (config.workload = "C2"): the Frenet-NMPC IRBFN forward over a dense 2^20-point state
lattice (BASELINE.json configs[1]): module shape R,K,D,O = 12,100,8,2,
gaussian basis, parameters of scripts/ckpts/dnmpc_12regions_frenet_l1_bigdata/checkpoint_9500
(carried by tests/golden/c2_frenet_12x100.npz), lattice ey 32 x epsi 32 x vx_car 16 x wz 8 x
delta 4 x vy_car 2 over the config's bounds, vx_goal and curv held (irbfn_vs_explicit_frenet.ipynb
cell 8 holds its other dims the same way).

A "step" is one forward pass over the whole lattice.  At N > 1 every rank owns its own 2^20
lattice (vx_goal differs per rank), no data-path collective: weak scaling.

  value     queries/s, inputs resident in HBM, one kernel launch per step, CUDA-event timed
  e2e       the same through the public host-buffer call (WCRBFNet.apply on pinned host memory):
            host->device copy of the queries and device->host copy of the outputs inside the
            timed region
  roofline  the kernel is bound by the SM FP32 pipe (arithmetic intensity ~800 flop/byte; see
            DESIGN.md): achieved = algorithmic FP32-pipe lane-operations / s against the lane-op
            peak measured in this run by csrc/fp32_peak.cu; the flop and HBM views ride along
  cpu_baseline  the C/OpenMP float32 oracle (oracle/irbfn_oracle.c) on the host cores

--impl reference times the CPU implementation of the path (the oracle port: JAX, Flax and
flax_rbf are not installable here) on the same workload.
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

FIXTURE = os.path.join(ROOT, "tests", "golden", "c2_frenet_12x100.npz")
LATTICE = (("ey", 0, 32), ("epsi", 6, 32), ("vx_car", 2, 16), ("wz", 5, 8), ("delta", 1, 4), ("vy_car", 3, 2))
B_LATTICE = 1 << 20
CPU_SAMPLE = 1 << 18      # stride sample of the lattice both CPU legs (cpu_baseline, --impl reference) time
METRIC = "irbfn_forward_queries_per_s"
UNIT = "queries/s"


def load_c2():
    z = np.load(FIXTURE)
    cfg = json.loads(str(z["cfg_json"]))
    p = {"rbf_list": {"centers": z["p/rbf_list/centers"], "log_sigs": z["p/rbf_list/log_sigs"]},
         "linear": {"kernel": z["p/linear/kernel"], "bias": z["p/linear/bias"]}}
    return cfg, p


def lattice(cfg, vx_goal=4.0, curv=0.0, jitter=0.0):
    """(2^20, 8) float32: columns (ey, delta, vx_car, vy_car, vx_goal, wz, epsi, curv)."""
    lo = [min(b) for b in cfg["lower_bounds"]]
    hi = [max(b) for b in cfg["upper_bounds"]]
    axes = [np.linspace(lo[c], hi[c], n) + jitter * (hi[c] - lo[c]) / n for _, c, n in LATTICE]
    grids = np.meshgrid(*axes, indexing="ij")
    x = np.empty((B_LATTICE, 8), np.float32)
    x[:, 4] = vx_goal
    x[:, 7] = curv
    for (_, c, _), g in zip(LATTICE, grids):
        x[:, c] = g.reshape(-1)
    return x


def _ncu_traffic():
    """dram__bytes_read.sum + dram__bytes_write.sum of one forward launch from the committed ncu --set full
    capture (profiles/traffic_fwd.json), or None."""
    try:
        t = json.load(open(os.path.join(ROOT, "profiles", "traffic_fwd.json")))
        return int(t["dram_bytes_read"]) + int(t["dram_bytes_write"])
    except Exception:
        return None


def algorithmic_counts(cfg):
    """Per-query algorithmic work."""
    R, K, D, O = cfg["num_regions"], cfg["num_kernels"], cfg["in_features"], cfg["out_features"]
    nb = sum(len(b) for b in cfg["lower_bounds"])
    S = len(cfg["activation_idx"])
    flops = R * K * (3 * D + 2) + (2 * K * O + O) + (10 * nb + R * (S - 1))
    lane_ops = R * K * (2 * D + 2) + K * O          # FP32-pipe thread-instructions, direct form
    mufu = R * K + 4 * nb
    hbm_bytes = 4 * (D + O)
    return dict(flops=flops, lane_ops=lane_ops, mufu=mufu, hbm_bytes=hbm_bytes)


class ClockSampler(threading.Thread):
    """Samples SM clock and throttle reasons through NVML while the timed region runs."""

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.stop_flag, self.samples, self.reasons, self.max_mhz = index, False, [], set(), None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception:
            self.nv = None

    def run(self):
        if self.nv is None:
            return
        nv = self.nv
        names = {"hw_slowdown": nv.nvmlClocksThrottleReasonHwSlowdown,
                 "hw_thermal_slowdown": nv.nvmlClocksThrottleReasonHwThermalSlowdown,
                 "sw_thermal_slowdown": nv.nvmlClocksThrottleReasonSwThermalSlowdown,
                 "sw_power_cap": nv.nvmlClocksThrottleReasonSwPowerCap}
        while not self.stop_flag:
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for k, bit in names.items():
                    if r & bit:
                        self.reasons.add(k)
            except Exception:
                pass
            time.sleep(0.005)

    def summary(self):
        return {"sm_mhz": float(np.median(self.samples)) if self.samples else None,
                "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons), "samples": len(self.samples)}


_REAL_STDOUT = None


def quiet_stdout():
    """Send everything libraries print on fd 1 (NCCL prints its version banner there when NCCL_DEBUG is set
    on the box) to stderr, so that stdout carries exactly one line: the JSON result written by emit()."""
    global _REAL_STDOUT
    if _REAL_STDOUT is None:
        sys.stdout.flush()
        _REAL_STDOUT = os.dup(1)
        os.dup2(2, 1)


def emit(obj):
    line = (json.dumps(obj) + "\n").encode()
    if _REAL_STDOUT is None:
        sys.stdout.write(line.decode())
        sys.stdout.flush()
    else:
        sys.stdout.flush()
        os.write(_REAL_STDOUT, line)


def run_reference(args):
    """CPU arm: the C/OpenMP oracle port on all host cores, bounded sample per step."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from oracle import c_oracle
    cfg, p = load_c2()
    sample = CPU_SAMPLE
    x = lattice(cfg)[:: B_LATTICE // sample].copy()
    # torchrun exports OMP_NUM_THREADS=1; the CPU arm uses every core this process may run on
    cores = len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1)
    run = lambda: c_oracle.forward(x, p["rbf_list"]["centers"], p["rbf_list"]["log_sigs"], p["linear"]["kernel"],
                                   p["linear"]["bias"], 0, cfg=cfg, nthreads=cores)
    for _ in range(args.warmup):
        run()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        run()
    dt = (time.perf_counter() - t0) / args.steps
    v = sample / dt
    emit({
        "impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt * 1e3, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": "C2", "shape": "R12_K100_D8_O2_gaussian", "queries_per_step": sample,
                   "note": "CPU arm: oracle/irbfn_oracle.c (float32, OpenMP) -- JAX/Flax/flax_rbf are not "
                           "installable here, so the reference's own JAX CPU path cannot run"},
        "cpu_baseline": {"value": v, "unit": UNIT, "cores": cores, "kind": "port",
                         "sample": "%d-point stride sample of the 2^20 lattice per step" % sample},
        "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    })


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--no-extras", action="store_true", help="skip the train-step / input-gradient side measurements")
    ap.add_argument("--no-other", action="store_true", help="skip the C1 / C2b / C5 side measurements")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3)
    quiet_stdout()
    if args.impl == "reference":
        return run_reference(args)

    import torch
    from irbfn_b200 import ops, parallel, peak
    from irbfn_b200.basis import gaussian
    from irbfn_b200.model import WCRBFNet, device_params

    rank, local_rank, world = parallel.init_distributed()
    dev = torch.device("cuda", local_rank)
    torch.cuda.set_device(dev)
    cfg, p_np = load_c2()
    net = WCRBFNet(cfg["in_features"], cfg["out_features"], cfg["num_kernels"], gaussian, cfg["num_regions"],
                   cfg["lower_bounds"], cfg["upper_bounds"], cfg["dimension_ranges"], cfg["activation_idx"],
                   cfg["delta"])
    params = device_params({"params": p_np}, dev)
    B = B_LATTICE
    counts = algorithmic_counts(cfg)

    # rotating input sets: 6 x 33.5 MB > 126 MB of L2, so no step finds its queries in cache
    NSETS = 6
    xs = [torch.from_numpy(lattice(cfg, vx_goal=4.0 + 0.25 * rank, jitter=0.1 * i)).to(dev) for i in range(NSETS)]
    ys = [torch.empty((B, cfg["out_features"]), device=dev) for _ in range(NSETS)]
    p = params["params"]
    gd = net._gate_pack(dev)
    fwd = lambda i: ops.forward(xs[i % NSETS], p["rbf_list"]["centers"], p["rbf_list"]["log_sigs"],
                                p["linear"]["kernel"], p["linear"]["bias"], gd=gd, basis=0, out=ys[i % NSETS])

    def timed(fn, steps, warmup):
        for i in range(warmup):
            fn(i)
        parallel.barrier()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for i in range(steps):
            fn(i)
        e1.record()
        torch.cuda.synchronize()
        parallel.barrier()
        return parallel.allreduce_max_scalar(e0.elapsed_time(e1) / steps, dev)

    # ---- device-resident throughput ----
    sampler = ClockSampler(local_rank)
    sampler.start()
    ms = timed(fwd, args.steps, args.warmup)
    sampler.stop_flag = True
    sampler.join()
    value = world * B / (ms * 1e-3)

    # ---- end to end through the public host-buffer call ----
    xh = torch.from_numpy(lattice(cfg, vx_goal=4.0 + 0.25 * rank)).pin_memory()
    yh = torch.empty((B, cfg["out_features"]), dtype=torch.float32).pin_memory()
    # apply_host(streamed="auto"): the chunk pipeline (a launch and a download per chunk) or ONE persistent launch whose
    # query tiles wait for their piece of the upload and which stores y straight into the pinned host buffer
    # (irbfn_forward_streamed) -- which of the two wins depends on the box's host link, so the call times both once
    # (outside the timed region: the warm-up steps) and keeps the faster; both fixed choices are reported beside it
    e2e_mode, e2e_errors = ["auto"], []

    def e2e_step(i):        # a launch that gave up on its upload is raised one call later: finish the run on chunks
        try:
            net.apply_host(params, xh, out=yh, streamed=e2e_mode[0])
        except RuntimeError as exc:
            e2e_errors.append(str(exc))
            e2e_mode[0] = False
            net.apply_host(params, xh, out=yh, streamed=False)
    ms_e2e = timed(e2e_step, args.steps, args.warmup)
    e2e = {"value": world * B / (ms_e2e * 1e-3), "unit": UNIT, "ms_per_step": ms_e2e,
           "h2d_bytes_per_step": int(xh.numel() * 4), "d2h_bytes_per_step": int(yh.numel() * 4),
           "ms_per_step_chunk_pipeline": timed(lambda i: net.apply_host(params, xh, out=yh, streamed=False),
                                               args.steps, args.warmup)}
    if e2e_errors:
        e2e["error_during_timed_region"] = e2e_errors[0]
    if ops.upload_streamed_available():
        e2e["auto_chose"] = "streamed" if net._const.get(("auto", str(dev), B)) else "chunks"
        gave_up = []

        def streamed_step(i):           # errors are kept out of timed(): its barriers must run on every rank
            try:
                net.apply_host(params, xh, out=yh, streamed=True)
            except RuntimeError as exc:  # under a profiler that serialises the kernel against its own uploads
                gave_up.append(str(exc))
        ms_streamed = timed(streamed_step, args.steps, args.warmup)
        try:
            net._const[("spipe", str(dev), None)].check()
        except RuntimeError as exc:
            gave_up.append(str(exc))
        e2e["ms_per_step_streamed_single_launch"] = None if gave_up else ms_streamed
        if gave_up:
            e2e["streamed_single_launch_error"] = gave_up[0]
    # the same sweep through the lattice entry point: the queries are generated on the device from the axis
    # values (a few hundred bytes up), the outputs still come down to pinned host memory every step.  Reported
    # beside the generic host-buffer number, which stays the e2e headline.
    lo_ = [min(b) for b in cfg["lower_bounds"]]
    hi_ = [max(b) for b in cfg["upper_bounds"]]
    axes_ = {c: np.linspace(lo_[c], hi_[c], n).astype(np.float32) for _, c, n in LATTICE}
    fixed_ = np.zeros(8, np.float32)
    fixed_[4], fixed_[7] = 4.0 + 0.25 * rank, 0.0

    def lattice_step(i):
        _, yl = net.apply_lattice(params, axes_, fixed_)
        yh.copy_(yl, non_blocking=True)
    ms_lat_dl = timed(lattice_step, args.steps, args.warmup)
    # the same with y stored by the kernel straight into the pinned host buffer (no download pass)
    ms_lat = timed(lambda i: net.apply_lattice(params, axes_, fixed_, out=yh), args.steps, args.warmup)
    e2e["lattice_api"] = {"value": world * B / (ms_lat * 1e-3), "ms_per_step": ms_lat,
                          "ms_per_step_with_download_pass": ms_lat_dl,
                          "h2d_bytes_per_step": int(sum(v.size for v in axes_.values()) * 4 + 32 + 64),
                          "d2h_bytes_per_step": int(yh.numel() * 4),
                          "note": "WCRBFNet.apply_lattice(out=pinned): lattice built on the device (irbfn_lattice_queries), y stored to host memory by the kernel"}

    out = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": "C2", "shape": "R12_K100_D8_O2_gaussian", "queries_per_step_per_gpu": B,
                   "params": "scripts/ckpts/dnmpc_12regions_frenet_l1_bigdata/checkpoint_9500 (tests/golden fixture)",
                   "l2": "%d rotating input sets (%.0f MB) > 126 MB L2" % (NSETS, NSETS * B * 32 / 1e6),
                   "parallelism": "dp%d (queries sharded, no collective)" % world},
        "e2e": e2e, "gpu_launches": args.steps, "clocks": sampler.summary(),
    }

    if rank == 0:
        # ---- roofline of irbfn_fwd_kernel (the only kernel of the step) ----
        peak_lane, _, _ = peak.fp32_microbench(2, iters=20000)    # FADD2+FFMA2: the distance loop's pair
        peak_ffma, _, _ = peak.fp32_microbench(0, iters=20000)
        _, peak_mufu, _ = peak.fp32_microbench(3, iters=5000)
        t = ms * 1e-3
        ach_lane = counts["lane_ops"] * B / t
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        except Exception:
            pass
        hbm_peak = peaks.get("hbm_gbs", 6650.0)
        out["roofline"] = {
            "bound": "fp32_pipe", "achieved": ach_lane / 1e12, "peak": peak_lane / 1e12, "unit": "T lane-op/s",
            "frac": ach_lane / peak_lane, "traffic": _ncu_traffic(),
            "kernel": "irbfn_fwd_kernel<8,2,gaussian>", "launch_ms": ms,
            "note": "FP32-pipe lane-operations: R*K*(2D+2) + K*O per query (direct-form distance, scale, blend, "
                    "readout); peak = FADD2+FFMA2 micro-benchmark measured in this run",
            "flops": {"achieved_tflops": counts["flops"] * B / t / 1e12, "peak_tflops": 2 * peak_ffma / 1e12,
                      "frac": counts["flops"] * B / t / (2 * peak_ffma)},
            "mufu": {"achieved": counts["mufu"] * B / t / 1e12, "peak": peak_mufu / 1e12,
                     "frac": counts["mufu"] * B / t / peak_mufu, "unit": "T op/s"},
            "hbm": {"achieved_gbs": counts["hbm_bytes"] * B / t / 1e9, "peak_gbs": hbm_peak,
                    "frac": counts["hbm_bytes"] * B / t / 1e9 / hbm_peak,
                    "peak_source": "measured" if "hbm_gbs" in peaks else "fallback"},
        }

        tail = {}
        if not args.no_extras:
            out["extras"], tail = extras(net, params, cfg, dev, world, args)

        if world == 1 and not args.no_cpu:
            from oracle import c_oracle
            sample = CPU_SAMPLE
            xs_cpu = xs[0][:: B // sample].cpu().numpy().copy()
            cores = len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1)
            run = lambda: c_oracle.forward(xs_cpu, p_np["rbf_list"]["centers"], p_np["rbf_list"]["log_sigs"],
                                           p_np["linear"]["kernel"], p_np["linear"]["bias"], 0, cfg=cfg, nthreads=cores)
            y_cpu = run()
            t0 = time.perf_counter()
            reps = 0
            while reps < 3 or time.perf_counter() - t0 < 10.0:
                run()
                reps += 1
            dt = (time.perf_counter() - t0) / reps
            y_gpu = ys[0][:: B // sample].cpu().numpy()
            out["cpu_baseline"] = {"value": sample / dt, "unit": UNIT, "cores": cores, "kind": "port",
                                   "sample": "%d-point stride sample of the 2^20 lattice, %d passes of "
                                             "oracle/irbfn_oracle.c (float32, OpenMP)" % (sample, reps)}
            out["parity_check"] = {"max_abs_err_over_max_abs": float(np.abs(y_gpu - y_cpu).max() / np.abs(y_cpu).max()),
                                   "against": "C oracle float32, same %d queries" % sample}
        out.update(tail)      # compact c3 / c5 / dp_parity objects LAST: they survive a truncated tail of the line
        if tail:
            # ... and one flat summary of them at the very end (a 1,500-character tail of the line always holds it)
            r4 = lambda v: None if v is None else float("%.4g" % v)
            c3, o10, c1s, c4, c5, dp = (tail.get(k) or {} for k in ("c3", "c3_o10_fullint", "c3_c1shape", "c4", "c5", "dp_parity"))
            sp = lambda d: (d.get("gate_sparse_step") or {}).get("ms_graph")
            out["summary"] = {
                "n_gpus": world, "fwd_ms": r4(ms), "e2e_ms": r4(ms_e2e),
                "c3_ms_graph": r4(c3.get("ms_graph")), "c3_ms_graph_nccl": r4(c3.get("ms_graph_nccl")),
                "c3_ms_graph_gate_sparse": r4(sp(c3)), "c3_ms_kernels_this_rank": r4(c3.get("ms_kernels_this_rank")),
                "c3_ms_nccl_allreduce_alone": r4(c3.get("ms_nccl_allreduce_alone")), "c3_allreduce": c3.get("allreduce"),
                "dp_parity_rel_l2": r4(dp.get("rel_l2_max_over_leaves")), "dp_parity_ok": dp.get("ok"),
                "c1shape_ms_graph": r4(c1s.get("ms_graph")), "o10_ms_graph": r4(o10.get("ms_graph")),
                "o10_ms_graph_gate_sparse": r4(sp(o10)), "c4_ms": r4(c4.get("ms_vjp_dx")),
                "c4_ms_gate_sparse": r4(c4.get("ms_vjp_dx_gate_sparse")), "c5_fwd_ms_per_2p24": r4(c5.get("fwd_ms_per_2p24")),
                "c5_fwd_ms_gate_sparse": r4(c5.get("fwd_gate_sparse_ms_per_2p24")),
                "c5_fwd_vjp_ms_per_2p22": r4(c5.get("fwd_vjp_ms_per_2p22")),
                "c5_fwd_vjp_ms_gate_sparse": r4(c5.get("fwd_vjp_gate_sparse_ms_per_2p22"))}
            pr = (out.get("extras") or {}).get("pred_step_latency") or {}
            out["summary"].update({"pred_step_b1_us_graph": r4(pr.get("kernel_us_in_cuda_graph_B1")),
                                   "pred_step_b1_us_apply_numpy": r4(pr.get("module_apply_numpy_us_wall_B1")),
                                   "pred_step_b400_us_graph": r4(pr.get("kernel_us_in_cuda_graph_B400")),
                                   "pred_step_b400_us_apply_numpy": r4(pr.get("module_apply_numpy_us_wall_B400"))})
        emit(out)
    elif not args.no_extras:
        extras(net, params, cfg, dev, world, args)   # collectives inside: every rank takes part
    if world > 1:
        # every rank is done: leave without running destructors that talk to the (soon gone) process-group
        # store -- torch symmetric-memory handles did hang dist.destroy_process_group() at exit on this pool
        import torch.distributed as dist
        torch.cuda.synchronize()
        dist.barrier()
        sys.stdout.flush()
        sys.stderr.flush()
        os._exit(0)


def extras(net, params, cfg, dev, world, args):
    """Side measurements that explain the headline: the data-parallel training step (C3) and the
    input-gradient sweep (C4).  Same timing discipline; reported under "extras"."""
    import torch
    from irbfn_b200 import ops, parallel
    from irbfn_b200.train import TrainState, grad_step, train_step
    rank = parallel.env_world()[0]
    res, tail = {}, {}
    p = params["params"]
    R, K, D, O = cfg["num_regions"], cfg["num_kernels"], cfg["in_features"], cfg["out_features"]
    lo = np.array([min(b) for b in cfg["lower_bounds"]], np.float32)
    hi = np.array([max(b) for b in cfg["upper_bounds"]], np.float32)

    def timed(fn, steps, warmup):
        for _ in range(warmup):
            fn()
        parallel.barrier()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(steps):
            fn()
        e1.record()
        torch.cuda.synchronize()
        return parallel.allreduce_max_scalar(e0.elapsed_time(e1) / steps, dev)

    # gate-sparse, region-binned forward on the headline lattice: same queries, same parameters,
    # every output within tol_abs of the dense kernel; the dense launch stays the headline and the roofline
    xl = torch.from_numpy(lattice(cfg, vx_goal=4.0 + 0.25 * rank)).to(dev)
    yd = ops.forward(xl, p["rbf_list"]["centers"], p["rbf_list"]["log_sigs"], p["linear"]["kernel"], p["linear"]["bias"],
                     gd=net._gate_pack(dev), basis=0)
    yscale = float(yd.abs().max())
    cnt = torch.zeros(2, dtype=torch.int64, device=dev)
    ysp = torch.empty_like(yd)
    run_sp = lambda: ops.forward_sparse(xl, p["rbf_list"]["centers"], p["rbf_list"]["log_sigs"], p["linear"]["kernel"],
                                        p["linear"]["bias"], gd=net._gate_pack(dev), basis=0, tol_abs=1e-6 * yscale,
                                        out=ysp, counters=cnt)
    ms_sp = timed(run_sp, args.steps, args.warmup)
    ms_dn = timed(lambda: ops.forward(xl, p["rbf_list"]["centers"], p["rbf_list"]["log_sigs"], p["linear"]["kernel"],
                                      p["linear"]["bias"], gd=net._gate_pack(dev), basis=0, out=yd), args.steps, args.warmup)
    sk, dn = (int(v) for v in cnt.cpu())
    res["forward_C2_gate_sparse"] = {"queries_per_gpu": B_LATTICE, "ms_sparse_incl_binning": ms_sp, "ms_dense": ms_dn,
                                     "speedup": ms_dn / ms_sp, "tol_abs_over_max_abs_y": 1e-6,
                                     "max_abs_diff_over_max_abs_y": float((ysp - yd).abs().max()) / yscale,
                                     "skipped_fraction": sk / max(1, sk + dn),
                                     "queries_per_s": world * B_LATTICE / (ms_sp * 1e-3)}
    del xl, yd, ysp

    # C3: 262,144 samples per step in total (strong scaling: the global batch is fixed)
    GB = 262144
    a, b = parallel.shard_range(GB, rank, world)
    g = torch.Generator().manual_seed(7)
    xall = torch.rand((GB, D), generator=g) * torch.from_numpy(hi - lo) + torch.from_numpy(lo)
    yall = torch.randn((GB, O), generator=g)
    x, y = xall[a:b].to(dev).contiguous(), yall[a:b].to(dev).contiguous()
    state = TrainState(net, params, lr=1e-3, max_grad_norm=1.0, device=dev)
    # the gradient all-reduce: one-shot kernel over NVLink peer memory fused with clip + Adam when the node allows
    # it (irbfn_fused_update); a second state keeps the plain NCCL path for comparison
    how = state.enable_peer_allreduce()
    ms_g = timed(lambda: grad_step(state, x, y, GB), args.steps, args.warmup)
    ms_t = timed(lambda: train_step(state, x, y, GB), args.steps, args.warmup)
    from irbfn_b200.train import GraphedTrainStep
    gstep = GraphedTrainStep(state, x.shape[0], GB, loss="l2").load(x, y)
    ms_graph = timed(lambda: gstep(), args.steps, args.warmup)
    ms_t_nccl = ms_graph_nccl = None
    if how.startswith("peer"):
        st_n = TrainState(net, params, lr=1e-3, max_grad_norm=1.0, device=dev)
        ms_t_nccl = timed(lambda: train_step(st_n, x, y, GB), args.steps, args.warmup)
        gstep_n = GraphedTrainStep(st_n, x.shape[0], GB, loss="l2").load(x, y)
        ms_graph_nccl = timed(lambda: gstep_n(), args.steps, args.warmup)
        del gstep_n, st_n
    # the same step gate-sparse (gate_tol = 2^-24): forward (irbfn_forward_sparse_gate: binned samples, regions skipped
    # per warp) and parameter gradients (irbfn_backward_sparse: a ballot per 32 samples) leave out the (sample, region)
    # pairs whose gate cannot move an output or a gradient
    GATE_TOL = 2.0 ** -24
    # (the forward stays dense here: on uniformly drawn samples at 12 regions its gate-sparse variant gains less than
    # its binning pass and the backward's own gate pass cost -- 0.527 against 0.514 ms; it pays at 27 regions, below)
    st_s = TrainState(net, params, lr=1e-3, max_grad_norm=1.0, device=dev, gate_tol=GATE_TOL)
    st_s.enable_peer_allreduce()
    ms_t_sp = timed(lambda: train_step(st_s, x, y, GB), args.steps, args.warmup)
    gstep_s = GraphedTrainStep(st_s, x.shape[0], GB, loss="l2").load(x, y)
    ms_graph_sp = timed(lambda: gstep_s(), args.steps, args.warmup)
    st_s.flat.copy_(state.flat)
    grad_step(st_s, x, y, GB)
    grad_step(state, x, y, GB)
    sp_dev = max(float((st_s.gviews[k].double() - state.gviews[k].double()).norm() / state.gviews[k].double().norm())
                 for k in state.gviews)
    del gstep_s, st_s
    # where the step's time goes: this rank's kernels alone (no collective: global batch = shard), the
    # NCCL all-reduce of the flat gradient alone, and the rest (launch gaps, rank skew)
    st_l = TrainState(net, params, lr=1e-3, max_grad_norm=1.0, device=dev)
    ms_local = timed(lambda: grad_step(st_l, x, y, x.shape[0]), args.steps, args.warmup)
    ms_ar = timed(lambda: parallel.allreduce_sum_(st_l._grad_ext), args.steps, args.warmup) if world > 1 else 0.0
    del st_l
    fl = R * K * (3 * D + 2) + 2 * K * O + R * K * (5 * D + 9) + 4 * K * O
    lane_step = R * K * (2 * D + 2) + K * O + R * K * (3 * D + 5) + 2 * K * O
    res["train_step_C3"] = {"global_batch": GB, "scaling": "strong", "ms_fwd_vjp_allreduce": ms_g,
                            "ms_with_clip_adam": ms_t, "ms_with_clip_adam_cuda_graph": ms_graph,
                            "samples_per_s": GB / (min(ms_t, ms_graph) * 1e-3),
                            "algorithmic_tflops": fl * GB / (ms_g * 1e-3) / 1e12,
                            "allreduce_bytes": int(state._grad_ext.numel() * 4), "allreduce": how,
                            "ms_with_clip_adam_nccl": ms_t_nccl, "ms_with_clip_adam_cuda_graph_nccl": ms_graph_nccl}
    tail["c3"] = {"shape": "R12_K100_D8_O2", "global_batch": GB, "n_gpus": world, "ms_eager": ms_t, "ms_graph": ms_graph,
                  "allreduce": how, "ms_graph_nccl": ms_graph_nccl,
                  "ms_kernels_this_rank": ms_local, "ms_nccl_allreduce_alone": ms_ar,
                  "ms_fwd_vjp_allreduce": ms_g, "allreduce_bytes": int(state._grad_ext.numel() * 4),
                  "lane_ops_T_per_s": lane_step * GB / (min(ms_t, ms_graph) * 1e-3) / 1e12,
                  "gate_sparse_step": {"gate_tol": GATE_TOL, "forward": "dense", "ms_eager": ms_t_sp,
                                       "ms_graph": ms_graph_sp,
                                      "grad_rel_l2_vs_dense_max_over_leaves": sp_dev}}

    # 8(e) result check on hardware: the NCCL-reduced gradient of the sharded step against the gradient of
    # the whole batch on one device (rank 0 holds the full batch too); per-leaf ||g - g1|| / ||g1||
    grad_step(state, x, y, GB)
    torch.cuda.synchronize()
    if rank == 0:
        full = TrainState(net, params, lr=1e-3, max_grad_norm=1.0, device=dev)
        full.flat.copy_(state.flat)       # the timed optimiser steps above moved the parameters
        if world > 1:
            grad_step(full, xall.to(dev), yall.to(dev), GB)
            got, how = state, "%s all-reduce of %d shard gradients vs the full batch on rank 0" % (how.split(":")[0], world)
        else:
            # one device: two half-batch shard gradients (means over the global batch) summed vs the full batch
            h = GB // 2
            halves = TrainState(net, params, lr=1e-3, max_grad_norm=1.0, device=dev)
            halves.flat.copy_(state.flat)
            acc = torch.zeros_like(halves.grad)
            for lo_, hi_ in ((0, h), (h, GB)):
                grad_step(halves, xall[lo_:hi_].to(dev), yall[lo_:hi_].to(dev), GB)   # world 1: the all-reduce is a no-op
                acc += halves.grad
            halves.grad.copy_(acc)
            grad_step(full, xall.to(dev), yall.to(dev), GB)
            got, how = halves, "two half-batch shard gradients summed vs the full batch (one device: no collective)"
        worst = 0.0
        for k in full.gviews:
            ref = full.gviews[k].double()
            worst = max(worst, float((got.gviews[k].double() - ref).norm() / ref.norm()))
        tail["dp_parity"] = {"n_gpus": world, "rel_l2_max_over_leaves": worst, "tol": 1e-4, "ok": bool(worst <= 1e-4),
                             "how": how}
        del full

    # the same step on the other shipped training shapes: the Cartesian C1 shape, and the
    # five-step Frenet trainer's shape (train_step_fullint needs O = 10; shipped 27 x 500 x 8 -> 10)
    for tag, fixture, loss in (("c3_c1shape", "c1_cartesian_1x1000.npz", "l2"),
                               ("c3_o10_fullint", "warmstart_27x500_o10.npz", "frenet_fullint")):
        tail[tag] = train_case(fixture, loss, GB, dev, rank, world, args, timed)

    # C4: input gradient for 1M queries per GPU (weak), dx of sum(y)
    B4 = 1000000
    x4 = (torch.rand((B4, D), generator=g) * torch.from_numpy(hi - lo) + torch.from_numpy(lo)).to(dev)
    dy4 = torch.ones((B4, O), device=dev)
    gd = net._gate_pack(dev)
    run4 = lambda: ops.backward(x4, p["rbf_list"]["centers"], p["rbf_list"]["log_sigs"], p["linear"]["kernel"],
                                dy4, gd=gd, basis=0, grad_mask=ops.GRAD_X)
    ms4 = timed(run4, max(3, args.steps // 2), 3)
    # gate-sparse: queries binned by their non-negligible gate intervals, regions skipped per warp (gate_tol = 2^-24)
    dx_dense = run4()["x"].clone()
    run4s = lambda: ops.backward(x4, p["rbf_list"]["centers"], p["rbf_list"]["log_sigs"], p["linear"]["kernel"],
                                 dy4, gd=gd, basis=0, grad_mask=ops.GRAD_X, gate_tol=2.0 ** -24)
    ms4s = timed(run4s, max(3, args.steps // 2), 3)
    dev4 = float((run4s()["x"] - dx_dense).norm() / dx_dense.norm())
    del dx_dense
    res["input_grad_C4"] = {"queries_per_gpu": B4, "scaling": "weak", "ms_vjp_dx": ms4,
                            "queries_per_s": world * B4 / (ms4 * 1e-3),
                            "algorithmic_tflops": (R * K * (5 * D + 7) + 2 * K * O) * B4 / (ms4 * 1e-3) / 1e12,
                            "gate_sparse": {"gate_tol": 2.0 ** -24, "ms_vjp_dx_incl_binning": ms4s,
                                            "rel_l2_vs_dense": dev4}}
    tail["c4"] = {"queries_per_gpu": B4, "ms_vjp_dx": ms4, "ms_vjp_dx_gate_sparse": ms4s, "n_gpus": world}
    # 8f rank 2: one-pass Jacobian dy/dx (B,O,D) on the C4 queries, against O one-hot VJP launches (jacrev)
    W_, b_ = p["linear"]["kernel"], p["linear"]["bias"]
    msj = timed(lambda: ops.jacobian(x4, p["rbf_list"]["centers"], p["rbf_list"]["log_sigs"], W_, gd=gd, basis=0),
                max(3, args.steps // 2), 3)
    msjs = timed(lambda: ops.jacobian(x4, p["rbf_list"]["centers"], p["rbf_list"]["log_sigs"], W_, gd=gd, basis=0,
                                      gate_tol=2.0 ** -24), max(3, args.steps // 2), 3)
    res["jacobian_8f2"] = {"queries_per_gpu": B4, "ms_one_pass": msj, "ms_one_pass_gate_sparse": msjs,
                           "ms_jacrev_equivalent": O * ms4,
                           "queries_per_s": world * B4 / (msj * 1e-3),
                           "lane_ops_T_per_s": (R * K * (2 * D + 3 + O * (D + 2))) * B4 / (msj * 1e-3) / 1e12}
    # 8f rank 1: the same C3 step with the reference's Frenet one-step dynamics loss fused in
    ms_d = timed(lambda: grad_step(state, x, y, GB, loss="frenet_oneint"), args.steps, args.warmup)
    res["train_step_C3_frenet_dynamics_loss"] = {"global_batch": GB, "ms_fwd_loss_vjp_allreduce": ms_d,
                                                 "samples_per_s": GB / (ms_d * 1e-3)}
    xl, yl, ypl = x4[:, :8].contiguous(), torch.randn((B4, 10), device=dev), torch.randn((B4, 10), device=dev)
    ms_l = timed(lambda: ops.dynamics_loss(1, xl, ypl, yl, state.dyn_params), args.steps, 3)
    res["dynamics_loss_fullint_8f1"] = {"samples": B4, "ms": ms_l, "hbm_GBps": B4 * 4 * (8 + 30) / (ms_l * 1e-3) / 1e9,
                                        "note": "five-step Frenet rollout x2 + adjoint, 152 B/sample"}
    # 8f rank 3: on-device input pipeline -- one gather launch per step instead of a host gather + H2D copy
    NT_ = 1 << 22
    tabx, taby = torch.rand((NT_, 8), device=dev), torch.rand((NT_, 10), device=dev)
    perm_ = torch.randperm(NT_, device=dev)[:GB]
    gx, gy = torch.empty((GB, 8), device=dev), torch.empty((GB, 10), device=dev)
    ms_ga = timed(lambda: ops.gather_rows(tabx, taby, perm_, x_out=gx, y_out=gy), args.steps, 3)
    res["gather_rows_8f3"] = {"rows": GB, "table_rows": NT_, "ms": ms_ga,
                              "hbm_GBps": GB * (2 * 72 + 8) / (ms_ga * 1e-3) / 1e9,
                              "note": "per row: 8 B index + 72 B read + 72 B written"}
    # 8f rank 4: the explicit side of the irbfn_vs_explicit sweep -- brute-force nearest table row
    NTAB = 1 << 18
    tab_ = (torch.rand((NTAB, D), device=dev) * torch.from_numpy(hi - lo).to(dev) + torch.from_numpy(lo).to(dev)).contiguous()
    qn = x4[: 1 << 20].contiguous()
    ms_nn = timed(lambda: ops.nearest_row(qn, tab_), 2, 1)
    res["nearest_row_8f4"] = {"queries_per_gpu": 1 << 20, "table_rows": NTAB, "ms": ms_nn,
                              "pairs_per_s": world * (1 << 20) * NTAB / (ms_nn * 1e-3),
                              "lane_ops_T_per_s": (1 << 20) * NTAB * (2 * D) / (ms_nn * 1e-3) / 1e12,
                              "timed_steps": 2, "warmup": 1}
    q400 = x4[:400].contiguous()
    ms_n4 = timed(lambda: ops.nearest_row(q400, tab_), 10, 3)
    res["nearest_row_400_queries"] = {"queries": 400, "table_rows": NTAB, "ms": ms_n4}
    # the notebook's sweep end to end on the device: 20 x 20 lattice -> network, nearest row, row fetch, errors
    from irbfn_b200.sweep import ExplicitTable, irbfn_vs_explicit, lattice_queries
    etab = ExplicitTable(tab_.cpu().numpy(), np.random.default_rng(5).normal(size=(NTAB, O)).astype(np.float32), dev)
    fixed = [0.0, 0.0, 1.076, 0.0, 4.0, 0.0, 0.0, 0.0][:D]
    qlat = torch.from_numpy(lattice_queries({0: np.linspace(0, 2, 20), D - 2: np.linspace(0, 1, 20)}, fixed)).to(dev)
    ms_sw = timed(lambda: irbfn_vs_explicit(net, params, etab, qlat, rank=0, world=1), 10, 3)
    res["sweep_400_lattice_vs_explicit"] = {"queries": 400, "table_rows": NTAB, "ms": ms_sw,
                                            "note": "irbfn_vs_explicit_frenet.ipynb cells 8-13 on one device"}
    del etab
    del tabx, taby, tab_

    # pred_step (src/irbfn_mpc/irbfn_planner.py:29-32): the planner evaluates ONE state per control step, the
    # notebooks 400.  (a) the kernel alone, replayed from a CUDA graph (irbfn_forward_path: the pair-parallel latency
    # kernel, one launch); (b) module.apply on a NumPy float64 batch, wall clock of the whole call with the result
    # back on the host (upload + kernel + download replayed as one captured graph per batch size).
    pred = {"shape": "C2 R12_K100_D8_O2", "path": {}}
    for Bp in (1, 400):
        xp = x4[:Bp].contiguous()
        yp = torch.empty((Bp, O), device=dev)
        fwd_p = lambda: ops.forward(xp, p["rbf_list"]["centers"], p["rbf_list"]["log_sigs"], p["linear"]["kernel"],
                                    p["linear"]["bias"], gd=net._gate_pack(dev), basis=0, out=yp)
        gp_, sp_ = torch.cuda.CUDAGraph(), torch.cuda.Stream(dev)
        with torch.cuda.stream(sp_):
            fwd_p()
            with torch.cuda.graph(gp_, stream=sp_):
                for _ in range(20):
                    fwd_p()
        torch.cuda.synchronize()
        pred["path"]["B%d" % Bp] = ops.forward_path(Bp, R, K, D, O)
        pred["kernel_us_in_cuda_graph_B%d" % Bp] = timed(gp_.replay, 20, 3) / 20 * 1e3
        xn = xp.cpu().numpy().astype(np.float64)
        for _ in range(20):
            net.apply(params, xn)
        t0 = time.perf_counter()
        for _ in range(200):
            net.apply(params, xn)
        pred["module_apply_numpy_us_wall_B%d" % Bp] = (time.perf_counter() - t0) / 200 * 1e6
    res["pred_step_latency"] = pred
    if not args.no_other:
        other = other_configs(dev, world, args, timed)
        tail["c5"] = other.pop("c5_summary")
        res.update(other)
    return res, tail


def train_case(fixture, loss, GB, dev, rank, world, args, timed):
    """One data-parallel optimiser step (forward, fused loss, VJP, all-reduce, clip + Adam) on a shipped shape."""
    import torch
    from irbfn_b200 import basis, parallel
    from irbfn_b200.model import WCRBFNet, device_params
    from irbfn_b200.train import GraphedTrainStep, TrainState, grad_step, train_step
    cfg, pf = _fixture(fixture)
    R, K, D, O = cfg["num_regions"], cfg["num_kernels"], cfg["in_features"], cfg["out_features"]
    net = WCRBFNet(D, O, K, getattr(basis, cfg["basis_func"]), R, cfg["lower_bounds"], cfg["upper_bounds"],
                   cfg["dimension_ranges"], cfg["activation_idx"], cfg["delta"])
    params = device_params({"params": {"rbf_list": {"centers": pf["centers"], "log_sigs": pf["log_sigs"]},
                                       "linear": {"kernel": pf["kernel"], "bias": pf["bias"]}}}, dev)
    lo = np.array([min(b) for b in cfg["lower_bounds"]] + [-1.0] * D, np.float32)[:D]
    hi = np.array([max(b) for b in cfg["upper_bounds"]] + [1.0] * D, np.float32)[:D]
    a, b = parallel.shard_range(GB, rank, world)
    g = torch.Generator().manual_seed(11)
    x = (torch.rand((GB, D), generator=g) * torch.from_numpy(hi - lo) + torch.from_numpy(lo))[a:b].to(dev).contiguous()
    y = torch.randn((GB, O), generator=g)[a:b].to(dev).contiguous()
    state = TrainState(net, params, lr=1e-3, max_grad_norm=1.0, device=dev)
    how = state.enable_peer_allreduce()
    steps = max(3, args.steps // 2)
    ms_g = timed(lambda: grad_step(state, x, y, GB, loss), steps, 3)
    ms_t = timed(lambda: train_step(state, x, y, GB, loss), steps, 3)
    gstep = GraphedTrainStep(state, x.shape[0], GB, loss=loss).load(x, y)
    ms_graph = timed(lambda: gstep(), steps, 3)
    lane = R * K * (2 * D + 2) + K * O + R * K * (3 * D + 5) + 2 * K * O
    sparse = None
    if R > 1:   # gate-sparse parameter gradients (irbfn_backward_sparse), as in c3
        yscale = float(net.apply(params, x).abs().max())
        st_s = TrainState(net, params, lr=1e-3, max_grad_norm=1.0, device=dev, gate_tol=2.0 ** -24,
                          fwd_tol_abs=1e-6 * yscale)
        st_s.enable_peer_allreduce()
        gstep_s = GraphedTrainStep(st_s, x.shape[0], GB, loss=loss).load(x, y)
        ms_graph_sp = timed(lambda: gstep_s(), steps, 3)
        st_s.flat.copy_(state.flat)
        # deviation of the gradients under the SMOOTH l2 loss: the full-integration loss clips (dynamics.py), and a 1e-6
        # change of y_pred -- fp32 rounding of any two correct forwards -- flips a clip for a few samples
        grad_step(st_s, x, y, GB, "l2")
        grad_step(state, x, y, GB, "l2")
        dev_sp = max(float((st_s.gviews[k].double() - state.gviews[k].double()).norm() / state.gviews[k].double().norm())
                     for k in state.gviews)
        sparse = {"gate_tol": 2.0 ** -24, "forward": "gate-sparse, tol_abs = 1e-6 max|y|", "ms_graph": ms_graph_sp,
                  "grad_rel_l2_vs_dense_max_over_leaves_l2_loss": dev_sp}
        del gstep_s, st_s
    return {"shape": "R%d_K%d_D%d_O%d" % (R, K, D, O), "loss": loss, "global_batch": GB, "n_gpus": world,
            "ms_eager": ms_t, "ms_graph": ms_graph, "ms_fwd_loss_vjp_allreduce": ms_g, "gate_sparse_step": sparse,
            "lane_ops_T_per_s": lane * GB / (ms_g * 1e-3) / 1e12,
            "allreduce_bytes": int(state._grad_ext.numel() * 4), "allreduce": how}


def _fixture(name):
    z = np.load(os.path.join(ROOT, "tests", "golden", name))
    cfg = json.loads(str(z["cfg_json"]))
    p = {"centers": z["p/rbf_list/centers"], "log_sigs": z["p/rbf_list/log_sigs"],
         "kernel": z["p/linear/kernel"], "bias": z["p/linear/bias"]}
    return cfg, p


def other_configs(dev, world, args, timed):
    """The remaining BASELINE.json configs, measured with the same discipline:
    C1 (4,096-query call, latency-bound), C2b (readout-heavy 1 x 1000 x 8 -> 10 on the 2^20 lattice)
    and C5 (64 local networks x 4,096 centres, 6-D, 16 outputs, 2^24 queries per GPU; parameters
    larger than shared memory, so the kernels walk (region, kernel) stages)."""
    import torch
    from irbfn_b200 import ops
    res = {}
    to = lambda a: torch.from_numpy(np.ascontiguousarray(a, dtype=np.float32)).to(dev)

    def module(cfg, p):
        gd = ops.GatePack(cfg["lower_bounds"], cfg["upper_bounds"], cfg["delta"], cfg["dimension_ranges"], dev)
        return gd, [to(p[k]) for k in ("centers", "log_sigs", "kernel", "bias")]

    # C1: one 4,096-query call on the Cartesian NMPC shape; launch latency dominates 1.5 us of math
    cfg, p = _fixture("c1_cartesian_1x1000.npz")
    gd, (c, ls, W, b) = module(cfg, p)
    lo = np.array([min(v) for v in cfg["lower_bounds"]], np.float32)
    hi = np.array([max(v) for v in cfg["upper_bounds"]], np.float32)
    x1 = to(np.random.default_rng(0).uniform(lo, hi, (4096, len(lo))))
    y1 = torch.empty((4096, cfg["out_features"]), device=dev)
    ms1 = timed(lambda: ops.forward(x1, c, ls, W, b, gd=gd, basis=0, out=y1), 200, 20)
    g = torch.cuda.CUDAGraph()
    s = torch.cuda.Stream(dev)
    with torch.cuda.stream(s):
        ops.forward(x1, c, ls, W, b, gd=gd, basis=0, out=y1)
        with torch.cuda.graph(g, stream=s):
            for _ in range(20):
                ops.forward(x1, c, ls, W, b, gd=gd, basis=0, out=y1)
    torch.cuda.synchronize()
    ms1g = timed(g.replay, 20, 3) / 20
    res["forward_C1"] = {"queries": 4096, "shape": "R1_K1000_D7_O2_gaussian", "us_per_call": ms1 * 1e3,
                         "us_per_call_in_cuda_graph": ms1g * 1e3, "queries_per_s": 4096 / (ms1 * 1e-3),
                         "queries_per_s_in_cuda_graph": 4096 / (ms1g * 1e-3)}
    if world == 1 and not args.no_cpu:
        # BASELINE config 1 names the CPU comparison: the same 4,096-query call through the C/OpenMP oracle port
        from oracle import c_oracle
        cores = len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1)
        x1h = x1.cpu().numpy()
        run1 = lambda: c_oracle.forward(x1h, p["centers"], p["log_sigs"], p["kernel"], p["bias"], 0, cfg=cfg, nthreads=cores)
        y1c = run1()
        t0 = time.perf_counter()
        reps = 0
        while reps < 20 or time.perf_counter() - t0 < 2.0:
            run1()
            reps += 1
        dt1 = (time.perf_counter() - t0) / reps
        res["forward_C1"]["cpu_baseline"] = {"value": 4096 / dt1, "unit": UNIT, "us_per_call": dt1 * 1e6, "cores": cores,
                                             "kind": "port", "sample": "the same 4,096 queries, %d calls of "
                                             "oracle/irbfn_oracle.c (float32, OpenMP)" % reps}
        res["forward_C1"]["parity_vs_cpu"] = float(np.abs(y1.cpu().numpy() - y1c).max() / np.abs(y1c).max())

    # C2b: readout share 43 %
    cfg, p = _fixture("c2b_frenet_1x1000_o10.npz")
    gd, (c, ls, W, b) = module(cfg, p)
    xb = to(np.random.default_rng(1).uniform([min(v) for v in cfg["lower_bounds"]],
                                             [max(v) for v in cfg["upper_bounds"]], (B_LATTICE, 8)))
    yb = torch.empty((B_LATTICE, 10), device=dev)
    msb = timed(lambda: ops.forward(xb, c, ls, W, b, gd=gd, basis=0, out=yb), args.steps, 3)
    lane = 1000 * (2 * 8 + 2) + 1000 * 10
    res["forward_C2b"] = {"queries_per_gpu": B_LATTICE, "shape": "R1_K1000_D8_O10_gaussian", "ms": msb,
                          "queries_per_s": world * B_LATTICE / (msb * 1e-3),
                          "lane_ops_T_per_s": lane * B_LATTICE / (msb * 1e-3) / 1e12}
    del xb, yb

    # C5: stress shape, synthetic parameters (seed 1234), 4 x 4 x 4 regions on dims 0-2 over [-1,1]
    R, K, D, O = 64, 4096, 6, 16
    rng = np.random.default_rng(1234)
    edges = np.linspace(-1.0, 1.0, 5)
    lbs = [list(edges[:-1])] * 3
    ubs = [list(edges[1:])] * 3
    ranges = [[i, j, k] for i in range(4) for j in range(4) for k in range(4)]
    gd5 = ops.GatePack(lbs, ubs, [100.0] * 3, ranges, dev)
    c5 = to(rng.uniform(-1, 1, (R, K, D)))
    ls5 = to(rng.uniform(0, 2, (R, K)))
    W5 = to(rng.standard_normal((K, O)))
    b5 = torch.zeros(O, device=dev)
    B5 = 1 << 24
    x5 = torch.rand((B5, D), device=dev) * 2 - 1
    y5 = torch.empty((B5, O), device=dev)
    ms5 = timed(lambda: ops.forward(x5, c5, ls5, W5, b5, gd=gd5, basis=0, out=y5), 2, 1)
    lane5 = R * K * (2 * D + 2) + K * O
    res["forward_C5"] = {"queries_per_gpu": B5, "shape": "R64_K4096_D6_O16_gaussian", "ms": ms5,
                         "queries_per_s": world * B5 / (ms5 * 1e-3),
                         "lane_ops_T_per_s": lane5 * B5 / (ms5 * 1e-3) / 1e12, "timed_steps": 2, "warmup": 1}
    # gate-sparse forward on the same 2^24 queries (delta = 100, non-overlapping regions: ~2 of 64 matter per query)
    y5s = torch.empty_like(y5)
    y5scale = float(y5.abs().max())
    cnt5 = torch.zeros(2, dtype=torch.int64, device=dev)
    ms5s = timed(lambda: ops.forward_sparse(x5, c5, ls5, W5, b5, gd=gd5, basis=0, tol_abs=1e-6 * y5scale, out=y5s,
                                            counters=cnt5), 3, 1)
    sk5, dn5 = (int(v) for v in cnt5.cpu())
    res["forward_C5_gate_sparse"] = {"queries_per_gpu": B5, "ms_sparse_incl_binning": ms5s, "ms_dense": ms5,
                                     "speedup": ms5 / ms5s, "tol_abs_over_max_abs_y": 1e-6,
                                     "max_abs_diff_over_max_abs_y": float((y5s - y5).abs().max()) / y5scale,
                                     "skipped_fraction": sk5 / max(1, sk5 + dn5),
                                     "queries_per_s": world * B5 / (ms5s * 1e-3)}
    del y5s
    B5b = 1 << 22
    dy5 = torch.randn((B5b, O), device=dev)
    flat = torch.empty(R * K * (D + 1) + K * O + O, device=dev)
    ms5b = timed(lambda: (ops.forward(x5[:B5b], c5, ls5, W5, b5, gd=gd5, basis=0, out=y5[:B5b]),
                          ops.backward(x5[:B5b], c5, ls5, W5, dy5, gd=gd5, basis=0, grad_mask=ops.GRAD_PARAMS,
                                       flat_out=flat)), 2, 1)
    res["fwd_vjp_C5"] = {"queries_per_gpu": B5b, "ms": ms5b, "queries_per_s": world * B5b / (ms5b * 1e-3),
                         "algorithmic_tflops": (R * K * (3 * D + 2) + R * K * (5 * D + 9) + 6 * K * O) * B5b
                                               / (ms5b * 1e-3) / 1e12, "timed_steps": 2, "warmup": 1}
    ms5bs = timed(lambda: (ops.forward_sparse(x5[:B5b], c5, ls5, W5, b5, gd=gd5, basis=0, tol_abs=1e-6 * y5scale,
                                              out=y5[:B5b]),
                           ops.backward(x5[:B5b], c5, ls5, W5, dy5, gd=gd5, basis=0, grad_mask=ops.GRAD_PARAMS,
                                        flat_out=flat, gate_tol=2.0 ** -24)), 2, 1)
    res["fwd_vjp_C5"]["ms_gate_sparse"] = ms5bs
    lane5b = lane5 + R * K * (3 * D + 5) + 2 * K * O
    res["c5_summary"] = {"shape": "R64_K4096_D6_O16", "n_gpus": world, "fwd_ms_per_2p24": ms5,
                         "fwd_gate_sparse_ms_per_2p24": ms5s, "fwd_gate_sparse_skipped_fraction": sk5 / max(1, sk5 + dn5),
                         "fwd_lane_ops_T_per_s": lane5 * B5 / (ms5 * 1e-3) / 1e12,
                         "fwd_vjp_ms_per_2p22": ms5b, "fwd_vjp_gate_sparse_ms_per_2p22": ms5bs, "fwd_vjp_lane_ops_T_per_s": lane5b * B5b / (ms5b * 1e-3) / 1e12}
    return res


if __name__ == "__main__":
    main()
