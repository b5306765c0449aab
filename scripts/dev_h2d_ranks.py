"""Per-rank host<->device bandwidth with 1 rank active and with ALL ranks active at once (torchrun, N ranks):
what bounds bench.py's `e2e` at 8 GPUs.  33.5 MB up + 8.4 MB down per step, as the e2e path moves them."""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist
from irbfn_b200 import parallel

rank, local_rank, world = parallel.init_distributed()
dev = torch.device("cuda", local_rank)
torch.cuda.set_device(dev)
N = 1 << 20
xh = torch.ones((N, 8)).pin_memory()
yh = torch.empty((N, 2)).pin_memory()
xd = torch.empty((N, 8), device=dev)
yd = torch.ones((N, 2), device=dev)
s1, s2 = torch.cuda.Stream(dev), torch.cuda.Stream(dev)


def timed(fn, n=20, w=3):
    for _ in range(w):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n


def h2d():
    xd.copy_(xh, non_blocking=True)


def d2h():
    yh.copy_(yd, non_blocking=True)


def both():
    with torch.cuda.stream(s1):
        xd.copy_(xh, non_blocking=True)
    with torch.cuda.stream(s2):
        yh.copy_(yd, non_blocking=True)
    torch.cuda.current_stream().wait_stream(s1)
    torch.cuda.current_stream().wait_stream(s2)


def both_split():            # upload cut in two halves on two streams (two copy engines)
    with torch.cuda.stream(s1):
        xd[: N // 2].copy_(xh[: N // 2], non_blocking=True)
    with torch.cuda.stream(s2):
        xd[N // 2:].copy_(xh[N // 2:], non_blocking=True)
        yh.copy_(yd, non_blocking=True)
    torch.cuda.current_stream().wait_stream(s1)
    torch.cuda.current_stream().wait_stream(s2)


res = {}
up, down = N * 32, N * 8
for name, fn, nbytes in (("h2d", h2d, up), ("d2h", d2h, down), ("h2d+d2h", both, up + down),
                         ("h2d(2 streams)+d2h", both_split, up + down)):
    # everyone at once
    parallel.barrier()
    ms_all = timed(fn)
    t = torch.tensor([ms_all], device=dev, dtype=torch.float64)
    worst = t.clone()
    if world > 1:
        dist.all_reduce(worst, op=dist.ReduceOp.MAX)
    # rank 0 alone
    parallel.barrier()
    ms_solo = timed(fn) if rank == 0 else 0.0
    parallel.barrier()
    res[name] = {"MB": nbytes / 1e6, "ms_all_ranks_max": float(worst), "GBps_per_rank_all_active": nbytes / float(worst) / 1e6,
                 "ms_rank0_alone": ms_solo, "GBps_rank0_alone": (nbytes / ms_solo / 1e6) if ms_solo else None}
if rank == 0:
    topo = os.popen("nvidia-smi topo -m 2>/dev/null | head -12").read()
    print(json.dumps({"world": world, "cpus": os.cpu_count(), "affinity": len(os.sched_getaffinity(0)), "table": res}))
    print(topo)
if world > 1:
    torch.cuda.synchronize()
    dist.barrier()
    sys.stdout.flush()
    os._exit(0)
