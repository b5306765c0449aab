"""H2D bandwidth experiments on the GPU box: torch pinned vs write-combined pinned vs NUMA placement."""
import ctypes as C, glob, os, sys, time
import numpy as np, torch

dev = torch.device("cuda:0")
torch.cuda.init()
N = 1 << 20
nbytes = N * 8 * 4
xd = torch.empty((N, 8), device=dev)
rt = C.CDLL("libcudart.so.12") if os.path.exists("/usr/local/cuda/lib64/libcudart.so.12") else C.CDLL("libcudart.so")
try:
    rt = C.CDLL("/usr/local/cuda/lib64/libcudart.so.12")
except OSError:
    pass

def timed(fn, n=20, w=5):
    for _ in range(w): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n

def host_alloc(flags):
    p = C.c_void_p()
    rc = rt.cudaHostAlloc(C.byref(p), C.c_size_t(nbytes), C.c_uint(flags))
    assert rc == 0, rc
    buf = (C.c_float * (N * 8)).from_address(p.value)
    t = torch.frombuffer(buf, dtype=torch.float32).view(N, 8)
    t.fill_(1.0)
    return t, p

print("cpus", os.cpu_count(), "affinity", len(os.sched_getaffinity(0)))
for node in sorted(glob.glob("/sys/devices/system/node/node*/cpulist")):
    print(node, open(node).read().strip())
os.system("nvidia-smi topo -m 2>/dev/null | head -8")
t0 = torch.ones((N, 8)).pin_memory()
ms = timed(lambda: xd.copy_(t0, non_blocking=True)); print("torch pinned      ms %.3f GB/s %.1f" % (ms, nbytes / ms / 1e6))
t1, _ = host_alloc(0)
print("is_pinned", t1.is_pinned())
ms = timed(lambda: xd.copy_(t1, non_blocking=True)); print("cudaHostAlloc dflt ms %.3f GB/s %.1f" % (ms, nbytes / ms / 1e6))
t2, _ = host_alloc(4)
ms = timed(lambda: xd.copy_(t2, non_blocking=True)); print("write-combined    ms %.3f GB/s %.1f" % (ms, nbytes / ms / 1e6))
t3, _ = host_alloc(1)
ms = timed(lambda: xd.copy_(t3, non_blocking=True)); print("portable          ms %.3f GB/s %.1f" % (ms, nbytes / ms / 1e6))
# NUMA placement by first touch under different affinities
full = os.sched_getaffinity(0)
for node in sorted(glob.glob("/sys/devices/system/node/node*/cpulist")):
    cpus = set()
    for part in open(node).read().strip().split(","):
        if "-" in part:
            a, b = part.split("-"); cpus |= set(range(int(a), int(b) + 1))
        elif part: cpus.add(int(part))
    cpus &= full
    if not cpus: continue
    os.sched_setaffinity(0, cpus)
    t, _ = host_alloc(0)
    ms = timed(lambda: xd.copy_(t, non_blocking=True)); print(os.path.basename(os.path.dirname(node)), "first touch ms %.3f GB/s %.1f" % (ms, nbytes / ms / 1e6))
    os.sched_setaffinity(0, full)
# two halves on two streams concurrently
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
def two():
    with torch.cuda.stream(s1): xd[: N // 2].copy_(t0[: N // 2], non_blocking=True)
    with torch.cuda.stream(s2): xd[N // 2:].copy_(t0[N // 2:], non_blocking=True)
    torch.cuda.current_stream().wait_stream(s1); torch.cuda.current_stream().wait_stream(s2)
ms = timed(two); print("two streams       ms %.3f GB/s %.1f" % (ms, nbytes / ms / 1e6))
