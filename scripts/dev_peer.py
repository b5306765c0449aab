"""Dev / verification of the one-shot NVLink all-reduce fused with clip + Adam (torchrun, N >= 2):
  * the same K optimiser steps through the NCCL path and through the peer-memory path from one initial state:
    parameters agree to float32 rounding, and are bit-identical across ranks on either path;
  * graph-replayed step time of both paths at the C3 shard size."""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import torch.distributed as dist
from irbfn_b200 import parallel
from irbfn_b200.basis import gaussian
from irbfn_b200.model import WCRBFNet, device_params
from irbfn_b200.train import GraphedTrainStep, TrainState, grad_step, train_step

rank, local_rank, world = parallel.init_distributed()
dev = torch.device("cuda", local_rank)
torch.cuda.set_device(dev)
GB = int(sys.argv[1]) if len(sys.argv) > 1 else 262144
z = np.load(os.path.join(os.path.dirname(__file__), "..", "tests", "golden", "c2_frenet_12x100.npz"))
cfg = json.loads(str(z["cfg_json"]))
params = {"params": {"rbf_list": {"centers": z["p/rbf_list/centers"], "log_sigs": z["p/rbf_list/log_sigs"]},
                     "linear": {"kernel": z["p/linear/kernel"], "bias": z["p/linear/bias"]}}}
net = WCRBFNet(cfg["in_features"], cfg["out_features"], cfg["num_kernels"], gaussian, cfg["num_regions"],
               cfg["lower_bounds"], cfg["upper_bounds"], cfg["dimension_ranges"], cfg["activation_idx"], cfg["delta"])
lo = np.array([min(b) for b in cfg["lower_bounds"]], np.float32)
hi = np.array([max(b) for b in cfg["upper_bounds"]], np.float32)
g = torch.Generator().manual_seed(7)
xall = torch.rand((GB, 8), generator=g) * torch.from_numpy(hi - lo) + torch.from_numpy(lo)
yall = torch.randn((GB, 2), generator=g)
a, b = parallel.shard_range(GB, rank, world)
x, y = xall[a:b].to(dev).contiguous(), yall[a:b].to(dev).contiguous()

s_nccl = TrainState(net, device_params(params, dev), lr=1e-3, device=dev)
s_peer = TrainState(net, device_params(params, dev), lr=1e-3, device=dev)
how = s_peer.enable_peer_allreduce()
if rank == 0:
    print("peer path:", how, flush=True)
assert how.startswith("peer"), how

# gradients: NCCL vs peer all-reduce
grad_step(s_nccl, x, y, GB)
grad_step(s_peer, x, y, GB)
torch.cuda.synchronize()
e = float((s_peer.grad - s_nccl.grad).norm() / s_nccl.grad.norm())
K = 25
for _ in range(K):
    ln = train_step(s_nccl, x, y, GB)
    lp = train_step(s_peer, x, y, GB)
torch.cuda.synchronize()
assert int(s_peer.peer_status.item()) == 0, "a peer timed out"
dp = float((s_peer.flat - s_nccl.flat).abs().max() / s_nccl.flat.abs().max())
# replicas identical across ranks?
def same_everywhere(t):
    ref = t.clone()
    dist.broadcast(ref, 0)
    ok = torch.tensor([1 if torch.equal(ref, t) else 0], device=dev)
    dist.all_reduce(ok, op=dist.ReduceOp.MIN)
    return bool(ok.item())
id_peer, id_nccl = same_everywhere(s_peer.flat), same_everywhere(s_nccl.flat)

def timed(fn, n=200, w=20):
    for _ in range(w):
        fn()
    parallel.barrier(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        fn()
    e1.record(); torch.cuda.synchronize()
    return parallel.allreduce_max_scalar(e0.elapsed_time(e1) / n * 1e3, dev)

t_eager_n = timed(lambda: train_step(s_nccl, x, y, GB))
t_eager_p = timed(lambda: train_step(s_peer, x, y, GB))
gn = GraphedTrainStep(s_nccl, x.shape[0], GB).load(x, y)
gp = GraphedTrainStep(s_peer, x.shape[0], GB).load(x, y)
t_graph_n = timed(lambda: gn())
t_graph_p = timed(lambda: gp())
id_peer2 = same_everywhere(s_peer.flat)
if rank == 0:
    print(json.dumps({"world": world, "global_batch": GB, "how": how, "grad_rel_l2_peer_vs_nccl": e,
                      "params_rel_max_after_%d_steps" % K: dp, "loss_nccl": float(ln), "loss_peer": float(lp),
                      "replicas_bit_identical": {"peer": id_peer, "nccl": id_nccl, "peer_after_graph": id_peer2},
                      "us_eager": {"nccl": t_eager_n, "peer": t_eager_p},
                      "us_graph": {"nccl": t_graph_n, "peer": t_graph_p}}), flush=True)
# teardown: symmetric-memory handles must go before the process group (their destructors talk to its store),
# and a hard exit after the last barrier keeps a stuck destructor from hanging the job
del gp, gn
s_peer.peer = None
torch.cuda.synchronize()
dist.barrier()
sys.stdout.flush()
os._exit(0)
