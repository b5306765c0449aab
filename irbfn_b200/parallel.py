"""Data-parallel plumbing for the IRBFN path: one process per GPU, queries sharded by rank,
parameters replicated, one all-reduce of one flat gradient buffer per training step
(SURVEY.md 8e).  Forward and input-gradient sweeps need no collective at all.

The reference is single-device (scripts/train_nmpc.py:43 pins CUDA_VISIBLE_DEVICES); the only
behaviour to preserve is "same gradients as one device on the concatenated batch".
"""
from __future__ import annotations

import os

import torch
import torch.distributed as dist


def env_world():
    """(rank, local_rank, world_size) from the torchrun environment (1 process -> 0, 0, 1)."""
    return (int(os.environ.get("RANK", "0")), int(os.environ.get("LOCAL_RANK", "0")),
            int(os.environ.get("WORLD_SIZE", "1")))


def init_distributed(backend: str = None):
    """Initialise torch.distributed from the environment if WORLD_SIZE > 1.  Returns
    (rank, local_rank, world_size).  NCCL on CUDA, gloo otherwise."""
    rank, local_rank, world = env_world()
    if world > 1 and not dist.is_initialized():
        if backend is None:
            backend = "nccl" if torch.cuda.is_available() else "gloo"
        if backend == "nccl":
            torch.cuda.set_device(local_rank)
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        kw = {"device_id": torch.device("cuda", local_rank)} if backend == "nccl" else {}
        dist.init_process_group(backend=backend, rank=rank, world_size=world, **kw)
    elif torch.cuda.is_available():
        torch.cuda.set_device(local_rank)
    return rank, local_rank, world


def shard_range(n: int, rank: int, world: int):
    """Contiguous, balanced [lo, hi) of n items for ``rank`` (the first n % world ranks get one
    more).  The union over ranks is [0, n) with no overlap, in rank order."""
    if world <= 0 or not (0 <= rank < world):
        raise ValueError("bad rank %d / world %d" % (rank, world))
    base, extra = divmod(n, world)
    lo = rank * base + min(rank, extra)
    return lo, lo + base + (1 if rank < extra else 0)


def allreduce_sum_(flat: torch.Tensor, group=None) -> torch.Tensor:
    """In-place sum of one flat buffer over all ranks (no-op for a single process)."""
    if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.all_reduce(flat, op=dist.ReduceOp.SUM, group=group)
    return flat


def allreduce_max_scalar(value: float, device=None) -> float:
    """Max of a python float over ranks (used for max-over-ranks timing)."""
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
        return float(value)
    t = torch.tensor([float(value)], dtype=torch.float64,
                     device=device if device is not None else ("cuda" if dist.get_backend() == "nccl" else "cpu"))
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def barrier():
    if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
        dist.barrier()


class PeerBuffer:
    """A flat fp32 buffer of this rank that every other rank of the node can read (and flag) directly over
    NVLink / NVSwitch: the memory side of ``ops.fused_update_`` (one-shot all-reduce fused with the optimiser
    update).  One process per GPU: the buffers are exchanged as CUDA IPC handles through
    the process group (torch's own ``_share_cuda_`` / ``_new_shared_cuda`` storage plumbing -- the mechanism
    torch.multiprocessing uses), falling back to torch symmetric memory.

    ``local``   this rank's buffer (n_total floats, then the kernel's flag words), zero-filled
    ``ptrs``    int64 device tensor: address of rank r's buffer in THIS process
    Construction is collective (every rank must call it) and raises if peer memory cannot be set up; callers
    fall back to NCCL."""

    def __init__(self, n_total: int, device, group=None):
        from . import _lib
        if not (dist.is_available() and dist.is_initialized()):
            raise RuntimeError("PeerBuffer needs an initialised process group")
        self.world, self.rank = dist.get_world_size(group), dist.get_rank(group)
        device = torch.device(device)
        nfloats = int(_lib.load().irbfn_fused_update_buffer_bytes(int(n_total), self.world)) // 4
        self._keep = []
        self.how = None
        err = None
        first, second = (self._init_symm, "symmetric-memory"), (self._init_ipc, "cuda-ipc")
        if os.environ.get("IRBFN_PEER_MEMORY", "") == "ipc":
            first, second = second, first
        try:
            first[0](nfloats, device, group)
            self.how = first[1]
        except Exception as e:           # noqa: BLE001 - any failure means: try the other mechanism
            err = e
        ok = torch.tensor([1 if self.how else 0], device=device)
        dist.all_reduce(ok, op=dist.ReduceOp.MIN, group=group)
        if int(ok.item()) == 0:
            self.how = None
            try:
                second[0](nfloats, device, group)
                self.how = second[1]
            except Exception as e:       # noqa: BLE001
                err = e
            ok = torch.tensor([1 if self.how else 0], device=device)
            dist.all_reduce(ok, op=dist.ReduceOp.MIN, group=group)
            if int(ok.item()) == 0:
                raise RuntimeError("peer memory unavailable on this node: %r" % (err,))
        self.n_total = int(n_total)
        self.local.zero_()
        torch.cuda.synchronize(device)
        dist.barrier(group)              # nobody flags a buffer that is not zeroed yet

    def _init_ipc(self, nfloats, device, group):
        # a dedicated cudaMalloc block (not a slice of the caching allocator's pool shared with other tensors)
        self.local = torch.empty(max(nfloats, 1 << 19), device=device, dtype=torch.float32)[:nfloats]
        handle = self.local.untyped_storage()._share_cuda_()
        offset_bytes = self.local.storage_offset() * 4
        gathered = [None] * self.world
        dist.all_gather_object(gathered, (handle, offset_bytes), group=group)
        addrs = []
        for r, (h, off) in enumerate(gathered):
            if r == self.rank:
                addrs.append(self.local.data_ptr())
                continue
            (dev_idx, ipc_handle, size_bytes, blk_off, ref_h, ref_off, ev_h, ev_sync) = h
            st = torch.UntypedStorage._new_shared_cuda(device.index, ipc_handle, size_bytes, blk_off, ref_h, ref_off,
                                                       ev_h, ev_sync)
            self._keep.append(st)
            addrs.append(st.data_ptr() + off)
            # kernels of THIS device dereference the peer's memory: peer access must be on in that direction
            import ctypes
            rt = ctypes.CDLL("libcudart.so.12")
            rc = rt.cudaDeviceEnablePeerAccess(int(dev_idx), 0)
            rt.cudaGetLastError()        # "already enabled" must not linger as the thread's last error
            if rc not in (0, 704):       # 704 = cudaErrorPeerAccessAlreadyEnabled
                raise RuntimeError("cudaDeviceEnablePeerAccess(%d) -> %d" % (dev_idx, rc))
        torch.cuda.synchronize(device)
        self.ptrs = torch.tensor(addrs, dtype=torch.int64, device=device)

    def _init_symm(self, nfloats, device, group):
        import torch.distributed._symmetric_memory as symm
        t = symm.empty(nfloats, dtype=torch.float32, device=device)
        h = symm.rendezvous(t, group if group is not None else dist.group.WORLD)
        self.local = t
        self._keep.append(h)
        self.ptrs = torch.tensor([int(v) for v in h.buffer_ptrs], dtype=torch.int64, device=device)
