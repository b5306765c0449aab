"""Host-buffer path: queries in host memory in, outputs in host memory out.

The batch is cut into chunks; chunk i's host->device copy, chunk i-1's kernel and chunk i-2's
device->host copy run concurrently on three CUDA streams ordered by events, out of pinned
staging memory.  Nothing is computed on the host.  This is the call bench.py's ``e2e`` times.
"""
from __future__ import annotations

import numpy as np
import torch

from . import ops


class HostPipeline:
    NBUF = 3

    def __init__(self, device, in_features: int, out_features: int, chunk: int = 1 << 17, copy_streams: int = 1,
                 run_streams: int = 1):
        self.device = torch.device(device)
        self.D, self.O, self.chunk = in_features, out_features, int(chunk)
        # Optionally cut each chunk's upload into `copy_streams` pieces on separate streams (separate DMA
        # engines).  Measured on this pool (scripts/dev_h2d.py): a lone 33.5 MB copy runs at
        # 25-55 GB/s depending on the box and two concurrent copies reach 55 GB/s, but inside the pipeline
        # the step is bound by the chunk kernels plus fill/drain, so 1 stream is the default.
        self.ncopy = max(1, int(copy_streams))
        self.s_ins = [torch.cuda.Stream(self.device) for _ in range(self.ncopy)]
        self.s_in = self.s_ins[0]
        # Chunk kernels alternate over `run_streams` streams: with chunks smaller than one wave of query tiles,
        # several chunk kernels share the device, so the pipeline fills after a short first copy and the SMs
        # never wait for a whole wave's upload.
        self.nrun = max(1, int(run_streams))
        self.NBUF = max(3, self.nrun + 2)
        self.s_runs = [torch.cuda.Stream(self.device) for _ in range(self.nrun)]
        self.s_run = self.s_runs[0]
        self.s_out = torch.cuda.Stream(self.device)
        self.xd = [torch.empty((self.chunk, self.D), device=self.device) for _ in range(self.NBUF)]
        self.yd = [torch.empty((self.chunk, self.O), device=self.device) for _ in range(self.NBUF)]
        self.ev_in = [[torch.cuda.Event() for _ in range(self.ncopy)] for _ in range(self.NBUF)]
        self.ev_run = [torch.cuda.Event() for _ in range(self.NBUF)]
        self.ev_out = [torch.cuda.Event() for _ in range(self.NBUF)]
        self._pin_in = None
        self._pin_out = None
        self._pin_in_busy = None      # recorded after the last upload that reads _pin_in
        self._graphs, self._keep = {}, {}

    def _pinned_in(self, x):
        """x as a pinned fp32 CPU tensor (zero-copy when it already is one)."""
        if isinstance(x, torch.Tensor) and x.is_pinned() and x.dtype == torch.float32 and x.is_contiguous():
            return x
        t = torch.as_tensor(np.ascontiguousarray(np.asarray(x), dtype=np.float32)) if not isinstance(x, torch.Tensor) \
            else x.float().contiguous()
        if self._pin_in_busy is not None:
            # a previous call that returned without synchronising (caller-provided ``out``) may still be
            # uploading from the staging buffer: do not overwrite it under the DMA engine
            self._pin_in_busy.synchronize()
            self._pin_in_busy = None
        if self._pin_in is None or self._pin_in.shape[0] < t.shape[0]:
            self._pin_in = torch.empty((t.shape[0], self.D), dtype=torch.float32).pin_memory()
        buf = self._pin_in[: t.shape[0]]
        buf.copy_(t)
        return buf

    def run(self, x, out, launch, graph: bool = False, graph_key=()):
        """launch(x_dev_chunk, y_dev_chunk) enqueues the kernel on the current stream."""
        xin = self._pinned_in(x)
        B = xin.shape[0]
        if xin.dim() != 2 or xin.shape[1] != self.D:
            raise ValueError("x must be (B, %d)" % self.D)
        ret_numpy = out is None
        if out is None:
            if self._pin_out is None or self._pin_out.shape[0] < B:
                self._pin_out = torch.empty((B, self.O), dtype=torch.float32).pin_memory()
            out = self._pin_out[:B]
        elif not (isinstance(out, torch.Tensor) and out.is_pinned() and tuple(out.shape) == (B, self.O)):
            raise ValueError("out must be a pinned CPU tensor of shape (B, %d)" % self.O)
        if graph:
            self._replay(xin, out, B, launch, graph_key)
        else:
            self._enqueue(xin, out, B, launch)
        if self._pin_in is not None and xin.data_ptr() == self._pin_in.data_ptr():
            ev = torch.cuda.Event()
            ev.record(torch.cuda.current_stream(self.device))     # the current stream waited for the pipeline's tail
            self._pin_in_busy = ev
        if ret_numpy:
            torch.cuda.current_stream(self.device).synchronize()
            return out.numpy().copy()
        return out

    def _replay(self, xin, out, B, launch, graph_key=()):
        """The whole chunk pipeline (uploads, kernels, downloads, their cross-stream dependencies) captured
        once per (input buffer, output buffer, batch) as a CUDA graph and replayed: one launch from the host
        instead of ~20 stream operations per chunk, which otherwise bound the pipeline on the CPU."""
        key = (xin.data_ptr(), out.data_ptr(), B) + tuple(graph_key)   # every pointer the graph bakes in
        g = self._graphs.get(key)
        if g is None:
            self._enqueue(xin, out, B, launch)          # warm-up: workspaces, lazy kernel attributes
            torch.cuda.current_stream(self.device).synchronize()
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g):
                self._enqueue(xin, out, B, launch)
            if len(self._graphs) >= 8:
                old = next(iter(self._graphs))
                self._graphs.pop(old)
                self._keep.pop(old, None)
            self._graphs[key] = g
            self._keep[key] = (xin, out)                # the captured host pointers must stay alive
        g.replay()

    def _enqueue(self, xin, out, B, launch):
        cur = torch.cuda.current_stream(self.device)
        for s in (*self.s_ins, *self.s_runs, self.s_out):
            s.wait_stream(cur)
        n = (B + self.chunk - 1) // self.chunk
        for i in range(n):
            b = i % self.NBUF
            lo, hi = i * self.chunk, min(B, (i + 1) * self.chunk)
            m = hi - lo
            piece = (m + self.ncopy - 1) // self.ncopy
            for k, s_in in enumerate(self.s_ins):
                a, e = min(m, k * piece), min(m, (k + 1) * piece)
                with torch.cuda.stream(s_in):
                    if i >= self.NBUF:
                        s_in.wait_event(self.ev_run[b])       # kernel that read xd[b] is done
                    if e > a:
                        self.xd[b][a:e].copy_(xin[lo + a:lo + e], non_blocking=True)
                    self.ev_in[b][k].record(s_in)
            s_run = self.s_runs[i % self.nrun]
            with torch.cuda.stream(s_run):
                for k in range(self.ncopy):
                    s_run.wait_event(self.ev_in[b][k])
                if i >= self.NBUF:
                    s_run.wait_event(self.ev_out[b])          # copy-out that read yd[b] is done
                launch(self.xd[b][:m], self.yd[b][:m])
                self.ev_run[b].record(s_run)
            with torch.cuda.stream(self.s_out):
                self.s_out.wait_event(self.ev_run[b])
                out[lo:hi].copy_(self.yd[b][:m], non_blocking=True)
                self.ev_out[b].record(self.s_out)
        cur.wait_stream(self.s_out)


class StreamedPipeline:
    """Host buffers in, host buffers out, ONE kernel launch (ops.forward_streamed).

    The persistent forward is launched first; the batch is then uploaded in pieces alternating over a few copy streams
    (one stream's per-copy set-up hides behind another's transfer), each piece followed on its stream by a stream-ordered
    store that sets the piece's flag; every query tile of the kernel waits for its piece's flag before it reads its
    rows.  Outputs are stored by the kernel straight into the pinned host buffer.  Compared with HostPipeline (one
    launch per chunk, a download pass) there is no per-chunk launch ramp / tail and no D2H stage.
    """

    def __init__(self, device, in_features: int, out_features: int, chunk: int = 1 << 15, copy_streams: int = 3):
        self.device = torch.device(device)
        self.D, self.O = in_features, out_features
        self.chunk = max(1024, (int(chunk) // 1024) * 1024)        # rows per upload piece / flag
        self.s_ins = [torch.cuda.Stream(self.device) for _ in range(max(1, int(copy_streams)))]
        self.ev_zero = torch.cuda.Event()
        self.ev_done = torch.cuda.Event()
        self.xd = None
        self.flags = None
        self.status = torch.zeros(1, dtype=torch.int32, device=self.device)
        self.status_host = torch.zeros(1, dtype=torch.int32).pin_memory()   # copied behind every launch
        self._pin_in = None
        self._pin_out = None

    def check(self, wait: bool = True) -> None:
        """Raise if an earlier call's kernel gave up on an upload piece (its result is then undefined).  The status
        word is sticky on the device -- later launches give up at their first wait -- until this clears it.  With
        wait=False only a call that has already finished is looked at (run() does that for the call before it)."""
        if wait:
            self.ev_done.synchronize()
        elif not self.ev_done.query():
            return
        if int(self.status_host[0]) != 0:
            self.status.zero_()
            self.status_host.zero_()
            raise RuntimeError("irbfn_forward_streamed: an upload piece never arrived (the kernel gave the launch up)")

    def _pinned_in(self, x):
        if isinstance(x, torch.Tensor) and x.is_pinned() and x.dtype == torch.float32 and x.is_contiguous():
            return x
        t = torch.as_tensor(np.ascontiguousarray(np.asarray(x), dtype=np.float32)) if not isinstance(x, torch.Tensor) \
            else x.float().contiguous()
        self.ev_done.synchronize()                  # an earlier call may still be uploading from the staging buffer
        if self._pin_in is None or self._pin_in.shape[0] < t.shape[0]:
            self._pin_in = torch.empty((t.shape[0], self.D), dtype=torch.float32).pin_memory()
        buf = self._pin_in[: t.shape[0]]
        buf.copy_(t)
        return buf

    def run(self, x, out, launch):
        """launch(x_dev, y_host, ready, queries_per_flag, status) enqueues ops.forward_streamed on the current stream."""
        self.check(wait=False)
        xin = self._pinned_in(x)
        B = xin.shape[0]
        if xin.dim() != 2 or xin.shape[1] != self.D:
            raise ValueError("x must be (B, %d)" % self.D)
        ret_numpy = out is None
        if out is None:
            if self._pin_out is None or self._pin_out.shape[0] < B:
                self._pin_out = torch.empty((B, self.O), dtype=torch.float32).pin_memory()
            out = self._pin_out[:B]
        elif not (isinstance(out, torch.Tensor) and out.is_pinned() and tuple(out.shape) == (B, self.O)):
            raise ValueError("out must be a pinned CPU tensor of shape (B, %d)" % self.O)
        if B == 0:
            return out.numpy().copy() if ret_numpy else out
        n = (B + self.chunk - 1) // self.chunk
        if self.xd is None or self.xd.shape[0] < B:
            self.xd = torch.empty((B, self.D), device=self.device)
        if self.flags is None or self.flags.numel() < n + 1:
            self.flags = torch.zeros(n + 1, dtype=torch.int32, device=self.device)
        cur = torch.cuda.current_stream(self.device)
        # the previous call's kernel has consumed its uploads (stream order), so flags and tile counter may be cleared;
        # no copy stream may set a flag before the clear
        self.flags[:n + 1].zero_()
        self.ev_zero.record(cur)
        launch(self.xd[:B], out, self.flags[:n + 1], self.chunk, self.status)
        for s in self.s_ins:
            s.wait_event(self.ev_zero)
        ops.upload_streamed(xin, self.xd[:B], self.flags[:n + 1], self.chunk, self.s_ins)
        self.status_host.copy_(self.status, non_blocking=True)
        self.ev_done.record(cur)                    # kernel finished => every upload finished and every y row stored
        if ret_numpy:
            self.check()
            return out.numpy().copy()
        return out


class PredStepPipeline:
    """The planner's call (pred_step, src/irbfn_mpc/irbfn_planner.py:29-32): a handful of queries in a NumPy array in,
    a NumPy array out, the caller waits for it.  Upload, kernel and download of a batch size are captured ONCE (on the
    second call with the same batch size and parameter buffers) as a CUDA graph over fixed pinned / device staging
    buffers; a call is then a copy into the pinned buffer, one graph launch and one stream synchronisation -- no
    per-call allocation, no event traffic, no chunking."""

    MAX_ROWS = 4096
    MAX_GRAPHS = 8

    def __init__(self, device, in_features: int, out_features: int):
        self.device = torch.device(device)
        self.D, self.O = in_features, out_features
        self.pin_in = torch.empty((self.MAX_ROWS, self.D), dtype=torch.float32).pin_memory()
        self.pin_out = torch.empty((self.MAX_ROWS, self.O), dtype=torch.float32).pin_memory()
        self.np_in, self.np_out = self.pin_in.numpy(), self.pin_out.numpy()
        self.xd = torch.empty((self.MAX_ROWS, self.D), device=self.device)
        self.yd = torch.empty((self.MAX_ROWS, self.O), device=self.device)
        self._graphs, self._seen = {}, set()

    def _enqueue(self, B, launch):
        self.xd[:B].copy_(self.pin_in[:B], non_blocking=True)
        launch(self.xd[:B], self.yd[:B])
        self.pin_out[:B].copy_(self.yd[:B], non_blocking=True)

    def run(self, x, launch, graph_key):
        """x (B,D) NumPy / CPU tensor, B <= MAX_ROWS; launch(x_dev, y_dev) enqueues the forward on the current stream;
        graph_key identifies the parameter buffers the captured launch reads."""
        x = x.numpy() if isinstance(x, torch.Tensor) else np.asarray(x)
        B = x.shape[0]
        if x.ndim != 2 or x.shape[1] != self.D or B > self.MAX_ROWS:
            raise ValueError("x must be (B <= %d, %d)" % (self.MAX_ROWS, self.D))
        if B == 0:
            return np.empty((0, self.O), np.float32)
        np.copyto(self.np_in[:B], x, casting="unsafe")            # float64 callers (NumPy states) narrow here
        cur = torch.cuda.current_stream(self.device)
        if torch.cuda.is_current_stream_capturing():
            raise RuntimeError("a host-buffer call waits for its result and cannot be part of a CUDA graph capture; "
                               "pass a CUDA tensor instead")
        g = self._graphs.get((B, graph_key))
        if g is None:
            self._enqueue(B, launch)                              # first use: also warms allocations / lazy loading
            cur.synchronize()
            first = self.np_out[:B].copy()
            if (B, graph_key) not in self._seen:
                # capture only what comes back: a caller that hands over fresh parameter buffers on every call
                # (a tree of NumPy leaves) would otherwise pay a capture per call
                if len(self._seen) >= 64:
                    self._seen.clear()
                self._seen.add((B, graph_key))
                return first
            if len(self._graphs) >= self.MAX_GRAPHS:
                self._graphs.pop(next(iter(self._graphs)))
            g = torch.cuda.CUDAGraph()
            side = torch.cuda.Stream(self.device)
            with torch.cuda.stream(side):
                with torch.cuda.graph(g, stream=side):
                    self._enqueue(B, launch)
            self._graphs[(B, graph_key)] = g
            return first
        g.replay()
        cur.synchronize()
        return self.np_out[:B].copy()
