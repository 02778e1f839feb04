"""Zero-copy data-parallel gradient engine for CPU ranks - the host-memory twin of data_parallel.py.

The reference is an allreduce between CPU processes whose use case is exactly this: summing the gradients of
data-parallel workers (allreduce_over_mpi/mpi_mod.hpp:1167-1221; it ships no training integration). Here every
gradient of a module is a view into one flat buffer inside the communicator's *registered* shared-memory region,
cut into buckets; when the last gradient of a bucket has been produced during backward, a worker thread all-reduces
that bucket IN PLACE (peers read the buffer itself: no copy-in, no copy-out, the 1/world_size average fused into
the reduction) while autograd keeps producing the next bucket - the C call releases the GIL.

    comm = ft.HostCommunicator(window_bytes=ft.HostDataParallel.bytes_needed(model))
    dp = ft.HostDataParallel(model, comm)           # broadcasts rank 0's parameters
    loss = loss_fn(dp(x), y); loss.backward()
    dp.synchronize()                                  # wait for the bucket collectives
    optimizer.step(); dp.zero_grad()
"""
import queue
import threading

import torch


def _pad(n, elem):  # every view starts on a 16-byte boundary (the registered-piece rule of HostCommunicator)
    per = 16 // elem
    return -(-n // per) * per


class _Bucket:
    __slots__ = ("flat", "params", "pending", "done", "launched")

    def __init__(self, flat, params):
        self.flat = flat
        self.params = params
        self.pending = len(params)
        self.done = threading.Event()
        self.launched = False


class HostDataParallel(torch.nn.Module):
    @staticmethod
    def bytes_needed(module):
        """Registered-region size that holds every trainable gradient of `module` (pass it as window_bytes)."""
        total = 0
        for p in module.parameters():
            if p.requires_grad:
                total += _pad(p.numel(), p.element_size()) * p.element_size()
        return total + 4096

    def __init__(self, module, comm, bucket_mb=4, average=True, widths=None, threads=1, broadcast=True):
        super().__init__()
        self.module = module
        self.comm = comm
        self.scale = 1.0 / comm.size if average else None
        self.widths = widths
        self.threads = threads
        self.buckets = []
        self._bucket_of = {}
        params = [p for p in module.parameters() if p.requires_grad]
        if self.bytes_needed(module) > comm.window_bytes:
            raise ValueError(f"gradients need {self.bytes_needed(module)} bytes of registered memory; create the "
                             f"communicator with window_bytes=HostDataParallel.bytes_needed(module)")
        if broadcast and comm.size > 1:
            self._broadcast_parameters()
        cap = int(bucket_mb * (1 << 20))
        off = 0  # byte offset inside the registered region: identical on every rank (same model, same order)
        by_dtype = {}
        for p in reversed(params):  # gradients become ready roughly in reverse registration order
            by_dtype.setdefault(p.dtype, []).append(p)
        for dtype, plist in by_dtype.items():
            cur, cur_bytes = [], 0
            for p in plist:
                nbytes = _pad(p.numel(), p.element_size()) * p.element_size()
                if cur and cur_bytes + nbytes > cap:
                    off = self._make_bucket(cur, dtype, off)
                    cur, cur_bytes = [], 0
                cur.append(p)
                cur_bytes += nbytes
            if cur:
                off = self._make_bucket(cur, dtype, off)
        for p in params:
            p.register_post_accumulate_grad_hook(self._on_grad_ready)
        self._work = queue.Queue()
        self._error = None
        self._worker = threading.Thread(target=self._run, name="flextree-host-dp", daemon=True)
        self._worker.start()

    # -- set-up ------------------------------------------------------------------------------------------
    def _broadcast_parameters(self):
        """Rank 0's values everywhere, through the library's own allreduce (the other ranks contribute zeros)."""
        with torch.no_grad():
            for t in list(self.module.parameters()) + list(self.module.buffers()):
                if not t.is_floating_point() or t.numel() == 0:
                    continue
                flat = t.detach().contiguous().view(-1).clone()
                if self.comm.rank != 0:
                    flat.zero_()
                self.comm.allreduce(flat)
                t.detach().copy_(flat.view_as(t))

    def _make_bucket(self, plist, dtype, off):
        elem = torch.empty(0, dtype=dtype).element_size()
        numel = sum(_pad(p.numel(), elem) for p in plist)
        flat = self.comm.registered_tensor(numel, dtype, off)
        flat.zero_()
        b = _Bucket(flat, plist)
        at = 0
        for p in plist:
            p.grad = torch.as_strided(flat, p.size(), p.stride(), storage_offset=flat.storage_offset() + at)
            at += _pad(p.numel(), elem)
            self._bucket_of[p] = b
        self.buckets.append(b)
        return off + numel * elem

    # -- autograd side -------------------------------------------------------------------------------------
    def _on_grad_ready(self, p):
        b = self._bucket_of[p]
        b.pending -= 1
        if b.pending == 0:
            self._launch(b)

    def _launch(self, b):
        b.launched = True
        self._work.put(b)

    def _run(self):  # worker thread: bucket collectives in launch order (the same on every rank)
        while True:
            b = self._work.get()
            if b is None:
                return
            try:
                self.comm.allreduce(b.flat, scale=self.scale, widths=self.widths, ctas=self.threads, registered=True)
            except Exception as e:  # surfaced by synchronize()
                self._error = e
            b.done.set()

    def synchronize(self):
        """Call after backward: launches any bucket whose gradients were not all produced (unused parameters) and
        waits for every bucket's collective."""
        for b in self.buckets:
            if not b.launched:
                self._launch(b)
        for b in self.buckets:
            b.done.wait()
            b.done.clear()
            b.pending = len(b.params)
            b.launched = False
        if self._error is not None:
            err, self._error = self._error, None
            raise err

    def zero_grad(self, set_to_none=False):
        for b in self.buckets:
            b.flat.zero_()

    def close(self):
        if self._worker.is_alive():
            self._work.put(None)
            self._worker.join()

    def forward(self, *args, **kwargs):
        return self.module(*args, **kwargs)
