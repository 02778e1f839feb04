"""Zero-copy data-parallel gradient engine.

``FlexTreeDataParallel`` keeps every gradient of a module inside a few flat buffers that live in
*symmetric* (registered, NVLS-capable) memory: ``param.grad`` is a view into its bucket, so backward
accumulates straight into the buffer the collective works on. When the last gradient of a bucket has
been produced, the bucket is all-reduced IN PLACE on a side stream - zero copies, the 1/world_size
average fused into the reduction, and (on >= 4 GPUs) the in-switch NVLS kernel that occupies 8 SMs,
so backward keeps the rest of the GPU.

    dp = ft.FlexTreeDataParallel(model, comm)         # broadcasts rank 0's parameters
    loss = loss_fn(dp(x), y); loss.backward()
    dp.synchronize()                                    # current stream waits for the bucket collectives
    optimizer.step(); dp.zero_grad()

This is the B200-first form of what the DDP comm hook (ddp.py) does through DDP's private buckets. The
reference has no training integration (its use case is this gradient allreduce:
allreduce_over_mpi/mpi_mod.hpp:1167-1221).
"""
import torch


class _Bucket:
    __slots__ = ("flat", "params", "pending", "event", "launched")

    def __init__(self, flat, params):
        self.flat = flat
        self.params = params
        self.pending = len(params)
        self.event = torch.cuda.Event()
        self.launched = False


class FlexTreeDataParallel(torch.nn.Module):
    def __init__(self, module, comm, bucket_mb=25, average=True, wire_dtype=None, max_ctas=32, broadcast=True):
        super().__init__()
        self.module = module
        self.comm = comm
        self.scale = 1.0 / comm.size if average else 1.0
        self.wire_dtype = wire_dtype
        self.stream = torch.cuda.Stream(device=comm.device)
        if max_ctas:
            comm.set_max_ctas(max_ctas)
        self.buckets = []
        self._next = 0  # index of the first bucket not launched in this iteration
        self._bucket_of = {}
        params = [p for p in module.parameters() if p.requires_grad]
        if broadcast and comm.size > 1:
            self._broadcast_parameters(params)
        cap = int(bucket_mb * (1 << 20))
        # gradients become ready roughly in reverse registration order
        by_dtype = {}
        for p in reversed(params):
            by_dtype.setdefault(p.dtype, []).append(p)
        for dtype, plist in by_dtype.items():
            cur, cur_bytes = [], 0
            for p in plist:
                nbytes = p.numel() * p.element_size()
                if cur and cur_bytes + nbytes > cap:
                    self._make_bucket(cur, dtype)
                    cur, cur_bytes = [], 0
                cur.append(p)
                cur_bytes += nbytes
            if cur:
                self._make_bucket(cur, dtype)
        for p in params:
            p.register_post_accumulate_grad_hook(self._on_grad_ready)

    # -- set-up ------------------------------------------------------------------------------------------
    def _broadcast_parameters(self, params):
        """Rank 0's values everywhere, through the library's own allreduce (others contribute zeros)."""
        with torch.no_grad():
            tensors = list(params) + list(self.module.buffers())
            for t in tensors:
                if not t.is_floating_point() or t.numel() == 0:
                    continue
                flat = t.detach().contiguous().view(-1).clone()
                if self.comm.rank != 0:
                    flat.zero_()
                self.comm.allreduce(flat)
                t.detach().copy_(flat.view_as(t))
        torch.cuda.synchronize(self.comm.device)

    def _make_bucket(self, plist, dtype):
        numel = sum(-(-p.numel() // 8) * 8 for p in plist)  # keep every view 16-byte aligned
        flat = self.comm.symm_empty(numel, dtype)
        flat.zero_()
        b = _Bucket(flat, plist)
        off = 0
        for p in plist:
            n = p.numel()
            # same strides as the parameter (e.g. channels_last conv weights): autograd's layout contract
            p.grad = torch.as_strided(flat, p.size(), p.stride(), storage_offset=flat.storage_offset() + off)
            off += -(-n // 8) * 8
            self._bucket_of[p] = b
        self.buckets.append(b)

    # -- autograd side -------------------------------------------------------------------------------------
    def _on_grad_ready(self, p):
        b = self._bucket_of[p]
        b.pending -= 1
        if b.pending == 0:
            self._launch_ready()

    def _launch_ready(self):
        """Collectives are matched across ranks by ISSUE ORDER (flags and epochs carry no bucket id), so buckets are
        launched strictly in index order: bucket i goes out only after buckets 0..i-1 have. A bucket that becomes
        ready early (data-dependent control flow can reorder gradient production between ranks) waits here for
        its predecessors; synchronize() flushes whatever is left, again in index order."""
        while self._next < len(self.buckets) and self.buckets[self._next].pending == 0:
            self._launch(self.buckets[self._next])
            self._next += 1

    def _launch(self, b):
        cur = torch.cuda.current_stream(self.comm.device)
        self.stream.wait_stream(cur)
        with torch.cuda.stream(self.stream):
            self.comm.allreduce(b.flat, scale=self.scale, wire_dtype=self.wire_dtype, stream=self.stream)
            b.event.record(self.stream)
        b.launched = True

    def synchronize(self):
        """Call after backward: launches any bucket whose gradients were not all produced (unused
        parameters) and makes the current stream wait for every bucket's collective."""
        cur = torch.cuda.current_stream(self.comm.device)
        for b in self.buckets:
            if not b.launched:
                self._launch(b)
            cur.wait_event(b.event)
            b.pending = len(b.params)
            b.launched = False
        self._next = 0
        # a watchdog that fired in one of this (or an earlier) iteration's collectives means wrong gradients: the
        # kernel mirrors its error word into mapped host memory, so this check costs no synchronisation
        self.comm.raise_if_failed()

    def zero_grad(self, set_to_none=False):
        for b in self.buckets:
            b.flat.zero_()

    def forward(self, *args, **kwargs):
        return self.module(*args, **kwargs)
