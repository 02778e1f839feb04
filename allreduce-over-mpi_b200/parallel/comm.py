"""Communicators over the native runtime (csrc/runtime/comm.cu).

* ``Communicator.from_process_group(pg)``: one process per GPU; torch.distributed is used only to move
  the memory handles at creation/growth (the ``ft_allgather_fn`` callback) - never for the collective.
* ``Communicator.emulated(n)``: n logical ranks on ONE GPU served by one cooperative launch (tests,
  1-GPU benchmark).

The reference's only "API" is MPI_Allreduce(sendbuf, recvbuf, count, datatype, op, comm)
(allreduce_over_mpi/mpi_mod.hpp:1167-1221); ``allreduce`` below is that call for torch tensors, plus
the fused scale/cast the DDP path needs.
"""
import ctypes as C
import os

import torch

from .. import _lib
from ..ops import constants as K

_AG_FN = C.CFUNCTYPE(C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t)
_NVTX = os.environ.get("FT_NVTX") == "1"


class CommOptions(C.Structure):
    _fields_ = [("window_bytes", C.c_size_t), ("ll_bytes", C.c_size_t), ("use_multicast", C.c_int),
                ("use_vmm", C.c_int), ("timeout_s", C.c_double)]


class CallInfo(C.Structure):
    _fields_ = [("algo", C.c_int), ("nstages", C.c_int), ("widths", C.c_int * 8), ("ctas", C.c_int),
                ("threads", C.c_int), ("tile_bytes", C.c_longlong), ("mode", C.c_int), ("vectorized", C.c_int),
                ("est_us", C.c_double)]

    def as_dict(self):
        return {"algo": K.ALGO_NAMES.get(self.algo, "?"), "widths": list(self.widths[:self.nstages]),
                "ctas": self.ctas, "threads": self.threads, "tile_bytes": self.tile_bytes,
                "mode": "registered" if self.mode else "staging", "vectorized": bool(self.vectorized),
                "est_us": self.est_us}


class FlexTreeError(RuntimeError):
    pass


def _check(L, rc, what):
    if rc != 0:
        raise FlexTreeError(f"{what} failed (code {rc}): {L.ft_last_error().decode()}")


class _RawCudaBuffer:
    """Minimal __cuda_array_interface__ holder so torch can wrap symmetric memory without copying."""

    def __init__(self, ptr, nbytes, owner):
        self.__cuda_array_interface__ = {"shape": (nbytes,), "typestr": "|u1", "data": (ptr, False), "version": 2}
        self._owner = owner


class Communicator:
    def __init__(self):
        self._L = _lib.load()
        if not _lib.has_cuda_lib():
            raise FlexTreeError("flextree: the CUDA library (allreduce-over-mpi_b200/lib/libflextree.so) is not built")
        self._h = C.c_void_p()
        self._fast = _lib.fast()     # compiled per-call path (no ctypes), None if not built
        self._raw_stream = getattr(torch._C, "_cuda_getCurrentRawStream", None)
        self._cb = None
        self.emulated_ranks = 0
        self.device = None
        self._symm = []

    # ---- construction ---------------------------------------------------------------------------------
    @staticmethod
    def _options(window_bytes=0, ll_bytes=0, multicast=-1, vmm=-1, timeout_s=-1.0):
        o = CommOptions()
        o.window_bytes, o.ll_bytes, o.use_multicast, o.use_vmm, o.timeout_s = window_bytes, ll_bytes, multicast, vmm, timeout_s
        return o

    @classmethod
    def from_process_group(cls, group=None, device=None, **opts):
        import torch.distributed as dist
        self = cls()
        rank, world = dist.get_rank(group), dist.get_world_size(group)
        if device is None:
            device = torch.cuda.current_device()
        self.device = torch.device("cuda", device if isinstance(device, int) else device.index)
        backend = dist.get_backend(group)

        def _allgather(_ctx, send, recv, nbytes):
            try:
                src = (C.c_ubyte * nbytes).from_address(send)
                t = torch.frombuffer(src, dtype=torch.uint8).clone()
                if "nccl" in backend:
                    t = t.to(self.device)
                out = [torch.empty_like(t) for _ in range(world)]
                dist.all_gather(out, t, group=group)
                flat = torch.cat(out).cpu().contiguous()
                C.memmove(recv, flat.data_ptr(), nbytes * world)
                return 0
            except Exception as e:  # never let an exception cross the C boundary
                import sys
                print(f"[flextree] bootstrap allgather failed: {e!r}", file=sys.stderr)
                return 1

        self._cb = _AG_FN(_allgather)
        o = cls._options(**opts)
        with torch.cuda.device(self.device):
            rc = self._L.ft_comm_create(C.byref(self._h), rank, world, self.device.index,
                                        C.cast(self._cb, C.c_void_p), None, C.cast(C.byref(o), C.c_void_p))
        _check(self._L, rc, "ft_comm_create")
        return self

    @classmethod
    def emulated(cls, nranks, device=0, **opts):
        self = cls()
        self.device = torch.device("cuda", device)
        o = cls._options(**opts)
        rc = self._L.ft_comm_create_emulated(C.byref(self._h), nranks, device, C.cast(C.byref(o), C.c_void_p))
        _check(self._L, rc, "ft_comm_create_emulated")
        self.emulated_ranks = nranks
        return self

    def close(self):
        if getattr(self, "_h", None) and self._h.value:
            self._L.ft_comm_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ---- properties -------------------------------------------------------------------------------------
    @property
    def rank(self):
        return self._L.ft_comm_rank(self._h)

    @property
    def size(self):
        return self._L.ft_comm_size(self._h)

    @property
    def has_multicast(self):
        return bool(self._L.ft_comm_has_multicast(self._h))

    def error(self):
        """Sticky device error word (0 = healthy). Synchronizes the device."""
        torch.cuda.synchronize(self.device)
        return int(self._L.ft_comm_error(self._h))

    def poll_error(self):
        """The same word without any synchronisation (mapped host mirror written by the failing kernel)."""
        if self._fast is not None and self._h.value:
            return self._fast.error_peek(self._h.value)
        return int(self._L.ft_comm_error_peek(self._h))

    def raise_if_failed(self):
        """Raise FlexTreeError if a device watchdog has fired on this communicator (non-blocking)."""
        e = self.poll_error()
        if e:
            raise FlexTreeError(f"flextree: device watchdog fired (error word {e:#x}: code {e & 0xff}, phase {(e >> 8) & 0xff}, "
                                f"waiting for rank {(e >> 16) & 0xff}, on rank {e >> 24}); the results of that collective are invalid")

    def set_tuning(self, key, value):
        """Kernel-variant / launch-shape knobs for scans (ft_comm_set_tuning): nvls_variant, nvls_flags, nvls_ctas,
        nvls_max_tile_bytes."""
        _check(self._L, self._L.ft_comm_set_tuning(self._h, key.encode(), int(value)), "ft_comm_set_tuning")

    def link_probe(self, mode, npeers, bytes_per_peer, ctas, stream=None):
        """Raw NVLink traffic (ft_link_probe): mode 'ldg_pull' | 'stg_push' | 'tma_pull' | 'tma_push'."""
        m = {"ldg_pull": 0, "stg_push": 1, "tma_pull": 2, "tma_push": 3}[mode] if isinstance(mode, str) else int(mode)
        _check(self._L, self._L.ft_link_probe(self._h, m, int(npeers), int(bytes_per_peer), int(ctas), self._stream(stream)),
               "ft_link_probe")

    def inject_fault(self, rank):
        """Tests/drills: logical rank `rank` plays dead in the following collectives (None = off)."""
        _check(self._L, self._L.ft_comm_inject_fault(self._h, -1 if rank is None else int(rank)), "ft_comm_inject_fault")

    def clear_error(self):
        _check(self._L, self._L.ft_comm_clear_error(self._h), "ft_comm_clear_error")

    def last_call(self):
        info = CallInfo()
        self._L.ft_comm_last_call(self._h, C.cast(C.byref(info), C.c_void_p))
        return info.as_dict()

    # ---- policy -------------------------------------------------------------------------------------------
    def set_widths(self, widths):
        if widths is None:
            _check(self._L, self._L.ft_comm_set_widths(self._h, None, 0), "ft_comm_set_widths")
        else:
            a = (C.c_int * len(widths))(*widths)
            _check(self._L, self._L.ft_comm_set_widths(self._h, a, len(widths)), "ft_comm_set_widths")

    def set_algo(self, algo):
        code = K.ALGOS[algo] if isinstance(algo, str) else int(algo)
        _check(self._L, self._L.ft_comm_set_algo(self._h, code), "ft_comm_set_algo")

    def set_launch(self, ctas=0, threads=0, tile_bytes=0):
        _check(self._L, self._L.ft_comm_set_launch(self._h, ctas, threads, tile_bytes), "ft_comm_set_launch")

    def trace_enable(self, on=True):
        """Record device timestamps at the phase boundaries of the staged kernel (SHOW_TIME parity)."""
        _check(self._L, self._L.ft_comm_trace_enable(self._h, int(bool(on))), "ft_comm_trace_enable")

    def trace(self, local_rank=0):
        """[(label, microseconds since kernel start)] for the last traced staged call."""
        torch.cuda.synchronize(self.device)
        buf = (C.c_ulonglong * 32)()
        n = self._L.ft_comm_trace_read(self._h, local_rank, buf, 32)
        if n <= 0:
            return []
        m = (n - 3) // 2
        labels = ["start", "window published"] + [f"reduce-scatter stage {k} done" for k in range(m)] + \
                 [f"all-gather stage {m - 1 - i} done" for i in range(m)] + ["exit"]
        return [(labels[i] if i < len(labels) else f"slot {i}", (buf[i] - buf[0]) / 1e3) for i in range(n)]

    def set_max_ctas(self, n):
        """Cap the SMs a collective may occupy (0 = whole GPU)."""
        _check(self._L, self._L.ft_comm_set_max_ctas(self._h, int(n)), "ft_comm_set_max_ctas")

    def set_tma(self, on):
        """P2P stages: True = TMA-staged tiles (cp.async.bulk ring), False = register-staged loads, None = default."""
        _check(self._L, self._L.ft_comm_set_tma(self._h, -1 if on is None else int(bool(on))), "ft_comm_set_tma")

    def set_link_model(self, model):
        _check(self._L, self._L.ft_comm_set_link_model(self._h, model._h if model is not None else None),
               "ft_comm_set_link_model")

    def reserve(self, nbytes):
        """Collective: make the staging window large enough for messages of nbytes (do this before
        capturing a CUDA graph)."""
        with torch.cuda.device(self.device):
            _check(self._L, self._L.ft_comm_reserve(self._h, nbytes), "ft_comm_reserve")

    # ---- symmetric tensors ----------------------------------------------------------------------------------
    def symm_empty(self, shape, dtype=torch.float32):
        """Collective: a tensor (a list of them for an emulated communicator) living in symmetric
        memory; all-reducing it in place takes the zero-copy path."""
        if isinstance(shape, int):
            shape = (shape,)
        numel = 1
        for s in shape:
            numel *= s
        nbytes = max(16, numel * torch.empty((), dtype=dtype).element_size())
        n = max(1, self.emulated_ranks)
        ptrs = (C.c_void_p * n)()
        with torch.cuda.device(self.device):
            _check(self._L, self._L.ft_symm_alloc(self._h, nbytes, ptrs), "ft_symm_alloc")
        outs = []
        for p in ptrs:
            raw = torch.as_tensor(_RawCudaBuffer(p, nbytes, self), device=self.device)
            outs.append(raw.view(dtype)[:numel].view(*shape))
        self._symm.append((list(ptrs), outs))
        return outs if self.emulated_ranks else outs[0]

    # ---- collectives ------------------------------------------------------------------------------------------
    def _stream(self, stream):
        s = stream if stream is not None else torch.cuda.current_stream(self.device)
        return C.c_void_p(s.cuda_stream)

    @staticmethod
    def _wire_code(dt, wire_dtype, op="sum"):
        if wire_dtype is None:
            return K.default_wire(dt, op)  # FT_WIRE, default native
        if isinstance(wire_dtype, str):
            if wire_dtype in ("fp8", "fp8-block16", "fp8_block"):
                return K.FP8_BLOCK  # opt-in block-scaled e4m3 wire
            wire_dtype = {"bf16": torch.bfloat16, "fp16": torch.float16, "fp32": torch.float32}[wire_dtype]
        return K.torch_dtype_code(wire_dtype)

    def allreduce(self, tensor, op="sum", out=None, scale=1.0, wire_dtype=None, stream=None):
        """In place (``out=None``) or out of place allreduce of a contiguous CUDA tensor.

        ``scale`` is fused into the last stage (``out = cast(sum * scale)``); ``wire_dtype`` (bf16/fp16) moves
        fp32 tensors in 16 bits."""
        if self.emulated_ranks:
            raise FlexTreeError("emulated communicator: use allreduce_multi")
        out = tensor if out is None else out
        if not (tensor.is_contiguous() and out.is_contiguous()):
            raise ValueError("flextree.allreduce needs contiguous tensors")
        if out.dtype != tensor.dtype or out.numel() != tensor.numel():
            raise ValueError("out must match the input dtype/size")
        dt = K.torch_dtype_code(tensor.dtype)
        wt = self._wire_code(dt, wire_dtype, op)
        opc = K.OPS[op] if isinstance(op, str) else int(op)
        # (the native call selects the communicator's device itself; no torch device context on the hot path)
        if _NVTX:  # FT_NVTX=1: one range per call for nsys / ncu timelines
            torch.cuda.nvtx.range_push(f"flextree.allreduce {tensor.numel() * tensor.element_size()} B")
        if self._fast is not None:
            # one vectorcall into the compiled module: plain integers in, result code out
            if stream is not None:
                sp = stream.cuda_stream
            elif self._raw_stream is not None:
                sp = self._raw_stream(self.device.index)
            else:
                sp = torch.cuda.current_stream(self.device).cuda_stream
            rc = self._fast.allreduce_ex(self._h.value, tensor.data_ptr(), out.data_ptr(), tensor.numel(), dt, wt, opc,
                                         float(scale), sp)
        else:
            rc = self._L.ft_allreduce_ex(tensor.data_ptr(), out.data_ptr(), tensor.numel(),
                                         dt, wt, opc, float(scale), self._h, self._stream(stream))
        if _NVTX:
            torch.cuda.nvtx.range_pop()
        if rc:
            _check(self._L, rc, "ft_allreduce")
        return out

    def allreduce_multi(self, tensors, op="sum", outs=None, scale=1.0, wire_dtype=None, stream=None):
        """Emulated communicator: one tensor per logical rank."""
        n = self.emulated_ranks
        if not n or len(tensors) != n:
            raise ValueError("need one tensor per emulated rank")
        outs = tensors if outs is None else outs
        dt = K.torch_dtype_code(tensors[0].dtype)
        wt = self._wire_code(dt, wire_dtype, op)
        opc = K.OPS[op] if isinstance(op, str) else int(op)
        s = (C.c_void_p * n)(*[t.data_ptr() for t in tensors])
        r = (C.c_void_p * n)(*[t.data_ptr() for t in outs])
        rc = self._L.ft_allreduce_multi(s, r, tensors[0].numel(), dt, wt, opc, float(scale), self._h,
                                        self._stream(stream))
        if rc:
            _check(self._L, rc, "ft_allreduce_multi")
        return outs

    def allreduce_host(self, tensors, outs=None, op="sum", scale=1.0, chunk_bytes=0):
        """Allreduce of HOST tensors (pinned for full speed): ``outs`` (default: in place) hold the result when the
        call returns. Chunks of the message are pipelined H2D -> in-place zero-copy allreduce -> D2H over three
        streams (csrc/runtime/host_pipe.cu). One tensor per logical rank for an emulated communicator, else a single
        tensor. This is the reference's calling convention: MPI_Allreduce on host vectors (benchmark.cpp:149-159)."""
        single = isinstance(tensors, torch.Tensor)
        ts = [tensors] if single else list(tensors)
        os_ = ts if outs is None else ([outs] if isinstance(outs, torch.Tensor) else list(outs))
        n = max(1, self.emulated_ranks)
        if len(ts) != n or len(os_) != n:
            raise ValueError("need one host tensor per logical rank")
        for t, o in zip(ts, os_):
            if t.is_cuda or o.is_cuda or not (t.is_contiguous() and o.is_contiguous()):
                raise ValueError("allreduce_host takes contiguous CPU tensors")
            if o.dtype != t.dtype or o.numel() != t.numel():
                raise ValueError("out must match the input dtype/size")
        dt = K.torch_dtype_code(ts[0].dtype)
        opc = K.OPS[op] if isinstance(op, str) else int(op)
        s = (C.c_void_p * n)(*[t.data_ptr() for t in ts])
        r = (C.c_void_p * n)(*[t.data_ptr() for t in os_])
        rc = self._L.ft_allreduce_host_multi(s, r, ts[0].numel(), dt, opc, float(scale), self._h, int(chunk_bytes))
        if rc:
            _check(self._L, rc, "ft_allreduce_host")
        return os_[0] if single else os_

    # ---- the halves ------------------------------------------------------------------------------------------------
    def reduce_scatter(self, tensor, out, op="sum", stream=None):
        """out (numel/size elements) = this rank's block of the reduction of `tensor` across ranks."""
        n = self.size
        if tensor.numel() != out.numel() * n or tensor.dtype != out.dtype:
            raise ValueError("reduce_scatter: input must hold size * out.numel() elements of out's dtype")
        opc = K.OPS[op] if isinstance(op, str) else int(op)
        with torch.cuda.device(self.device):
            rc = self._L.ft_reduce_scatter(C.c_void_p(tensor.data_ptr()), C.c_void_p(out.data_ptr()), out.numel(),
                                           K.torch_dtype_code(out.dtype), opc, self._h, self._stream(stream))
        _check(self._L, rc, "ft_reduce_scatter")
        return out

    def all_gather(self, tensor, out, stream=None):
        """out (size * numel elements) = the blocks of all ranks in rank order."""
        n = self.size
        if out.numel() != tensor.numel() * n or tensor.dtype != out.dtype:
            raise ValueError("all_gather: out must hold size * input.numel() elements of the input dtype")
        with torch.cuda.device(self.device):
            rc = self._L.ft_all_gather(C.c_void_p(tensor.data_ptr()), C.c_void_p(out.data_ptr()), tensor.numel(),
                                       K.torch_dtype_code(out.dtype), self._h, self._stream(stream))
        _check(self._L, rc, "ft_all_gather")
        return out

    def reduce_scatter_multi(self, tensors, outs, op="sum", stream=None):
        n = self.emulated_ranks
        opc = K.OPS[op] if isinstance(op, str) else int(op)
        s = (C.c_void_p * n)(*[t.data_ptr() for t in tensors])
        r = (C.c_void_p * n)(*[t.data_ptr() for t in outs])
        with torch.cuda.device(self.device):
            rc = self._L.ft_reduce_scatter_multi(s, r, outs[0].numel(), K.torch_dtype_code(outs[0].dtype), opc, self._h,
                                                 self._stream(stream))
        _check(self._L, rc, "ft_reduce_scatter_multi")
        return outs

    def all_gather_multi(self, tensors, outs, stream=None):
        n = self.emulated_ranks
        s = (C.c_void_p * n)(*[t.data_ptr() for t in tensors])
        r = (C.c_void_p * n)(*[t.data_ptr() for t in outs])
        with torch.cuda.device(self.device):
            rc = self._L.ft_all_gather_multi(s, r, tensors[0].numel(), K.torch_dtype_code(outs[0].dtype), self._h,
                                             self._stream(stream))
        _check(self._L, rc, "ft_all_gather_multi")
        return outs
