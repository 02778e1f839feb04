"""Multi-process CPU communicator (csrc/host/host_comm.cc): one process per rank, shared-memory windows,
the same algorithm templates as the GPU kernels. Works under any torch.distributed backend (gloo on
machines without GPUs) - this is what the CPU multi-process tests drive, and it is a usable CPU allreduce
in its own right (the reference is exactly that, over MPI: allreduce_over_mpi/mpi_mod.hpp:1167-1221)."""
import ctypes as C

import numpy as np
import torch

from .. import _lib
from ..ops import constants as K

_AG_FN = C.CFUNCTYPE(C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t)


class HostCommunicator:
    def __init__(self, group=None, window_bytes=8 << 20, ll_bytes=1 << 20, max_phantoms=0):
        """``max_phantoms`` > 0 makes room for "-k" trees: ``allreduce(x, widths=[2, 2, 2], lonely=-1)`` on 7 ranks runs
        the 8-position tree with one phantom rank (hosted by rank 0, contributing the identity) - the shapes
        cost_model/ChooseWidth.h:16-22 proposes for prime world sizes."""
        import torch.distributed as dist
        self._L = L = _lib.load()
        vp, i32, i64 = C.c_void_p, C.c_int, C.c_longlong
        L.ft_host_comm_create_ex.restype = i32
        L.ft_host_comm_create_ex.argtypes = [C.POINTER(vp), i32, i32, vp, vp, C.c_size_t, C.c_size_t, i32]
        self.max_phantoms = int(max_phantoms)
        L.ft_host_comm_destroy.argtypes = [vp]
        L.ft_host_comm_registered.restype = vp
        L.ft_host_comm_registered.argtypes = [vp]
        L.ft_host_comm_error.restype = C.c_uint
        L.ft_host_comm_error.argtypes = [vp]
        L.ft_host_comm_allreduce.restype = i64
        L.ft_host_comm_allreduce.argtypes = [vp, vp, vp, i64, i32, i32, i32, i32, C.POINTER(i32), i32, i32, i64, i32,
                                             C.c_double, i32, C.c_ulonglong, i32, i32]
        self.rank, self.size = dist.get_rank(group), dist.get_world_size(group)
        world = self.size

        def _allgather(_ctx, send, recv, nbytes):
            try:
                src = (C.c_ubyte * nbytes).from_address(send)
                t = torch.frombuffer(src, dtype=torch.uint8).clone()
                if "nccl" in dist.get_backend(group):
                    t = t.cuda()
                out = [torch.empty_like(t) for _ in range(world)]
                dist.all_gather(out, t, group=group)
                flat = torch.cat(out).cpu().contiguous()
                C.memmove(recv, flat.data_ptr(), nbytes * world)
                return 0
            except Exception as e:
                import sys
                print(f"[flextree] host bootstrap allgather failed: {e!r}", file=sys.stderr)
                return 1

        self._cb = _AG_FN(_allgather)
        self._h = vp()
        self._ll_bytes = ll_bytes
        self._create(window_bytes)

    def _create(self, window_bytes):
        L, vp = self._L, C.c_void_p
        rc = L.ft_host_comm_create_ex(C.byref(self._h), self.rank, self.size, C.cast(self._cb, vp), None, window_bytes,
                                      self._ll_bytes, self.max_phantoms)
        if rc:
            raise RuntimeError(f"ft_host_comm_create failed ({rc}): {L.ft_last_error().decode()}")
        self.window_bytes = window_bytes

    def reserve(self, nbytes):
        """Grow-only staging, like the reference's flextree_register_the_buffer (mpi_mod.hpp:932-950): make room for
        messages of ``nbytes`` (wire type). Collective when it has to grow (every rank sees the same sizes); views
        returned by registered() before a growth are invalid afterwards. allreduce() calls this itself."""
        need = nbytes + nbytes // 16 + 32 * self.size + 4096  # blocks padded to 16 B + the scale plane of the layout
        if need <= self.window_bytes:
            return
        if getattr(self, "_views_out", False):
            raise RuntimeError(f"a {nbytes}-byte message needs a larger window, and growing it would invalidate the registered "
                               "views handed out earlier: create the communicator with a larger window_bytes")
        win = self.window_bytes
        while win < need:
            win *= 2
        self._L.ft_host_comm_destroy(self._h)
        self._h = C.c_void_p()
        self._create(win)

    def close(self):
        if getattr(self, "_h", None) and self._h.value:
            self._L.ft_host_comm_destroy(self._h)
            self._h = C.c_void_p()

    def registered(self, count, dtype, offset_bytes=0):
        """numpy view of (a piece of) this rank's registered (zero-copy) region. Pieces must start at a multiple of
        16 bytes and at the same offset on every rank; allreduce(view, registered=True) works on them in place."""
        if offset_bytes % 16:
            raise ValueError("registered views start at multiples of 16 bytes")
        ptr = self._L.ft_host_comm_registered(self._h)
        nbytes = count * np.dtype(dtype).itemsize
        if offset_bytes + nbytes > self.window_bytes:
            raise ValueError(f"registered region is {self.window_bytes} bytes (HostCommunicator(window_bytes=...))")
        self._views_out = True
        return np.frombuffer((C.c_char * nbytes).from_address(ptr + offset_bytes), dtype=dtype, count=count)

    def registered_tensor(self, numel, dtype=torch.float32, offset_bytes=0):
        """The same as a torch tensor sharing the memory (fp32/fp64/ints; bf16/fp16 through a uint16 view)."""
        if dtype in (torch.bfloat16, torch.float16):
            return torch.from_numpy(self.registered(numel, np.uint16, offset_bytes)).view(dtype)
        npdt = torch.empty(0, dtype=dtype).numpy().dtype
        return torch.from_numpy(self.registered(numel, npdt, offset_bytes))

    def error(self):
        return int(self._L.ft_host_comm_error(self._h))

    def _call(self, send_ptr, recv_ptr, count, dcode, iocode, op, algo, widths, ctas, tile_bytes, mode, scale, lonely, phases,
              timeout_s):
        widths = list(widths) if widths is not None else [self.size - lonely]
        w = (C.c_int * len(widths))(*widths)
        if mode != K.MODE_REGISTERED:  # a registered call works on a view of the current window: it cannot move
            self.reserve(count * K.DTYPE_SIZE[dcode])
        rc = self._L.ft_host_comm_allreduce(
            self._h, send_ptr, recv_ptr, count, dcode, iocode, K.OPS[op] if isinstance(op, str) else op,
            K.HOST_ALGOS[algo] if isinstance(algo, str) else algo, w, len(widths), ctas, tile_bytes, mode,
            1.0 if scale is None else float(scale), 0 if scale is None else 1, int(timeout_s * 1e9), lonely, phases)
        if rc:
            raise RuntimeError(f"host allreduce failed: code {rc}")

    @staticmethod
    def _as_numpy(x):
        if isinstance(x, torch.Tensor):
            if x.dtype == torch.bfloat16:
                return x.view(torch.int16).numpy().view(np.uint16), K.BFLOAT16
            a = x.numpy()
            return a, K.NUMPY_DTYPE[a.dtype]
        return x, K.NUMPY_DTYPE[x.dtype]

    def allreduce(self, x, out=None, op="sum", algo="stages_bulk", widths=None, ctas=1, tile_bytes=262144, scale=None,
                  wire=None, registered=False, lonely=0, timeout_s=30.0):
        """In-place (or into ``out``) allreduce of a numpy array / CPU tensor.

        ``algo="stages_bulk"`` (default) runs the staged algorithm with the CPU's bulk tiles
        (csrc/host/cpu_backend.h; falls back to the per-vector text where they do not apply), ``"stages"`` the exact
        text of the GPU kernels; ``ctas`` is the number of threads this rank uses. Defaults are the best measured
        shape for MB-sized messages with one rank per core (bench/cpu_compare.py)."""
        a, code = self._as_numpy(x)
        o, _ = self._as_numpy(out) if out is not None else (a, code)
        dcode = code if wire is None else wire
        self._call(a.ctypes.data, o.ctypes.data, a.size, dcode, code, op, algo, widths, ctas, tile_bytes,
                   K.MODE_REGISTERED if registered else K.MODE_STAGING, scale, lonely, 0, timeout_s)
        return out if out is not None else x

    def reduce_scatter(self, x, out, op="sum", widths=None, ctas=1, tile_bytes=262144):
        a, code = self._as_numpy(x)
        o, _ = self._as_numpy(out)
        self._call(a.ctypes.data, o.ctypes.data, a.size, code, code, op, "stages_bulk", widths, ctas, tile_bytes,
                   K.MODE_STAGING, None, 0, 1, 30.0)
        return out

    def all_gather(self, x, out, widths=None, ctas=1, tile_bytes=262144):
        a, code = self._as_numpy(x)
        o, _ = self._as_numpy(out)
        self._call(a.ctypes.data, o.ctypes.data, o.size, code, code, "sum", "stages_bulk", widths, ctas, tile_bytes,
                   K.MODE_STAGING, None, 0, 2, 30.0)
        return out
