"""CPU world: N logical ranks in this process, one thread per "CTA", running the same algorithm
templates as the CUDA kernels (csrc/core/staged_core.h via csrc/host). Used by the CPU tests as the
functional model of the GPU data plane; also a usable CPU allreduce (the reference is a CPU library,
allreduce_over_mpi/mpi_mod.hpp:1167-1221)."""
import ctypes as C

import numpy as np

from .. import _lib
from ..ops import constants as K


class HostWorld:
    def __init__(self, nranks, window_bytes=1 << 20, ll_bytes=1 << 18):
        self._L = _lib.load()
        self.n = nranks
        self._h = self._L.ft_host_world_create(nranks, window_bytes, ll_bytes)
        if not self._h:
            raise ValueError("bad world size")
        self.window_bytes = window_bytes

    def set_jitter(self, permille):
        """Tests: random 0-200 us sleeps at flag publications / acquisitions / syncs of every CPU "CTA" (process-wide)."""
        self._L.ft_host_set_jitter(int(permille))

    def close(self):
        if self._h:
            self._L.ft_host_world_destroy(self._h)
            self._h = None

    __del__ = close

    def registered(self, rank, count, dtype):
        """numpy view of rank's registered (zero-copy) region."""
        ptr = self._L.ft_host_registered(self._h, rank)
        nbytes = count * np.dtype(dtype).itemsize
        buf = (C.c_char * nbytes).from_address(ptr)
        return np.frombuffer(buf, dtype=dtype, count=count)

    def epoch(self, rank=0):
        return int(self._L.ft_host_epoch(self._h, rank))

    def error(self, rank):
        return int(self._L.ft_host_error(self._h, rank))

    def clear_error(self):
        self._L.ft_host_clear_error(self._h)

    def allreduce(self, send, recv=None, op="sum", algo="stages", widths=None, ctas=2, tile_bytes=4096,
                  mode=K.MODE_STAGING, scale=None, dtype_code=None, io_dtype_code=None,
                  timeout_s=20.0, skip_rank=-1, lonely=0, phases=0, count=None):
        """send/recv: lists of numpy arrays (one per rank). recv=None -> in place."""
        n = self.n
        recv = send if recv is None else recv
        count = send[0].size if count is None else count
        if dtype_code is None:
            dtype_code = K.NUMPY_DTYPE[send[0].dtype]
        if io_dtype_code is None:
            io_dtype_code = dtype_code
        vp = C.c_void_p * n
        s = vp(*[a.ctypes.data for a in send])
        r = vp(*[a.ctypes.data for a in recv])
        widths = list(widths) if widths is not None else [n - lonely]
        w = (C.c_int * len(widths))(*widths)
        rc = self._L.ft_host_allreduce(
            self._h, s, r, count, dtype_code, io_dtype_code, K.OPS[op] if isinstance(op, str) else op,
            K.HOST_ALGOS[algo] if isinstance(algo, str) else algo, w, len(widths), ctas, tile_bytes, mode,
            1.0 if scale is None else float(scale), 0 if scale is None else 1,
            int(timeout_s * 1e9), skip_rank, lonely, phases)
        return rc
