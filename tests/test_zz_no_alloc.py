"""ft_allreduce_emulated_noalloc: the allocation-free entry point (static __device__ scratch).

It exists for callers that may not allocate device memory at the point of the call - plugins and
callbacks running inside another framework's loop, services whose GPU memory is entirely owned by a
pool, latency-critical loops that cannot risk an allocation failing or synchronising the device. This
test drives it through the C ABI (ctypes) exactly as such a caller would, and checks that it returns
the same bits as the regular entry point (ft_allreduce_multi on an emulated communicator, one-shot
algorithm) on two consecutive calls - the second call re-uses the scratch array that the first one left
dirty. The file name sorts last on purpose: a run that stops at the first failure has then already
executed every other test. It must also pass on its own in a fresh process."""
import ctypes as C

import numpy as np
import pytest

torch = pytest.importorskip("torch")
import flextree as ft  # noqa: E402
from flextree.ops import constants as K  # noqa: E402

pytestmark = pytest.mark.gpu


def _ptrs(tensors):
    return (C.c_void_p * len(tensors))(*[t.data_ptr() for t in tensors])


def test_noalloc_matches_regular_entry_point_twice():
    L = ft._lib.load()
    assert ft._lib.has_cuda_lib()
    n = 8
    torch.cuda.set_device(0)
    comm = ft.Communicator.emulated(n, 0)   # the EXISTING path, for the expected values only
    comm.set_algo("oneshot")
    stream = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    g = torch.Generator(device="cuda").manual_seed(2024)
    for call, (count, dtype, code) in enumerate([(4096, torch.float32, K.FLOAT), (1237, torch.float32, K.FLOAT)]):
        xs = [torch.randn(count, device="cuda", dtype=dtype, generator=g) for _ in range(n)]
        want = [torch.empty_like(x) for x in xs]
        comm.allreduce_multi(xs, outs=want)
        got = [torch.full_like(x, float("nan")) for x in xs]
        torch.cuda.synchronize()
        rc = L.ft_allreduce_emulated_noalloc(_ptrs(xs), _ptrs(got), n, count, code, K.SUM, stream)
        assert rc == 0, L.ft_last_error().decode()
        torch.cuda.synchronize()
        for r in range(n):
            # same kernel, same summation order: the tolerance of the other tests (exact equality) applies
            assert torch.equal(got[r], want[r]), f"call {call} rank {r}"
        ref = torch.stack(xs).double().sum(0)
        assert torch.allclose(got[0].double(), ref, rtol=1e-5, atol=1e-5)
    # in place + another dtype/op on the same (now twice-used) scratch
    xs = [torch.randint(-1000, 1000, (513,), device="cuda", dtype=torch.int32, generator=g) for _ in range(4)]
    want = torch.stack(xs).amax(0)
    rc = L.ft_allreduce_emulated_noalloc(None, _ptrs(xs), 4, 513, K.INT32, K.MAX, stream)
    assert rc == 0, L.ft_last_error().decode()
    torch.cuda.synchronize()
    for x in xs:
        assert torch.equal(x, want)
    comm.close()


def test_noalloc_world_size_may_change_between_calls():
    """A call with fewer ranks must not leave the idle ranks' epoch words behind: 4 -> 8 -> 2 -> 8 -> 3 ranks,
    every result bit-identical to the regular entry point, and no watchdog error recorded."""
    L = ft._lib.load()
    torch.cuda.set_device(0)
    stream = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    g = torch.Generator(device="cuda").manual_seed(77)
    L.ft_noalloc_error(1)
    for n, count in [(4, 1000), (8, 4096), (2, 33), (8, 777), (3, 2048), (8, 4096)]:
        comm = ft.Communicator.emulated(n, 0)
        comm.set_algo("oneshot")
        xs = [torch.randn(count, device="cuda", generator=g) for _ in range(n)]
        want = [torch.empty_like(x) for x in xs]
        comm.allreduce_multi(xs, outs=want)
        got = [torch.full_like(x, float("nan")) for x in xs]
        torch.cuda.synchronize()
        rc = L.ft_allreduce_emulated_noalloc(_ptrs(xs), _ptrs(got), n, count, K.FLOAT, K.SUM, stream)
        assert rc == 0, L.ft_last_error().decode()
        torch.cuda.synchronize()
        for r in range(n):
            assert torch.equal(got[r], want[r]), f"n={n} rank {r}"
        comm.close()
    assert L.ft_noalloc_error(0) == 0


def test_noalloc_rejects_what_the_static_scratch_cannot_hold():
    L = ft._lib.load()
    torch.cuda.set_device(0)
    stream = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    xs = [torch.zeros(8192, device="cuda") for _ in range(2)]          # 32 KiB > 16 KiB limit
    rc = L.ft_allreduce_emulated_noalloc(None, _ptrs(xs), 2, 8192, K.FLOAT, K.SUM, stream)
    assert rc == 1 and "FT_NOALLOC_MAX_BYTES" in L.ft_last_error().decode()
    ys = [torch.zeros(16, device="cuda") for _ in range(9)]            # 9 ranks > 8
    rc = L.ft_allreduce_emulated_noalloc(None, _ptrs(ys), 9, 16, K.FLOAT, K.SUM, stream)
    assert rc == 1 and "FT_NOALLOC_MAX_RANKS" in L.ft_last_error().decode()
    rc = L.ft_allreduce_emulated_noalloc(None, _ptrs(xs), 2, 16, K.FLOAT, K.BAND, stream)
    assert rc == 3                                                      # float BAND: unsupported
    torch.cuda.synchronize()
