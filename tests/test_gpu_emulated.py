"""GPU tests on ONE device: N logical ranks served by one cooperative launch of the real sm_100a
kernels (same device code as the multi-GPU path; every "peer" pointer lands in local HBM).
Oracle: the independent numpy model in refimpl.py (plain fp32/fp64 PyTorch-free reference)."""
import os

import numpy as np
import pytest

os.environ.setdefault("FT_TIMEOUT_S", "5")

torch = pytest.importorskip("torch")
import flextree as ft  # noqa: E402
from refimpl import block_size, round_bf16, staged_allreduce  # noqa: E402

pytestmark = pytest.mark.gpu

RNG = np.random.default_rng(7)


@pytest.fixture(scope="module")
def comms():
    cache = {}

    def get(n):
        if n not in cache:
            cache[n] = ft.Communicator.emulated(n, 0, window_bytes=8 << 20)
        return cache[n]

    yield get
    for c in cache.values():
        c.close()


def dev(a):
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


def run(comm, arrays, out_of_place=False, **kw):
    ts = [dev(a) for a in arrays]
    outs = [torch.full_like(t, 0) for t in ts] if out_of_place else None
    res = comm.allreduce_multi(ts, outs=outs, **kw)
    torch.cuda.synchronize()
    assert comm.error() == 0, hex(comm.error())
    return [t.cpu().numpy() for t in res], [t.cpu().numpy() for t in ts]


WIDTHS = {2: [[2]], 4: [[4], [2, 2]], 6: [[6], [2, 3], [3, 2]], 8: [[8], [2, 4], [4, 2], [2, 2, 2]]}


@pytest.mark.parametrize("n", [2, 4, 6, 8])
@pytest.mark.parametrize("count", [1, 7, 35, 1001, 65536 + 3])
def test_fp32_sum_bit_exact(comms, n, count):
    c = comms(n)
    c.set_algo("stages")
    xs = [RNG.standard_normal(count).astype(np.float32) for _ in range(n)]
    B = block_size(count, n, 4)
    for widths in WIDTHS[n]:
        c.set_widths(widths)
        want = staged_allreduce(xs, widths, block=B)
        got, src = run(c, xs, out_of_place=True)
        for r in range(n):
            np.testing.assert_array_equal(got[r], want[r], err_msg=f"{widths} rank {r}")
            np.testing.assert_array_equal(src[r], xs[r])
        got, _ = run(c, xs)
        for r in range(n):
            np.testing.assert_array_equal(got[r], want[r])
        info = c.last_call()
        assert info["algo"] == "stages" and info["widths"] == widths
    c.set_widths(None)
    c.set_algo("auto")


@pytest.mark.parametrize("dtype", [np.uint8, np.int8, np.int16, np.int32, np.int64, np.float64, np.bool_])
@pytest.mark.parametrize("op", ["sum", "max", "min", "band", "bor"])
def test_dtype_op_matrix(comms, dtype, op):
    n, count = 6, 4099
    c = comms(n)
    dt = np.dtype(dtype)
    if dt.kind == "f":
        xs = [RNG.standard_normal(count).astype(dt) for _ in range(n)]
    elif dt.kind == "b":
        xs = [RNG.integers(0, 2, count).astype(dt) for _ in range(n)]
    else:
        info = np.iinfo(dt)
        xs = [RNG.integers(info.min, info.max, count, dtype=dt, endpoint=True) for _ in range(n)]
    c.set_algo("stages")
    c.set_widths([2, 3])
    if dt.kind == "f" and op in ("band", "bor"):
        with pytest.raises(ft.FlexTreeError):
            run(c, xs, op=op)
        return
    got, _ = run(c, xs, op=op)
    npop = {"sum": np.add, "max": np.maximum, "min": np.minimum, "band": np.bitwise_and, "bor": np.bitwise_or}[op]
    if dt.kind == "b":
        st = np.stack(xs)
        want = st.any(0) if op in ("sum", "max", "bor") else st.all(0)
    elif dt.kind == "f":
        want = staged_allreduce(xs, [2, 3], op=npop, block=block_size(count, n, dt.itemsize))[0]
    else:
        with np.errstate(over="ignore"):
            want = xs[0].copy()
            for x in xs[1:]:
                want = npop(want, x)
    for r in range(n):
        np.testing.assert_array_equal(got[r], want)
    c.set_widths(None)
    c.set_algo("auto")


def test_bf16_and_fp16(comms):
    n, count, widths = 8, 40001, [2, 4]
    c = comms(n)
    c.set_algo("stages")
    c.set_widths(widths)
    xs = [RNG.standard_normal(count).astype(np.float32) for _ in range(n)]
    for tdt, rnd in ((torch.bfloat16, round_bf16), (torch.float16, lambda a: a.astype(np.float16).astype(np.float32))):
        ts = [torch.from_numpy(x).cuda().to(tdt) for x in xs]
        ins = [t.float().cpu().numpy() for t in ts]
        c.allreduce_multi(ts)
        torch.cuda.synchronize()
        assert c.error() == 0
        want = staged_allreduce(ins, widths, block=block_size(count, n, 2), round_fn=rnd)[0]
        for r in range(n):
            np.testing.assert_array_equal(ts[r].float().cpu().numpy(), want)
    c.set_widths(None)
    c.set_algo("auto")


def test_fp32_grads_bf16_wire_fused_scale(comms):
    n, count = 4, 123457
    c = comms(n)
    c.set_algo("stages")
    xs = [RNG.standard_normal(count).astype(np.float32) for _ in range(n)]
    ts = [dev(x) for x in xs]
    c.allreduce_multi(ts, scale=1.0 / n, wire_dtype=torch.bfloat16)
    torch.cuda.synchronize()
    assert c.error() == 0
    want = staged_allreduce([round_bf16(x) for x in xs], [4], block=block_size(count, n, 2), round_fn=round_bf16,
                            final_scale=1.0 / n)[0]
    for r in range(n):
        np.testing.assert_array_equal(ts[r].cpu().numpy(), want)
    ref = torch.stack([dev(x) for x in xs]).double().mean(0).float().cpu().numpy()
    assert np.max(np.abs(ts[0].cpu().numpy() - ref)) < 0.03
    c.set_algo("auto")


@pytest.mark.parametrize("n", [2, 8])
def test_oneshot_and_ring(comms, n):
    c = comms(n)
    for count in (1, 33, 5000):
        xs = [RNG.standard_normal(count).astype(np.float32) for _ in range(n)]
        exact = np.sum(np.stack(xs).astype(np.float64), 0)
        acc = xs[0].copy()
        for x in xs[1:]:
            acc = acc + x
        c.set_algo("oneshot")
        got, _ = run(c, xs, out_of_place=True)
        for r in range(n):
            np.testing.assert_array_equal(got[r], acc)  # ascending-rank order on every rank
        c.set_algo("ring")
        got, _ = run(c, xs)
        for r in range(n):
            np.testing.assert_allclose(got[r], exact, rtol=1e-5, atol=1e-5)
            np.testing.assert_array_equal(got[r], got[0])
    c.set_algo("auto")


def test_auto_selection_small_is_oneshot_large_is_stages(comms):
    c = comms(8)
    xs = [RNG.standard_normal(256).astype(np.float32) for _ in range(8)]
    run(c, xs)
    assert c.last_call()["algo"] == "oneshot"
    xs = [RNG.standard_normal(1 << 18).astype(np.float32) for _ in range(8)]
    got, _ = run(c, xs)
    assert c.last_call()["algo"] in ("twoshot", "oneshot")     # 1 MiB: still an LL size
    np.testing.assert_allclose(got[0], np.sum(np.stack(xs).astype(np.float64), 0), rtol=1e-5, atol=1e-5)
    xs = [RNG.standard_normal(1 << 22).astype(np.float32) for _ in range(8)]   # 16 MiB: bandwidth kernels
    got, _ = run(c, xs)
    assert c.last_call()["algo"] == "stages"
    np.testing.assert_allclose(got[3], np.sum(np.stack(xs).astype(np.float64), 0), rtol=1e-4, atol=1e-4)


def test_registered_zero_copy_in_place(comms):
    n, count = 8, 300000
    c = comms(n)
    c.set_algo("stages")
    bufs = c.symm_empty(count, torch.float32)
    xs = [RNG.standard_normal(count).astype(np.float32) for _ in range(n)]
    for widths in ([8], [2, 4]):
        c.set_widths(widths)
        for it in range(3):
            for r in range(n):
                bufs[r].copy_(dev(xs[r]))
            c.allreduce_multi(bufs)
            torch.cuda.synchronize()
            assert c.error() == 0
            assert c.last_call()["mode"] == "registered"
            want = staged_allreduce(xs, widths, block=block_size(count, n, 4))[0]
            for r in range(n):
                np.testing.assert_array_equal(bufs[r].cpu().numpy(), want)
    c.set_widths(None)
    c.set_algo("auto")


def test_odd_prefix_of_symmetric_tensor_does_not_touch_the_rest(comms):
    """A ragged view of a registered buffer must not take the zero-copy path (full-vector stores would spill
    past the view)."""
    n = 4
    c = comms(n)
    c.set_algo("stages")
    bufs = c.symm_empty(4096, torch.float32)
    for r in range(n):
        bufs[r].fill_(float(r + 1))
    views = [b[:1001] for b in bufs]
    c.allreduce_multi(views)
    torch.cuda.synchronize()
    assert c.error() == 0 and c.last_call()["mode"] == "staging"
    for r in range(n):
        assert torch.all(bufs[r][:1001] == 10.0) and torch.all(bufs[r][1001:] == float(r + 1))
    views = [b[:1024] for b in bufs]   # whole vectors: zero-copy
    c.allreduce_multi(views)
    torch.cuda.synchronize()
    assert c.last_call()["mode"] == "registered"
    for r in range(n):
        assert torch.all(bufs[r][:1001] == 40.0) and torch.all(bufs[r][1024:] == float(r + 1))
    c.set_algo("auto")


def test_unaligned_views_use_scalar_path(comms):
    n, count = 4, 10001
    c = comms(n)
    c.set_algo("stages")
    xs = [RNG.standard_normal(count).astype(np.float32) for _ in range(n)]
    base = [torch.zeros(count + 1, device="cuda") for _ in range(n)]
    views = [b[1:] for b in base]  # 4-byte aligned only
    for v, x in zip(views, xs):
        v.copy_(dev(x))
    c.allreduce_multi(views)
    torch.cuda.synchronize()
    assert c.error() == 0 and c.last_call()["vectorized"] is False
    want = staged_allreduce(xs, [4], block=block_size(count, n, 4, vectorized=False))[0]
    for r in range(n):
        np.testing.assert_array_equal(views[r].cpu().numpy(), want)
    c.set_algo("auto")


def test_cuda_graph_capture_and_replay(comms):
    """Epochs live on the device, so a captured allreduce can be replayed without host resets."""
    n, count = 4, 50000
    c = comms(n)
    c.set_algo("stages")
    c.reserve(count * 4)
    ts = [torch.ones(count, device="cuda") * (r + 1) for r in range(n)]
    outs = [torch.zeros(count, device="cuda") for _ in range(n)]
    c.allreduce_multi(ts, outs=outs)  # warm-up outside capture
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    s = torch.cuda.Stream()
    with torch.cuda.stream(s):
        g.capture_begin()
        c.allreduce_multi(ts, outs=outs, stream=s)
        g.capture_end()
    for it in range(20):
        for r in range(n):
            ts[r].fill_(float(r + 1 + it))
        g.replay()
        torch.cuda.synchronize()
        want = float(sum(r + 1 + it for r in range(n)))
        for r in range(n):
            assert torch.all(outs[r] == want), (it, r)
    assert c.error() == 0
    c.set_algo("auto")


@pytest.mark.parametrize("n,widths", [(2, [2]), (4, [2, 2]), (8, [8]), (8, [2, 4]), (6, [3, 2])])
def test_tma_staged_tiles_match_register_path(comms, n, widths):
    """The TMA variant (cp.async.bulk ring + mbarriers) must give the same bits as the LDG variant."""
    c = comms(n)
    c.set_algo("stages")
    c.set_widths(widths)
    for count, dt in ((4096, np.float32), (1 << 20, np.float32), (300000, np.float32), (65536 + 8, np.int32)):
        if np.dtype(dt).kind == "f":
            xs = [RNG.standard_normal(count).astype(dt) for _ in range(n)]
        else:
            xs = [RNG.integers(-10**6, 10**6, count, dtype=dt) for _ in range(n)]
        c.set_tma(False)
        ref, _ = run(c, xs, out_of_place=True)
        c.set_tma(True)
        got, src = run(c, xs, out_of_place=True)
        got2, _ = run(c, xs)
        for r in range(n):
            np.testing.assert_array_equal(got[r], ref[r], err_msg=f"{widths} count {count} rank {r}")
            np.testing.assert_array_equal(got2[r], ref[r])
            np.testing.assert_array_equal(src[r], xs[r])
    # bf16 wire + fused scale on fp32 buffers, and registered buffers, through the TMA kernel
    xs = [RNG.standard_normal(40000).astype(np.float32) for _ in range(n)]
    c.set_tma(False)
    ts = [dev(x) for x in xs]
    c.allreduce_multi(ts, scale=0.5, wire_dtype=torch.bfloat16)
    c.set_tma(True)
    ts2 = [dev(x) for x in xs]
    c.allreduce_multi(ts2, scale=0.5, wire_dtype=torch.bfloat16)
    torch.cuda.synchronize()
    assert c.error() == 0
    for a_, b_ in zip(ts, ts2):
        assert torch.equal(a_, b_)
    bufs = c.symm_empty(100000, torch.float32)
    for it in range(2):
        for r in range(n):
            bufs[r].copy_(dev(xs[r][:100000] if xs[r].size >= 100000 else np.resize(xs[r], 100000)))
        expect = [b.clone() for b in bufs]
        c.set_tma(False)
        c.allreduce_multi(expect)
        c.set_tma(True)
        c.allreduce_multi(bufs)
        torch.cuda.synchronize()
        assert c.error() == 0 and c.last_call()["mode"] == "registered"
        for a_, b_ in zip(expect, bufs):
            assert torch.equal(a_, b_)
    c.set_tma(None)
    c.set_widths(None)
    c.set_algo("auto")


@pytest.mark.parametrize("n,widths", [(7, [2, 3]), (5, [2, 2]), (8, [3, 2])])
def test_lonely_ranks(n, widths):
    """A tree smaller than the world: the remaining ranks are "lonely" (reference "+k" trees)."""
    c = ft.Communicator.emulated(n, 0, window_bytes=4 << 20)
    c.set_algo("stages")
    c.set_widths(widths)
    T = int(np.prod(widths))
    for count in (3, 1001, 200000):
        xs = [RNG.standard_normal(count).astype(np.float32) for _ in range(n)]
        got, src = run(c, xs, out_of_place=True)
        want = staged_allreduce(xs[:T], widths, block=block_size(count, T, 4))[0]
        for x in xs[T:]:
            want = want + x
        for r in range(n):
            np.testing.assert_array_equal(got[r], want, err_msg=f"count {count} rank {r}")
            np.testing.assert_array_equal(src[r], xs[r])
    ints = [RNG.integers(-1000, 1000, 5000, dtype=np.int32) for _ in range(n)]
    got, _ = run(c, ints, op="max")
    for r in range(n):
        np.testing.assert_array_equal(got[r], np.max(np.stack(ints), 0))
    c.close()


def test_fp8_block_scaled_wire_matches_cpu_model(comms):
    """Opt-in fp8 wire on the GPU: same bytes as the CPU model of the same templates, error within bounds."""
    n, count, widths = 8, 100000, [2, 4]
    c = comms(n)
    c.set_algo("stages")
    c.set_widths(widths)
    xs = [(RNG.standard_normal(count) * 3).astype(np.float32) for _ in range(n)]
    ts = [dev(x) for x in xs]
    c.allreduce_multi(ts, scale=1.0 / n, wire_dtype="fp8")
    torch.cuda.synchronize()
    assert c.error() == 0
    w = ft.HostWorld(n, window_bytes=1 << 20)
    bufs = []
    for x in xs:
        raw = np.empty(count * 4 + 64, dtype=np.uint8)
        off = (-raw.ctypes.data) % 16
        v = raw[off:off + count * 4].view(np.float32)
        v[:] = x
        bufs.append(v)
    info = c.last_call()
    assert w.allreduce(bufs, None, widths=widths, ctas=info["ctas"], tile_bytes=info["tile_bytes"],
                       dtype_code=ft.FP8_BLOCK, io_dtype_code=ft.FLOAT, scale=1.0 / n) == 0
    for r in range(n):
        np.testing.assert_array_equal(ts[r].cpu().numpy(), bufs[r])
    exact = np.mean(np.stack(xs).astype(np.float64), 0)
    rel = np.sqrt(np.mean((ts[0].cpu().numpy() - exact) ** 2)) / np.sqrt(np.mean(exact ** 2))
    assert rel < 0.15, rel
    # bf16 buffers too
    tb = [dev(x).to(torch.bfloat16) for x in xs]
    c.allreduce_multi(tb, scale=1.0 / n, wire_dtype="fp8")
    torch.cuda.synchronize()
    assert c.error() == 0
    rel = np.sqrt(np.mean((tb[0].float().cpu().numpy() - exact) ** 2)) / np.sqrt(np.mean(exact ** 2))
    assert rel < 0.15, rel
    c.set_widths(None)
    c.set_algo("auto")


@pytest.mark.parametrize("n,widths,per_rank", [(8, [2, 4], 4096), (8, [8], 1000), (4, [2, 2], 65536), (6, [3, 2], 37)])
def test_reduce_scatter_and_all_gather(comms, n, widths, per_rank):
    c = comms(n)
    c.set_algo("stages")
    c.set_widths(widths)
    total = per_rank * n
    xs = [torch.randn(total, device="cuda") for _ in range(n)]
    outs = [torch.zeros(per_rank, device="cuda") for _ in range(n)]
    c.reduce_scatter_multi(xs, outs)
    torch.cuda.synchronize()
    assert c.error() == 0
    full = torch.stack(xs).double().sum(0)
    for r in range(n):
        assert torch.allclose(outs[r].double(), full[r * per_rank:(r + 1) * per_rank], rtol=1e-5, atol=1e-5)
    gathered = [torch.zeros(total, device="cuda") for _ in range(n)]
    c.all_gather_multi(outs, gathered)
    torch.cuda.synchronize()
    assert c.error() == 0
    want = torch.cat(outs)
    for r in range(n):
        assert torch.equal(gathered[r], want)
    c.set_widths(None)
    c.set_algo("auto")


def test_phase_trace(comms):
    c = comms(8)
    c.set_algo("stages")
    c.set_widths([2, 4])
    c.trace_enable(True)
    xs = [RNG.standard_normal(1 << 18).astype(np.float32) for _ in range(8)]
    run(c, xs)
    tr = c.trace(3)
    labels = [l for l, _ in tr]
    assert labels == ["start", "window published", "reduce-scatter stage 0 done", "reduce-scatter stage 1 done",
                      "all-gather stage 1 done", "all-gather stage 0 done", "exit"]
    times = [t for _, t in tr]
    assert times == sorted(times) and times[-1] > 0
    c.trace_enable(False)
    c.set_widths(None)
    c.set_algo("auto")


def test_single_rank_fast_path():
    c = ft.Communicator.emulated(1, 0)
    x = torch.arange(1000, device="cuda", dtype=torch.float32)
    y = torch.zeros_like(x)
    c.allreduce_multi([x], outs=[y])
    torch.cuda.synchronize()
    assert torch.equal(x, y)
    c.allreduce_multi([x], outs=[y], scale=0.5)
    torch.cuda.synchronize()
    assert torch.equal(y, x * 0.5)
    c.close()


def test_native_library_is_loaded():
    """Guard against a silent fallback: the CUDA .so must be the one mapped into this process."""
    assert ft._lib.has_cuda_lib()
    maps = open("/proc/self/maps").read()
    assert "libflextree.so" in maps


def test_fired_watchdog_is_visible_on_the_next_call():
    """A dead rank (fault injection) makes its peers time out; the kernel mirrors the error word into mapped host
    memory, so the NEXT collective call fails loudly instead of queueing onto a broken communicator, and
    clear_error() re-arms it."""
    n, count = 4, 4096
    c = ft.Communicator.emulated(n, 0, window_bytes=1 << 20, timeout_s=0.05)
    try:
        ts = [torch.ones(count, device="cuda") for _ in range(n)]
        for algo in ("stages", "oneshot"):
            c.set_algo(algo)
            c.allreduce_multi(ts)
            torch.cuda.synchronize()
            assert c.poll_error() == 0 and c.error() == 0
            c.inject_fault(2)
            c.allreduce_multi(ts)            # rank 2 plays dead: launch succeeds, peers hit the watchdog
            torch.cuda.synchronize()
            e = c.poll_error()               # host-mapped mirror: no device round trip
            assert e != 0 and (e & 0xff) == 1 and ((e >> 16) & 0xff) == 2, hex(e)
            assert c.error() == e or c.error() != 0
            c.inject_fault(None)
            with pytest.raises(ft.FlexTreeError, match="watchdog"):
                c.allreduce_multi(ts)        # refused: an earlier call failed
            with pytest.raises(ft.FlexTreeError, match="watchdog"):
                c.raise_if_failed()
            c.clear_error()
            ts = [torch.full((count,), float(r + 1), device="cuda") for r in range(n)]
            c.allreduce_multi(ts)            # healthy again (epochs stayed aligned)
            torch.cuda.synchronize()
            assert c.error() == 0 and torch.all(ts[0] == 10.0) and torch.all(ts[3] == 10.0)
    finally:
        c.close()


@pytest.mark.parametrize("algo,registered,tma", [("oneshot", False, None), ("stages", True, True), ("stages", True, False),
                                                 ("stages", False, True), ("ring", True, None), ("twoshot", False, None)])
def test_cuda_graph_replay_every_path(comms, algo, registered, tma):
    """Capture + 20 replays for the paths with their own replay hazards: the one-shot / two-shot LL lines alternate
    parity halves by epoch, registered mode ends in an exit barrier, the TMA ring's mbarrier phases restart in
    every launch."""
    n = 4
    count = 2048 if algo in ("oneshot", "twoshot") else 60000
    c = comms(n)
    if algo == "twoshot" and "twoshot" not in ft.ops.constants.ALGOS:
        pytest.skip("two-shot kernel not built")
    c.set_algo(algo)
    c.set_widths([2, 2] if algo == "stages" else None)
    if tma is not None:
        c.set_tma(tma)
    c.reserve(count * 4)
    ts = c.symm_empty(count, torch.float32) if registered else [torch.empty(count, device="cuda") for _ in range(n)]
    for r in range(n):
        ts[r].fill_(1.0)
    c.allreduce_multi(ts)  # warm-up outside capture: resolves the plan
    torch.cuda.synchronize()
    assert c.last_call()["mode"] == ("registered" if registered else "staging")
    g = torch.cuda.CUDAGraph()
    s = torch.cuda.Stream()
    with torch.cuda.stream(s):
        g.capture_begin()
        c.allreduce_multi(ts, stream=s)
        g.capture_end()
    for it in range(20):
        for r in range(n):
            ts[r].fill_(float(r + 1 + it))
        g.replay()
        torch.cuda.synchronize()
        want = float(sum(r + 1 + it for r in range(n)))
        for r in range(n):
            assert torch.all(ts[r] == want), (algo, it, r)
    assert c.error() == 0
    c.set_algo("auto")
    c.set_widths(None)
    c.set_tma(None)


def test_allreduce_host_pipelined(comms):
    """Host-resident message (the reference's calling convention): chunked H2D -> in-place allreduce -> D2H over three
    streams; result equals the fp64 sum of the host inputs, in place and out of place, odd sizes, several chunk sizes."""
    n = 4
    c = comms(n)
    c.set_algo("auto")
    g = torch.Generator().manual_seed(5)
    for count, chunk in ((1 << 20, 1 << 20), (300001, 256 << 10), (5000, 0), ((1 << 20) + 4, 1 << 20)):
        xs = [torch.randn(count, generator=g).pin_memory() for _ in range(n)]
        want = torch.stack(xs).double().sum(0)
        outs = [torch.empty(count).pin_memory() for _ in range(n)]
        c.allreduce_host(xs, outs=outs, chunk_bytes=chunk)
        for r in range(n):
            assert torch.allclose(outs[r].double(), want, rtol=1e-5, atol=1e-5), (count, r)
            assert torch.equal(outs[r], outs[0])
        c.allreduce_host(xs, scale=0.25, chunk_bytes=chunk)          # in place, fused scale
        assert torch.allclose(xs[2].double(), want * 0.25, rtol=1e-5, atol=1e-5)
    assert c.error() == 0


@pytest.mark.parametrize("variant", [0, 1, 2, 3, 4, 5, 6])
@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_nvls_protocol_emulated_every_variant(comms, variant, dtype):
    """The NVLS kernels (flags, tiles, every inner-loop variant incl. the warp-specialised one) with the switch played
    by a loop over the N windows: registered buffers -> dedicated kernel (kernels/nvls_kernel.cu), private buffers ->
    the role-split general path. Several tiles per CTA, ragged sizes, back-to-back calls."""
    n = 4
    c = comms(n)
    c.set_algo("nvls")
    c.set_tuning("nvls_variant", variant)
    tol = dict(rtol=1e-5, atol=1e-5) if dtype == torch.float32 else dict(rtol=2e-2, atol=2e-2)
    g = torch.Generator(device="cuda").manual_seed(11 + variant)
    try:
        for count, ctas, tile in ((1 << 18, 2, 16 << 10), (300000, 3, 8 << 10), (4096, 0, 0), (1 << 20, 4, 0)):
            count -= count % 8
            c.set_launch(ctas, 0, tile)
            sym = c.symm_empty(count, dtype)
            xs = [torch.randn(count, device="cuda", generator=g).to(dtype) for _ in range(n)]
            want = torch.stack([x.double() for x in xs]).sum(0)
            for rep in range(3):
                for s_, x in zip(sym, xs):
                    s_.copy_(x)
                c.allreduce_multi(sym)
                torch.cuda.synchronize()
                assert c.error() == 0, hex(c.error())
                info = c.last_call()
                assert info["algo"] == "nvls" and info["mode"] == "registered", info
                for r in range(n):
                    assert torch.allclose(sym[r].double(), want, **tol), (variant, count, rep, r)
                    assert torch.equal(sym[r], sym[0])
            priv = [x.clone() for x in xs]       # private buffers: copy-in / reduce / copy-out roles
            outs = [torch.empty_like(x) for x in xs]
            c.allreduce_multi(priv, outs=outs, scale=0.5)
            torch.cuda.synchronize()
            assert c.error() == 0 and c.last_call()["mode"] == "staging"
            for r in range(n):
                assert torch.allclose(outs[r].double(), want * 0.5, **tol)
                assert torch.equal(priv[r], xs[r])
    finally:
        c.set_launch(0, 0, 0)
        c.set_tuning("nvls_variant", -1)
        c.set_algo("auto")


def test_nvls_registered_graph_replay_emulated(comms):
    n, count = 4, 1 << 16
    c = comms(n)
    c.set_algo("nvls")
    c.set_launch(2, 0, 8 << 10)
    ts = c.symm_empty(count, torch.float32)
    for t in ts:
        t.fill_(1.0)
    c.allreduce_multi(ts)
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    s = torch.cuda.Stream()
    with torch.cuda.stream(s):
        g.capture_begin()
        c.allreduce_multi(ts, stream=s)
        g.capture_end()
    for it in range(20):
        for r in range(n):
            ts[r].fill_(float(r + 1 + it))
        g.replay()
        torch.cuda.synchronize()
        want = float(sum(r + 1 + it for r in range(n)))
        for r in range(n):
            assert torch.all(ts[r] == want), (it, r)
    assert c.error() == 0
    c.set_launch(0, 0, 0)
    c.set_algo("auto")


@pytest.mark.parametrize("n,widths", [(2, [2]), (4, [4]), (4, [2, 2]), (8, [8]), (8, [2, 4]), (6, [3, 2])])
def test_cluster_pair_per_tile_matches_reference(comms, n, widths):
    """Thread-block clusters: a pair of CTAs owns a tile, splits the stage's sources and combines the partial sums
    through distributed shared memory (kernels/cluster_tiles.cuh). Same values as the one-CTA kernels up to the fp32
    rounding of a different (but fixed) summation order; integers-in-float inputs must match exactly."""
    c = comms(n)
    c.set_algo("stages")
    c.set_widths(widths)
    c.set_tma(True)
    c.set_tuning("cluster", 1)
    try:
        g = torch.Generator(device="cuda").manual_seed(3)
        for count in (4096, 300000, (1 << 20) + 8):
            xs = [torch.randint(-50, 50, (count,), device="cuda", generator=g).float() for _ in range(n)]
            want = torch.stack(xs).sum(0)
            outs = [torch.empty_like(x) for x in xs]
            c.allreduce_multi(xs, outs=outs)             # private buffers, out of place
            torch.cuda.synchronize()
            assert c.error() == 0, hex(c.error())
            for r in range(n):
                assert torch.equal(outs[r], want), (widths, count, r)
            sym = c.symm_empty(count, torch.float32)     # registered, in place, random data, fused scale
            ys = [torch.randn(count, device="cuda", generator=g) for _ in range(n)]
            for s_, y in zip(sym, ys):
                s_.copy_(y)
            c.allreduce_multi(sym, scale=0.5)
            torch.cuda.synchronize()
            assert c.error() == 0 and c.last_call()["mode"] == "registered"
            ref = torch.stack(ys).double().sum(0) * 0.5
            for r in range(n):
                assert torch.allclose(sym[r].double(), ref, rtol=1e-5, atol=1e-5)
                assert torch.equal(sym[r], sym[0])
        zs = [torch.randn(65536, device="cuda", generator=g).bfloat16() for _ in range(n)]   # bf16 wire
        ref = torch.stack([z.double() for z in zs]).sum(0)
        c.allreduce_multi(zs)
        torch.cuda.synchronize()
        assert c.error() == 0
        assert torch.allclose(zs[0].double(), ref, rtol=5e-2, atol=1e-1) and torch.equal(zs[0], zs[n - 1])  # bf16 is rounded once per stage
    finally:
        c.set_tuning("cluster", 0)
        c.set_tma(None)
        c.set_widths(None)
        c.set_algo("auto")


def test_ft_topo_environment_forms(monkeypatch):
    """FT_TOPO as the reference reads it (mpi_mod.hpp:882-929) plus the tree-size suffixes: widths, ring, a lonely rank
    ('2,3+1' on 7 ranks), errors for a wrong product and for a phantom tree on a communicator that cannot host one."""
    n, count = 7, 20000
    c = ft.Communicator.emulated(n, 0, window_bytes=1 << 20)
    try:
        xs = [torch.full((count,), float(r + 1), device="cuda") for r in range(n)]
        want = float(sum(range(1, n + 1)))
        for topo, algo, widths in (("7", "stages", [7]), ("2,3+1", "stages", [2, 3]), ("1", "ring", [7]), ("", None, None)):
            monkeypatch.setenv("FT_TOPO", topo)
            c.set_algo("auto")              # bumps the plan generation: FT_TOPO is read when a plan is built
            ys = [x.clone() for x in xs]
            c.allreduce_multi(ys)
            torch.cuda.synchronize()
            assert c.error() == 0
            for y in ys:
                assert torch.all(y == want), topo
            if algo:
                info = c.last_call()
                assert info["algo"] == algo and info["widths"] == widths, (topo, info)
        for bad in ("2,3", "2,4-1"):        # product != 7; '-k' trees need one process per GPU
            monkeypatch.setenv("FT_TOPO", bad)
            c.set_algo("auto")
            with pytest.raises(ft.FlexTreeError):
                c.allreduce_multi([x.clone() for x in xs])
    finally:
        monkeypatch.delenv("FT_TOPO", raising=False)
        c.close()
