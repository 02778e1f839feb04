"""The algorithm templates of csrc/core/staged_core.h, executed on CPU threads (one per "CTA"),
checked against an independent numpy model. These are the same templates the CUDA kernels instantiate."""
import numpy as np
import pytest

import flextree as ft
from flextree import HostWorld
from flextree.ops import constants as K

from refimpl import block_size, from_bf16_bits, round_bf16, staged_allreduce, to_bf16_bits

RNG = np.random.default_rng(1234)
NP_OPS = {"sum": np.add, "max": np.maximum, "min": np.minimum, "band": np.bitwise_and, "bor": np.bitwise_or}


def make_inputs(n, count, dtype):
    dt = np.dtype(dtype)
    if dt.kind == "f":
        return [RNG.standard_normal(count).astype(dt) for _ in range(n)]
    if dt.kind == "b":
        return [RNG.integers(0, 2, count).astype(dt) for _ in range(n)]
    info = np.iinfo(dt)
    return [RNG.integers(info.min, info.max, count, dtype=dt, endpoint=True) for _ in range(n)]


def aligned_empty(count, dtype, misalign=0):
    raw = np.empty(count * np.dtype(dtype).itemsize + 64, dtype=np.uint8)
    off = (-raw.ctypes.data) % 16 + misalign
    return raw[off:off + count * np.dtype(dtype).itemsize].view(dtype)


def aligned_copy(a, misalign=0):
    out = aligned_empty(a.size, a.dtype, misalign)
    out[:] = a
    return out


WIDTHS = {2: [[2]], 4: [[4], [2, 2]], 6: [[6], [2, 3], [3, 2]], 8: [[8], [2, 4], [4, 2], [2, 2, 2]],
          12: [[2, 3, 2], [3, 4]], 20: [[20], [4, 5]]}


@pytest.mark.parametrize("n", [2, 4, 6, 8, 12])
@pytest.mark.parametrize("count", [1, 3, 5, 7, 35, 1000, 1001, 4099])
def test_fp32_sum_bit_exact_all_factorizations(n, count):
    """fp32 SUM must be bit-identical to the order-aware model (own first, peers ascending), on every
    rank, for every width vector, in place and out of place, incl. count < N."""
    inputs = make_inputs(n, count, np.float32)
    B = block_size(count, n, 4)
    with_world = HostWorld(n, window_bytes=1 << 16)
    for widths in WIDTHS[n]:
        want = staged_allreduce(inputs, widths, block=B)
        send = [aligned_copy(x) for x in inputs]
        recv = [aligned_empty(count, np.float32) for _ in range(n)]
        for r in recv:
            r[:] = np.nan
        assert with_world.allreduce(send, recv, widths=widths, ctas=3, tile_bytes=256) == 0
        for r in range(n):
            np.testing.assert_array_equal(recv[r], want[r], err_msg=f"{widths} rank {r}")
            np.testing.assert_array_equal(send[r], inputs[r])  # sendbuf untouched
        buf = [aligned_copy(x) for x in inputs]  # in place
        assert with_world.allreduce(buf, None, widths=widths, ctas=2, tile_bytes=512) == 0
        for r in range(n):
            np.testing.assert_array_equal(buf[r], want[r])
    with_world.close()


@pytest.mark.parametrize("dtype", [np.uint8, np.int8, np.uint16, np.int16, np.int32, np.uint32, np.int64,
                                   np.uint64, np.float64, np.bool_])
@pytest.mark.parametrize("op", ["sum", "max", "min", "band", "bor"])
def test_dtype_op_matrix(dtype, op):
    n, count, widths = 6, 1001, [2, 3]
    dt = np.dtype(dtype)
    if dt.kind == "f" and op in ("band", "bor"):
        w = HostWorld(n)
        x = make_inputs(n, 8, dtype)
        assert w.allreduce([aligned_copy(a) for a in x], None, op=op, widths=widths) == -7  # unsupported
        return
    inputs = make_inputs(n, count, dtype)
    w = HostWorld(n, window_bytes=1 << 16)
    buf = [aligned_copy(x) for x in inputs]
    assert w.allreduce(buf, None, op=op, widths=widths, ctas=2, tile_bytes=128) == 0
    if dt.kind == "b":
        stack = np.stack(inputs)
        want = stack.any(0) if op in ("sum", "max", "bor") else stack.all(0)
    elif dt.kind == "f":
        want = staged_allreduce(inputs, widths, op=NP_OPS[op], block=block_size(count, n, dt.itemsize))[0]
    else:
        with np.errstate(over="ignore"):
            want = inputs[0].copy()
            for x in inputs[1:]:
                want = NP_OPS[op](want, x)
    for r in range(n):
        np.testing.assert_array_equal(buf[r], want, err_msg=f"rank {r}")


@pytest.mark.parametrize("misalign", [4, 8, 2])
def test_unaligned_buffers_take_scalar_path(misalign):
    n, count = 4, 777
    dtype = np.float32 if misalign % 4 == 0 else np.int16
    inputs = make_inputs(n, count, dtype)
    w = HostWorld(n)
    send = [aligned_copy(x, misalign) for x in inputs]
    recv = [aligned_empty(count, dtype, misalign) for _ in range(n)]
    assert w.allreduce(send, recv, widths=[2, 2], ctas=2, tile_bytes=64) == 0
    if np.dtype(dtype).kind == "f":
        want = staged_allreduce(inputs, [2, 2], block=block_size(count, n, 4, vectorized=False))[0]
    else:
        with np.errstate(over="ignore"):
            want = inputs[0] + inputs[1] + inputs[2] + inputs[3]
    for r in range(n):
        np.testing.assert_array_equal(recv[r], want)


def test_half_and_bf16_wire():
    n, count, widths = 8, 2049, [2, 4]
    # fp16 through numpy
    x16 = [RNG.standard_normal(count).astype(np.float16) for _ in range(n)]
    w = HostWorld(n, window_bytes=1 << 16)
    buf = [aligned_copy(x) for x in x16]
    assert w.allreduce(buf, None, widths=widths, ctas=2, tile_bytes=256) == 0
    want = staged_allreduce(x16, widths, block=block_size(count, n, 2),
                            round_fn=lambda a: a.astype(np.float16).astype(np.float32))[0].astype(np.float16)
    for r in range(n):
        np.testing.assert_array_equal(buf[r], want)
    # bf16 carried as uint16 bit patterns
    xf = [RNG.standard_normal(count).astype(np.float32) for _ in range(n)]
    xb = [to_bf16_bits(x) for x in xf]
    buf = [aligned_copy(x) for x in xb]
    assert w.allreduce(buf, None, widths=widths, ctas=2, tile_bytes=256, dtype_code=K.BFLOAT16) == 0
    want = staged_allreduce([from_bf16_bits(b) for b in xb], widths, block=block_size(count, n, 2),
                            round_fn=round_bf16)[0]
    for r in range(n):
        np.testing.assert_array_equal(from_bf16_bits(buf[r]), want)


def test_fp32_io_bf16_wire_with_scale():
    """DDP shape: fp32 gradients, bf16 on the wire, fused 1/N scale + cast back to fp32."""
    n, count, widths = 4, 1234, [4]
    xf = [RNG.standard_normal(count).astype(np.float32) for _ in range(n)]
    w = HostWorld(n, window_bytes=1 << 16)
    buf = [aligned_copy(x) for x in xf]
    assert w.allreduce(buf, None, widths=widths, ctas=2, tile_bytes=256, dtype_code=K.BFLOAT16,
                       io_dtype_code=K.FLOAT, scale=1.0 / n) == 0
    want = staged_allreduce([round_bf16(x) for x in xf], widths, block=block_size(count, n, 2),
                            round_fn=round_bf16, final_scale=1.0 / n)[0]
    for r in range(n):
        np.testing.assert_array_equal(buf[r], want)
    exact = np.mean(np.stack(xf).astype(np.float64), 0)
    assert np.max(np.abs(buf[0] - exact)) < 0.03


@pytest.mark.parametrize("n,widths", [(2, [2]), (8, [8]), (8, [2, 4]), (8, [2, 2, 2])])
def test_fp8_block_scaled_wire(n, widths):
    """Opt-in fp8 wire: e4m3 values + a power-of-two scale per 16 elements, re-quantised at every stage.
    fp32 user buffers in and out, fused 1/N. The error must stay within the e4m3 resolution (2^-4 relative
    to the largest magnitude of a 16-block) times a small factor per stage."""
    count = 4096 + 32
    xf = [RNG.standard_normal(count).astype(np.float32) * np.float32(10.0 ** RNG.integers(-3, 3)) for _ in range(n)]
    w = HostWorld(n, window_bytes=1 << 16)
    buf = [aligned_copy(x) for x in xf]
    assert w.allreduce(buf, None, widths=widths, ctas=2, tile_bytes=256, dtype_code=K.FP8_BLOCK,
                       io_dtype_code=K.FLOAT, scale=1.0 / n) == 0
    exact = np.mean(np.stack(xf).astype(np.float64), 0)
    for r in range(1, n):
        np.testing.assert_array_equal(buf[r], buf[0])  # every rank decodes the same bytes
    blocks = exact[:count // 16 * 16].reshape(-1, 16)
    # scale of the comparison: the largest partial-sum magnitude a block can see, bounded by sum |x_i| / n
    bound = np.stack([np.abs(x.astype(np.float64)) for x in xf]).sum(0)[:count // 16 * 16].reshape(-1, 16).max(1) / n
    err = np.abs(buf[0][:count // 16 * 16].astype(np.float64).reshape(-1, 16) - blocks).max(1)
    stages = len(widths)
    assert np.all(err <= bound * 0.0625 * (1 + stages)), float((err / bound).max())
    rel_rms = np.sqrt(np.mean((buf[0] - exact) ** 2)) / np.sqrt(np.mean(exact ** 2))
    assert rel_rms < 0.06 * stages + 0.03, rel_rms
    # NVLS / one-shot / registered have no block-scaled variant
    assert w.allreduce(buf, None, algo="oneshot", dtype_code=K.FP8_BLOCK, io_dtype_code=K.FLOAT) == -9


def test_fp8_conversions():
    import ctypes as C
    L = ft._lib.load()
    # every e4m3 code round-trips; monotone; saturation and ties
    vals = []
    for code in range(256):
        f = L.ft_e4m3_to_float(code)
        if code & 0x7f == 0x7f:
            assert np.isnan(f)
            continue
        vals.append((code, f))
        assert L.ft_float_to_e4m3(C.c_float(f)) == (code if code != 0x80 else 0x80), hex(code)
    pos = sorted(f for c, f in vals if c < 0x80)
    assert pos[0] == 0.0 and pos[1] == 2.0 ** -9 and pos[-1] == 448.0
    assert L.ft_float_to_e4m3(C.c_float(1e9)) == 0x7e and L.ft_float_to_e4m3(C.c_float(-1e9)) == 0xfe
    assert L.ft_float_to_e4m3(C.c_float(2.0 ** -10)) == 0          # tie -> even (0)
    assert L.ft_float_to_e4m3(C.c_float(2.0 ** -10 * 1.01)) == 1   # above the tie -> smallest subnormal
    assert L.ft_float_to_e4m3(C.c_float(1.0625)) == L.ft_float_to_e4m3(C.c_float(1.0))  # tie -> even mantissa
    for amax in (0.0, 1e-30, 1.0, 447.0, 448.0, 449.0, 3e38):
        s = L.ft_fp8_block_scale(C.c_float(amax))
        if amax > 0:
            scaled = amax * 2.0 ** (127 - s)
            assert scaled <= 448.0 * (1 + 1e-6) and (scaled > 224.0 or s in (0, 254)), (amax, s, scaled)


def test_scale_epilogue_fp32():
    n, count = 6, 333
    xf = make_inputs(n, count, np.float32)
    w = HostWorld(n)
    buf = [aligned_copy(x) for x in xf]
    assert w.allreduce(buf, None, widths=[3, 2], scale=0.25) == 0
    want = staged_allreduce(xf, [3, 2], block=block_size(count, n, 4), final_scale=0.25)[0]
    np.testing.assert_array_equal(buf[3], want)


@pytest.mark.parametrize("widths", [[8], [2, 4], [2, 2, 2]])
def test_registered_in_place_zero_copy(widths):
    n, count = 8, 5000
    xf = make_inputs(n, count, np.float32)
    w = HostWorld(n, window_bytes=1 << 16)
    views = [w.registered(r, count, np.float32) for r in range(n)]
    for it in range(3):  # back-to-back calls exercise the exit barrier / epoch logic
        for r in range(n):
            views[r][:] = xf[r]
        assert w.allreduce(views, None, widths=widths, ctas=4, tile_bytes=512, mode=K.MODE_REGISTERED) == 0
        want = staged_allreduce(xf, widths, block=block_size(count, n, 4))[0]
        for r in range(n):
            np.testing.assert_array_equal(views[r], want)
    assert w.epoch(0) == 3 and w.epoch(n - 1) == 3


@pytest.mark.parametrize("algo,widths", [("stages", [2, 2]), ("stages", [4]), ("nvls", None), ("ring", None)])
def test_registered_in_place_with_fused_scale(algo, widths):
    """The scale is applied when a block becomes final, so zero-copy (registered) buffers can use it too."""
    n, count = 4, 3000
    xf = make_inputs(n, count, np.float32)
    w = HostWorld(n, window_bytes=1 << 16)
    views = [w.registered(r, count, np.float32) for r in range(n)]
    for r in range(n):
        views[r][:] = xf[r]
    assert w.allreduce(views, None, algo=algo, widths=widths, ctas=3, tile_bytes=256, mode=K.MODE_REGISTERED,
                       scale=0.25) == 0
    exact = np.sum(np.stack(xf).astype(np.float64), 0) * 0.25
    for r in range(n):
        np.testing.assert_allclose(views[r], exact, rtol=1e-5, atol=1e-6)
        np.testing.assert_array_equal(views[r], views[0])


def test_back_to_back_sizes_and_plans_reuse_flags():
    """Different sizes, tilings, CTA counts and plans in sequence: epochs keep flags monotone and the
    parity windows never collide."""
    n = 6
    w = HostWorld(n, window_bytes=1 << 18)
    for i, (count, widths, ctas, tile) in enumerate([(10, [6], 1, 64), (40000, [2, 3], 5, 1024), (3, [3, 2], 2, 64),
                                                     (777, [6], 4, 128), (40000, [3, 2], 3, 4096), (1, [2, 3], 1, 64)]):
        xf = make_inputs(n, count, np.float32)
        buf = [aligned_copy(x) for x in xf]
        assert w.allreduce(buf, None, widths=widths, ctas=ctas, tile_bytes=tile) == 0
        want = staged_allreduce(xf, widths, block=block_size(count, n, 4))[0]
        for r in range(n):
            np.testing.assert_array_equal(buf[r], want, err_msg=f"call {i} rank {r}")
    assert w.epoch(0) == 6


def test_width_20_single_stage():
    """The reference caps a stage at 20-way reduce (mpi_mod.hpp:245); make sure 20 works (batched peers)."""
    n, count = 20, 2000
    xf = make_inputs(n, count, np.float32)
    w = HostWorld(n, window_bytes=1 << 16)
    for widths in ([20], [4, 5]):
        buf = [aligned_copy(x) for x in xf]
        assert w.allreduce(buf, None, widths=widths, ctas=1, tile_bytes=512) == 0
        want = staged_allreduce(xf, widths, block=block_size(count, n, 4))[0]
        for r in range(n):
            np.testing.assert_array_equal(buf[r], want)


@pytest.mark.parametrize("n", [2, 3, 5, 8])
@pytest.mark.parametrize("count", [1, 6, 1000, 4097])
def test_ring(n, count):
    """Ring mode (FT_TOPO containing 1; reference ring_allreduce mpi_mod.hpp:1113-1163)."""
    xf = make_inputs(n, count, np.float32)
    w = HostWorld(n, window_bytes=1 << 16)
    send = [aligned_copy(x) for x in xf]
    recv = [aligned_empty(count, np.float32) for _ in range(n)]
    assert w.allreduce(send, recv, algo="ring", ctas=2, tile_bytes=256) == 0
    exact = np.sum(np.stack(xf).astype(np.float64), 0)
    for r in range(n):
        np.testing.assert_allclose(recv[r], exact, rtol=1e-5, atol=1e-5)
        np.testing.assert_array_equal(recv[r], recv[0])  # all ranks bit-identical
    ints = make_inputs(n, count, np.int32)
    buf = [aligned_copy(x) for x in ints]
    assert w.allreduce(buf, None, algo="ring", ctas=3, tile_bytes=128) == 0
    with np.errstate(over="ignore"):
        want = np.sum(np.stack(ints), 0, dtype=np.int32)
    for r in range(n):
        np.testing.assert_array_equal(buf[r], want)


@pytest.mark.parametrize("n", [2, 4, 8])
@pytest.mark.parametrize("dtype", [np.float32, np.float16, np.int32, np.uint8, np.float64])
def test_oneshot_ll(n, dtype):
    w = HostWorld(n, ll_bytes=1 << 18)
    for count in (1, 5, 1023):
        x = make_inputs(n, count, dtype)
        send = [aligned_copy(a) for a in x]
        recv = [aligned_empty(count, dtype) for _ in range(n)]
        assert w.allreduce(send, recv, algo="oneshot", ctas=2) == 0
        if np.dtype(dtype).kind == "f":
            acc = x[0].astype(np.float64 if dtype == np.float64 else np.float32)
            for a in x[1:]:
                acc = acc + a.astype(acc.dtype)
            want = acc.astype(dtype)
        else:
            with np.errstate(over="ignore"):
                want = np.sum(np.stack(x), 0, dtype=dtype)
        for r in range(n):
            np.testing.assert_array_equal(recv[r], want, err_msg=f"count {count} rank {r}")
    assert w.allreduce([aligned_copy(a) for a in make_inputs(n, 1 << 16, np.float32)], None, algo="oneshot") == -5


@pytest.mark.parametrize("n", [2, 3, 4, 8])
@pytest.mark.parametrize("dtype", [np.float32, np.float16, np.int32, np.uint8, np.float64])
def test_twoshot_ll_equals_oneshot(n, dtype):
    """Two-shot LL (block p of every message to rank p, reduce, broadcast) against the one-shot: same ascending
    summation order, so the same bits - for ragged counts, counts smaller than N (empty blocks), in place, and
    several CTAs."""
    LL = 1 << 20
    w = HostWorld(n, window_bytes=8 << 20, ll_bytes=LL)
    for count in (1, 5, 1023, 4096):
        x = make_inputs(n, count, dtype)
        one = [aligned_empty(count, dtype) for _ in range(n)]
        assert w.allreduce([aligned_copy(a) for a in x], one, algo="oneshot", ctas=2) == 0
        two = [aligned_copy(a) for a in x]  # in place
        assert w.allreduce(two, None, algo="twoshot", ctas=3) == 0
        for r in range(n):
            np.testing.assert_array_equal(two[r], one[r], err_msg=f"count {count} rank {r}")
    # the two-shot holds ~N/2 times the one-shot's message in the same LL buffer
    big = LL // 16 // n * 8 // 4 + 64  # fp32 elements: a little more than one one-shot region
    xs = make_inputs(n, big, np.float32)
    assert w.allreduce([aligned_copy(a) for a in xs], None, algo="oneshot") == -5
    if n > 2:
        bufs = [aligned_copy(a) for a in xs]
        assert w.allreduce(bufs, None, algo="twoshot", ctas=2) == 0
        acc = xs[0].copy()
        for a in xs[1:]:
            acc = acc + a
        np.testing.assert_array_equal(bufs[0], acc)
    assert w.allreduce([aligned_copy(a) for a in make_inputs(n, LL, np.float32)], None, algo="twoshot") == -5


def test_twoshot_ll_fused_scale_and_bf16_wire():
    n, count = 4, 2000
    xf = make_inputs(n, count, np.float32)
    w = HostWorld(n, ll_bytes=1 << 18)
    bufs = [aligned_copy(x) for x in xf]
    assert w.allreduce(bufs, None, algo="twoshot", ctas=2, scale=0.25) == 0
    acc = xf[0].copy()
    for a in xf[1:]:
        acc = acc + a
    for b in bufs:
        np.testing.assert_array_equal(b, acc * np.float32(0.25))
    # fp32 data on a bf16 wire: inputs and the reduced block are rounded to bf16, every rank gets the same bits
    bufs = [aligned_copy(x) for x in xf]
    assert w.allreduce(bufs, None, algo="twoshot", ctas=2, dtype_code=K.BFLOAT16, io_dtype_code=K.FLOAT, scale=0.25) == 0
    acc = round_bf16(xf[0]).astype(np.float32)
    for a in xf[1:]:
        acc = acc + round_bf16(a)
    want = round_bf16(acc * np.float32(0.25))
    for b in bufs:
        np.testing.assert_array_equal(b, want)


@pytest.mark.parametrize("mode", [K.MODE_STAGING, K.MODE_REGISTERED])
def test_nvls_protocol_with_emulated_switch(mode):
    n, count = 8, 3001
    xf = make_inputs(n, count, np.float32)
    w = HostWorld(n, window_bytes=1 << 16)
    if mode == K.MODE_REGISTERED:
        bufs = [w.registered(r, count, np.float32) for r in range(n)]
        for r in range(n):
            bufs[r][:] = xf[r]
    else:
        bufs = [aligned_copy(x) for x in xf]
    for _ in range(2):
        assert w.allreduce(bufs, None, algo="nvls", ctas=3, tile_bytes=256, mode=mode) == 0
        acc = xf[0].copy()
        for a in xf[1:]:
            acc = acc + a
        for r in range(n):
            np.testing.assert_array_equal(bufs[r], acc)
        xf = [acc.copy() for _ in range(n)]


@pytest.mark.parametrize("n,widths,lonely", [(7, [2, 3], 1), (5, [2, 2], 1), (8, [2, 3], 2), (3, [2], 1), (9, [2, 2, 2], 1)])
def test_lonely_ranks(n, widths, lonely):
    """World size = tree + lonely ranks (the reference's "+k" trees, stubbed out there: mpi_mod.hpp:981-1099)."""
    w = HostWorld(n, window_bytes=1 << 16)
    for count in (1, 5, 1000, 4099):
        ints = make_inputs(n, count, np.int32)
        send = [aligned_copy(x) for x in ints]
        recv = [aligned_empty(count, np.int32) for _ in range(n)]
        assert w.allreduce(send, recv, widths=widths, lonely=lonely, ctas=2, tile_bytes=256) == 0
        with np.errstate(over="ignore"):
            want = np.sum(np.stack(ints), 0, dtype=np.int32)
        for r in range(n):
            np.testing.assert_array_equal(recv[r], want, err_msg=f"count {count} rank {r}")
            np.testing.assert_array_equal(send[r], ints[r])
        xf = make_inputs(n, count, np.float32)
        buf = [aligned_copy(x) for x in xf]
        assert w.allreduce(buf, None, widths=widths, lonely=lonely, ctas=3, tile_bytes=128) == 0
        # tree part in the staged order, then the lonely ranks in ascending order
        T = n - lonely
        tree = staged_allreduce(xf[:T], widths, block=block_size(count, T, 4))[0]
        for x in xf[T:]:
            tree = tree + x
        for r in range(n):
            np.testing.assert_array_equal(buf[r], tree, err_msg=f"fp32 count {count} rank {r}")
    assert w.allreduce([aligned_copy(x) for x in make_inputs(n, 8, np.float32)], None, widths=widths, lonely=lonely,
                       mode=K.MODE_REGISTERED) == -8   # lonely trees use private windows only


@pytest.mark.parametrize("n,widths", [(4, [4]), (6, [2, 3]), (8, [2, 2, 2])])
@pytest.mark.parametrize("per_rank", [4, 37, 1024])
def test_reduce_scatter_and_all_gather_halves(n, widths, per_rank):
    """The two halves of the algorithm as collectives of their own."""
    w = HostWorld(n, window_bytes=1 << 16)
    total = per_rank * n
    xs = make_inputs(n, total, np.int32)
    send = [aligned_copy(x) for x in xs]
    recv = [aligned_empty(per_rank, np.int32) for _ in range(n)]
    assert w.allreduce(send, recv, widths=widths, ctas=2, tile_bytes=128, phases=1, count=total) == 0
    with np.errstate(over="ignore"):
        full = np.sum(np.stack(xs), 0, dtype=np.int32)
    for r in range(n):
        np.testing.assert_array_equal(recv[r], full[r * per_rank:(r + 1) * per_rank], err_msg=f"rs rank {r}")
        np.testing.assert_array_equal(send[r], xs[r])
    blocks = [aligned_copy(x[:per_rank]) for x in make_inputs(n, per_rank, np.float32)]
    outs = [aligned_empty(total, np.float32) for _ in range(n)]
    assert w.allreduce(blocks, outs, widths=widths, ctas=3, tile_bytes=64, phases=2, count=total) == 0
    want = np.concatenate(blocks)
    for r in range(n):
        np.testing.assert_array_equal(outs[r], want, err_msg=f"ag rank {r}")
    # RS followed by AG == allreduce
    xf = make_inputs(n, total, np.float32)
    part = [aligned_empty(per_rank, np.float32) for _ in range(n)]
    assert w.allreduce([aligned_copy(x) for x in xf], part, widths=widths, phases=1, count=total) == 0
    allr = [aligned_empty(total, np.float32) for _ in range(n)]
    assert w.allreduce(part, allr, widths=widths, phases=2, count=total) == 0
    ref = [aligned_copy(x) for x in xf]
    assert w.allreduce(ref, None, widths=widths) == 0
    vec = (per_rank * 4) % 16 == 0
    if vec:  # same block split -> same summation order -> same bits
        np.testing.assert_array_equal(allr[0], ref[0])
    else:
        np.testing.assert_allclose(allr[0], ref[0], rtol=1e-5, atol=1e-5)


@pytest.mark.parametrize("n,ctas", [(2, 1), (4, 2), (4, 5), (8, 8), (8, 3)])
def test_nvls_private_buffers_role_split(n, ctas):
    """NVLS with private buffers: copy CTAs and reduce CTAs meet through per-tile flags (incl. self flags)."""
    w = HostWorld(n, window_bytes=1 << 20)
    for count, tile in ((40000, 1024), (1001, 64), (3, 64), (65536, 4096)):
        xf = make_inputs(n, count, np.float32)
        send = [aligned_copy(x) for x in xf]
        recv = [aligned_empty(count, np.float32) for _ in range(n)]
        for it in range(2):  # twice: per-tile flag slots are reused with the next epoch
            assert w.allreduce(send, recv, algo="nvls", ctas=ctas, tile_bytes=tile, scale=0.5) == 0
            acc = xf[0].copy()
            for a in xf[1:]:
                acc = acc + a
            acc = acc * np.float32(0.5)
            for r in range(n):
                np.testing.assert_array_equal(recv[r], acc, err_msg=f"count {count} rank {r}")
                np.testing.assert_array_equal(send[r], xf[r])


def test_watchdog_fires_instead_of_hanging():
    """Fault injection: one rank never joins the collective. The others must report a timeout."""
    n = 4
    w = HostWorld(n)
    xf = make_inputs(n, 100, np.float32)
    bufs = [aligned_copy(x) for x in xf]
    rc = w.allreduce(bufs, None, widths=[2, 2], ctas=1, timeout_s=0.2, skip_rank=2)
    assert rc > 0 and (rc & 0xFF) == 1  # ERR_TIMEOUT
    assert any(w.error(r) for r in range(n))


def test_conversions_match_numpy():
    import ctypes as C
    L = ft._lib.load()
    vals = np.concatenate([RNG.standard_normal(2000).astype(np.float32) * 10.0 ** RNG.integers(-9, 6, 2000),
                           np.array([0.0, -0.0, np.inf, -np.inf, 65504.0, 65520.0, 65519.9, 5.96e-8, 2.98e-8,
                                     2.9802322e-8, 6.1e-5, 6.0975552e-05, 1e-10, 70000.0], dtype=np.float32)])
    for v in vals:
        h = L.ft_float_to_f16(C.c_float(v))
        assert h == int(np.float32(v).astype(np.float16).view(np.uint16)), v
        assert L.ft_float_to_bf16(C.c_float(v)) == int(to_bf16_bits(np.array([v], np.float32))[0]), v
    for bits in list(range(0, 65536, 7)) + [0x7c00, 0xfc00, 0x0001, 0x03ff, 0x0400]:
        f = L.ft_f16_to_float(bits)
        want = np.array([bits], np.uint16).view(np.float16).astype(np.float32)[0]
        assert (np.isnan(f) and np.isnan(want)) or f == want, hex(bits)


# ---- bulk tiles (csrc/host/cpu_backend.h): the CPU counterpart of the TMA-staged tiles -------------------------
# Same hook as the TMA kernel (StagedCta::run<true>), selected with algo 5. They must give the bits of the
# per-vector text in every configuration they accept and fall back to it silently everywhere else.

def _both(n, make, **kw):
    """Run the same call with the per-vector text and with the bulk tiles; return the two result lists."""
    out = []
    for algo in ("stages", "stages_bulk"):
        w = HostWorld(n, window_bytes=1 << 18)
        send, recv = make()
        assert w.allreduce(send, recv, algo=algo, **kw) == 0
        out.append(recv if recv is not None else send)
        w.close()
    return out


@pytest.mark.parametrize("dtype", [np.uint8, np.int16, np.int32, np.int64, np.float32, np.float64, np.float16, np.bool_])
@pytest.mark.parametrize("op", ["sum", "max", "band"])
def test_bulk_tiles_dtype_op_matrix_bit_identical(dtype, op):
    dt = np.dtype(dtype)
    if dt.kind == "f" and op == "band":
        pytest.skip("no bitwise ops on floats")
    n, widths = 6, [2, 3]
    count = 6 * 1024 + 16 * 5  # a multiple of 16 elements: whole 16-byte vectors for every dtype
    inputs = make_inputs(n, count, dtype)
    a, b = _both(n, lambda: ([aligned_copy(x) for x in inputs], None), op=op, widths=widths, ctas=2, tile_bytes=512)
    for r in range(n):
        np.testing.assert_array_equal(a[r], b[r], err_msg=f"rank {r}")
    if dt.kind != "f":  # and against numpy (floats: bit-exact model is covered by the per-vector tests)
        if dt.kind == "b":
            want = np.stack(inputs).any(0) if op in ("sum", "max") else np.stack(inputs).all(0)
        else:
            with np.errstate(over="ignore"):
                want = inputs[0].copy()
                for x in inputs[1:]:
                    want = NP_OPS[op](want, x)
        np.testing.assert_array_equal(b[0], want)


@pytest.mark.parametrize("n,widths", [(2, [2]), (4, [2, 2]), (8, [8]), (8, [2, 2, 2]), (12, [3, 4]), (20, [20])])
@pytest.mark.parametrize("ctas,tile", [(1, 1 << 20), (3, 256), (2, 4096)])
def test_bulk_tiles_fp32_bit_exact_vs_model(n, widths, ctas, tile):
    count = 4 * 1237  # not a multiple of n * 4: short and empty tail blocks
    inputs = make_inputs(n, count, np.float32)
    want = staged_allreduce(inputs, widths, block=block_size(count, n, 4))
    w = HostWorld(n, window_bytes=1 << 18)
    send = [aligned_copy(x) for x in inputs]
    recv = [aligned_empty(count, np.float32) for _ in range(n)]
    assert w.allreduce(send, recv, algo="stages_bulk", widths=widths, ctas=ctas, tile_bytes=tile) == 0
    for r in range(n):
        np.testing.assert_array_equal(recv[r], want[r], err_msg=f"rank {r}")
        np.testing.assert_array_equal(send[r], inputs[r])  # sendbuf preserved
    # back to back, in place, other parity of the window
    assert w.allreduce(send, None, algo="stages_bulk", widths=widths, ctas=ctas, tile_bytes=tile) == 0
    for r in range(n):
        np.testing.assert_array_equal(send[r], want[r])


@pytest.mark.parametrize("widths", [[8], [2, 4], [2, 2, 2]])
def test_bulk_tiles_registered_zero_copy_and_scale(widths):
    n, count = 8, 4096
    inputs = make_inputs(n, count, np.float32)
    res = []
    for algo in ("stages", "stages_bulk"):
        w = HostWorld(n, window_bytes=1 << 18)
        bufs = [w.registered(r, count, np.float32) for r in range(n)]
        for b, x in zip(bufs, inputs):
            b[:] = x
        assert w.allreduce(bufs, None, algo=algo, widths=widths, ctas=2, tile_bytes=512, mode=K.MODE_REGISTERED,
                           scale=0.125) == 0
        res.append([b.copy() for b in bufs])
        w.close()
    for r in range(n):
        np.testing.assert_array_equal(res[0][r], res[1][r])
    np.testing.assert_allclose(res[1][0], np.mean(np.stack(inputs), 0), rtol=1e-5, atol=1e-6)


def test_bulk_tiles_fp32_io_bf16_wire_with_scale():
    n, count, widths = 4, 2048, [2, 2]
    xf = make_inputs(n, count, np.float32)
    kw = dict(widths=widths, ctas=2, tile_bytes=256, dtype_code=K.BFLOAT16, io_dtype_code=K.FLOAT, scale=1.0 / n)
    a, b = _both(n, lambda: ([aligned_copy(x) for x in xf], [aligned_empty(count, np.float32) for _ in range(n)]), **kw)
    for r in range(n):
        np.testing.assert_array_equal(a[r], b[r])
    want = staged_allreduce([round_bf16(x) for x in xf], widths, block=block_size(count, n, 2), round_fn=round_bf16,
                            final_scale=1.0 / n)[0]
    np.testing.assert_array_equal(b[0], want)


def test_bulk_tiles_fall_back_where_they_do_not_apply():
    n, widths = 4, [2, 2]
    # ragged message (not whole 16-byte vectors), unaligned buffers, lonely ranks: algo 5 silently runs the per-vector text
    for count, misalign in [(1001, 0), (1024, 4)]:
        inputs = make_inputs(n, count, np.float32)
        a, b = _both(n, lambda: ([aligned_copy(x, misalign) for x in inputs], None), widths=widths, ctas=2, tile_bytes=256)
        for r in range(n):
            np.testing.assert_array_equal(a[r], b[r])
    inputs = make_inputs(5, 1024, np.float32)
    a, b = _both(5, lambda: ([aligned_copy(x) for x in inputs], None), widths=[2, 2], ctas=2, tile_bytes=256, lonely=1)
    for r in range(5):
        np.testing.assert_array_equal(a[r], b[r])


@pytest.mark.parametrize("n,widths", [(4, [4]), (6, [2, 3])])
def test_bulk_tiles_reduce_scatter_and_all_gather_halves(n, widths):
    per_rank = 256
    total = per_rank * n
    xs = make_inputs(n, total, np.float32)
    w = HostWorld(n, window_bytes=1 << 18)
    part = [aligned_empty(per_rank, np.float32) for _ in range(n)]
    assert w.allreduce([aligned_copy(x) for x in xs], part, algo="stages_bulk", widths=widths, ctas=2, tile_bytes=128,
                       phases=1, count=total) == 0
    full = [aligned_empty(total, np.float32) for _ in range(n)]
    assert w.allreduce(part, full, algo="stages_bulk", widths=widths, ctas=2, tile_bytes=128, phases=2, count=total) == 0
    ref = [aligned_copy(x) for x in xs]
    assert w.allreduce(ref, None, widths=widths) == 0
    for r in range(n):
        np.testing.assert_array_equal(full[r], ref[r])


def test_randomized_delays_stress():
    """random sleeps before every flag publication, after every acquisition and at every CTA sync, over
    back-to-back calls that change size, plan, algorithm and buffer kind: a missing wait, a window reused too early
    or a leaky exit barrier shows up as a wrong sum. (Mutation-checked: removing any single wait of staged_core.h
    that an in-process world can observe makes this test fail.)"""
    n = 6
    w = HostWorld(n, window_bytes=1 << 16, ll_bytes=1 << 16)
    w.set_jitter(150)
    try:
        rng = np.random.default_rng(7)
        # (algorithm, widths, lonely ranks, registered buffers?, threads per rank)
        plans = [("stages", [2, 3], 0, False, 2), ("stages_bulk", [3, 2], 0, True, 1), ("stages", [6], 0, True, 3),
                 ("ring", None, 0, False, 2), ("nvls", None, 0, False, 2), ("oneshot", None, 0, False, 1),
                 ("stages_bulk", [6], 0, False, 3), ("nvls", None, 0, True, 2), ("nvls", None, 0, False, 5),
                 ("stages", [5], 1, False, 2), ("ring", None, 0, True, 1), ("stages", [2, 2], 2, False, 3),
                 ("nvls", None, 0, False, 8), ("twoshot", None, 0, False, 2)]
        for it in range(2 * len(plans)):
            algo, widths, lonely, registered, ctas = plans[it % len(plans)]
            count = int(rng.integers(1, 3000)) if algo not in ("oneshot", "twoshot") else int(rng.integers(1, 600))
            xs = [rng.integers(-1000, 1000, count).astype(np.int32) for _ in range(n)]
            if registered:
                bufs = [w.registered(r, count, np.int32) for r in range(n)]
                for b, x in zip(bufs, xs):
                    b[:] = x
            else:
                bufs = [aligned_copy(x) for x in xs]
            dt_kw = {}
            if algo == "nvls":  # the in-switch reduce exists for floats; small integers in fp32 are exact
                xs = [x.astype(np.float32) for x in xs]
                if registered:
                    bufs = [w.registered(r, count, np.float32) for r in range(n)]
                    for b, x in zip(bufs, xs):
                        b[:] = x
                else:
                    bufs = [aligned_copy(x) for x in xs]
            tile = [64, 256, 1024][it % 3]
            if algo == "nvls" and not registered:
                tile = 1024  # one flag slot per tile: keep ntiles <= FT_MAX_CTAS
            rc = w.allreduce(bufs, None, algo=algo, widths=widths, ctas=ctas, tile_bytes=tile, lonely=lonely,
                             mode=K.MODE_REGISTERED if registered else K.MODE_STAGING, **dt_kw)
            assert rc == 0, (it, algo, rc)
            want = np.sum(np.stack(xs), 0, dtype=xs[0].dtype)
            for r in range(n):
                np.testing.assert_array_equal(bufs[r], want, err_msg=f"iteration {it} {algo} {widths} rank {r}")
    finally:
        w.set_jitter(0)
        w.close()
