"""Property tests (hypothesis): random world size, random ordered factorization, random count and tiling."""
import numpy as np
from hypothesis import HealthCheck, given, settings, strategies as st

import flextree as ft
from flextree import HostWorld

from refimpl import block_size, staged_allreduce


@st.composite
def plans(draw):
    n = draw(st.integers(2, 12))
    widths = draw(st.sampled_from(ft.enumerate_widths(n)))
    count = draw(st.integers(1, 3000))
    ctas = draw(st.integers(1, 4))
    tile = draw(st.sampled_from([16, 64, 256, 4096]))
    return n, widths, count, ctas, tile


@settings(max_examples=40, deadline=None, suppress_health_check=[HealthCheck.too_slow])
@given(plans(), st.integers(0, 2**31 - 1))
def test_any_factorization_any_count(plan, seed):
    n, widths, count, ctas, tile = plan
    rng = np.random.default_rng(seed)
    xs = [rng.integers(-2**31, 2**31 - 1, count, dtype=np.int32) for _ in range(n)]
    w = HostWorld(n, window_bytes=1 << 16)
    bufs = []
    for x in xs:
        raw = np.empty(x.nbytes + 64, dtype=np.uint8)
        off = (-raw.ctypes.data) % 16
        v = raw[off:off + x.nbytes].view(np.int32)
        v[:] = x
        bufs.append(v)
    assert w.allreduce(bufs, None, widths=widths, ctas=ctas, tile_bytes=tile) == 0
    with np.errstate(over="ignore"):
        want = np.sum(np.stack(xs), 0, dtype=np.int32)
    for b in bufs:
        np.testing.assert_array_equal(b, want)
    w.close()


@settings(max_examples=25, deadline=None, suppress_health_check=[HealthCheck.too_slow])
@given(plans(), st.integers(0, 2**31 - 1))
def test_fp32_matches_order_aware_model(plan, seed):
    n, widths, count, ctas, tile = plan
    rng = np.random.default_rng(seed)
    xs = [rng.standard_normal(count).astype(np.float32) for _ in range(n)]
    w = HostWorld(n, window_bytes=1 << 16)
    bufs = []
    for x in xs:
        raw = np.empty(x.nbytes + 64, dtype=np.uint8)
        off = (-raw.ctypes.data) % 16
        v = raw[off:off + x.nbytes].view(np.float32)
        v[:] = x
        bufs.append(v)
    assert w.allreduce(bufs, None, widths=widths, ctas=ctas, tile_bytes=tile) == 0
    want = staged_allreduce(xs, widths, block=block_size(count, n, 4))[0]
    for b in bufs:
        np.testing.assert_array_equal(b, want)
    w.close()


@given(st.integers(1, 400))
def test_enumeration_matches_count(n):
    lists = ft.enumerate_widths(n)
    assert len(lists) == ft.count_factorizations(n)
    for wl in lists:
        p = 1
        for x in wl:
            p *= x
        assert p == n or (n == 1 and wl == [])


@settings(max_examples=40, deadline=None, suppress_health_check=[HealthCheck.too_slow])
@given(plans(), st.integers(0, 2**31 - 1), st.sampled_from([np.float32, np.float64, np.int16, np.uint8]),
       st.booleans(), st.booleans())
def test_bulk_tiles_equal_per_vector_text(plan, seed, dtype, registered, scaled):
    """The CPU's bulk tiles (algo 5, the TMA hook) against the per-vector text of the GPU kernels: same bits for any
    plan, count (ragged ones fall back), tiling, dtype, private or registered buffers, with or without the scale."""
    n, widths, count, ctas, tile = plan
    rng = np.random.default_rng(seed)
    dt = np.dtype(dtype)
    if dt.kind == "f":
        xs = [rng.standard_normal(count).astype(dt) for _ in range(n)]
    else:
        info = np.iinfo(dt)
        xs = [rng.integers(info.min, info.max, count, dtype=dt, endpoint=True) for _ in range(n)]
    scale = 0.25 if (scaled and dt.kind == "f") else None
    res = []
    for algo in ("stages", "stages_bulk"):
        w = HostWorld(n, window_bytes=1 << 16)
        bufs = []
        for r, x in enumerate(xs):
            if registered:
                v = w.registered(r, count, dt)
            else:
                raw = np.empty(x.nbytes + 64, dtype=np.uint8)
                off = (-raw.ctypes.data) % 16
                v = raw[off:off + x.nbytes].view(dt)
            v[:] = x
            bufs.append(v)
        mode = ft.ops.constants.MODE_REGISTERED if registered else ft.ops.constants.MODE_STAGING
        assert w.allreduce(bufs, None, algo=algo, widths=widths, ctas=ctas, tile_bytes=tile, mode=mode, scale=scale) == 0
        res.append([b.copy() for b in bufs])
        w.close()
    for a, b in zip(*res):
        np.testing.assert_array_equal(a, b)


@st.composite
def algo_calls(draw):
    n = draw(st.integers(2, 9))
    algo = draw(st.sampled_from(["ring", "nvls", "oneshot", "stages+lonely", "halves"]))
    count = draw(st.integers(1, 2500))
    ctas = draw(st.integers(1, 6))
    tile = draw(st.sampled_from([64, 256, 1024, 4096]))
    registered = draw(st.booleans())
    return n, algo, count, ctas, tile, registered


@settings(max_examples=60, deadline=None, suppress_health_check=[HealthCheck.too_slow])
@given(algo_calls(), st.integers(0, 2**31 - 1))
def test_other_algorithms_any_shape(call, seed):
    """Ring, NVLS (emulated switch; one-role, role-split and registered variants), one-shot LL, lonely trees and the
    reduce-scatter / all-gather halves for random world sizes, counts, thread counts and tilings (exact: integers,
    or small integers carried in fp32 for the float-only NVLS path)."""
    n, algo, count, ctas, tile, registered = call
    rng = np.random.default_rng(seed)
    K = ft.ops.constants
    w = HostWorld(n, window_bytes=1 << 16, ll_bytes=1 << 16)
    try:
        dt = np.float32 if algo == "nvls" else np.int32
        xs = [rng.integers(-500, 500, count).astype(dt) for _ in range(n)]

        def buffers(src):
            out = []
            for r, x in enumerate(src):
                if registered and algo in ("ring", "nvls"):
                    v = w.registered(r, x.size, x.dtype)
                else:
                    raw = np.empty(x.nbytes + 64, dtype=np.uint8)
                    off = (-raw.ctypes.data) % 16
                    v = raw[off:off + x.nbytes].view(x.dtype)
                v[:] = x
                out.append(v)
            return out

        want = np.sum(np.stack(xs), 0, dtype=dt)
        if algo in ("ring", "nvls"):
            mode = K.MODE_REGISTERED if registered else K.MODE_STAGING
            if algo == "nvls" and not registered:
                tile = max(tile, 1024)  # one flag slot per tile
            bufs = buffers(xs)
            assert w.allreduce(bufs, None, algo=algo, ctas=ctas, tile_bytes=tile, mode=mode) == 0
            for b in bufs:
                np.testing.assert_array_equal(b, want)
        elif algo == "oneshot":
            xs = [x[:500] for x in xs]
            bufs = buffers(xs)
            assert w.allreduce(bufs, None, algo="oneshot", ctas=ctas, tile_bytes=tile) == 0
            for b in bufs:
                np.testing.assert_array_equal(b, np.sum(np.stack(xs), 0, dtype=dt))
        elif algo == "stages+lonely":
            lonely = 1 if n < 4 else int(rng.integers(1, 3))
            tree = n - lonely
            if tree < 1:
                return
            widths = ft.enumerate_widths(tree)[int(rng.integers(0, len(ft.enumerate_widths(tree))))] if tree > 1 else [1]
            if tree == 1:
                return
            bufs = buffers(xs)
            assert w.allreduce(bufs, None, widths=widths, lonely=lonely, ctas=ctas, tile_bytes=tile) == 0
            for b in bufs:
                np.testing.assert_array_equal(b, want)
        else:  # halves
            per = max(1, count // n)
            total = per * n
            xs = [x[:total] if x.size >= total else np.resize(x, total) for x in xs]
            widths = ft.enumerate_widths(n)[int(rng.integers(0, len(ft.enumerate_widths(n))))]
            send = buffers(xs)
            part = buffers([np.zeros(per, dt) for _ in range(n)])
            assert w.allreduce(send, part, widths=widths, ctas=ctas, tile_bytes=tile, phases=1, count=total) == 0
            full = np.sum(np.stack(xs), 0, dtype=dt)
            for r in range(n):
                np.testing.assert_array_equal(part[r], full[r * per:(r + 1) * per])
            out = buffers([np.zeros(total, dt) for _ in range(n)])
            assert w.allreduce(part, out, widths=widths, ctas=ctas, tile_bytes=tile, phases=2, count=total) == 0
            for r in range(n):
                np.testing.assert_array_equal(out[r], full)
    finally:
        w.close()
