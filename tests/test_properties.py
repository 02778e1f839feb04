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
