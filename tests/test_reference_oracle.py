"""This library against the UNMODIFIED reference (mpi_mod.hpp run through the in-repo MPI shim) on the same
random data: integers and fp64/fp32 sums must match bit for bit whenever the block split coincides (the
summation order - own value first, then group peers in ascending rank - is the same by design)."""
import os
import subprocess

import numpy as np
import pytest

import flextree as ft
from flextree import HostWorld

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BIN = os.path.join(ROOT, "baseline", "_ref", "bin")
RNG = np.random.default_rng(99)


def _have_ref():
    if not os.path.exists(os.path.join(BIN, "ref_oracle")):
        subprocess.run(["bash", os.path.join(ROOT, "baseline", "build_ref.sh")], capture_output=True)
    return os.path.exists(os.path.join(BIN, "ref_oracle"))


def _aligned(a):
    raw = np.empty(a.nbytes + 64, dtype=np.uint8)
    off = (-raw.ctypes.data) % 16
    v = raw[off:off + a.nbytes].view(a.dtype)
    v[:] = a
    return v


def run_reference(tmp_path, xs, topo, dt, op, inplace=False):
    n = len(xs)
    for r, x in enumerate(xs):
        x.tofile(tmp_path / f"in.{r}.bin")
    env = dict(os.environ, FT_TOPO=topo, OMP_NUM_THREADS="2")
    cmd = [os.path.join(BIN, "ftmpirun"), "-np", str(n), os.path.join(BIN, "ref_oracle"), str(tmp_path), str(xs[0].size), dt, op]
    if inplace:
        cmd.append("inplace")
    p = subprocess.run(cmd, capture_output=True, text=True, env=env, timeout=120)
    assert p.returncode == 0, p.stdout + p.stderr
    return [np.fromfile(tmp_path / f"out.{r}.bin", dtype=xs[0].dtype) for r in range(n)]


@pytest.mark.parametrize("n,topo,widths", [(4, "2,2", [2, 2]), (6, "2,3", [2, 3]), (6, "3,2", [3, 2]), (6, "", [6]),
                                           (8, "2,4", [2, 4]), (8, "4,2", [4, 2]), (8, "2,2,2", [2, 2, 2]), (8, "8", [8])])
def test_bit_exact_against_unmodified_reference(tmp_path, n, topo, widths):
    if not _have_ref():
        pytest.skip("reference sources not available")
    w = HostWorld(n, window_bytes=1 << 18)
    # counts whose ceil(count/N) block is already a multiple of 16 bytes: identical block split
    for count, dt, name in ((4 * n * 25, np.float32, "f32"), (4 * n * 7, np.int32, "i32"), (2 * n * 33, np.float64, "f64")):
        if dt == np.int32:
            xs = [RNG.integers(-2**31, 2**31 - 1, count, dtype=np.int32) for _ in range(n)]
        else:
            xs = [(RNG.standard_normal(count) * 10.0 ** RNG.integers(-3, 4)).astype(dt) for _ in range(n)]
        ref = run_reference(tmp_path, xs, topo, name, "sum")
        bufs = [_aligned(x) for x in xs]
        assert w.allreduce(bufs, None, widths=widths, ctas=2, tile_bytes=256) == 0
        for r in range(n):
            np.testing.assert_array_equal(bufs[r], ref[r], err_msg=f"{name} {widths} rank {r}")
    # BAND + in place + a count that is NOT a multiple of N (tail blocks) - integers are order-free
    count = 1001
    xs = [RNG.integers(0, 2**31 - 1, count, dtype=np.int32) for _ in range(n)]
    ref = run_reference(tmp_path, xs, topo, "i32", "band", inplace=True)
    bufs = [_aligned(x) for x in xs]
    assert w.allreduce(bufs, None, op="band", widths=widths, ctas=3, tile_bytes=128) == 0
    for r in range(n):
        np.testing.assert_array_equal(bufs[r], ref[r])


def test_ring_mode_against_reference(tmp_path):
    if not _have_ref():
        pytest.skip("reference sources not available")
    n, count = 5, 1000
    xs = [RNG.integers(-10**6, 10**6, count, dtype=np.int32) for _ in range(n)]
    ref = run_reference(tmp_path, xs, "1", "i32", "sum")     # FT_TOPO containing 1 -> ring
    w = HostWorld(n, window_bytes=1 << 16)
    bufs = [_aligned(x) for x in xs]
    assert w.allreduce(bufs, None, algo="ring", ctas=2, tile_bytes=256) == 0
    for r in range(n):
        np.testing.assert_array_equal(bufs[r], ref[r])
    assert ft.parse_topo("1", n) == [1]
