"""Source-level drop-in (compat/mpi_mod.hpp): programs written against the reference's header run on this
library's CPU data plane with no source change. Two programs are used: the reference's own, unmodified
benchmark.cpp (STANDALONE_TEST -> MPI_Allreduce_FT) and tests/cpp/dropin_user.cc (plain MPI code whose
MPI_Allreduce calls are captured, allreduce_over_mpi/mpi_mod.hpp:1167-1172). Both are built by
baseline/build_ref.sh against the in-repo MPI subset and run under its launcher."""
import os
import re
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BIN = os.path.join(ROOT, "baseline", "_ref", "bin")


@pytest.fixture(scope="module")
def bins():
    need = ["allreduce_over_mpi", "allreduce_over_mpi_dropin", "dropin_user", "ftmpirun"]
    stale = any(not os.path.exists(os.path.join(BIN, b)) for b in need)
    hdr = os.path.join(ROOT, "compat", "mpi_mod.hpp")
    lib = os.path.join(ROOT, "build", "libflextree_host.so")
    if not os.path.exists(lib):
        subprocess.run(["make", "-C", ROOT, "host"], capture_output=True)
    if not stale:
        built = os.path.getmtime(os.path.join(BIN, "allreduce_over_mpi_dropin"))
        stale = built < os.path.getmtime(hdr) or built < os.path.getmtime(lib)
    if stale:
        r = subprocess.run(["bash", os.path.join(ROOT, "baseline", "build_ref.sh")], capture_output=True, text=True)
        if r.returncode != 0 or not os.path.exists(os.path.join(BIN, "allreduce_over_mpi_dropin")):
            pytest.skip("reference sources or compiler not available: " + r.stderr[-200:])
    return BIN


def run(bins, np_, exe, args=(), topo=None, cwd=None):
    env = dict(os.environ)
    env.pop("FT_TOPO", None)
    if topo is not None:
        env["FT_TOPO"] = topo
    return subprocess.run([os.path.join(bins, "ftmpirun"), "-np", str(np_), os.path.join(bins, exe), *args],
                          capture_output=True, text=True, env=env, cwd=cwd, timeout=300)


def check_lines(out):
    return sorted(re.findall(r"CHECK \d+: [^\n]*", out))


@pytest.mark.parametrize("np_,topo", [(6, None), (6, "2,3"), (6, "3,2"), (8, "2,2,2"), (8, "2 4"), (5, "1"), (7, None), (1, None)])
def test_reference_benchmark_unmodified_on_this_data_plane(bins, tmp_path, np_, topo):
    """Same benchmark.cpp, both headers: identical CHECK lines, for every tree shape, the ring and N = 1."""
    ours = run(bins, np_, "allreduce_over_mpi_dropin", ["--size", "35", "--repeat", "3"], topo, tmp_path)
    ref = run(bins, np_, "allreduce_over_mpi", ["--size", "35", "--repeat", "3"], topo, tmp_path)
    assert ours.returncode == 0, ours.stderr[-400:]
    assert ref.returncode == 0, ref.stderr[-400:]
    a, b = check_lines(ours.stdout + ours.stderr), check_lines(ref.stdout + ref.stderr)
    assert len(a) == np_ and a == b


def test_reference_benchmark_large_message_and_time_file(bins, tmp_path):
    r = run(bins, 6, "allreduce_over_mpi_dropin", ["--size", "1000000", "--repeat", "5", "--to-file", "--tag", "dropin"],
            "2,3", tmp_path)
    assert r.returncode == 0, r.stderr[-400:]
    files = [f for f in os.listdir(tmp_path) if f.startswith("dropin.6.1000000.2-3-")]
    assert len(files) == 1 and ".ar_test." in files[0]  # the reference's file naming (benchmark.cpp:194-213)
    times = [float(v) for v in open(os.path.join(tmp_path, files[0])).read().split()]
    assert len(times) == 5 and all(0 < t < 5 for t in times)
    assert "average time" in r.stdout + r.stderr


def test_invalid_topo_exits_like_the_reference(bins, tmp_path):
    r = run(bins, 6, "allreduce_over_mpi_dropin", ["--size", "35"], "2,2", tmp_path)
    assert r.returncode != 0
    assert "invalid FT_TOPO" in r.stdout + r.stderr


@pytest.mark.parametrize("np_,topo", [(1, None), (4, None), (6, "2,3"), (6, "3,2"), (7, None), (5, "1"), (8, "2,2,2")])
def test_plain_mpi_program_is_captured(bins, tmp_path, np_, topo):
    """MPI_Allreduce calls of an ordinary MPI program: dtypes/ops of the reference's table, in place and out of place
    (sendbuf preserved), count < N, and a message that forces the grow-only window to be re-created."""
    r = run(bins, np_, "dropin_user", [], topo, tmp_path)
    assert r.returncode == 0, (r.stdout + r.stderr)[-400:]
    assert "DROPIN_USER_OK" in r.stdout
