"""Real multi-GPU tests: one process per GPU through torch.distributed.run (NCCL only for bootstrap and
as the oracle). Skipped unless >= 2 GPUs are visible."""
import os
import subprocess
import sys

import pytest

pytestmark = [pytest.mark.gpu, pytest.mark.multigpu]
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _launch(nproc, port, script="mp_worker.py"):
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={nproc}",
           "--master-addr", "127.0.0.1", "--master-port", str(port), os.path.join(ROOT, "tests", script)]
    return subprocess.run(cmd, capture_output=True, text=True, timeout=900, cwd=ROOT)


def _ngpu():
    import torch
    return torch.cuda.device_count()


def test_two_ranks():
    p = _launch(2, 29611)
    assert p.returncode == 0 and "MP_WORKER_OK" in p.stdout, p.stdout[-3000:] + p.stderr[-3000:]


def test_all_ranks():
    n = _ngpu()
    if n <= 2:
        pytest.skip("covered by test_two_ranks")
    p = _launch(n, 29612)
    assert p.returncode == 0 and "MP_WORKER_OK" in p.stdout, p.stdout[-3000:] + p.stderr[-3000:]


def test_watchdog_on_missing_rank():
    """Failure detection: a rank that never shows up makes the others time out with an error word."""
    p = _launch(2, 29613, "mp_fault_worker.py")
    assert p.returncode == 0 and "MP_FAULT_OK" in p.stdout, p.stdout[-3000:] + p.stderr[-3000:]


def test_ft_bench_multi_device():
    """The C benchmark with the reference's CLI: one process, one communicator per GPU (peer access)."""
    exe = os.path.join(ROOT, "build", "ft_bench")
    if not os.path.exists(exe):
        pytest.skip("build/ft_bench not built")
    env = dict(os.environ, FT_TOPO="2")
    p = subprocess.run([exe, "--np", "2", "--size", "1000000", "--repeat", "3"], capture_output=True, text=True,
                       timeout=300, env=env, cwd=ROOT)
    assert p.returncode == 0, p.stdout[-2000:] + p.stderr[-2000:]
    lines = [ln for ln in p.stdout.splitlines() if ln.startswith("CHECK")]
    # in-place repeats compound like the reference benchmark: data[i] = i * 2^3
    assert lines == [f"CHECK {r}: " + " ".join(str(i * 8) for i in range(9, 20)) + " " for r in range(2)], lines
    p = subprocess.run([exe, "--np", "2", "--size", "1000000", "--repeat", "2", "--comm-type", "nccl"],
                       capture_output=True, text=True, timeout=300, cwd=ROOT)
    assert p.returncode == 0 and "CHECK 1: 36 40" in p.stdout, p.stdout[-2000:] + p.stderr[-2000:]


def _build_dropin_gpu_user():
    exe = os.path.join(ROOT, "build", "dropin_gpu_user")
    libdir = os.path.join(ROOT, "allreduce-over-mpi_b200", "lib")
    cuda = os.environ.get("CUDA_HOME", "/usr/local/cuda")
    cmd = ["/usr/bin/g++", "-std=c++17", "-O2", "-w", "-I", os.path.join(ROOT, "compat"), "-I", os.path.join(ROOT, "include"),
           "-I", os.path.join(ROOT, "baseline", "mpi_shim"), "-I", os.path.join(cuda, "include"),
           os.path.join(ROOT, "tests", "cpp", "dropin_gpu_user.cc"), os.path.join(ROOT, "baseline", "mpi_shim", "mpi_shim.cc"),
           "-o", exe, "-L", libdir, "-lflextree", "-Wl,-rpath," + libdir, "-L", os.path.join(cuda, "lib64"), "-lcudart",
           "-Wl,-rpath," + os.path.join(cuda, "lib64"), "-lpthread", "-lrt"]
    r = subprocess.run(cmd, capture_output=True, text=True)
    assert r.returncode == 0, r.stderr[-2000:]
    launcher = os.path.join(ROOT, "build", "ftmpirun")
    r = subprocess.run(["/usr/bin/g++", "-std=c++17", "-O2", os.path.join(ROOT, "baseline", "mpi_shim", "ftmpirun.cc"), "-o", launcher],
                       capture_output=True, text=True)
    assert r.returncode == 0, r.stderr[-2000:]
    return launcher, exe


def test_mpi_allreduce_on_device_memory_through_the_dropin_header():
    """compat/mpi_mod.hpp: MPI_Allreduce on cudaMalloc'ed buffers is captured and runs on the GPU kernels (and host
    buffers in the same program on the CPU plane): `ftmpirun -np 2 dropin_gpu_user`."""
    launcher, exe = _build_dropin_gpu_user()
    for topo in (None, "2"):
        env = dict(os.environ)
        env.pop("FT_TOPO", None)
        if topo:
            env["FT_TOPO"] = topo
        p = subprocess.run([launcher, "-np", "2", exe], capture_output=True, text=True, timeout=300, env=env, cwd=ROOT)
        assert p.returncode == 0, p.stdout[-2000:] + p.stderr[-2000:]
        assert [ln for ln in p.stdout.splitlines() if ln.startswith("DROPIN_GPU")] == ["DROPIN_GPU 0/2 OK", "DROPIN_GPU 1/2 OK"]


def test_ft_bench_under_ftrun_one_process_per_gpu():
    """The C benchmark as the reference runs it: a launcher starts one process per GPU (ftrun = mpirun stand-in); this
    bootstrap reaches VMM + multicast, so --algo nvls on symmetric buffers works without Python."""
    exe, launcher = os.path.join(ROOT, "build", "ft_bench"), os.path.join(ROOT, "build", "ftrun")
    if not (os.path.exists(exe) and os.path.exists(launcher)):
        pytest.skip("tools not built")
    for extra in (["--algo", "auto"], ["--algo", "nvls", "--symmetric"], ["--comm-type", "nccl"]):
        p = subprocess.run([launcher, "-np", "2", exe, "--size", "1000000", "--repeat", "3", *extra], capture_output=True, text=True,
                           timeout=300, cwd=ROOT)
        assert p.returncode == 0, " ".join(extra) + p.stdout[-2000:] + p.stderr[-2000:]
        lines = [ln for ln in p.stdout.splitlines() if ln.startswith("CHECK")]
        assert lines == [f"CHECK {r}: " + " ".join(str(i * 8) for i in range(9, 20)) + " " for r in range(2)], (extra, lines)
