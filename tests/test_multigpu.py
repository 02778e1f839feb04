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
