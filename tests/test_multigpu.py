"""Real multi-GPU tests: one process per GPU through torch.distributed.run (NCCL only for bootstrap and
as the oracle). Skipped unless >= 2 GPUs are visible."""
import os
import subprocess
import sys

import pytest

pytestmark = [pytest.mark.gpu, pytest.mark.multigpu]
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _launch(nproc, port):
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={nproc}",
           "--master-addr", "127.0.0.1", "--master-port", str(port), os.path.join(ROOT, "tests", "mp_worker.py")]
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
