"""Real multi-process runs on the CPU (torch.distributed, gloo): bootstrap through the all-gather callback,
shared-memory windows, the kernels' algorithm templates executing in separate processes."""
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.parametrize("nproc,port", [(2, 29811), (4, 29812), (6, 29813)])
def test_host_communicator_under_torchrun(nproc, port):
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={nproc}",
           "--master-addr", "127.0.0.1", "--master-port", str(port), os.path.join(ROOT, "tests", "mp_cpu_worker.py")]
    env = dict(os.environ, OMP_NUM_THREADS="1", FT_ALLOW_HOST_ONLY="1")
    p = subprocess.run(cmd, capture_output=True, text=True, timeout=600, cwd=ROOT, env=env)
    assert p.returncode == 0 and "MP_CPU_WORKER_OK" in p.stdout, p.stdout[-3000:] + p.stderr[-3000:]


@pytest.mark.parametrize("nproc,port", [(2, 29821), (4, 29822)])
def test_host_data_parallel_matches_gloo_averaged_gradients(nproc, port):
    """ft.HostDataParallel: gradients live in the registered shared-memory region, bucket collectives overlap
    backward on a worker thread; 4 SGD steps equal to plain backward + gloo all_reduce / N."""
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={nproc}",
           "--master-addr", "127.0.0.1", "--master-port", str(port), os.path.join(ROOT, "tests", "mp_cpu_dp_worker.py")]
    env = dict(os.environ, OMP_NUM_THREADS="1", FT_ALLOW_HOST_ONLY="1")
    p = subprocess.run(cmd, capture_output=True, text=True, timeout=600, cwd=ROOT, env=env)
    assert p.returncode == 0 and "MP_CPU_DP_WORKER_OK" in p.stdout, p.stdout[-3000:] + p.stderr[-3000:]
