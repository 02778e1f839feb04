"""CPU tests of the command-line tools and of the reference-baseline harness (no GPU needed)."""
import json
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _build_tools():
    subprocess.run(["make", "-C", ROOT, "build/ft_plan"], capture_output=True, check=False)


def test_ft_plan_cli():
    exe = os.path.join(ROOT, "build", "ft_plan")
    if not os.path.exists(exe):
        _build_tools()
    if not os.path.exists(exe):
        pytest.skip("ft_plan not built")
    out = subprocess.run([exe, "--candidates", "8"], capture_output=True, text=True).stdout.split()
    assert out == ["2*2*2", "2*4", "4*2", "1*8", "8*1", "3*3-1", "1*9-1", "9*1-1", "1*7+1", "7*1+1"]
    out = subprocess.run([exe, "--legacy", "8"], capture_output=True, text=True).stdout
    assert "should be: 1*8" in out and "0.666" in out
    out = subprocess.run([exe, "--count", "41160"], capture_output=True, text=True).stdout.strip()
    assert out == "24924"
    out = subprocess.run([exe, "--print-ops", "--widths", "2,4", "--rank", "5"], capture_output=True, text=True).stdout
    assert "peer 4: blocks 0 2 4 6" in out and "peer 7: blocks 7" in out
    out = subprocess.run([exe, "--sweep", "--max", "16"], capture_output=True, text=True).stdout.splitlines()
    assert out[0].startswith("n,structures") and out[16].startswith("16,9,") and ",2*8," in out[16]
    out = subprocess.run([exe, "--choose", "8", "--bytes", "67108864", "--nvls", "--registered"], capture_output=True, text=True).stdout
    assert "choice for N=8" in out and "nvls" in out.splitlines()[-1]


def test_fd_channel_scm_rights(tmp_path):
    """The fd-passing channel used by the VMM symmetric-memory bootstrap, between forked processes."""
    exe = tmp_path / "fd_channel_test"
    r = subprocess.run(["/usr/bin/g++" if os.path.exists("/usr/bin/g++") else "g++", "-std=c++17", "-O1", "-o", str(exe),
                        os.path.join(ROOT, "tests", "cpp", "fd_channel_test.cc")], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    for n in ("2", "8"):
        p = subprocess.run([str(exe), n], capture_output=True, text=True, timeout=60)
        assert p.returncode == 0 and "ok" in p.stdout, p.stdout + p.stderr


def test_reference_baseline_runs_unmodified():
    """baseline/_ref: the reference's own benchmark.cpp + mpi_mod.hpp against the in-repo MPI/glog shims."""
    ref_bin = os.path.join(ROOT, "baseline", "_ref", "bin", "allreduce_over_mpi")
    if not os.path.exists(ref_bin):
        r = subprocess.run(["bash", os.path.join(ROOT, "baseline", "build_ref.sh")], capture_output=True, text=True)
        if r.returncode != 0:
            pytest.skip("reference sources not available here: " + r.stderr[-200:])
    run = os.path.join(ROOT, "baseline", "_ref", "bin", "ftmpirun")
    env = dict(os.environ, FT_TOPO="2,3", OMP_NUM_THREADS="2")
    p = subprocess.run([run, "-np", "6", ref_bin, "--size", "35"], capture_output=True, text=True, env=env, timeout=120,
                       cwd=str(os.path.join(ROOT, "build")) if os.path.isdir(os.path.join(ROOT, "build")) else ROOT)
    checks = [ln for ln in p.stdout.splitlines() if ln.startswith("CHECK")]
    assert len(checks) == 6 and all(ln.split(": ")[1].split() == [str(v) for v in range(54, 120, 6)] for ln in checks), p.stdout
    # the sources in baseline/_ref/src are byte-identical to the mounted reference when it is available
    src = "/root/reference/allreduce_over_mpi/mpi_mod.hpp"
    if os.path.exists(src):
        assert open(src, "rb").read() == open(os.path.join(ROOT, "baseline", "_ref", "src", "allreduce_over_mpi", "mpi_mod.hpp"), "rb").read()


def test_bench_reference_arm_json():
    p = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--gpus", "2", "--steps", "2",
                        "--warmup", "1", "--bytes", str(1 << 22)], capture_output=True, text=True, timeout=300,
                       env=dict(os.environ, OMP_NUM_THREADS="2"))
    assert p.returncode == 0, p.stderr[-2000:]
    d = json.loads(p.stdout.strip().splitlines()[-1])
    assert d["impl"] == "reference"
    if "unavailable" not in d:
        assert d["n_gpus"] == 2 and d["value"] > 0 and d["config"]["ranks"] == 2 and d["check"].startswith("CHECK 0")


def test_public_header_is_c99_and_cxx11(tmp_path):
    """include/flextree.h must stay usable from plain C (the reference's entry point is a C signature)."""
    src = tmp_path / "hdr.c"
    src.write_text('#define STANDALONE_TEST\n#include "flextree.h"\nint main(void) { ft_comm_options_t o; (void)o; return (int)FT_SUCCESS; }\n')
    inc = os.path.join(ROOT, "include")
    r = subprocess.run(["gcc", "-std=c99", "-Wall", "-Wextra", "-pedantic", "-Werror", f"-I{inc}", "-fsyntax-only", str(src)],
                       capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    r = subprocess.run(["g++", "-std=c++11", "-Wall", "-Wextra", "-Werror", f"-I{inc}", "-fsyntax-only", "-x", "c++", str(src)],
                       capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    cuda_inc = "/usr/local/cuda/include"
    if os.path.isdir(cuda_inc):  # the C example compiles against the header
        r = subprocess.run(["gcc", "-std=c99", f"-I{inc}", f"-I{cuda_inc}", "-fsyntax-only",
                            os.path.join(ROOT, "examples", "c_local_allreduce.c")], capture_output=True, text=True)
        assert r.returncode == 0, r.stderr


def test_launcher_front_end(tmp_path):
    """python -m flextree.launch: the single-host stand-in for mpirun + hostfile (allreduce_over_mpi/Makefile:8-24)."""
    from flextree import launch
    cmd = launch.build_command(3, "x.py", ["--a", "1"], port=12345)
    assert cmd[1:4] == ["-m", "torch.distributed.run", "--nnodes=1"] and "--nproc-per-node=3" in cmd
    assert cmd[cmd.index("--master-addr") + 1] == "127.0.0.1" and cmd[-3:] == ["x.py", "--a", "1"]
    script = tmp_path / "w.py"
    # one os.write per rank: two ranks share the pipe, and print() would emit its fields in separate writes
    script.write_text("import os\nos.write(1, ('RANK %s %s %s\\n' % (os.environ['RANK'], os.environ['WORLD_SIZE'], "
                      "os.environ.get('FT_TOPO'))).encode())\n")
    p = subprocess.run([sys.executable, "-m", "flextree.launch", "-np", "2", "--cpu", "--topo", "2", str(script)],
                       capture_output=True, text=True, cwd=ROOT, timeout=300)
    assert p.returncode == 0, p.stderr[-500:]
    assert "RANK 0 2 2" in p.stdout and "RANK 1 2 2" in p.stdout


def test_ftrun_launcher_and_env_bootstrap(tmp_path):
    """`ftrun -np 4 prog`: the C launcher + unix-socket all-gather that bootstraps C programs (no Python, no MPI)."""
    import flextree._lib as L
    if not os.path.exists(L.CUDA_LIB):
        pytest.skip("libflextree.so not built")
    subprocess.check_call(["make", "-C", ROOT, "-j4", "tools"], stdout=subprocess.DEVNULL)
    exe = str(tmp_path / "env_boot_test")
    libdir = os.path.dirname(L.CUDA_LIB)
    subprocess.check_call(["/usr/bin/g++", "-std=c++17", "-O1", "-I", os.path.join(ROOT, "include"), "-o", exe,
                           os.path.join(ROOT, "tests", "cpp", "env_boot_test.cc"), "-L", libdir, "-lflextree",
                           "-Wl,-rpath," + libdir])
    p = subprocess.run([os.path.join(ROOT, "build", "ftrun"), "-np", "4", "--gpus", "2", exe], capture_output=True, text=True, timeout=120)
    assert p.returncode == 0, p.stdout + p.stderr
    lines = [ln for ln in p.stdout.splitlines() if ln.startswith("ENV_BOOT_OK")]
    assert lines == [f"ENV_BOOT_OK {r}/4 local {r % 2}" for r in range(4)], p.stdout
    # a failing rank takes the job down with its exit code
    p = subprocess.run([os.path.join(ROOT, "build", "ftrun"), "-np", "2", "sh", "-c", "test $FT_RANK = 0 || exit 7; sleep 30"],
                       capture_output=True, text=True, timeout=60)
    assert p.returncode == 7


def test_ft_plan_topo_mode():
    """`ft_plan --topo TEXT --ranks N`: how FT_TOPO is read, incl. the '+k' / '-k' suffixes of the reference's listings."""
    exe = os.path.join(ROOT, "build", "ft_plan")
    if not os.path.exists(exe):
        pytest.skip("ft_plan not built")
    run = lambda *a: subprocess.run([exe, *a], capture_output=True, text=True)
    assert run("--topo", "2,4-1", "--ranks", "7").stdout.strip() == "tree 2*4-1: 8 positions on 7 ranks, 0 lonely, 1 phantom"
    assert run("--topo", "2 3+1", "--ranks", "7").stdout.strip() == "tree 2*3+1: 6 positions on 7 ranks, 1 lonely, 0 phantom"
    assert run("--topo", "1", "--ranks", "5").stdout.strip() == "ring over 5 ranks"
    bad = run("--topo", "2,3", "--ranks", "7")
    assert bad.returncode == 1 and "invalid FT_TOPO" in bad.stderr
