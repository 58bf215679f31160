"""The single-node MPI stand-in the reference arm is built against (baseline/mpi_shim): its
self-test on host buffers - collectives, tag matching out of posting order, probe/get_count,
zero-length messages, ring wrap - with 1, 2 and 4 ranks as processes.  No GPU needed (device
buffers take the CUDA-IPC path, which the reference arm itself exercises on the GPU boxes)."""
import os
import shutil
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SHIM = os.path.join(ROOT, "baseline", "mpi_shim")
CUDA = os.environ.get("CUDA_HOME", "/usr/local/cuda")


@pytest.fixture(scope="module")
def selftest(tmp_path_factory):
    if not shutil.which("g++") or not os.path.exists(os.path.join(CUDA, "include", "cuda_runtime.h")):
        pytest.skip("needs g++ and the CUDA headers")
    d = tmp_path_factory.mktemp("shim")
    lib, exe = str(d / "libmpi.so"), str(d / "selftest")
    subprocess.run(["g++", "-std=c++17", "-O1", "-fPIC", "-shared", f"-I{SHIM}/include", f"-I{CUDA}/include",
                    f"{SHIM}/src/mpi_shim.cpp", "-o", lib, f"-L{CUDA}/lib64", "-lcudart", "-lrt", "-lpthread",
                    f"-Wl,-rpath,{CUDA}/lib64"], check=True)
    subprocess.run(["g++", "-std=c++17", "-O1", f"-I{SHIM}/include", f"{SHIM}/test/selftest.cpp", "-o", exe,
                    f"-L{d}", "-lmpi", f"-Wl,-rpath,{d}"], check=True)
    return exe


@pytest.mark.timeout(180)
@pytest.mark.parametrize("ranks", [1, 2, 4])
def test_mpi_stand_in_self_test(selftest, ranks):
    job = f"pytest{os.getpid()}_{ranks}"
    procs = [subprocess.Popen([selftest], env=dict(os.environ, TZMPI_RANK=str(r), TZMPI_SIZE=str(ranks),
                                                   TZMPI_JOB=job, CUDA_VISIBLE_DEVICES=""),
                              stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
             for r in range(ranks)]
    outs = []
    try:
        for p in procs:
            out, _ = p.communicate(timeout=120)
            outs.append(out)
    finally:
        for p in procs:
            if p.poll() is None:
                p.kill()
    assert all(p.returncode == 0 for p in procs), outs
    assert f"selftest ok ({ranks} ranks)" in outs[0]
