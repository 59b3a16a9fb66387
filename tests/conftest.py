import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")
    # a fresh checkout has no built libraries (they are git-ignored): build them once, where a
    # compiler exists (nvcc cross-compiles sm_100a without a GPU); on the GPU box the prebuilt
    # files travel with the snapshot and this is a no-op
    import shutil
    import subprocess
    lib = os.path.join(ROOT, "sxmc_b200", "libsxmc_b200.so")
    shim = os.path.join(ROOT, "sxmc_b200", "libsxmc_cxx.so")
    runner = os.path.join(ROOT, "cxx", "tests", "test_class_api")
    sxmc_bin = os.path.join(ROOT, "cxx", "bin", "sxmc")
    nvcc = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    if not os.path.exists(lib) and os.path.exists(nvcc):
        subprocess.run(["make", "-j4", "-C", os.path.join(ROOT, "sxmc_b200", "csrc")], check=False,
                       stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
    if os.path.exists(lib) and not (os.path.exists(shim) and os.path.exists(runner) and os.path.exists(sxmc_bin)) and shutil.which("g++"):
        subprocess.run(["make", "-C", os.path.join(ROOT, "cxx")], check=False, stdout=subprocess.DEVNULL,
                       stderr=subprocess.DEVNULL)


@pytest.fixture(scope="session")
def oracle_built():
    import oracle
    oracle.build()
    return oracle
