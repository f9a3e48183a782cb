"""The C ABI from a host that is neither Python nor torch: tests/c_abi/abi_demo.cpp is built with nvcc,
dlopen()s libia_b200.so, allocates with the CUDA runtime and checks K2 / K1 (both entry points) bit for
bit against a scalar C restatement."""

import os
import shutil
import subprocess

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_plain_c_host_drives_the_abi(tmp_path):
    import torch

    if not torch.cuda.is_available():
        pytest.skip("needs a GPU")
    nvcc = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    if not os.path.exists(nvcc):
        pytest.skip("nvcc not available on this box")
    from introspective_awareness_b200 import _lib

    exe = str(tmp_path / "abi_demo")
    src = os.path.join(ROOT, "tests", "c_abi", "abi_demo.cpp")
    build = subprocess.run([nvcc, "-O1", "-Xcompiler", "-ffp-contract=off", "-o", exe, src, "-ldl"], capture_output=True, text=True)
    assert build.returncode == 0, build.stderr[-2000:]
    run = subprocess.run([exe, _lib.LIB_PATH], capture_output=True, text=True, timeout=120)
    assert run.returncode == 0, run.stdout + run.stderr
    assert "abi_demo ok" in run.stdout
