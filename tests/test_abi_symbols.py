"""CPU-only checks of the C-ABI boundary: the library builds/loads, exports every symbol the
public header declares, and rejects bad arguments before touching a GPU."""

import ctypes
import os
import subprocess

import pytest

from introspective_awareness_b200 import _lib


@pytest.fixture(scope="module")
def lib():
    if not os.path.exists(_lib.LIB_PATH):
        from introspective_awareness_b200.build import build

        build()
    return _lib.load()


def test_header_symbols_are_exported(lib):
    declared = _lib.declared_symbols()
    assert len(declared) >= 8 and "ia_inject_rmsnorm" in declared and "ia_gather_meandiff" in declared
    exported = subprocess.run(["nm", "-D", "--defined-only", _lib.LIB_PATH], capture_output=True, text=True).stdout
    exported = {line.split()[-1] for line in exported.splitlines() if " T " in line}
    for name in declared:
        assert name in exported, f"{name} declared in include/ia_b200.h but not exported"
        assert hasattr(lib, name)
    # nothing but the declared C symbols leaks out of the library
    assert {s for s in exported if s.startswith("ia_")} == set(declared)
    # every declared function has a ctypes signature in the binding
    for name in declared:
        assert name in _lib._SIGNATURES, name


def test_version_and_error_strings(lib):
    assert lib.ia_version() == 1
    assert lib.ia_error_string(0) == b"ok"
    for code in (-1, -2, -3, -4, -5, -6):
        assert lib.ia_error_string(code) not in (b"ok", b"unknown error")


def test_argument_errors_return_before_any_launch(lib):
    buf = (ctypes.c_char * 4096)()
    p = ctypes.addressof(buf)
    p = (p + 15) & ~15
    # NULL pointers
    assert lib.ia_inject_rmsnorm(0, p, 0, 0, 0, 0, p, 1e-6, 1, 1, 128, 0, 0, 0, 0, 0, 0) == -1
    # H not a multiple of 4
    assert lib.ia_inject_rmsnorm(p, p, 0, 0, 0, 0, p, 1e-6, 1, 1, 126, 0, 0, 0, 0, 0, 0) == -2
    # misaligned pointer
    assert lib.ia_inject_rmsnorm(p + 2, p, 0, 0, 0, 0, p, 1e-6, 1, 1, 128, 0, 0, 0, 0, 0, 0) == -4
    # unknown variant
    assert lib.ia_inject_rmsnorm(p, p, 0, 0, 0, 0, p, 1e-6, 1, 1, 128, 0, 0, 0, 7, 0, 0) == -3
    # steering rows given but no table
    assert lib.ia_inject_rmsnorm(p, p, 0, p, p, p, p, 1e-6, 1, 1, 128, 0, 0, 0, 0, 0, 0) == -1
    # zero rows is a no-op
    assert lib.ia_inject_rmsnorm(p, p, 0, 0, 0, 0, p, 1e-6, 0, 1, 128, 0, 0, 0, 0, 0, 0) == 0
    # K1: workspace query and too-small workspace
    need = lib.ia_meandiff_workspace_bytes(32, 50, 100, 4096, 0)
    assert need > 0 and need % 256 == 0
    assert lib.ia_meandiff_workspace_bytes(32, 50, 100, 4100, 0) == -2
    assert lib.ia_gather_meandiff(p, 8, 0, 0, 50, p, 8, 0, 0, 100, 32, 4096, 0, 0, 0, p, 0, p, 16, 0, 0) == -5
    assert lib.ia_gather_meandiff(p, 8, 0, 0, 50, p, 8, 0, 0, 100, 32, 4096, 9, 0, 0, p, 0, p, need, 0, 0) == -3
    assert lib.ia_gather_meandiff(0, 8, 0, 0, 50, p, 8, 0, 0, 100, 32, 4096, 0, 0, 0, p, 0, p, need, 0, 0) == -1


def test_product_path_refuses_cpu_tensors(lib):
    import torch

    from introspective_awareness_b200 import kernels

    h = torch.zeros(2, 1, 128, dtype=torch.bfloat16)
    w = torch.ones(128, dtype=torch.bfloat16)
    with pytest.raises(RuntimeError, match="no CPU path"):
        kernels.inject_rmsnorm(h, w, 1e-6, "llama")
    with pytest.raises(RuntimeError, match="no CPU path"):
        kernels.gather_meandiff([h], [h], "baseline")


def test_missing_library_fails_loudly(monkeypatch, tmp_path):
    monkeypatch.setattr(_lib, "_lib", None)
    monkeypatch.setattr(_lib, "LIB_PATH", str(tmp_path / "libia_b200.so"))
    with pytest.raises(_lib.KernelLibraryError, match="no CPU fallback"):
        _lib.load()
