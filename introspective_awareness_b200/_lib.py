"""ctypes binding of libia_b200.so (include/ia_b200.h).  No fallback: if the library is missing
or a call fails, an exception is raised — there is no CPU or eager-torch path for the hot ops."""

from __future__ import annotations

import ctypes
import os
import re
from ctypes import c_char_p, c_float, c_int, c_int32, c_int64, c_void_p

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "libia_b200.so")
HEADER_PATH = os.path.join(os.path.dirname(HERE), "include", "ia_b200.h")

IA_BF16, IA_F16, IA_F32 = 0, 1, 2
IA_NORM_LLAMA, IA_NORM_GEMMA = 0, 1
IA_STEER_ALL, IA_RSQRT_IEEE, IA_IMPL_TMA, IA_IMPL_REGS, IA_IMPL_ADD_STAGED, IA_IMPL_STREAM = 1, 2, 4, 8, 16, 32
IA_MODE_BASELINE, IA_MODE_SIMPLE, IA_MODE_RAW, IA_MODE_POSNEG = 0, 1, 2, 3


class KernelLibraryError(RuntimeError):
    pass


_SIGNATURES = {
    "ia_version": (c_int, []),
    "ia_error_string": (c_char_p, [c_int]),
    "ia_inject_rmsnorm": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_float,
                                  c_int64, c_int64, c_int64, c_int32, c_int32, c_int32, c_int32, c_int32, c_void_p]),
    "ia_gemma2_tail_inject_rmsnorm": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p,
                                              c_void_p, c_void_p, c_float, c_int64, c_int64, c_int64, c_int32,
                                              c_int32, c_int32, c_int32, c_void_p]),
    "ia_inject_rmsnorm_tap": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_float,
                                      c_int64, c_int64, c_int64, c_int32, c_int32, c_int32, c_int32, c_int32,
                                      c_void_p, c_void_p, c_int32, c_void_p]),
    "ia_add_inject_rmsnorm_tap": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p,
                                          c_float, c_int64, c_int64, c_int64, c_int32, c_int32, c_int32, c_int32, c_int32,
                                          c_void_p, c_void_p, c_int32, c_void_p]),
    "ia_gemma2_tail_inject_rmsnorm_tap": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p,
                                                  c_void_p, c_void_p, c_float, c_int64, c_int64, c_int64, c_int32,
                                                  c_int32, c_int32, c_int32, c_void_p, c_void_p, c_int32, c_void_p]),
    "ia_meandiff_workspace_bytes": (c_int64, [c_int64, c_int64, c_int64, c_int64, c_int32]),
    "ia_gather_meandiff": (c_int, [c_void_p, c_int64, c_void_p, c_void_p, c_int64, c_void_p, c_int64, c_void_p,
                                   c_void_p, c_int64, c_int64, c_int64, c_int32, c_int32, c_void_p, c_void_p,
                                   c_void_p, c_void_p, c_int64, c_int32, c_void_p]),
    "ia_gather_segsum": (c_int, [c_void_p, c_int64, c_void_p, c_void_p, c_int64, c_int64, c_int64, c_void_p,
                                 c_void_p, c_void_p, c_int64, c_int32, c_void_p]),
    "ia_meandiff_finalize": (c_int, [c_void_p, c_int64, c_void_p, c_int64, c_int64, c_int64, c_int32, c_int32,
                                     c_void_p, c_void_p, c_void_p, c_void_p, c_int64, c_int32, c_void_p]),
    "ia_meandiff_compact": (c_int, [c_void_p, c_int64, c_void_p, c_int64, c_int64, c_int64, c_int32, c_int32, c_void_p, c_void_p,
                                    c_void_p, c_void_p, c_int64, c_int32, c_void_p]),
    "ia_segsum_compact": (c_int, [c_void_p, c_int64, c_int64, c_int64, c_void_p, c_void_p, c_int64, c_int32, c_void_p]),
    "ia_gather_meandiff_static": (c_int, [c_void_p, c_int64, c_void_p, c_void_p, c_int64, c_void_p, c_int64, c_void_p,
                                          c_void_p, c_int64, c_int64, c_int64, c_int32, c_int32, c_void_p, c_void_p,
                                          c_void_p, c_int32, c_void_p]),
    "ia_rope": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int64, c_int64, c_int64, c_int64,
                        c_int32, c_void_p]),
    "ia_add_inject_rmsnorm": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_float,
                                      c_int64, c_int64, c_int64, c_int32, c_int32, c_int32, c_int32, c_int32, c_void_p]),
    "ia_rope_append": (c_int, [c_void_p, c_void_p, c_void_p, c_int64, c_int64, c_int64, c_void_p, c_void_p, c_void_p, c_void_p,
                               c_void_p, c_int64, c_int64, c_int64, c_int64, c_int64, c_int64, c_int64, c_int64, c_int64,
                               c_int32, c_void_p]),
    "ia_silu_mul": (c_int, [c_void_p, c_void_p, c_void_p, c_int64, c_int64, c_int64, c_int64, c_int32, c_void_p]),
    "ia_act_mul_table": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_int64, c_int64, c_int64, c_int64, c_int32, c_void_p]),
}
IA_STATIC_SCRATCH_BYTES = 4 * 1024 * 1024

_lib = None


def declared_symbols(header_path: str = HEADER_PATH):
    """Every function the public header declares with IA_API."""
    text = open(header_path).read()
    return re.findall(r"IA_API\s+[\w\s\*]+?\b(ia_\w+)\s*\(", text)


def load() -> ctypes.CDLL:
    """Load the kernel library, failing loudly when it has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise KernelLibraryError(
            f"{LIB_PATH} not found: build it with `python -m introspective_awareness_b200.build` "
            "(or __graft_entry__.build()).  There is no CPU fallback for the steering/extraction kernels.")
    lib = ctypes.CDLL(LIB_PATH)
    for name, (res, args) in _SIGNATURES.items():
        try:
            fn = getattr(lib, name)
        except AttributeError:
            continue  # optional entry points are checked by the symbol test against the header
        fn.restype = res
        fn.argtypes = args
    if lib.ia_version() != 1:
        raise KernelLibraryError(f"libia_b200.so ABI version {lib.ia_version()} != 1")
    _lib = lib
    return lib


def check(rc: int, what: str) -> None:
    if rc != 0:
        msg = load().ia_error_string(int(rc))
        raise KernelLibraryError(f"{what} failed with code {rc}: {msg.decode() if msg else '?'}")
