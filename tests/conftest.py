import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


def load_golden(name):
    return dict(np.load(os.path.join(GOLDEN, name), allow_pickle=False))


def widen(arr, dtype_name):
    """Stored bit pattern -> float32 values."""
    from oracle import kernels_np as K

    if dtype_name == "bf16":
        return K.bf16_to_f32(arr)
    return np.asarray(arr, dtype=np.float32)


def ulp_diff_bf16(a_f32, b_f32):
    """Distance in bf16 units-in-the-last-place between two bf16-valued float32 arrays."""
    from oracle import kernels_np as K

    def key(x):
        b = K.bf16_bits(x).astype(np.int32)
        return np.where(b & 0x8000, 0x8000 - b, b)  # monotone integer order

    return np.abs(key(a_f32) - key(b_f32))


@pytest.fixture(scope="session")
def golden_dir():
    return GOLDEN
