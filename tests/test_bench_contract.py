"""The driver-facing contract of bench.py: one JSON line with the agreed keys.  The reference arm runs on
the CPU (tiny architecture here); our arm needs the GPU."""

import json
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BASE_KEYS = {"metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better", "scaling",
             "vs_baseline", "dtype", "data", "config", "e2e", "gpu_launches", "cpu_baseline"}


def _run(args, timeout):
    out = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), *args], capture_output=True, text=True,
                         timeout=timeout, cwd=ROOT)
    assert out.returncode == 0, out.stderr[-2000:]
    lines = out.stdout.splitlines()
    assert len(lines) == 1 and lines[0].startswith("{"), out.stdout[-2000:]  # ONE JSON line and nothing else on stdout
    return json.loads(lines[0])


def test_stdout_carries_only_the_json_line():
    """Whatever a library writes to file descriptor 1 while the bench runs (NCCL prints its version banner there when the
    box sets NCCL_DEBUG=VERSION) must not reach the driver's pipe."""
    code = ("import importlib.util, os\n"
            "spec = importlib.util.spec_from_file_location('bench', 'bench.py'); b = importlib.util.module_from_spec(spec)\n"
            "spec.loader.exec_module(b)\n"
            "b._stdout_for_the_json_line_only()\n"
            "os.write(1, b'NCCL version 2.28.9+cuda12.9\\n'); print('library chatter')\n"
            "b.emit({'ok': 1})\n")
    out = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, timeout=300, cwd=ROOT)
    assert out.returncode == 0, out.stderr[-2000:]
    assert out.stdout == '{"ok": 1}\n' and "NCCL version" in out.stderr and "library chatter" in out.stderr


def test_reference_arm_line_on_cpu():
    d = _run(["--impl", "reference", "--arch", "tiny-llama", "--steps", "1", "--warmup", "0", "--cpu-rows", "2"], 600)
    assert BASE_KEYS <= set(d) and d["impl"] == "reference"
    assert d["unit"] == "trials/s" and d["higher_is_better"] is True and d["value"] > 0
    assert d["e2e"] == {"value": d["value"], "unit": "trials/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
    assert d["cpu_baseline"]["kind"] in ("reference", "port") and d["cpu_baseline"]["value"] == d["value"] and d["cpu_baseline"]["cores"] >= 1
    assert "workload" in d["config"] and d["gpu_launches"] == 0


@pytest.mark.gpu
def test_our_arm_line_on_gpu():
    import torch

    if not torch.cuda.is_available():
        pytest.skip("needs a GPU")
    d = _run(["--arch", "tiny-llama", "--steps", "2", "--warmup", "3"], 900)
    assert BASE_KEYS | {"roofline", "clocks"} <= set(d) and "impl" not in d
    assert d["n_gpus"] == 1 and d["steps"] == 2 and d["warmup"] >= 3 and d["scaling"] == "weak" and d["vs_baseline"] is None
    assert d["value"] > 0 and d["e2e"]["value"] > 0 and d["e2e"]["h2d_bytes_per_step"] > 0 and d["e2e"]["d2h_bytes_per_step"] > 0
    r = d["roofline"]
    assert r["bound"] == "hbm" and r["unit"] == "GB/s" and r["peak"] > 0 and r["launches_timed"] == 2
    assert abs(r["frac"] - r["achieved"] / r["peak"]) < 1e-9
    n = d["gpu_launches_by_kernel"]
    layers = 4
    # per step: prompt pass + 31 graph-replayed decode steps, each 2 norms per layer + final norm through K2 ...
    assert n["k2_launches"] >= 2 * 32 * (2 * layers + 1) and n["rope_launches"] >= 2 * 32 * layers and n["k1_launches"] == 2
    assert d["gpu_launches"] == sum(n.values())
    assert d["cpu_baseline"]["kind"] in ("reference", "port") and d["cpu_baseline"]["value"] > 0
    assert set(d["clocks"]) >= {"sm_mhz", "sm_max_mhz", "reasons"}
