"""In-tree build of libia_b200.so (sm_100a only) with plain nvcc.

`python -m introspective_awareness_b200.build` or `build()` from `__graft_entry__.py`.
nvcc cross-compiles without a GPU; the resulting .so sits next to this file so that it travels
with the source tree (it is git-ignored, not gpurun-ignored).
"""

from __future__ import annotations

import os
import shutil
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
OBJ = os.path.join(HERE, "build")
LIB = os.path.join(HERE, "libia_b200.so")
SOURCES = ["ia_abi.cu", "ia_inject_rmsnorm.cu", "ia_inject_rmsnorm_tma.cu", "ia_inject_rmsnorm_stream.cu", "ia_meandiff.cu", "ia_rope.cu", "ia_mlp_act.cu"]
NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-std=c++17",
    "-Xcompiler", "-fPIC", "-Xcompiler", "-fvisibility=hidden",
]


def _nvcc() -> str:
    exe = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    if not os.path.exists(exe):
        raise RuntimeError("nvcc not found; cannot build libia_b200.so")
    return exe


def _deps(src: str):
    yield os.path.join(CSRC, src)
    yield os.path.join(CSRC, "ia_common.cuh")
    yield os.path.join(CSRC, "ia_k2_common.cuh")
    yield os.path.join(os.path.dirname(HERE), "include", "ia_b200.h")


def _stale(target: str, deps) -> bool:
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.exists(d) and os.path.getmtime(d) > t for d in deps)


def build(force: bool = False, verbose: bool = True) -> str:
    os.makedirs(OBJ, exist_ok=True)
    nvcc = _nvcc()
    sources = [s for s in SOURCES if os.path.exists(os.path.join(CSRC, s))]
    jobs = []
    for src in sources:
        obj = os.path.join(OBJ, src.replace(".cu", ".o"))
        if force or _stale(obj, _deps(src)):
            jobs.append((src, [nvcc, *NVCC_FLAGS, "-c", os.path.join(CSRC, src), "-o", obj]))

    def run(job):
        src, cmd = job
        if verbose:
            print("[ia_b200] nvcc", src, flush=True)
        res = subprocess.run(cmd, capture_output=True, text=True)
        if res.returncode != 0:
            raise RuntimeError(f"nvcc failed for {src}:\n{res.stdout}\n{res.stderr}")
        return src

    if jobs:
        with ThreadPoolExecutor(max_workers=min(4, len(jobs))) as pool:
            list(pool.map(run, jobs))
    objs = [os.path.join(OBJ, s.replace(".cu", ".o")) for s in sources]
    if force or jobs or _stale(LIB, objs):
        cmd = [nvcc, "-shared", "-o", LIB, *objs, "-gencode", "arch=compute_100a,code=sm_100a"]
        res = subprocess.run(cmd, capture_output=True, text=True)
        if res.returncode != 0:
            raise RuntimeError(f"link failed:\n{res.stdout}\n{res.stderr}")
        if verbose:
            print("[ia_b200] linked", LIB, flush=True)
    return LIB


if __name__ == "__main__":
    build(force="--force" in sys.argv)
