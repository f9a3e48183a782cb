"""ORACLE — test infrastructure.  Recipe for `oracle/_ref/`: the REAL reference, compiled where it lies.

The reference's hot path is three pure-Python modules.  This compiles them from `/root/reference` (read-only;
nothing is copied into the repository's history) to sourceless byte code under `oracle/_ref/` —

    /root/reference/model_utils.py    -> oracle/_ref/model_utils.bin
    /root/reference/vector_utils.py   -> oracle/_ref/vector_utils.bin
    /root/reference/steering_utils.py -> oracle/_ref/steering_utils.bin
    /root/reference/eval_utils.py, detect_injected_thoughts.py -> likewise (the driver, for the drop-in test)

(`.pyc` content under another suffix: the snapshot that carries the tree to the GPU box leaves `*.pyc` files behind) —
which is git-ignored but travels to the GPU box with the tree, like the built `libia_b200.so`.  There,
`oracle/real_ref.py` imports it so that `bench.py`'s reference arm, its `cpu_baseline` and its `gpu_hook_baseline`
time the reference's own `ModelWrapper.generate_batch_with_multi_steering` / `extract_concept_vectors_batch`
(kind "reference") instead of the restatement in `oracle/ref_hooks.py` (kind "port", the fallback when `_ref` is absent
or was compiled by another Python version).  Run by `__graft_entry__.build()` when `/root/reference` is present.
"""

from __future__ import annotations

import os
import py_compile
import sys

REFERENCE = "/root/reference"
HERE = os.path.dirname(os.path.abspath(__file__))
OUT = os.path.join(HERE, "_ref")
MODULES = ("model_utils", "vector_utils", "steering_utils", "eval_utils", "detect_injected_thoughts")


def build(verbose: bool = True) -> bool:
    if not all(os.path.exists(os.path.join(REFERENCE, m + ".py")) for m in MODULES):
        return False
    os.makedirs(OUT, exist_ok=True)
    for m in MODULES:
        py_compile.compile(os.path.join(REFERENCE, m + ".py"), cfile=os.path.join(OUT, m + ".bin"), doraise=True, optimize=0)
    with open(os.path.join(OUT, "BUILT_WITH"), "w") as f:
        f.write(f"python {sys.version_info.major}.{sys.version_info.minor}; compiled from {REFERENCE} by oracle/build_ref.py\n")
    if verbose:
        print(f"[oracle] compiled {', '.join(MODULES)} from {REFERENCE} into {OUT}")
    return True


if __name__ == "__main__":
    sys.exit(0 if build() else 1)
