"""ORACLE — test infrastructure.  The REAL reference (`oracle/_ref/`, built by `oracle/build_ref.py`) on token ids.

`bench.py` works on synthetic token ids; the reference's API takes strings and tokenizes them itself.  `IdTokenizer`
closes that gap without touching the reference: a prompt is the ids written out as decimal numbers, and the tokenizer
reads them back (left padding, pad id 0, the mapping-with-`.to()` the reference expects; no chat template, so the
reference's helpers take their plain-text fallback and whatever words that adds are ignored).  `wrap()` builds the
reference's own `ModelWrapper` around an existing HF model the way `oracle/make_golden.py` does (its `__init__` needs the
Hub).  Everything that then runs — hooks, `model.generate`, the vector arithmetic — is the reference's own code.

Only `tests/`, `__graft_entry__.smoke()` and `bench.py`'s baseline legs may import this.
"""

from __future__ import annotations

import os
import re
import sys
from typing import List, Optional, Sequence

import torch

from introspective_awareness_b200.stub_tokenizer import StubTokenizer

REF_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "_ref")
_INT = re.compile(r"\d+")
_modules = None


def _load_one(name: str):
    """Import oracle/_ref/<name>.bin (byte code compiled from the reference's <name>.py) as module `name`."""
    from importlib.machinery import SourcelessFileLoader
    from importlib.util import module_from_spec, spec_from_file_location

    path = os.path.join(REF_DIR, name + ".bin")
    spec = spec_from_file_location(name, path, loader=SourcelessFileLoader(name, path))
    mod = module_from_spec(spec)
    sys.modules[name] = mod  # vector_utils does `from model_utils import ...`
    spec.loader.exec_module(mod)
    return mod


def load():
    """(model_utils, vector_utils) of the real reference, or None when oracle/_ref is missing or unusable
    (e.g. compiled by another Python version)."""
    global _modules
    if _modules is not None:
        return _modules or None
    _modules = False
    if not os.path.exists(os.path.join(REF_DIR, "model_utils.bin")):
        return None
    saved = {k: sys.modules.get(k) for k in ("model_utils", "vector_utils")}
    try:
        _modules = (_load_one("model_utils"), _load_one("vector_utils"))
    except Exception:
        _modules = False
    finally:
        for k, v in saved.items():  # leave the names as we found them; the reference modules stay reachable through us
            if v is not None:
                sys.modules[k] = v
            else:
                sys.modules.pop(k, None)
    return _modules or None


class IdTokenizer(StubTokenizer):
    """Prompts are token ids written as decimal numbers; anything that is not a number is ignored."""

    def encode(self, text: str) -> List[int]:
        return [int(t) for t in _INT.findall(text)]


def ids_to_prompts(ids: torch.Tensor, mask: torch.Tensor) -> List[str]:
    return [" ".join(str(int(t)) for t in row[m.bool()]) for row, m in zip(ids, mask)]


def wrap(model, device: str, dtype: torch.dtype, tokenizer=None, model_name: str = "synthetic"):
    """The reference's ModelWrapper around `model` (reference model_utils.py:56-137, minus the Hub download)."""
    mu, _ = load()
    w = mu.ModelWrapper.__new__(mu.ModelWrapper)
    w.model_name, w.hf_path = model_name, model_name
    w.device, w.dtype = device, dtype
    w.model = model
    w.tokenizer = tokenizer if tokenizer is not None else IdTokenizer(vocab_size=model.config.vocab_size)
    w.hooks = []
    w.n_layers = w._get_n_layers()
    return w


def concept_vectors(w, c_ids, c_mask, b_ids, b_mask, layer_idx: int):
    """reference vector_utils.extract_concept_vectors_batch on id prompts -> [C, H] in the model dtype (CPU)."""
    _, vu = load()
    cw, bw = ids_to_prompts(c_ids, c_mask), ids_to_prompts(b_ids, b_mask)
    vecs = vu.extract_concept_vectors_batch(w, cw, bw, layer_idx=layer_idx, extraction_method="baseline", template="{word}")
    return torch.stack([vecs[k] for k in cw])


def steered_generate(w, ids, mask, layer_idx: int, vectors: Sequence[torch.Tensor], strength: float,
                     start_positions: Optional[Sequence[int]], max_new_tokens: int) -> List[str]:
    """reference ModelWrapper.generate_batch_with_multi_steering on id prompts (greedy)."""
    return w.generate_batch_with_multi_steering(
        prompts=ids_to_prompts(ids, mask), layer_idx=layer_idx, steering_vectors=list(vectors), strength=strength,
        max_new_tokens=max_new_tokens, temperature=0.0, steering_start_positions=list(start_positions))


def load_driver(bind: dict):
    """The reference's sweep driver (detect_injected_thoughts.py, compiled into oracle/_ref) imported with the names
    `model_utils` / `vector_utils` / `steering_utils` bound to the modules in `bind` — the real reference's own, or this
    package's (INTEGRATION.md level 1: the import swap).  Plotting / judge dependencies that are not installed
    (matplotlib, openai, dotenv) are stubbed: the sweep itself uses none of them with --no-llm-judge.  Returns the
    driver module, or None when oracle/_ref does not hold it."""
    import types

    if not os.path.exists(os.path.join(REF_DIR, "detect_injected_thoughts.bin")) or load() is None:
        return None
    saved = {k: sys.modules.get(k) for k in ("model_utils", "vector_utils", "steering_utils", "eval_utils", "matplotlib",
                                             "matplotlib.pyplot", "seaborn", "openai", "dotenv")}
    try:
        for name in ("matplotlib", "matplotlib.pyplot", "seaborn", "openai", "dotenv"):
            try:
                __import__(name)
            except Exception:
                stub = types.ModuleType(name)
                stub.use = lambda *a, **k: None
                stub.load_dotenv = lambda *a, **k: None
                stub.AsyncOpenAI = stub.OpenAI = object
                sys.modules[name] = stub
        if "matplotlib.pyplot" in sys.modules:
            sys.modules["matplotlib"].pyplot = sys.modules["matplotlib.pyplot"]
        for k, mod in bind.items():
            sys.modules[k] = mod
        _load_one("eval_utils")
        drv = _load_one("detect_injected_thoughts")
        sys.modules.pop("detect_injected_thoughts", None)
        for name in ("create_sweep_plots", "create_cross_model_comparison_plots", "extract_example_transcripts"):
            setattr(drv, name, lambda *a, **k: None)  # reporting only
        return drv
    finally:
        for k, v in saved.items():
            if v is not None:
                sys.modules[k] = v
            else:
                sys.modules.pop(k, None)


def load_steering_utils():
    """The real reference's steering_utils (needs `model_utils` importable while it loads)."""
    mods = load()
    if mods is None:
        return None
    saved = sys.modules.get("model_utils")
    sys.modules["model_utils"] = mods[0]
    try:
        su = _load_one("steering_utils")
        sys.modules.pop("steering_utils", None)
        return su
    finally:
        if saved is not None:
            sys.modules["model_utils"] = saved
        else:
            sys.modules.pop("model_utils", None)
