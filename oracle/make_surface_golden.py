"""ORACLE — test infrastructure.  Generates `tests/golden/surface_*.npz` by running the REAL reference.

Run in the build container only (`python oracle/make_surface_golden.py`): it imports
`/root/reference/{model_utils,vector_utils}.py` unmodified and calls every single-concept / utility
function of the kept call surface that `make_golden.py` does not reach through the batch entry point:

    extract_concept_vector               reference vector_utils.py:63-111
    extract_concept_vector_with_baseline reference vector_utils.py:114-183
    extract_concept_vector_simple        reference vector_utils.py:186-251
    extract_concept_vector_no_baseline   reference vector_utils.py:254-307
    cosine_similarity                    reference vector_utils.py:310-328
    analyze_vector_underspecification    reference vector_utils.py:597-643

on a seeded random-init tiny Llama (fp32, CPU) with the stub tokenizer.  For each call the fixture
records the token ids the reference handed to its model (so the GPU test can check that our wrappers
build the same prompts — the templates differ per function, `"Tell me about {word}"` vs `"{word}"`,
control prompt `"The"`) and the value the reference returned.
"""

from __future__ import annotations

import os
import sys

sys.dont_write_bytecode = True
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, "/root/reference")

import numpy as np
import torch

import vector_utils as ref_vu  # noqa: E402  (the reference)

from introspective_awareness_b200.stub_tokenizer import StubTokenizer  # noqa: E402
from introspective_awareness_b200.synthetic import build_model  # noqa: E402
from oracle.make_golden import GOLD, ref_wrapper  # noqa: E402

from oracle.surface_cases import CALLS, UNDERSPEC, func_of, materialize  # noqa: E402

ARCH, SEED, LAYER = "tiny-llama", 21, 2


class RecordingTokenizer(StubTokenizer):
    """Stub tokenizer that keeps every batch it encodes."""

    def __init__(self, *a, **k):
        super().__init__(*a, **k)
        self.batches = []

    def __call__(self, prompts, **kw):
        enc = super().__call__(prompts, **kw)
        if kw.get("padding"):
            self.batches.append((enc["input_ids"].clone(), enc["attention_mask"].clone()))
        return enc


def main():
    model = build_model(ARCH, dtype=torch.float32, seed=SEED)
    tok = RecordingTokenizer(vocab_size=model.config.vocab_size)
    w = ref_wrapper(model, tok, torch.float32)
    out = dict(arch=ARCH, seed=np.int32(SEED), layer_idx=np.int32(LAYER))
    for tag, kw in CALLS:
        tok.batches.clear()
        vec = getattr(ref_vu, func_of(tag))(w, layer_idx=LAYER, **materialize(kw, ref_vu.get_baseline_words))
        assert vec.dim() == 1 and vec.dtype == torch.float32
        out[f"{tag}__vec"] = vec.numpy()
        for i, (ids, mask) in enumerate(tok.batches):
            out[f"{tag}__ids{i}"] = ids.numpy()
            out[f"{tag}__mask{i}"] = mask.numpy()
        out[f"{tag}__nbatches"] = np.int32(len(tok.batches))
        print(tag, func_of(tag), "batches", [tuple(b[0].shape) for b in tok.batches], "norm", float(vec.norm()))
    a, b = torch.from_numpy(out["with_baseline__vec"]), torch.from_numpy(out["simple__vec"])
    out["cosine_ab"] = np.float64(ref_vu.cosine_similarity(a, b))
    out["cosine_aa"] = np.float64(ref_vu.cosine_similarity(a, a))
    out["cosine_bf16"] = np.float64(ref_vu.cosine_similarity(a.to(torch.bfloat16), b.to(torch.bfloat16)))
    related = UNDERSPEC["related"]
    sims = ref_vu.analyze_vector_underspecification(w, UNDERSPEC["target"], related, LAYER,
                                                    baseline_words=ref_vu.get_baseline_words(UNDERSPEC["n_baseline"]))
    out["underspec_related"] = np.array(related)
    out["underspec_values"] = np.array([sims[r] for r in related], np.float64)
    print("cosine", out["cosine_ab"], "underspecification", sims)
    np.savez_compressed(os.path.join(GOLD, "surface_tiny_llama_f32.npz"), **out)


if __name__ == "__main__":
    main()
