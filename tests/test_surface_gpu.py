"""The single-concept and utility functions of the kept call surface, EXECUTED on the GPU (SURVEY.md §8 rows
a3, a4, a5): extract_concept_vector, extract_concept_vector_with_baseline / _simple / _no_baseline,
cosine_similarity, save / load_concept_vector, analyze_vector_underspecification.

Anchors:
  * tests/golden/surface_tiny_llama_f32.npz — what the REAL reference functions (reference vector_utils.py:63-111,
    114-307, 310-381, 597-643) returned on the CPU for a seeded tiny Llama, and the token ids they handed to their
    model (written by oracle/make_surface_golden.py).  Ours must hand the same ids to the runtime (the templates
    differ per function: "Tell me about {word}" vs "{word}", control prompt "The") and return the same values
    within fp32 CPU-vs-GPU rounding (1e-4 relative L2; north_star's bound is 1e-3).
  * the restated reference on the SAME device and model (oracle/ref_hooks.py) in bf16: same bits.
"""

import numpy as np
import pytest
import torch

from conftest import load_golden
from oracle import ref_hooks
from oracle.surface_cases import CALLS, UNDERSPEC, func_of, materialize

pytestmark = pytest.mark.gpu


def _dev():
    if not torch.cuda.is_available():
        pytest.skip("needs a GPU")
    torch.backends.cuda.matmul.allow_tf32 = False
    torch.backends.cudnn.allow_tf32 = False
    return "cuda:0"


class _Spy:
    """Records every batch our wrapper encodes."""

    def __init__(self, w):
        self.w, self.batches, self._real = w, [], w._encode

    def __enter__(self):
        def encode(prompts):
            ids, mask = self._real(prompts)
            self.batches.append((ids.clone(), mask.clone()))
            return ids, mask

        self.w._encode = encode
        return self

    def __exit__(self, *exc):
        self.w._encode = self._real


def _rel(a, b):
    a, b = a.float().flatten(), b.float().flatten()
    return ((a - b).norm() / b.norm()).item()


def test_single_concept_functions_match_the_real_reference():
    from introspective_awareness_b200 import vector_utils as vu
    from introspective_awareness_b200.model_utils import ModelWrapper

    g = load_golden("surface_tiny_llama_f32.npz")
    layer = int(g["layer_idx"])
    w = ModelWrapper.from_arch(str(g["arch"]), device=_dev(), dtype=torch.float32, seed=int(g["seed"]))
    got = {}
    for tag, kw in CALLS:
        with _Spy(w) as spy:
            vec = getattr(vu, func_of(tag))(w, layer_idx=layer, **materialize(kw, vu.get_baseline_words))
        assert vec.device.type == "cpu" and vec.dtype == torch.float32 and vec.shape == (w.engine.hidden,)
        assert len(spy.batches) == int(g[f"{tag}__nbatches"]), tag
        for i, (ids, mask) in enumerate(spy.batches):  # the same prompts, tokenised and padded the same way
            assert np.array_equal(ids.numpy(), g[f"{tag}__ids{i}"]), (tag, i)
            assert np.array_equal(mask.numpy(), g[f"{tag}__mask{i}"]), (tag, i)
        want = torch.from_numpy(g[f"{tag}__vec"])
        assert _rel(vec, want) <= 1e-4, (tag, _rel(vec, want))
        got[tag] = vec
    # cosine similarity: python float, in the vectors' own dtype, eps in the denominator (reference :310-328)
    a, b = got["with_baseline"], got["simple"]
    assert isinstance(vu.cosine_similarity(a, b), float)
    assert abs(vu.cosine_similarity(a, b) - float(g["cosine_ab"])) <= 1e-4
    assert abs(vu.cosine_similarity(a, a) - float(g["cosine_aa"])) <= 1e-6
    ra, rb = torch.from_numpy(g["with_baseline__vec"]), torch.from_numpy(g["simple__vec"])
    assert vu.cosine_similarity(ra, rb) == float(g["cosine_ab"])  # same inputs: the same float
    assert vu.cosine_similarity(ra.to(torch.bfloat16), rb.to(torch.bfloat16)) == float(g["cosine_bf16"])
    assert vu.cosine_similarity(ra.reshape(2, -1), rb.reshape(-1, 2)) == float(g["cosine_ab"])  # flattened first
    # underspecification analysis: one batched extraction instead of one pair of forwards per concept
    related = [str(r) for r in g["underspec_related"]]
    assert related == UNDERSPEC["related"]
    sims = vu.analyze_vector_underspecification(w, UNDERSPEC["target"], related, layer,
                                                baseline_words=vu.get_baseline_words(UNDERSPEC["n_baseline"]))
    assert list(sims) == related and all(isinstance(v, float) for v in sims.values())
    for r, want in zip(related, g["underspec_values"]):
        assert abs(sims[r] - float(want)) <= 1e-4, (r, sims[r], float(want))


def test_single_concept_functions_bf16_same_device(tmp_path):
    """bf16 on the GPU against the restated reference hooks on the same model: identical bits; then the file
    round trip of save_concept_vector / load_concept_vector (reference :331-381: bare torch.save + JSON side-car)."""
    from introspective_awareness_b200 import vector_utils as vu
    from introspective_awareness_b200.model_utils import ModelWrapper

    w = ModelWrapper.from_arch("tiny-qwen2", device=_dev(), dtype=torch.bfloat16, seed=31)
    layer = 1
    tok = w.tokenizer

    def acts(prompts, token_idx=-1):
        enc = tok(prompts, return_tensors="pt", padding=True, truncation=True)
        return ref_hooks.capture(w.model, enc["input_ids"], enc["attention_mask"], layer, token_idx)

    chat = lambda text: vu._chat_prompt(w, text)
    base_words = vu.get_baseline_words(9)
    cases = [
        (vu.extract_concept_vector(w, ["LOUD ONE", "LOUD TWO WORDS"], ["quiet one", "quiet", "soft spoken words"], layer),
         ref_hooks.concept_vectors(acts(["LOUD ONE", "LOUD TWO WORDS"]), acts(["quiet one", "quiet", "soft spoken words"]), "posneg")[0]),
        (vu.extract_concept_vector_with_baseline(w, "Dust", base_words, layer),
         ref_hooks.concept_vectors(acts([chat("Tell me about Dust")]), acts([chat(f"Tell me about {b}") for b in base_words]), "baseline")[0]),
        (vu.extract_concept_vector_simple(w, "Dust", layer, normalize=True),
         ref_hooks.concept_vectors(acts([chat("Dust")]), acts([chat("The")]), "simple", normalize=True)[0]),
        (vu.extract_concept_vector_no_baseline(w, "Dust", layer, token_idx=-2),
         ref_hooks.concept_vectors(acts([chat("Dust")], -2), None, "no_baseline")[0]),
    ]
    for k, (ours, ref) in enumerate(cases):
        assert ours.dtype == torch.bfloat16 and ours.device.type == "cpu"
        assert _rel(ours, ref) <= 1e-3, (k, _rel(ours, ref))
        assert torch.equal(ours, ref), k
    vec = cases[1][0]
    path = tmp_path / "vectors" / "layer_0.50" / "Dust.pt"
    vu.save_concept_vector(vec, path, metadata={"concept": "Dust", "layer": layer})
    back, meta = vu.load_concept_vector(path)
    assert torch.equal(back, vec) and meta == {"concept": "Dust", "layer": layer}
    assert torch.equal(torch.load(path), vec)  # a bare tensor, as the reference's plotting tools expect
    vu.save_concept_vector(vec, tmp_path / "bare.pt")
    assert vu.load_concept_vector(tmp_path / "bare.pt")[1] is None
