"""The REAL reference, compiled into oracle/_ref by oracle/build_ref.py, run on token ids through oracle/real_ref.py.

CPU (here): the bridge itself, and the restatement `oracle/ref_hooks.py` pinned against the real
`ModelWrapper.generate_batch_with_multi_steering` / `vector_utils.extract_concept_vectors_batch` at test time (the
golden generator pins it when the fixtures are made).  GPU (`-m gpu`, where oracle/_ref travelled with the tree): this
package against the real reference on the SAME device and model — vectors and returned strings identical.
"""

import os
import sys

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _real_ref():
    from oracle import build_ref, real_ref

    if not os.path.exists(os.path.join(real_ref.REF_DIR, "model_utils.bin")):
        build_ref.build(verbose=False)  # the build container has /root/reference; the GPU box gets the built files
    if real_ref.load() is None:
        pytest.skip("oracle/_ref not built (no /root/reference here)")
    return real_ref


@pytest.mark.parametrize("arch,dtype", [("tiny-llama", torch.float32), ("tiny-gemma2", torch.bfloat16), ("tiny-qwen2", torch.bfloat16)])
def test_restated_hooks_equal_the_real_reference_on_cpu(arch, dtype):
    from introspective_awareness_b200.synthetic import build_model, synthetic_prompt_ids
    from oracle import ref_hooks

    real_ref = _real_ref()
    m = build_model(arch, dtype=dtype, seed=3)
    w = real_ref.wrap(m, "cpu", dtype)
    V = m.config.vocab_size
    c_ids, c_mask, _ = synthetic_prompt_ids(4, 10, 5, V, seed=2)
    b_ids, b_mask, _ = synthetic_prompt_ids(7, 10, 5, V, seed=3)
    layer = 2
    vecs = real_ref.concept_vectors(w, c_ids, c_mask, b_ids, b_mask, layer)
    mine = ref_hooks.concept_vectors(ref_hooks.capture(m, c_ids, c_mask, layer), ref_hooks.capture(m, b_ids, b_mask, layer), "baseline")
    assert vecs.dtype == dtype and torch.equal(vecs, mine)
    ids, mask, lens = synthetic_prompt_ids(5, 16, 10, V, seed=1)
    t0 = (lens - 4).tolist()
    texts = real_ref.steered_generate(w, ids, mask, layer, [vecs[i % 4] for i in range(5)], 4.0, t0, 6)
    ref = ref_hooks.steered_generate(m, ids, mask, layer, [vecs[i % 4] for i in range(5)], 4.0, t0, 6)
    pads = (16 - lens).tolist()
    for i, text in enumerate(texts):
        # the reference slices at the unpadded length: pad_i prompt tokens lead the new ones
        full = torch.cat([ids[i], ref["new_ids"][i]])
        assert text == w.tokenizer.decode(full[int(lens[i]):], skip_special_tokens=True).strip(), (i, pads[i])


@pytest.mark.gpu
@pytest.mark.parametrize("arch,layer,rows", [("tiny-llama", 1, 12), ("tiny-gemma2", 2, 12), ("llama-8b-4layer", 2, 48)])
def test_against_the_real_reference_on_the_same_gpu(arch, layer, rows):
    """north_star's parity statement with the real reference as the other side: same GPU, same model object, bf16;
    `extract_concept_vectors_batch` and `generate_batch_with_multi_steering` of this package against the reference's
    own (oracle/_ref).  Vectors: north_star <= 1e-3 relative L2 — identical.  Generation: identical greedy tokens for
    32 steps — the returned strings (which carry every token id, `compat_slice` reproducing the reference's slicing)
    are identical, for each strength in turn (the reference takes one scalar strength per call)."""
    if not torch.cuda.is_available():
        pytest.skip("needs a GPU")
    from introspective_awareness_b200 import vector_utils as vu
    from introspective_awareness_b200.model_utils import ModelWrapper
    from introspective_awareness_b200.synthetic import synthetic_prompt_ids

    real_ref = _real_ref()
    torch.manual_seed(99)
    tiny = arch.startswith("tiny")
    ours = ModelWrapper.from_arch(arch, device="cuda:0", dtype=torch.bfloat16, seed=21 if tiny else None)
    V = ours.model.config.vocab_size
    ours.tokenizer = real_ref.IdTokenizer(vocab_size=V)
    ours.compat_slice = True
    ref = real_ref.wrap(ours.model, "cuda:0", torch.bfloat16)
    c_ids, c_mask, _ = synthetic_prompt_ids(10, 20, 8, V, seed=2)
    b_ids, b_mask, _ = synthetic_prompt_ids(25, 20, 8, V, seed=3)
    cw, bw = real_ref.ids_to_prompts(c_ids, c_mask), real_ref.ids_to_prompts(b_ids, b_mask)
    want = real_ref.concept_vectors(ref, c_ids, c_mask, b_ids, b_mask, layer)
    got = vu.extract_concept_vectors_batch(ours, cw, bw, layer_idx=layer, template="{word}")
    got = torch.stack([got[k] for k in cw])
    rel = ((got.float() - want.float()).norm(dim=1) / want.float().norm(dim=1)).max().item()
    assert rel <= 1e-3, rel
    assert torch.equal(got, want)
    ids, mask, lens = synthetic_prompt_ids(rows, 64, 48, V, seed=1)
    prompts = real_ref.ids_to_prompts(ids, mask)
    t0 = (lens - 20).tolist()
    vectors = [want[i % 10] for i in range(rows)]
    for strength in (2.0, 8.0):
        theirs = real_ref.steered_generate(ref, ids, mask, layer, vectors, strength, t0, 32)
        mine = ours.generate_batch_with_multi_steering(prompts, layer, vectors, strength=strength, max_new_tokens=32,
                                                       temperature=0.0, steering_start_positions=t0)
        assert mine == theirs, strength
    # per-row strengths in ONE batch of ours == one reference call per strength
    alphas = [[2.0, 8.0][i % 2] for i in range(rows)]
    mine = ours.generate_batch_with_multi_steering(prompts, layer, vectors, strength=alphas, max_new_tokens=32, temperature=0.0,
                                                   steering_start_positions=t0)
    by_strength = {a: real_ref.steered_generate(ref, ids, mask, layer, vectors, a, t0, 32) for a in (2.0, 8.0)}
    assert mine == [by_strength[a][i] for i, a in enumerate(alphas)]
