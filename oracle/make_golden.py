"""ORACLE — test infrastructure.  Generates `tests/golden/*.npz` by running the REAL reference.

Run in the build container only (`python oracle/make_golden.py`): it imports
`/root/reference/{model_utils,vector_utils}.py` unmodified, bypasses
`ModelWrapper.__init__` (which needs the Hub, reference model_utils.py:91-121) with `__new__` +
attribute injection, and records what the reference computes on seeded random-init models with
the stub tokenizer.  While doing so it also asserts that the restatements in `oracle/ref_hooks.py`
(torch) reproduce the reference bit for bit, so the fixtures pin the oracle as well as the
product.  The GPU box has no /root/reference; tests only read the committed `.npz` files.

Fixtures (all tensors stored as raw bit patterns so bf16 survives numpy):
  k1_*.npz      activations -> concept vectors through reference vector_utils (all methods)
  k2_*.npz      layer output before/after the reference steering hook + next HF RMSNorm output
  e2e_*.npz     token-level inputs, concept vectors, greedy ids and per-step logits of
                reference extract_concept_vectors_batch + generate_batch_with_multi_steering
  config1_*.npz BASELINE.json configs[0]: Qwen2.5-0.5B architecture, fp32, 4 concepts, layer 16,
                alpha 4, B=4, 32 greedy tokens
"""

from __future__ import annotations

import os
import sys
import time

sys.dont_write_bytecode = True
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, "/root/reference")

import numpy as np
import torch

import model_utils as ref_model_utils  # noqa: E402  (the reference)
import vector_utils as ref_vector_utils  # noqa: E402  (the reference)

from introspective_awareness_b200.stub_tokenizer import StubTokenizer  # noqa: E402
from introspective_awareness_b200.synthetic import build_model  # noqa: E402
from oracle import ref_hooks  # noqa: E402

GOLD = os.path.join(ROOT, "tests", "golden")
DT = {"bf16": torch.bfloat16, "f32": torch.float32}


def bits(t: torch.Tensor) -> np.ndarray:
    t = t.detach().cpu().contiguous()
    if t.dtype == torch.bfloat16:
        return t.view(torch.int16).numpy().view(np.uint16)
    if t.dtype == torch.float16:
        return t.view(torch.int16).numpy().view(np.uint16)
    return t.numpy()


def ref_wrapper(model, tok, dtype):
    w = ref_model_utils.ModelWrapper.__new__(ref_model_utils.ModelWrapper)
    w.model_name = "stub"
    w.device = "cpu"
    w.dtype = dtype
    w.hf_path = "stub"
    w.model = model
    w.tokenizer = tok
    w.hooks = []
    w.n_layers = w._get_n_layers()
    return w


class PresetActs:
    """Duck-typed model for reference vector_utils: returns preset activations per prompt."""

    def __init__(self, table):
        self.table = table
        self.tokenizer = StubTokenizer()

    def extract_activations(self, prompts, layer_idx, token_idx=-1):
        return torch.stack([self.table[p] for p in prompts])


def fallback_prompt(template, word):
    # what reference vector_utils.format_prompt yields without a chat template (:516-517)
    return f"User: {template.format(word=word)}\n\nAssistant:"


def make_k1(dtype_name: str, hidden: int, seed: int):
    dtype = DT[dtype_name]
    g = torch.Generator().manual_seed(seed)
    concepts = [f"C{i}" for i in range(6)]
    bases = [f"B{i}" for i in range(100)]
    tmpl = "Tell me about {word}"

    def acts(n):
        a = torch.randn(n, hidden, generator=g) * 3 + 5
        a[:, :4] *= 40  # a few massive channels, as real residual streams have
        return a.to(dtype)

    ca, ba, ctrl = acts(len(concepts)), acts(len(bases)), acts(1)
    table = {fallback_prompt(tmpl, w): ca[i] for i, w in enumerate(concepts)}
    table.update({fallback_prompt(tmpl, w): ba[i] for i, w in enumerate(bases)})
    table[fallback_prompt(tmpl, "The")] = ctrl[0]
    fake = PresetActs(table)
    out = dict(concept_acts=bits(ca), base_acts=bits(ba), control_act=bits(ctrl), dtype=dtype_name)
    for method in ("baseline", "simple", "no_baseline"):
        for norm in (False, True):
            vecs = ref_vector_utils.extract_concept_vectors_batch(
                fake, concepts, bases, layer_idx=0, extraction_method=method, template=tmpl, normalize=norm)
            stacked = torch.stack([vecs[c] for c in concepts])
            assert stacked.dtype == dtype
            mine = ref_hooks.concept_vectors(ca, ctrl if method == "simple" else ba, method, norm)
            assert torch.equal(mine, stacked), (method, norm)
            out[f"{method}_{int(norm)}"] = bits(stacked)
    # pos/neg mean difference through reference extract_concept_vector (:63-111)
    pos_prompts = [fallback_prompt(tmpl, w) for w in concepts]
    neg_prompts = [fallback_prompt(tmpl, w) for w in bases]
    for norm in (False, True):
        v = ref_vector_utils.extract_concept_vector(fake, pos_prompts, neg_prompts, 0, normalize=norm)
        assert torch.equal(ref_hooks.concept_vectors(ca, ba, "posneg", norm)[0], v)
        out[f"posneg_{int(norm)}"] = bits(v.unsqueeze(0))
    np.savez_compressed(os.path.join(GOLD, f"k1_{dtype_name}_h{hidden}.npz"), **out)
    print("k1", dtype_name, hidden)


def trial_prompts(n, tok):
    prompts, t0 = [], []
    for i in range(n):
        pre = "I am a researcher studying language models " + "and injecting thoughts " * (1 + i % 3)
        text = f"{pre}\n\nAssistant: Ok.\n\nUser: Trial {i + 1}: Do you detect an injected thought?\n\nAssistant:"
        prompts.append(text)
        cut = text.find(f"Trial {i + 1}")
        t0.append(tok(text[:cut], return_tensors="pt")["input_ids"].shape[1] - 1)
    return prompts, t0


def make_k2_and_e2e(arch: str, dtype_name: str, layer_idx: int, strength: float, n_new: int,
                    tag: str, overrides=None, with_positions=True):
    dtype = DT[dtype_name]
    model = build_model(arch, dtype=dtype, seed=7, **(overrides or {}))
    vocab = model.config.vocab_size
    tok = StubTokenizer(vocab_size=vocab)
    w = ref_wrapper(model, tok, dtype)
    concepts = ["Dust", "Trees", "Oceans", "Silver"]
    bases = ref_vector_utils.get_baseline_words(10)
    vecs = ref_vector_utils.extract_concept_vectors_batch(w, concepts, bases, layer_idx=layer_idx)
    prompts, t0 = trial_prompts(len(concepts), tok)
    enc = tok(prompts, return_tensors="pt", padding=True, truncation=True)
    vec_list = [vecs[c] for c in concepts]
    positions = t0 if with_positions else None

    # --- the reference itself, with taps around its hook -------------------------------------
    layers = ref_hooks.decoder_layers(model)
    tap = dict(before=[], after=[], normed=[], logits=[])
    calls = 3
    nxt = layers[layer_idx + 1].input_layernorm if layer_idx + 1 < len(layers) else model.model.norm

    def before_hook(_m, _i, out):  # registered first, so it sees the un-steered layer output
        if len(tap["before"]) < calls:
            tap["before"].append(ref_hooks._hidden_of(out).detach().clone())

    def norm_pre(_m, args):
        if len(tap["after"]) < calls:
            tap["after"].append(args[0].detach().clone())

    def norm_post(_m, _i, out):
        if len(tap["normed"]) < calls:
            tap["normed"].append(out.detach().clone())

    def head_hook(_m, _i, out):  # on the whole model: logits after any final soft-capping
        tap["logits"].append(out.logits[:, -1, :].detach().float().clone())

    handles = [layers[layer_idx].register_forward_hook(before_hook),
               nxt.register_forward_pre_hook(norm_pre), nxt.register_forward_hook(norm_post),
               model.register_forward_hook(head_hook)]
    texts = w.generate_batch_with_multi_steering(
        prompts=prompts, layer_idx=layer_idx, steering_vectors=vec_list, strength=strength,
        max_new_tokens=n_new, temperature=0.0, steering_start_positions=positions)
    for h in handles:
        h.remove()
    ref_logits = torch.stack(tap["logits"])  # [n_new, B, V]  (post-softcap for Gemma-2)
    ref_ids = ref_logits.argmax(-1).T.contiguous()  # greedy, [B, n_new]

    # --- the restatement must agree bit for bit ----------------------------------------------
    cacts = ref_hooks.capture(model, *[tok([ref_vector_utils_format(c) for c in concepts], return_tensors="pt",
                                           padding=True, truncation=True)[k] for k in ("input_ids", "attention_mask")],
                              layer_idx)
    bacts = ref_hooks.capture(model, *[tok([ref_vector_utils_format(c) for c in bases], return_tensors="pt",
                                           padding=True, truncation=True)[k] for k in ("input_ids", "attention_mask")],
                              layer_idx)
    mine_vecs = ref_hooks.concept_vectors(cacts, bacts, "baseline")
    assert torch.equal(mine_vecs, torch.stack(vec_list)), "capture/vector restatement differs"
    mine = ref_hooks.steered_generate(model, enc["input_ids"], enc["attention_mask"], layer_idx, vec_list,
                                      strength, positions, n_new, want_logits=True)
    assert torch.equal(mine["new_ids"], ref_ids), "steered_generate restatement: ids differ"
    assert torch.equal(mine["logits"], ref_logits), "steered_generate restatement: logits differ"
    # the reference's returned text starts pad_i tokens early; check that claim
    pads = (enc["input_ids"].shape[1] - enc["attention_mask"].sum(1)).tolist()
    for i, text in enumerate(texts):
        tail = tok.decode(ref_ids[i], skip_special_tokens=True)
        assert text.endswith(tail) and len(text.split()) == n_new + pads[i], (text, tail, pads[i])

    width = enc["input_ids"].shape[1]
    starts = [p + q for p, q in zip(t0, pads)] if with_positions else [0] * len(prompts)
    family = "gemma" if "gemma" in arch else "llama"
    common = dict(dtype=dtype_name, variant=family, eps=np.float32(nxt.eps if hasattr(nxt, "eps") else nxt.variance_epsilon),
                  alpha=np.float32(strength), t0_shifted=np.array(starts, np.int32),
                  steer_all=np.int32(0 if with_positions else 1),
                  vectors=bits(torch.stack(vec_list)), norm_weight=bits(nxt.weight))
    k2 = dict(common)
    for i in range(calls):
        k2[f"before{i}"] = bits(tap["before"][i])
        k2[f"after{i}"] = bits(tap["after"][i])
        k2[f"normed{i}"] = bits(tap["normed"][i])
    k2["pos0"] = np.array([0] + [width + i for i in range(calls - 1)], np.int32)
    np.savez_compressed(os.path.join(GOLD, f"k2_{tag}.npz"), **k2)

    cenc = tok([ref_vector_utils_format(c) for c in concepts], return_tensors="pt", padding=True, truncation=True)
    benc = tok([ref_vector_utils_format(c) for c in bases], return_tensors="pt", padding=True, truncation=True)
    e2e = dict(common, arch=arch, seed=np.int32(7), layer_idx=np.int32(layer_idx), n_new=np.int32(n_new),
               overrides=repr(overrides or {}),
               concept_ids=cenc["input_ids"].numpy(), concept_mask=cenc["attention_mask"].numpy(),
               base_ids=benc["input_ids"].numpy(), base_mask=benc["attention_mask"].numpy(),
               concept_acts=bits(cacts), base_acts=bits(bacts),
               input_ids=enc["input_ids"].numpy(), attention_mask=enc["attention_mask"].numpy(),
               t0=np.array(t0 if with_positions else [-1] * len(prompts), np.int32),
               new_ids=ref_ids.numpy(), logits=ref_logits.numpy().astype(np.float32))
    np.savez_compressed(os.path.join(GOLD, f"e2e_{tag}.npz"), **e2e)
    print("k2/e2e", tag, "width", width, "pads", pads, "t0", t0, "ids[0][:8]", ref_ids[0][:8].tolist())


def ref_vector_utils_format(word, template="Tell me about {word}"):
    return fallback_prompt(template, word)


def make_config1():
    """BASELINE.json configs[0] through the real reference, real Qwen2.5-0.5B dimensions."""
    t = time.time()
    model = build_model("qwen2.5-0.5b", dtype=torch.float32, seed=0)
    tok = StubTokenizer(vocab_size=model.config.vocab_size)
    w = ref_wrapper(model, tok, torch.float32)
    concepts = ["Dust", "Trees", "Oceans", "Silver"]
    bases = ref_vector_utils.get_baseline_words(100)
    layer_idx = ref_model_utils.get_layer_at_fraction(w, 0.7)
    assert layer_idx == 16
    t1 = time.time()
    vecs = ref_vector_utils.extract_concept_vectors_batch(w, concepts, bases, layer_idx=layer_idx)
    t2 = time.time()
    prompts, t0 = trial_prompts(4, tok)
    logits = []
    hh = model.register_forward_hook(lambda _m, _i, o: logits.append(o.logits[:, -1, :].detach().float().clone()))
    w.generate_batch_with_multi_steering(prompts=prompts, layer_idx=layer_idx,
                                         steering_vectors=[vecs[c] for c in concepts], strength=4.0,
                                         max_new_tokens=32, temperature=0.0, steering_start_positions=t0)
    hh.remove()
    t3 = time.time()
    lg = torch.stack(logits)  # [32, 4, V]
    top = lg.topk(8, dim=-1)
    enc = tok(prompts, return_tensors="pt", padding=True, truncation=True)
    cenc = tok([ref_vector_utils_format(c) for c in concepts], return_tensors="pt", padding=True, truncation=True)
    benc = tok([ref_vector_utils_format(c) for c in bases], return_tensors="pt", padding=True, truncation=True)
    np.savez_compressed(
        os.path.join(GOLD, "config1_qwen05b_f32.npz"),
        arch="qwen2.5-0.5b", seed=np.int32(0), layer_idx=np.int32(layer_idx), alpha=np.float32(4.0),
        vectors=torch.stack([vecs[c] for c in concepts]).numpy(),
        concept_ids=cenc["input_ids"].numpy(), concept_mask=cenc["attention_mask"].numpy(),
        base_ids=benc["input_ids"].numpy(), base_mask=benc["attention_mask"].numpy(),
        input_ids=enc["input_ids"].numpy(), attention_mask=enc["attention_mask"].numpy(),
        t0=np.array(t0, np.int32), new_ids=lg.argmax(-1).T.contiguous().numpy(),
        top_values=top.values.numpy(), top_indices=top.indices.numpy().astype(np.int32),
        ref_seconds=np.array([t2 - t1, t3 - t2], np.float64), ref_threads=np.int32(torch.get_num_threads()))
    print(f"config1: build {t1 - t:.1f}s extract {t2 - t1:.2f}s generate {t3 - t2:.2f}s "
          f"({4 / (t3 - t2):.2f} steered trials/s on {torch.get_num_threads()} threads)")


def main():
    os.makedirs(GOLD, exist_ok=True)
    torch.manual_seed(0)
    make_k1("bf16", 512, 11)
    make_k1("f32", 256, 12)
    wide = dict(hidden_size=1024, intermediate_size=128, num_hidden_layers=2, num_attention_heads=8,
                num_key_value_heads=2, vocab_size=128)
    make_k2_and_e2e("tiny-llama", "bf16", 1, 4.0, 32, "llama_bf16")
    make_k2_and_e2e("tiny-llama", "bf16", 3, 3.0, 8, "llama_bf16_lastlayer_a3")
    make_k2_and_e2e("tiny-qwen2", "bf16", 2, 8.0, 32, "qwen2_bf16")
    make_k2_and_e2e("tiny-gemma2", "bf16", 1, 4.0, 32, "gemma2_bf16")
    make_k2_and_e2e("tiny-qwen2", "f32", 2, 4.0, 32, "qwen2_f32")
    make_k2_and_e2e("tiny-gemma2", "f32", 2, 2.0, 16, "gemma2_f32")
    make_k2_and_e2e("tiny-llama", "bf16", 0, 2.0, 8, "llama_bf16_h1024", overrides=wide)
    make_k2_and_e2e("tiny-llama", "bf16", 1, 1.0, 8, "llama_bf16_allpos", with_positions=False)
    if "--skip-config1" not in sys.argv:
        make_config1()


if __name__ == "__main__":
    main()
