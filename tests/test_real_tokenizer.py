"""SURVEY.md §8(f)3, the part that can be done offline: the host side of the path driven by a REAL Hugging Face
fast tokenizer with a real Jinja chat template (tests/real_tokenizer.py: Llama-3 style with a BOS token, header
tokens and a system turn; Gemma style with no system role), not the stub tokenizer.

With the reference mounted (build container): the real reference `steering_utils` / `vector_utils` and the real
sweep driver against ours on the same scripted model — identical templated prompts through the
`apply_chat_template` branches (steering_utils.py:255-268, vector_utils.py:506-517, the driver's :1833-1840 /
:2004-2017), identical t0 with a BOS token present (steering_utils.py:270-287), the system turn dropped for Gemma
(`filter_messages_for_model`, detect_injected_thoughts.py:84-99).  Everywhere: structural checks of t0, left
padding with pad = eos (model_utils.py:97) and the pad shift (model_utils.py:822-825); on the GPU an end-to-end
steered generation with that tokenizer against the hook path.
"""

import os
import sys

import pytest
import torch

from real_tokenizer import build_fast_tokenizer

REF = "/root/reference"
HAVE_REF = os.path.exists(os.path.join(REF, "steering_utils.py"))


class _Holder:
    """The least a prompt builder needs from a model."""

    def __init__(self, tokenizer, name="llama_8b"):
        self.tokenizer, self.model_name = tokenizer, name


@pytest.mark.parametrize("style", ["llama3", "gemma"])
def test_tokenizer_behaves_like_the_registry_models(style):
    tok = build_fast_tokenizer(style)
    from introspective_awareness_b200.model_utils import ModelWrapper  # noqa: F401  (sets padding side / pad token)

    tok.padding_side = "left"
    if tok.pad_token is None:
        tok.pad_token = tok.eos_token
    enc = tok(["Tell me about Dust", "Tell me about the Oceans and the Trees"], return_tensors="pt", padding=True, truncation=True)
    ids, mask = enc["input_ids"], enc["attention_mask"]
    assert mask[0, 0] == 0 and mask[1, 0] == 1 and ids[0, 0] == tok.pad_token_id          # left padding
    assert ids[1, 0] == tok.bos_token_id and ids[0][mask[0].bool()][0] == tok.bos_token_id  # BOS on every row
    assert tok.decode(ids[0], skip_special_tokens=True) == "Tell me about Dust"
    msgs = [{"role": "system", "content": ""}, {"role": "user", "content": "hello"}]
    if style == "gemma":
        with pytest.raises(Exception, match="System role not supported"):
            tok.apply_chat_template(msgs, tokenize=False, add_generation_prompt=True)
        text = tok.apply_chat_template(msgs[1:], tokenize=False, add_generation_prompt=True)
        assert text == "<bos><start_of_turn>user\nhello<end_of_turn>\n<start_of_turn>model\n"
    else:
        text = tok.apply_chat_template(msgs, tokenize=False, add_generation_prompt=True)
        assert text.startswith("<|begin_of_text|><|start_header_id|>system<|end_header_id|>\n\n<|eot_id|>")
        assert text.endswith("<|start_header_id|>assistant<|end_header_id|>\n\n")


@pytest.mark.parametrize("style,name", [("llama3", "llama_8b"), ("gemma", "gemma2_9b")])
def test_t0_is_the_token_before_the_trial_marker(style, name):
    """t0 = len(tokens of the text before "Trial N") - 1 (reference steering_utils.py:270-287), counted with the BOS
    the tokenizer prepends: position t0 of the tokenised prompt is the last token before the marker, and the
    marker's first token sits at t0 + 1 — whatever the template puts in front."""
    from introspective_awareness_b200 import grid, steering_utils as su

    tok = build_fast_tokenizer(style)
    m = _Holder(tok, name)
    for n, forced in ((1, False), (12, False), (27, True)):
        prompt = grid.driver_prompt(m, n, forced)
        t0 = su.steering_start_position(m, prompt, n, missing=0)
        full = tok(prompt, return_tensors="pt")["input_ids"][0]
        before = tok(prompt[: prompt.find(f"Trial {n}")], return_tensors="pt")["input_ids"][0]
        assert t0 == len(before) - 1 and t0 > 5
        assert torch.equal(full[: t0 + 1], before)                      # the prefix tokenises the same inside the prompt
        assert tok.decode(full[t0 + 1: t0 + 4]).startswith("Trial")     # the marker starts right after t0
        assert full[0] == tok.bos_token_id
        if style == "gemma":
            assert "system" not in prompt                               # the system turn was filtered out
        if forced:
            assert prompt.endswith(grid.DRIVER_FORCED_PREFILL) and "What injected thought do you notice?" in prompt


def _load_ref(name):
    import importlib.util

    sys.dont_write_bytecode = True
    spec = importlib.util.spec_from_file_location(f"_ref_{name}", os.path.join(REF, f"{name}.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


@pytest.fixture
def _ref_modules(monkeypatch):
    monkeypatch.syspath_prepend(REF)
    for k in ("model_utils", "vector_utils", "steering_utils"):
        monkeypatch.delitem(sys.modules, k, raising=False)
    yield _load_ref("vector_utils"), _load_ref("steering_utils")
    for name in ("model_utils", "vector_utils", "steering_utils", "eval_utils"):
        mod = sys.modules.get(name)
        if mod is not None and str(getattr(mod, "__file__", "")).startswith(REF):
            del sys.modules[name]


@pytest.mark.skipif(not HAVE_REF, reason="reference not mounted")
def test_runners_and_extraction_prompts_equal_the_reference_with_a_real_chat_template(_ref_modules):
    import introspective_awareness_b200 as ia

    ref_vu, ref_su = _ref_modules

    class Spy:
        def __init__(self, tok):
            self.tokenizer, self.model_name, self.calls = tok, "llama_8b", []

        def generate_with_steering(self, **kw):
            self.calls.append(("steer", kw["prompt"], kw["steering_start_pos"], kw["layer_idx"], kw["strength"]))
            return "r"

        def generate(self, **kw):
            self.calls.append(("plain", kw["prompt"]))
            return "r"

        def generate_batch(self, **kw):
            self.calls.append(("plain_batch", tuple(kw["prompts"])))
            return ["r"] * len(kw["prompts"])

        # reference vector_utils surface
        def extract_activations(self, prompts, layer_idx, token_idx=-1):
            self.calls.append(("acts", tuple(prompts), layer_idx, token_idx))
            return torch.zeros(len(prompts), 4)

        # this package's duck type for the same request
        def extract_concept_vectors(self, concept_prompts, baseline_prompts, layer_indices, method="baseline", normalize=False,
                                    token_idx=-1, **kw):
            for layer in layer_indices:
                self.calls.append(("acts", tuple(concept_prompts), layer, token_idx))
                if baseline_prompts is not None:
                    self.calls.append(("acts", tuple(baseline_prompts), layer, token_idx))
            n_out = len(concept_prompts)
            return torch.zeros(len(layer_indices), n_out, 4)

    tok = build_fast_tokenizer("llama3")
    v = torch.zeros(4)
    a, b = Spy(tok), Spy(tok)
    for su, vu, spy in ((ref_su, ref_vu, a), (ia.steering_utils, ia.vector_utils, b)):
        su.run_steered_introspection_test(spy, "Dust", v, 3, strength=4.0, trial_number=7)
        su.run_forced_noticing_test(spy, "Dust", v, 3, strength=2.0, trial_number=12)
        su.run_unsteered_introspection_test(spy, "Dust", trial_number=5)
        su.run_unsteered_introspection_test_batch(spy, "Dust", trial_numbers=[1, 12])
        vu.extract_concept_vectors_batch(spy, ["Dust", "Trees"], vu.get_baseline_words(5), layer_idx=2)
        vu.extract_concept_vectors_batch(spy, ["Dust"], vu.get_baseline_words(5), layer_idx=1, extraction_method="simple")
        vu.extract_concept_vectors_batch(spy, ["Dust"], [], layer_idx=1, extraction_method="no_baseline", template="{word}")
    gen = lambda calls: [c for c in calls if c[0] != "acts"]
    acts = lambda calls: sorted(c for c in calls if c[0] == "acts")  # "simple" asks for the control prompt first there, last here
    assert gen(a.calls) == gen(b.calls)
    assert acts(a.calls) == acts(b.calls)
    steer = a.calls[0]
    assert steer[1].startswith("<|begin_of_text|><|start_header_id|>system") and steer[2] > 5  # templated, t0 past the explanation
    assert any(c[1][0].endswith("Tell me about Dust<|eot_id|><|start_header_id|>assistant<|end_header_id|>\n\n") for c in acts(a.calls))
    assert any(c[1] == ("<|begin_of_text|><|start_header_id|>user<|end_header_id|>\n\nDust<|eot_id|><|start_header_id|>assistant<|end_header_id|>\n\n",)
               for c in acts(a.calls))  # template="{word}"
    # Gemma's template rejects the system turn the runners always send: same failure on both sides (the driver
    # filters the turn out itself, see the sweep test below)
    gtok = build_fast_tokenizer("gemma")
    for su in (ref_su, ia.steering_utils):
        with pytest.raises(Exception, match="System role not supported"):
            su.run_steered_introspection_test(Spy(gtok), "Dust", v, 3, strength=4.0, trial_number=7)


@pytest.mark.skipif(not HAVE_REF, reason="reference not mounted")
@pytest.mark.parametrize("style,name", [("llama3", "llama_8b"), ("gemma", "gemma2_9b")])
def test_sweep_equals_the_real_drivers_sweep_with_a_real_tokenizer(monkeypatch, tmp_path, style, name):
    """The whole sweep comparison of tests/test_reference_sweep_dropin.py — the real driver's main() against
    run_sweep on a scripted model — with the real tokenizer and chat template in place of the stub."""
    import test_reference_sweep_dropin as T

    tok = build_fast_tokenizer(style)

    class RealTokModel(T.ScriptedModel):
        def __init__(self, chat, model_name):
            super().__init__(chat, model_name)
            self.tokenizer = tok

    monkeypatch.setattr(T, "ScriptedModel", RealTokModel)
    T.test_sweep_equals_the_real_drivers_sweep(monkeypatch, tmp_path, True, name, 5, 4)


@pytest.mark.gpu
@pytest.mark.parametrize("style,name", [("llama3", "llama_8b"), ("gemma", "gemma2_9b")])
def test_steered_generation_with_a_real_tokenizer_on_gpu(style, name):
    """End to end on the GPU with the real tokenizer: templated trial prompts (different trial numbers -> ragged
    lengths -> left padding), per-row t0 with the BOS counted, `generate_batch_with_multi_steering` against the
    hook path on the same token ids; text decoded with special tokens skipped and rows ended by the EOS token."""
    if not torch.cuda.is_available():
        pytest.skip("needs a GPU")
    from introspective_awareness_b200 import grid, steering_utils as su, vector_utils as vu
    from introspective_awareness_b200.model_utils import ModelWrapper
    from oracle import ref_hooks

    tok = build_fast_tokenizer(style)
    w = ModelWrapper.from_arch("tiny-llama", device="cuda:0", dtype=torch.bfloat16, seed=17, tokenizer=tok, vocab_size=len(tok))
    w.model_name = name
    assert tok.padding_side == "left" and tok.pad_token_id is not None
    w.model.generation_config.eos_token_id = [tok.eos_token_id, tok.convert_tokens_to_ids("<|end_of_text|>" if style == "llama3" else "<end_of_turn>")]
    w.model.generation_config.pad_token_id = tok.pad_token_id
    concepts = ["Dust", "Trees", "Oceans", "Silver"]
    vecs = vu.extract_concept_vectors_batch(w, concepts, vu.get_baseline_words(8), layer_idx=2)
    trials = [1, 7, 12, 30]
    prompts = [grid.driver_prompt(w, n, forced=(n == 30)) for n in trials]
    t0 = [su.steering_start_position(w, p, n, missing=0) for p, n in zip(prompts, trials)]
    texts = w.generate_batch_with_multi_steering(prompts, 2, [vecs[c] for c in concepts], strength=4.0, max_new_tokens=16,
                                                 temperature=0.0, steering_start_positions=t0)
    enc = tok(prompts, return_tensors="pt", padding=True, truncation=True)
    assert (enc["attention_mask"].sum(1) < enc["input_ids"].shape[1]).any()          # some rows are left-padded
    ref = ref_hooks.steered_generate(w.model, enc["input_ids"], enc["attention_mask"], 2, [vecs[c] for c in concepts], 4.0, t0, 16,
                                     pad_token_id=tok.pad_token_id)
    want = [tok.decode(r, skip_special_tokens=True).strip() for r in ref["new_ids"]]
    assert texts == want
    assert all("<|" not in t and "<start_of_turn>" not in t for t in texts)
