"""Drop-in check against the REAL reference driver (build container only; skipped where
/root/reference is absent, e.g. on the GPU box): with `model_utils` / `vector_utils` /
`steering_utils` pointed at this package (level 1), `detect_injected_thoughts.py`
imports cleanly — every name it pulls from those modules exists — and its own helpers keep working."""

import importlib.util
import os
import sys
import types

import pytest

REF = "/root/reference"


@pytest.fixture(autouse=True)
def _no_reference_modules_left_behind():
    yield
    for name in ("model_utils", "vector_utils", "steering_utils", "eval_utils"):
        mod = sys.modules.get(name)
        if mod is not None and str(getattr(mod, "__file__", "")).startswith(REF):
            del sys.modules[name]


@pytest.mark.skipif(not os.path.exists(os.path.join(REF, "detect_injected_thoughts.py")), reason="reference not mounted")
def test_reference_driver_imports_against_this_package(monkeypatch):
    import introspective_awareness_b200 as ia

    monkeypatch.setattr(sys, "dont_write_bytecode", True)
    # third-party modules the driver needs only for plotting / the OpenAI judge (out of scope, not installed)
    for name in ("matplotlib", "matplotlib.pyplot", "seaborn", "openai", "dotenv"):
        if name not in sys.modules:
            try:
                __import__(name)
            except Exception:
                stub = types.ModuleType(name)
                stub.use = lambda *a, **k: None
                stub.load_dotenv = lambda *a, **k: None
                stub.AsyncOpenAI = stub.OpenAI = object
                monkeypatch.setitem(sys.modules, name, stub)
    if "matplotlib" in sys.modules and "matplotlib.pyplot" in sys.modules:
        sys.modules["matplotlib"].pyplot = sys.modules["matplotlib.pyplot"]
    monkeypatch.setitem(sys.modules, "model_utils", ia.model_utils)
    monkeypatch.setitem(sys.modules, "vector_utils", ia.vector_utils)
    monkeypatch.setitem(sys.modules, "steering_utils", ia.steering_utils)
    monkeypatch.syspath_prepend(REF)  # for the reference's own eval_utils
    spec = importlib.util.spec_from_file_location("ref_driver", os.path.join(REF, "detect_injected_thoughts.py"))
    drv = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(drv)
    # the driver is now bound to OUR implementations
    assert drv.extract_concept_vectors_batch is ia.vector_utils.extract_concept_vectors_batch
    assert drv.load_model is ia.model_utils.load_model and drv.ModelWrapper is ia.model_utils.ModelWrapper
    assert drv.run_forced_noticing_test_batch is ia.steering_utils.run_forced_noticing_test_batch
    assert drv.check_concept_mentioned("I sense dust", "Dust") is True
    # constants this package's defaults were taken from
    assert len(drv.DEFAULT_TEST_CONCEPTS) == 50 and drv.DEFAULT_STRENGTH_SWEEP == [1.0, 2.0, 4.0, 8.0]
    assert drv.get_baseline_words(drv.DEFAULT_N_BASELINE) == ia.vector_utils.get_baseline_words(100)
    msgs = [{"role": "system", "content": ""}, {"role": "user", "content": "x"}]
    assert drv.filter_messages_for_model(msgs, "gemma2_9b") == msgs[1:]
    # our grid applies the same filter, from the same list, before templating
    from introspective_awareness_b200 import grid

    assert grid.MODELS_WITHOUT_SYSTEM_ROLE == drv.MODELS_WITHOUT_SYSTEM_ROLE
    for name in ("gemma2_9b", "llama_8b", "gemma3_27b", "qwen_32b"):
        assert grid.filter_messages_for_model(msgs, name) == drv.filter_messages_for_model(msgs, name)
    sys.modules.pop("eval_utils", None)


def _load_ref(name):
    """A reference module under a private name (with the reference's own model_utils as its dependency)."""
    sys.dont_write_bytecode = True
    spec = importlib.util.spec_from_file_location(f"_ref_{name}", os.path.join(REF, f"{name}.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


@pytest.mark.skipif(not os.path.exists(os.path.join(REF, "steering_utils.py")), reason="reference not mounted")
def test_prompts_t0_and_word_lists_equal_the_reference(monkeypatch):
    """Host-side parity with the real reference: identical word lists and concept pairs, and for every
    runner the same prompt text and steering start position reach the model."""
    import torch

    import introspective_awareness_b200 as ia
    from introspective_awareness_b200.stub_tokenizer import StubTokenizer

    monkeypatch.syspath_prepend(REF)
    for k in ("model_utils", "vector_utils", "steering_utils"):
        monkeypatch.delitem(sys.modules, k, raising=False)
    ref_vu, ref_su = _load_ref("vector_utils"), _load_ref("steering_utils")
    assert ia.vector_utils.DEFAULT_BASELINE_WORDS == ref_vu.DEFAULT_BASELINE_WORDS
    assert ia.vector_utils.CONCEPT_PAIRS == ref_vu.CONCEPT_PAIRS

    class Spy:
        """Duck-typed model that records what the runners hand to the generate methods."""
        def __init__(self, chat):
            self.tokenizer = StubTokenizer(vocab_size=4096, chat_template=chat)
            self.model_name = "spy"
            self.calls = []

        def generate_with_steering(self, **kw):
            self.calls.append(("steer", kw["prompt"], kw["steering_start_pos"], kw["layer_idx"], kw["strength"]))
            return "r"

        def generate(self, **kw):
            self.calls.append(("plain", kw["prompt"]))
            return "r"

        def generate_batch(self, **kw):
            self.calls.append(("plain_batch", tuple(kw["prompts"])))
            return ["r"] * len(kw["prompts"])

    v = torch.zeros(8)
    for chat in (False, True):
        a, b = Spy(chat), Spy(chat)
        for su, spy in ((ref_su, a), (ia.steering_utils, b)):
            su.run_steered_introspection_test(spy, "Dust", v, 3, strength=4.0, trial_number=7)
            su.run_forced_noticing_test(spy, "Dust", v, 3, strength=2.0, trial_number=2)
            su.run_unsteered_introspection_test(spy, "Dust", trial_number=5)
            su.run_unsteered_introspection_test_batch(spy, "Dust", trial_numbers=[1, 12])
        assert a.calls == b.calls, chat
        assert a.calls[0][2] is not None and a.calls[0][2] > 0
    # prompt factories
    for args in (("Dust", 3, True, False), ("Dust", 1, True, True), ("Dust", 1, False, True)):
        p, q = ref_su.create_introspection_test_prompt(*args), ia.steering_utils.create_introspection_test_prompt(*args)
        assert (p.system_prompt, p.user_prompt, p.prefill) == (q.system_prompt, q.user_prompt, q.prefill)
        tok = StubTokenizer()
        assert p.format_for_model(tok) == q.format_for_model(tok)
    for fn in ("create_style_detection_prompt", "create_abstract_concept_prompt"):
        p, q = getattr(ref_su, fn)("writing in all caps", 4), getattr(ia.steering_utils, fn)("writing in all caps", 4)
        assert (p.system_prompt, p.user_prompt, p.prefill) == (q.system_prompt, q.user_prompt, q.prefill)
    for text, word in (("I notice dust", "Dust"), ("trees!", "Tree"), ("a box", "Boxes"), ("industry", "dust"), ("Foxes run", "fox")):
        assert ref_su.check_concept_mentioned(text, word) == ia.steering_utils.check_concept_mentioned(text, word)
    for resp in ("Yes, I do", "No.", "maybe yes no", "yes yes no", ""):
        assert ref_su.extract_yes_no_answer(resp) == ia.steering_utils.extract_yes_no_answer(resp)
