"""CPU tests of the host-side mirror of the reference interface: prompt building, t0, grid
scheduling, the unsteered runtime path (pure torch, no kernels) and the refusal of the steered /
extraction paths to run without CUDA."""

import inspect

import pytest
import torch

from introspective_awareness_b200 import grid, model_utils, steering_utils as su, vector_utils as vu
from introspective_awareness_b200.model_utils import ModelWrapper, get_layer_at_fraction
from introspective_awareness_b200.stub_tokenizer import StubTokenizer


@pytest.fixture(scope="module")
def wrapper():
    return ModelWrapper.from_arch("tiny-llama", device="cpu", dtype=torch.float32, seed=3)


def test_reference_call_surface_is_present():
    """Names, positional order and defaults the reference driver relies on."""
    def params(fn):
        return [(p.name, p.default) for p in inspect.signature(fn).parameters.values()]

    assert params(vu.extract_concept_vectors_batch) == [
        ("model", inspect._empty), ("concept_words", inspect._empty), ("baseline_words", inspect._empty),
        ("layer_idx", inspect._empty), ("extraction_method", "baseline"), ("template", "Tell me about {word}"),
        ("token_idx", -1), ("normalize", False)]
    assert params(vu.extract_concept_vector)[-2:] == [("token_idx", -1), ("normalize", False)]
    assert params(vu.extract_concept_vector_simple)[3:5] == [("control_prompt", "The"), ("template", "{word}")]
    assert params(ModelWrapper.generate_batch_with_multi_steering)[1:] == [
        ("prompts", inspect._empty), ("layer_idx", inspect._empty), ("steering_vectors", inspect._empty),
        ("strength", 1.0), ("max_new_tokens", 512), ("temperature", 0.0), ("steering_start_positions", None),
        ("debug", False), ("generation_kwargs", inspect._empty)]
    assert params(ModelWrapper.generate_with_steering)[4:8] == [
        ("strength", 1.0), ("max_new_tokens", 512), ("temperature", 0.0), ("steering_start_pos", None)]
    assert params(ModelWrapper.extract_activations)[1:] == [("prompts", inspect._empty), ("layer_idx", inspect._empty),
                                                            ("token_idx", -1)]
    assert params(su.run_steered_introspection_test)[4:] == [
        ("strength", 8.0), ("trial_number", 1), ("max_new_tokens", 100), ("temperature", 1.0), ("is_first_trial", False)]
    for name in ("run_steered_introspection_test", "run_unsteered_introspection_test",
                 "run_steered_introspection_test_batch", "run_unsteered_introspection_test_batch",
                 "run_forced_noticing_test", "run_forced_noticing_test_batch", "calculate_detection_accuracy",
                 "check_concept_mentioned", "IntrospectionPrompt", "create_introspection_test_prompt"):
        assert hasattr(su, name), name  # everything detect_injected_thoughts.py:44-49 imports
    for name in ("create_concept_prompts", "create_contrastive_pair", "cosine_similarity", "save_concept_vector",
                 "load_concept_vector", "DEFAULT_BASELINE_WORDS", "CONCEPT_PAIRS", "get_baseline_words",
                 "get_concept_pair", "analyze_vector_underspecification", "extract_concept_vector_with_baseline",
                 "extract_concept_vector_no_baseline"):
        assert hasattr(vu, name), name


def test_word_lists():
    words = vu.get_baseline_words()
    assert len(words) == 100 and words[0] == "Desks" and words[-1] == "Enigmas"
    assert words.count("Butterflies") == 2  # reference vector_utils.py:398 and :402
    assert vu.get_baseline_words(10) == words[:10]
    with pytest.raises(ValueError, match="Unknown concept pair"):
        vu.get_concept_pair("nope")
    assert vu.create_concept_prompts("Dust", ["Sand"]) == (["Tell me about Dust"], ["Tell me about Sand"])


def test_stub_tokenizer_prefix_property_and_left_padding():
    tok = StubTokenizer(vocab_size=512)
    text = "alpha beta\n\nUser: Trial 3: Do you detect?\n\nAssistant:"
    cut = text.find("Trial 3")
    full = tok(text, return_tensors="pt")["input_ids"][0].tolist()
    pre = tok(text[:cut], return_tensors="pt")["input_ids"][0].tolist()
    assert full[: len(pre)] == pre
    enc = tok(["a b c", "a"], return_tensors="pt", padding=True, truncation=True)
    assert enc["attention_mask"].tolist() == [[1, 1, 1], [0, 0, 1]]
    assert enc["input_ids"][1, :2].tolist() == [0, 0]
    assert not hasattr(tok, "apply_chat_template") and hasattr(StubTokenizer(chat_template=True), "apply_chat_template")


def test_trial_prompt_and_t0(wrapper):
    prompt, t0 = su.build_trial(wrapper, 7)
    # plain-text fallback layout of reference steering_utils.py:263-268
    assert prompt.startswith("I am an interpretability researcher") and prompt.endswith("\n\nAssistant:")
    assert "Assistant: Ok.\n\nUser: Trial 7: Do you detect an injected thought?" in prompt
    ids = wrapper.tokenizer(prompt, return_tensors="pt")["input_ids"][0]
    before = wrapper.tokenizer(prompt[: prompt.find("Trial 7")], return_tensors="pt")["input_ids"][0]
    assert t0 == len(before) - 1 and 0 < t0 < len(ids)
    forced, _ = su.build_trial(wrapper, 2, forced=True)
    assert "Please complete this sentence" in forced and "identify what I inject" in forced
    assert su.steering_start_position(wrapper, "no marker here", 1) is None
    assert su.steering_start_position(wrapper, "no marker here", 1, missing=0) == 0
    chat = ModelWrapper("x", device="cpu", dtype=torch.float32, model=wrapper.model,
                        tokenizer=StubTokenizer(vocab_size=512, chat_template=True))
    p2, t2 = su.build_trial(chat, 1)
    assert "<|user|>" in p2 and p2.endswith("<|assistant|>\n\n") and t2 is not None


def test_scoring_helpers():
    assert su.check_concept_mentioned("I notice dust everywhere", "Dust")
    assert su.check_concept_mentioned("many trees", "Tree") and su.check_concept_mentioned("a tree", "Trees")
    assert su.check_concept_mentioned("boxes!", "box") and not su.check_concept_mentioned("industry", "dust")
    assert su.extract_yes_no_answer("Yes, I do.") is True and su.extract_yes_no_answer("No. Nothing") is False
    assert su.extract_yes_no_answer("maybe") is None
    res = [{"injected": True, "detected": True}, {"injected": False, "detected": True},
           {"injected": True, "detected": None, "response": ""}]
    assert su.calculate_detection_accuracy(res) == 0.5
    assert su.calculate_false_positive_rate([{"injected": False, "response": "yes."}, {"injected": False, "response": "no."}]) == 0.5


def test_layer_fraction_and_registry(wrapper):
    assert wrapper.n_layers == 4
    assert [get_layer_at_fraction(wrapper, f) for f in (0.0, 0.4, 0.7, 0.99, 1.0, 2.0)] == [0, 1, 2, 3, 3, 3]
    assert wrapper.get_layer_module(2) is wrapper.model.model.layers[2]
    assert model_utils.MODEL_NAME_MAP["llama_8b"] == "meta-llama/Llama-3.1-8B-Instruct"


def test_unsteered_generation_matches_hf_generate(wrapper):
    """The runtime's own layer loop + static cache against stock `model.generate` (control trials,
    reference model_utils.py:508-560); padded batch, greedy."""
    prompts = ["one two three four five six", "seven eight", "nine ten eleven twelve"]
    enc = wrapper.tokenizer(prompts, return_tensors="pt", padding=True, truncation=True)
    with torch.no_grad():
        ref = wrapper.model.generate(**enc, max_new_tokens=9, do_sample=False, pad_token_id=0)
    texts = wrapper.generate_batch(prompts, max_new_tokens=9, temperature=0.0)
    width = enc["input_ids"].shape[1]
    want = [wrapper.tokenizer.decode(ref[i, width:], skip_special_tokens=True).strip() for i in range(3)]
    assert texts == want
    # reference string slicing: starts pad_i prompt tokens early
    wrapper.compat_slice = True
    try:
        compat = wrapper.generate_batch(prompts, max_new_tokens=9, temperature=0.0)
    finally:
        wrapper.compat_slice = False
    lens = enc["attention_mask"].sum(1).tolist()
    assert compat == [wrapper.tokenizer.decode(ref[i, lens[i]:], skip_special_tokens=True).strip() for i in range(3)]
    assert wrapper.generate(prompts[0], max_new_tokens=4) == wrapper.tokenizer.decode(
        wrapper.model.generate(**wrapper.tokenizer(prompts[0], return_tensors="pt"), max_new_tokens=4, do_sample=False,
                               pad_token_id=0)[0, 6:], skip_special_tokens=True)


def test_eos_stops_rows_like_hf(wrapper):
    enc = wrapper.tokenizer(["a b c d", "e f"], return_tensors="pt", padding=True, truncation=True)
    with torch.no_grad():
        free = wrapper.model.generate(**enc, max_new_tokens=6, do_sample=False, pad_token_id=0)
    eos = int(free[0, enc["input_ids"].shape[1] + 2])  # make the 3rd new token of row 0 the EOS
    wrapper.model.generation_config.eos_token_id = eos
    try:
        with torch.no_grad():
            ref = wrapper.model.generate(**enc, max_new_tokens=6, do_sample=False, pad_token_id=0)
        got = wrapper.steered_generate_ids(enc["input_ids"], enc["attention_mask"], None, None, None, None, None, 6)
    finally:
        wrapper.model.generation_config.eos_token_id = None
    new = ref[:, enc["input_ids"].shape[1]:]
    assert torch.equal(got[:, : new.shape[1]].cpu(), new)
    assert (got[:, new.shape[1]:] == 0).all()


def test_eos_list_second_id_fires_like_hf(wrapper):
    """generation_config.eos_token_id may be a list (Llama-3.x-Instruct: [128001, 128008, 128009], the turn ends
    with the LAST of them): a row must stop on any id of the set, as HF `generate` stops it."""
    enc = wrapper.tokenizer(["a b c d", "e f"], return_tensors="pt", padding=True, truncation=True)
    width = enc["input_ids"].shape[1]
    with torch.no_grad():
        free = wrapper.model.generate(**enc, max_new_tokens=8, do_sample=False, pad_token_id=0)
    fires = int(free[0, width + 2])
    never = int(max(set(range(2, wrapper.model.config.vocab_size)) - set(free.flatten().tolist())))
    wrapper.model.generation_config.eos_token_id = [never, fires]
    try:
        with torch.no_grad():
            ref = wrapper.model.generate(**enc, max_new_tokens=8, do_sample=False, pad_token_id=0)
        got = wrapper.steered_generate_ids(enc["input_ids"], enc["attention_mask"], None, None, None, None, None, 8)
    finally:
        wrapper.model.generation_config.eos_token_id = None
    new = ref[:, width:]
    assert torch.equal(got[:, : new.shape[1]].cpu(), new)
    assert (got[:, new.shape[1]:] == 0).all()
    assert (got[0, 3:] == 0).all() and int(got[0, 2]) == fires  # row 0 ended at its third token


def test_repetition_penalty_and_generation_config_defaults_like_hf(wrapper):
    """Qwen2.5-Instruct ships repetition_penalty 1.05 (it applies to greedy decoding too); whatever the reference's
    `generate` call does not pass comes from the checkpoint's generation_config (reference model_utils.py:828-847)."""
    enc = wrapper.tokenizer(["a b c d a b", "e f e f"], return_tensors="pt", padding=True, truncation=True)
    gc_ = wrapper.model.generation_config
    old = (gc_.repetition_penalty, gc_.top_p, gc_.top_k, gc_.do_sample, gc_.temperature)
    gc_.repetition_penalty = 1.8
    try:
        with torch.no_grad():
            ref = wrapper.model.generate(**enc, max_new_tokens=10, pad_token_id=0)
            plain = wrapper.model.generate(**enc, max_new_tokens=10, pad_token_id=0, repetition_penalty=1.0)
        got = wrapper.steered_generate_ids(enc["input_ids"], enc["attention_mask"], None, None, None, None, None, 10)
        assert torch.equal(got.cpu(), ref[:, enc["input_ids"].shape[1]:])
        assert not torch.equal(ref, plain)  # the penalty really changed the greedy path
        gc_.top_p, gc_.top_k = 0.8, 20
        assert wrapper._generation_settings(0.0, {"top_k": 5}) == dict(
            do_sample=False, temperature=0.0, top_k=20, top_p=0.8, repetition_penalty=1.8)  # kwargs ignored at T=0 (:833)
        assert wrapper._generation_settings(0.7, {"top_k": 5}) == dict(
            do_sample=True, temperature=0.7, top_k=5, top_p=0.8, repetition_penalty=1.8)
        gc_.do_sample, gc_.temperature = True, 0.6  # Llama-3.1-Instruct's shipped config: temperature 0 still samples
        assert wrapper._generation_settings(0.0, {})["do_sample"] is True
        assert wrapper._generation_settings(0.0, {})["temperature"] == 0.6
    finally:
        gc_.repetition_penalty, gc_.top_p, gc_.top_k, gc_.do_sample, gc_.temperature = old


def test_hot_path_has_no_cpu_fallback(wrapper):
    v = torch.zeros(wrapper.engine.hidden)
    with pytest.raises(RuntimeError, match="no CPU path"):
        wrapper.generate_with_steering("a b c", 1, v, strength=2.0, max_new_tokens=2)
    with pytest.raises(RuntimeError, match="no CPU path"):
        vu.extract_concept_vectors_batch(wrapper, ["Dust"], ["Sand", "Clay"], layer_idx=1)
    with pytest.raises(RuntimeError, match="no CPU path"):
        wrapper.extract_activations(["a b c"], 1)
    with pytest.raises(ValueError, match="Unknown extraction method"):
        vu.extract_concept_vectors_batch(wrapper, ["Dust"], ["Sand"], layer_idx=1, extraction_method="bogus")
    with pytest.raises(AssertionError, match="one steering vector per prompt"):
        wrapper.generate_batch_with_multi_steering(["a", "b"], 1, [v])


def test_grid_rows_and_dedupe():
    concepts = [f"c{i}" for i in range(50)]
    rows = grid.layer_trials(concepts, [1, 2, 4, 8], n_injection=1)
    assert len(rows) == 200 and {r.strength for r in rows} == {1.0, 2.0, 4.0, 8.0}  # BASELINE configs[1]
    full = grid.layer_trials(concepts, [1, 2, 4, 8], n_injection=15, n_control=15, n_forced=15, dedupe_control=False)
    assert len(full) == 4 * 50 * 45  # the reference's per-layer work (detect_injected_thoughts.py:1792-1799)
    dedup = grid.layer_trials(concepts, [1, 2, 4, 8], n_injection=15, n_control=15, n_forced=15)
    assert len(dedup) == 4 * 50 * 30 + 50 * 15
    assert [len(c) for c in grid.chunk(rows, 64)] == [64, 64, 64, 8]
    assert rows[0] == grid.Trial("c0", 1.0, 1, "injection")
    # numbered as the driver numbers them: control n_inj + t (:1917), forced n_inj + n_ctl + t (:1984)
    few = grid.layer_trials(["a"], [2.0], n_injection=2, n_control=3, n_forced=2)
    assert [(r.kind, r.trial) for r in few] == [("injection", 1), ("injection", 2), ("forced_injection", 6),
                                                ("forced_injection", 7), ("control", 3), ("control", 4), ("control", 5)]


def test_sweep_writers_use_the_reference_wire_format(tmp_path):
    """results.json / results.csv / vectors layout the reference's re-judge, plotting and transcript
    tools read (detect_injected_thoughts.py:1651, :1780-1787, :2136-2157; eval_utils.py:910-914)."""
    import csv
    import json

    from introspective_awareness_b200 import sweep

    rows = [
        {"concept": "Dust", "trial": 1, "response": "I notice dust", "injected": True, "layer": 22, "layer_fraction": 0.7,
         "strength": 4.0, "detected": True, "trial_type": "injection"},
        {"concept": "Dust", "trial": 1, "response": "nothing", "injected": False, "layer": 22, "layer_fraction": 0.7,
         "strength": 4.0, "detected": False, "trial_type": "control"},
        {"concept": "Dust", "trial": 1, "response": "about trees", "injected": True, "layer": 22, "layer_fraction": 0.7,
         "strength": 4.0, "detected": False, "trial_type": "forced_injection"},
    ]
    d = sweep.write_config(tmp_path / "llama_8b", 0.7, 22, 4.0, rows, temperature=1.0, max_tokens=100)
    assert d.name == "layer_0.70_strength_4.0"
    blob = json.load(open(d / "results.json"))
    assert set(blob) == {"results", "metrics", "n_samples"} and blob["n_samples"] == 3
    assert blob["results"][0] == rows[0]
    m = blob["metrics"]
    assert m["detection_hit_rate"] == 1.0 and m["detection_false_alarm_rate"] == 0.0
    # the three judge-only figures are written as 0 when the judge is off, as the reference driver does (:2108-2120)
    assert m["detection_accuracy"] == 0 and m["identification_accuracy_given_claim"] == 0
    assert m["combined_detection_and_identification_rate"] == 0
    assert list(m)[:6] == ["detection_hit_rate", "detection_false_alarm_rate", "detection_accuracy",
                           "identification_accuracy_given_claim", "combined_detection_and_identification_rate",
                           "forced_identification_accuracy"]
    assert m["forced_identification_accuracy"] == 0.0 and m["layer_idx"] == 22 and m["strength"] == 4.0
    got = list(csv.DictReader(open(d / "results.csv")))
    assert len(got) == 3 and got[0]["concept"] == "Dust" and got[1]["trial_type"] == "control"
    sweep.save_vectors(tmp_path / "llama_8b", 0.7, {"Dust": torch.arange(4.0)})
    v = torch.load(tmp_path / "llama_8b" / "vectors" / "layer_0.70" / "Dust.pt")
    assert torch.equal(v, torch.arange(4.0))


def test_sampling_distribution_matches_hf_warpers():
    """temperature > 0 (the reference's default, detect_injected_thoughts.py:72): our sampler draws from
    exactly the distribution HF's default sampling pipeline builds (temperature, then top-k 50)."""
    from transformers.generation.logits_process import TemperatureLogitsWarper, TopKLogitsWarper

    from introspective_awareness_b200.engine import Session

    g = torch.Generator().manual_seed(0)
    logits = torch.randn(5, 300, generator=g) * 3
    from transformers.generation.logits_process import TopPLogitsWarper

    for temp, k, p in ((1.0, 50, None), (0.7, 50, None), (1.3, None, None), (0.6, 50, 0.9), (1.0, 20, 0.8), (1.0, None, 0.5)):
        scores = TemperatureLogitsWarper(temp)(None, logits.clone())
        if k is not None:
            scores = TopKLogitsWarper(top_k=k)(None, scores)
        if p is not None:
            scores = TopPLogitsWarper(top_p=p)(None, scores)
        want = torch.softmax(scores, dim=-1)
        got = Session.sampling_probs(logits, temp, k, p)
        assert torch.allclose(got, want, atol=1e-7) and ((got > 0) == (want > 0)).all()


def test_empty_inputs(wrapper):
    assert wrapper.generate_batch([], max_new_tokens=3) == []
    assert wrapper.generate_batch_with_multi_steering([], 1, []) == []
    assert vu.extract_concept_vectors_batch(wrapper, [], ["Sand"], layer_idx=1) == {}
    assert grid.layer_trials([], [1.0], n_injection=3) == []


def test_fused_linear_shares_storage_and_keeps_module_outputs():
    """`engine._FusedLinear`: q/k/v (or gate/up) as one GEMM over concatenated weights.  The modules' own
    parameters become row slices of the concatenated buffer — nothing stored twice, values unchanged, so the
    stock HF forward (the reference hook path) still computes what it computed before."""
    import torch

    from introspective_awareness_b200.engine import _FusedLinear
    from introspective_awareness_b200.synthetic import build_model

    model = build_model("tiny-qwen2", dtype=torch.float32, device="cpu", seed=5)  # Qwen2: q/k/v carry biases
    att = model.model.layers[0].self_attn
    mods = [att.q_proj, att.k_proj, att.v_proj]
    x = torch.randn(3, 7, model.config.hidden_size, generator=torch.Generator().manual_seed(0))
    with torch.no_grad():
        before = [m(x) for m in mods]
        weights_before = [m.weight.detach().clone() for m in mods]
        fused = _FusedLinear(mods)
        after = [m(x) for m in mods]
        parts = fused(x)
    base = fused.W.untyped_storage().data_ptr()
    for m, w0, b0, a0, part in zip(mods, weights_before, before, after, parts):
        assert m.weight.untyped_storage().data_ptr() == base and m.bias.untyped_storage().data_ptr() == fused.b.untyped_storage().data_ptr()
        assert torch.equal(m.weight, w0) and torch.equal(b0, a0)          # same values, same module output
        assert part.shape == a0.shape and torch.allclose(part, a0, rtol=1e-5, atol=1e-6)
    assert sum(fused.sizes) == fused.W.shape[0] and set(k for k in model.state_dict() if "layers.0.self_attn" in k) >= {
        "model.layers.0.self_attn.q_proj.weight", "model.layers.0.self_attn.k_proj.bias"}


def test_gemma2_sliding_window_steps_match_hf_generate():
    """Cached steps past Gemma-2's sliding window: the sliding layers see the last `sliding_window` keys only, as HF's
    sliding cache hands them over (tiny stand-in: window 64; prompts of 50-60 tokens, 24 new tokens cross it)."""
    from introspective_awareness_b200.model_utils import ModelWrapper
    from introspective_awareness_b200.synthetic import synthetic_prompt_ids

    w = ModelWrapper.from_arch("tiny-gemma2", device="cpu", dtype=torch.float32, seed=5)
    assert w.engine.sliding_window == 64
    ids, mask, _ = synthetic_prompt_ids(3, 60, 50, w.model.config.vocab_size, seed=8)
    with torch.no_grad():
        ref = w.model.generate(input_ids=ids, attention_mask=mask, max_new_tokens=24, do_sample=False, pad_token_id=0,
                               output_logits=True, return_dict_in_generate=True)
    got = w.steered_generate_ids(ids, mask, None, None, None, None, None, 24, keep_logits=True)
    ours = torch.stack(w.last_session.step_logits)
    theirs = torch.stack([l.float() for l in ref.logits])
    assert torch.equal(got, ref.sequences[:, 60:])
    assert (ours - theirs).abs().max().item() <= 1e-4


def test_sessions_are_kept_least_recently_used_first_out(wrapper):
    """The runtime keeps the static buffers of the last few batch shapes (a sweep alternates between a handful);
    a fourth shape evicts the least recently used one; `_session = None` drops them all."""
    eng = wrapper.engine
    eng._session = None
    made = {}
    for width in (5, 6, 7):
        made[width] = eng.session(2, width, 4, padded=True)
    assert list(k[1] for k in eng._sessions) == [5, 6, 7] and eng._session is made[7]
    assert eng.session(2, 5, 4, padded=True) is made[5]          # a hit moves the shape to the most recent end
    eng.session(2, 8, 4, padded=True)                           # the least recently used one (6) goes
    assert list(k[1] for k in eng._sessions) == [7, 5, 8]
    assert eng.session(2, 6, 4, padded=True) is not made[6]
    eng._session = None
    assert not eng._sessions and eng._session is None


def test_steer_spec_per_layer_rows():
    from introspective_awareness_b200.engine import SteerSpec

    table = torch.zeros(3, 8)
    one = SteerSpec(2, table, torch.tensor([0, 1], dtype=torch.int32), torch.ones(2), None)
    assert one.at(2) is one and one.at(1) is None
    rows = torch.full((4, 2), -1, dtype=torch.int32)
    rows[2] = torch.tensor([0, 1])
    every = SteerSpec(None, table, rows, torch.ones(2), None)
    for i in range(4):
        s = every.at(i)
        assert s.layer_idx == i and torch.equal(s.row_vec, rows[i]) and s.vtable is table
