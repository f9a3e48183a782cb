"""The sweep of this package against the REAL reference driver's sweep (build container only; skipped where
/root/reference is absent): `detect_injected_thoughts.main()` is run unmodified in sweep mode against a scripted
model, `sweep.run_sweep` against the same scripted model, and what reached the model (prompt text, steering start
positions, vectors, strengths, layers) and what was written (`results.json`: every result row in order, the metric
names and values; `results.csv`; the vector files) must be the same.

This pins the host-side half of SURVEY.md §8 rows a11 / f2: trial numbering (control trials are numbered
n_injection + t, forced trials n_injection + n_control + t: detect_injected_thoughts.py:1917, :1984), the driver's
own forced-identification prompt with its assistant prefill (:1985-2017, not steering_utils' forced-noticing
prompt), `trial_type == "forced_injection"` (:2056) and the no-judge metric fallback (:2108-2120).
"""

import importlib.util
import json
import os
import sys
import types
import zlib

import pytest
import torch

REF = "/root/reference"
H = 8


def _crc(*parts) -> int:
    return zlib.crc32("|".join(str(p) for p in parts).encode())


class ScriptedModel:
    """A model whose answers are a pure function of what it is asked: both drivers can be compared call by call."""

    def __init__(self, chat: bool, name: str = "spy_model"):
        from introspective_awareness_b200.stub_tokenizer import StubTokenizer

        self.tokenizer = StubTokenizer(vocab_size=4096, chat_template=chat)
        self.model_name = name
        self.hf_path = name
        self.n_layers = 10
        self.device, self.dtype = "cpu", torch.float32
        self.seen = []          # (kind, prompt, layer, strength, t0, vector checksum)
        self._prompts = None
        self._texts = None

    # ---- what the reference's vector_utils / driver call --------------------------------------
    def extract_activations(self, prompts, layer_idx, token_idx=-1):
        rows = []
        for p in prompts:
            g = torch.Generator().manual_seed(_crc(p, layer_idx, token_idx) % (2 ** 31))
            rows.append(torch.randn(H, generator=g))
        return torch.stack(rows)

    def _answer(self, prompt, layer, strength, t0, vec):
        key = _crc(prompt, layer, strength, t0, None if vec is None else [round(float(x), 5) for x in vec])
        self.seen.append(("steer" if vec is not None else "plain", prompt, layer, strength, t0,
                          None if vec is None else round(float(vec.sum()), 5)))
        return f"reply {key % 9973} " + ("about dust and trees" if key % 3 == 0 else "nothing unusual")

    def generate_batch_with_multi_steering(self, prompts, layer_idx, steering_vectors, strength=1.0, max_new_tokens=512,
                                           temperature=0.0, steering_start_positions=None, debug=False, **kw):
        assert len(prompts) == len(steering_vectors) == len(steering_start_positions)
        return [self._answer(p, layer_idx, float(strength), t, v) for p, v, t in zip(prompts, steering_vectors, steering_start_positions)]

    def generate_batch(self, prompts, max_new_tokens=512, temperature=0.0, **kw):
        return [self._answer(p, None, None, None, None) for p in prompts]

    def cleanup(self):
        pass

    # ---- what this package's grid / vector_utils call on its ModelWrapper -----------------------
    def extract_concept_vectors(self, concept_prompts, baseline_prompts, layer_indices, method="baseline", normalize=False,
                                token_idx=-1, fp32=False, sharded=False):
        out = []
        for layer in layer_indices:
            c = self.extract_activations(concept_prompts, layer, token_idx)
            if method == "baseline":  # reference vector_utils.py:541-548
                c = torch.stack([row - self.extract_activations(baseline_prompts, layer, token_idx).mean(dim=0) for row in c])
            out.append(c)
        return torch.stack(out)

    def _encode(self, prompts):
        self._prompts = list(prompts)
        n = len(prompts)
        return torch.zeros(n, 1, dtype=torch.long), torch.ones(n, 1, dtype=torch.long)

    def steered_generate_ids(self, ids, mask, layer_idx, table, row_vec, row_alpha, starts, max_new_tokens, temperature=0.0, **kw):
        self._texts = []
        for p, vi, a, t in zip(self._prompts, row_vec, row_alpha, starts):
            if vi < 0:
                self._texts.append(self._answer(p, None, None, None, None))
            else:
                self._texts.append(self._answer(p, layer_idx, float(a), t, table[vi]))
        return torch.arange(len(self._texts))[:, None]

    def _decode_rows(self, new_ids, ids, mask):
        return [self._texts[int(i)] for i in new_ids[:, 0]]


def _load_driver(monkeypatch):
    monkeypatch.setattr(sys, "dont_write_bytecode", True)
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
    for k in ("model_utils", "vector_utils", "steering_utils", "eval_utils"):
        monkeypatch.delitem(sys.modules, k, raising=False)
    monkeypatch.syspath_prepend(REF)
    spec = importlib.util.spec_from_file_location("ref_driver_sweep", os.path.join(REF, "detect_injected_thoughts.py"))
    drv = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(drv)
    return drv


@pytest.fixture(autouse=True)
def _no_reference_modules_left_behind():
    yield
    for name in ("model_utils", "vector_utils", "steering_utils", "eval_utils"):
        mod = sys.modules.get(name)
        if mod is not None and str(getattr(mod, "__file__", "")).startswith(REF):
            del sys.modules[name]


@pytest.mark.skipif(not os.path.exists(os.path.join(REF, "detect_injected_thoughts.py")), reason="reference not mounted")
@pytest.mark.parametrize("chat,model_name,n_trials,batch", [(False, "spy_model", 5, 4), (True, "spy_model", 5, 4),
                                                            (True, "gemma2_9b", 5, 4), (True, "spy_model", 4, 7)])
def test_sweep_equals_the_real_drivers_sweep(monkeypatch, tmp_path, chat, model_name, n_trials, batch):
    from introspective_awareness_b200 import sweep, vector_utils as vu

    concepts, fractions, strengths = ["Dust", "Trees", "Oceans"], [0.3, 0.8], [2.0, 8.0]
    n_base, max_tokens = 6, 7

    # ---- the reference driver, unmodified, in sweep mode ---------------------------------------
    drv = _load_driver(monkeypatch)
    ref_model = ScriptedModel(chat, model_name)
    monkeypatch.setattr(drv, "load_model", lambda **kw: ref_model)
    for name in ("create_sweep_plots", "create_cross_model_comparison_plots", "extract_example_transcripts"):
        monkeypatch.setattr(drv, name, lambda *a, **k: None)
    ref_out = tmp_path / "ref"
    monkeypatch.setattr(sys, "argv", ["detect_injected_thoughts.py", "-m", model_name, "-c", *concepts, "-nb", str(n_base),
                                      "-ls", *map(str, fractions), "-ss", *map(str, strengths), "-nt", str(n_trials),
                                      "-t", "0.0", "-mt", str(max_tokens), "-bs", str(batch), "-od", str(ref_out), "-nlj", "-ow",
                                      "-d", "cpu", "-dt", "float32"])
    drv.main()

    # ---- this package's sweep ------------------------------------------------------------------
    our_model = ScriptedModel(chat, model_name)
    ours_out = tmp_path / "ours"
    got = sweep.run_sweep(our_model, concepts, vu.get_baseline_words(n_base), fractions, strengths, n_trials=n_trials,
                          max_new_tokens=max_tokens, temperature=0.0, batch_size=batch, output_dir=str(ours_out))
    assert set(got) == {(f, a) for f in fractions for a in strengths}

    # every steered call reached the model with the same prompt, layer, strength, t0 and vector; control prompts too
    # (ours asks once per layer instead of once per (layer, strength): compare as sets)
    assert {s for s in ref_model.seen if s[0] == "steer"} == {s for s in our_model.seen if s[0] == "steer"}
    assert {s for s in ref_model.seen if s[0] == "plain"} == {s for s in our_model.seen if s[0] == "plain"}
    assert len([s for s in our_model.seen if s[0] == "steer"]) == len([s for s in ref_model.seen if s[0] == "steer"])
    n_ctl = n_trials - n_trials // 2
    assert len([s for s in our_model.seen if s[0] == "plain"]) == len(fractions) * len(concepts) * n_ctl
    assert len([s for s in ref_model.seen if s[0] == "plain"]) == len(fractions) * len(strengths) * len(concepts) * n_ctl

    for f in fractions:
        for a in strengths:
            name = f"layer_{f:.2f}_strength_{a}"
            ref_blob = json.load(open(ref_out / model_name / name / "results.json"))
            our_blob = json.load(open(ours_out / model_name / name / "results.json"))
            assert our_blob["n_samples"] == ref_blob["n_samples"] == len(concepts) * (n_trials // 2 * 2 + n_ctl)
            assert our_blob["results"] == ref_blob["results"], name     # every row, every field, in the driver's order
            assert our_blob["metrics"] == ref_blob["metrics"], name     # names, values, and the zeros of the no-judge fallback
            assert list(our_blob["metrics"]) == list(ref_blob["metrics"])
            assert got[(f, a)] == ref_blob["results"]
            ref_csv = open(ref_out / model_name / name / "results.csv").read().splitlines()
            our_csv = open(ours_out / model_name / name / "results.csv").read().splitlines()
            assert our_csv == ref_csv, name
        for c in concepts:
            rv = torch.load(ref_out / model_name / "vectors" / f"layer_{f:.2f}" / f"{c}.pt")
            ov = torch.load(ours_out / model_name / "vectors" / f"layer_{f:.2f}" / f"{c}.pt")
            assert torch.equal(rv, ov), (f, c)
    kinds = [r["trial_type"] for r in got[(0.3, 2.0)]]
    assert set(kinds) == {"injection", "control", "forced_injection"}
