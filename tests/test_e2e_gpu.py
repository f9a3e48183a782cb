"""End-to-end parity of the B200 runtime on the GPU.

north_star's bar: concept vectors within 1e-3 relative L2 of the reference; steered next-token
logits within 2e-2 max-abs of the reference hook path with identical greedy tokens for the first
32 steps, on synthetic prompts and random-init weights.

Two anchors:
  (a) the restated reference hook path (oracle/ref_hooks.py: HF `generate` + forward hook, pinned
      bit-for-bit to the real reference by oracle/make_golden.py) run on the SAME GPU and dtype;
  (b) the committed outputs of the real reference run on the CPU (tests/golden/e2e_*.npz,
      config1_*.npz).  fp32 cases must reproduce the reference's greedy ids exactly; bf16 cases
      differ from a CPU run by GEMM rounding, so there the bound is the 2e-2 logit tolerance on the
      first step (later steps may legitimately follow a different greedy path).
"""

import numpy as np
import pytest
import torch

from conftest import load_golden, widen
from oracle import ref_hooks

pytestmark = pytest.mark.gpu

DT = {"bf16": torch.bfloat16, "f32": torch.float32}


def _dev():
    if not torch.cuda.is_available():
        pytest.skip("needs a GPU")
    torch.backends.cuda.matmul.allow_tf32 = False
    torch.backends.cudnn.allow_tf32 = False
    return "cuda:0"


def _wrapper(arch, dtype, seed, attention="auto", **overrides):
    from introspective_awareness_b200.model_utils import ModelWrapper

    return ModelWrapper.from_arch(arch, device=_dev(), dtype=dtype, seed=seed, attention=attention, **overrides)


def _ours(w, g_ids, g_mask, layer, vec_table, alphas, t0, n_new):
    ids = torch.from_numpy(g_ids)
    mask = torch.from_numpy(g_mask)
    out = w.steered_generate_ids(ids, mask, layer, vec_table, list(range(vec_table.shape[0])), alphas, t0, n_new,
                                 temperature=0.0, keep_logits=True)
    return out.cpu().numpy(), torch.stack(w.last_session.step_logits).cpu()


@pytest.mark.parametrize("name", ["e2e_qwen2_f32.npz", "e2e_gemma2_f32.npz", "e2e_llama_bf16.npz", "e2e_qwen2_bf16.npz",
                                  "e2e_gemma2_bf16.npz", "e2e_llama_bf16_lastlayer_a3.npz", "e2e_llama_bf16_h1024.npz",
                                  "e2e_llama_bf16_allpos.npz"])
def test_against_real_reference_outputs(name):
    g = load_golden(name)
    dtn = str(g["dtype"])
    overrides = eval(str(g["overrides"]), {"__builtins__": {}})
    w = _wrapper(str(g["arch"]), DT[dtn], int(g["seed"]), **overrides)
    layer = int(g["layer_idx"])
    # extraction through K1 (two forwards + one fused launch), same token ids the reference saw
    v32, vr, _ = w.engine.extract_vectors(torch.from_numpy(g["concept_ids"]), torch.from_numpy(g["concept_mask"]),
                                          torch.from_numpy(g["base_ids"]), torch.from_numpy(g["base_mask"]), [layer])
    want = torch.from_numpy(widen(g["vectors"], dtn))
    got = vr[0].float().cpu()
    rel = ((got - want).norm(dim=1) / want.norm(dim=1)).max().item()
    # bf16 fixtures were computed on a CPU: its GEMM rounding differs from the GPU's upstream of K1, hence 2e-2 here; the
    # 1e-3 bound of north_star (in fact: identical bits) is asserted where both sides run on the same device —
    # test_same_device_extraction_against_the_hook_path and tests/test_real_reference_bridge.py
    assert rel <= (1e-3 if dtn == "f32" else 2e-2), rel
    # steering with the REFERENCE's vectors, so this half does not inherit extraction differences
    table = want.to(DT[dtn])
    t0 = None if bool(g["steer_all"]) else g["t0"].tolist()
    n_new = int(g["n_new"])
    ids, logits = _ours(w, g["input_ids"], g["attention_mask"], layer, table, [float(g["alpha"])] * table.shape[0], t0, n_new)
    ref_logits = torch.from_numpy(g["logits"])
    if dtn == "f32":
        assert (logits - ref_logits).abs().max().item() <= 2e-3
        assert np.array_equal(ids, g["new_ids"])
    else:
        assert (logits[0] - ref_logits[0]).abs().max().item() <= 2e-2
        same = (ids == g["new_ids"]).all(axis=1)
        # rows that followed the same greedy path must agree on every step's logits
        for r in np.nonzero(same)[0]:
            assert (logits[:, r] - ref_logits[:, r]).abs().max().item() <= 4e-2


def test_config1_qwen05b_fp32_matches_real_reference():
    """BASELINE.json configs[0]: what the reference produced on the CPU, reproduced on the B200."""
    g = load_golden("config1_qwen05b_f32.npz")
    w = _wrapper("qwen2.5-0.5b", torch.float32, int(g["seed"]))
    layer = int(g["layer_idx"])
    assert layer == 16
    _, vr, _ = w.engine.extract_vectors(torch.from_numpy(g["concept_ids"]), torch.from_numpy(g["concept_mask"]),
                                        torch.from_numpy(g["base_ids"]), torch.from_numpy(g["base_mask"]), [layer])
    want = torch.from_numpy(g["vectors"])
    rel = ((vr[0].float().cpu() - want).norm(dim=1) / want.norm(dim=1)).max().item()
    assert rel <= 1e-3, rel
    ids, logits = _ours(w, g["input_ids"], g["attention_mask"], layer, vr[0], [float(g["alpha"])] * 4, g["t0"].tolist(), 32)
    assert np.array_equal(ids, g["new_ids"])
    top_idx = torch.from_numpy(g["top_indices"]).long()
    got_top = torch.gather(logits, 2, top_idx)
    assert (got_top - torch.from_numpy(g["top_values"])).abs().max().item() <= 2e-3


def _hook_path_vs_ours(arch, layer, rows, t_max, t_min, n_new, alphas, seed, n_concepts=50, attention="hf", **overrides):
    from introspective_awareness_b200.synthetic import synthetic_prompt_ids

    if seed is None:
        torch.manual_seed(1234)
    w = _wrapper(arch, torch.bfloat16, seed, attention=attention, **overrides)
    V = w.model.config.vocab_size
    ids, mask, lens = synthetic_prompt_ids(rows, t_max, t_min, V, seed=1)
    gen = torch.Generator().manual_seed(5)
    table = (torch.randn(n_concepts, w.engine.hidden, generator=gen) * 0.5).to(torch.bfloat16)
    t0 = (lens - min(28, t_min - 1)).tolist()
    row_vec = [i % n_concepts for i in range(rows)]
    row_alpha = [alphas[(i // n_concepts) % len(alphas)] for i in range(rows)]
    ours = w.steered_generate_ids(ids, mask, layer, table, row_vec, row_alpha, t0, n_new, temperature=0.0, keep_logits=True)
    ours = ours.cpu()
    our_logits = torch.stack(w.last_session.step_logits).cpu()
    # reference hook path: one call per distinct alpha with the SAME batch
    ref_ids = torch.zeros_like(ours)
    ref_logits = torch.zeros_like(our_logits)
    for a in sorted(set(row_alpha)):
        out = ref_hooks.steered_generate(w.model, ids, mask, layer, [table[j] for j in row_vec], a, t0, n_new,
                                         want_logits=True)
        sel = torch.tensor([x == a for x in row_alpha])
        ref_ids[sel] = out["new_ids"][sel]
        ref_logits[:, sel] = out["logits"][:, sel]
    return ours, our_logits, ref_ids, ref_logits


@pytest.mark.parametrize("arch,layer", [("tiny-llama", 1), ("tiny-qwen2", 2), ("tiny-gemma2", 1), ("tiny-llama", 3)])
def test_same_device_hook_path_tiny(arch, layer):
    ours, ol, ref, rl = _hook_path_vs_ours(arch, layer, rows=24, t_max=40, t_min=30, n_new=32, alphas=[1.0, 3.0, 8.0],
                                           seed=11, n_concepts=8)
    assert (ol - rl).abs().max().item() <= 2e-2
    assert torch.equal(ours, ref)


def test_same_device_hook_path_llama_8b_dims():
    """Real Llama-3.1-8B hidden size / heads / vocabulary (4 layers), BASELINE configs[1] row count:
    200 rows = 50 concepts x 4 strengths in ONE batch vs four reference calls."""
    ours, ol, ref, rl = _hook_path_vs_ours("llama-8b-4layer", 2, rows=200, t_max=128, t_min=112, n_new=32,
                                           alphas=[1.0, 2.0, 4.0, 8.0], seed=3)
    diff = (ol - rl).abs().max().item()
    same_rows = (ours == ref).all(dim=1).float().mean().item()
    print(f"max |logit diff| = {diff:.4g}; rows with identical 32-token greedy paths: {same_rows:.3f}")
    assert diff <= 2e-2
    assert torch.equal(ours, ref)


@pytest.mark.parametrize("arch,layer,attention", [("gemma-2-9b-4layer", 1, "hf"), ("gemma-2-9b-4layer", 3, "hf"),
                                                  ("gemma-2-9b-4layer", 1, "gqa"), ("qwen2.5-32b-2layer", 0, "gqa")])
def test_same_device_hook_path_gemma2_9b_and_qwen_32b_dims(arch, layer, attention):
    """BASELINE configs[3]/[4] shapes: Gemma-2 (post+pre norm fusion variant, incl. injection at the
    LAST layer, fused with model.norm) and Qwen2.5-32B (H = 5120) against the hook path on the same GPU.

    Every kernel of ours is on the path (K2 for all norms, the Gemma-2 tail variant, fused rope; for Qwen the
    engine-driven block) and the logits must satisfy north_star's 2e-2 — in fact they are bit-identical, with the
    model's stock attention function and with the grouped-query mode alike (at head_dim = 256 that is the folded
    form, which runs the same fused kernel as HF's repeat_kv path with K/V read in place)."""
    # seed=None: Hugging Face's own initialisation, on the GPU (both paths share the model object)
    ours, ol, ref, rl = _hook_path_vs_ours(arch, layer, rows=96, t_max=64, t_min=48, n_new=32,
                                           alphas=[1.0, 2.0, 4.0, 8.0], seed=None, n_concepts=24, attention=attention)
    diff = (ol - rl).abs()
    same_rows = (ours == ref).all(dim=1).float().mean().item()
    print(f"{arch} L={layer} {attention}: max |logit diff| = {diff.max().item():.4g}; identical 32-token rows: {same_rows:.3f}")
    assert diff.max().item() == 0.0
    assert torch.equal(ours, ref)


@pytest.mark.parametrize("attention", ["gqa", "hf"])
def test_benched_configuration_is_bit_identical_to_the_hook_path(attention):
    """bench.py's configuration — Llama-3.1-8B dimensions, 200 rows = 50 concepts x 4 strengths, prompt 112..128,
    32 greedy tokens, and the DEFAULT runtime mode ("auto" = grouped-query SDPA, what `ModelWrapper` / `load_model`
    give a caller who passes nothing) — against the STOCK hook path on the same GPU: north_star asks for
    <= 2e-2 and identical greedy tokens; the logits are in fact the same bits on every step of every row."""
    ours, ol, ref, rl = _hook_path_vs_ours("llama-8b-4layer", 2, rows=200, t_max=128, t_min=112, n_new=32,
                                           alphas=[1.0, 2.0, 4.0, 8.0], seed=3, attention=attention)
    same = (ours == ref).all(dim=1)
    diff = (ol - rl).abs().max().item()
    print(f"attention={attention}: max |logit diff| over 32 steps = {diff:.4g}; rows with identical 32-token greedy paths: "
          f"{same.float().mean().item():.3f}")
    assert torch.equal(ours, ref)
    assert diff == 0.0


def test_default_runtime_mode_is_the_benched_one():
    from introspective_awareness_b200.model_utils import ModelWrapper, load_model

    w = ModelWrapper.from_arch("tiny-llama", device=_dev(), dtype=torch.bfloat16, seed=1)
    assert w.engine.attention == "gqa" and w.engine.native_block
    w2 = load_model("arch:tiny-qwen2", device=_dev(), dtype="bfloat16")
    assert w2.engine.attention == "gqa"


@pytest.mark.parametrize("arch,layers", [("llama-8b-4layer", [2]), ("llama-8b-4layer", [0, 1, 3]), ("gemma-2-9b-4layer", [1, 3]),
                                         ("qwen2.5-32b-2layer", [0, 1]), ("tiny-gemma2", [0, 2, 3])])
def test_same_device_extraction_against_the_hook_path(arch, layers):
    """Concept vectors end to end — extraction forwards with K2's capture tap, then K1 — against the restated
    reference on the SAME GPU and model: `ref_hooks.capture` (forward hook on the stock HF forward, one full
    forward per layer, `.cpu()`) + `ref_hooks.concept_vectors` (torch CPU arithmetic), bf16, BASELINE's 50 concept
    + 100 baseline prompts with ragged lengths (left padding).  north_star: <= 1e-3 relative L2; captured
    activations and vectors are in fact the same bits."""
    from introspective_awareness_b200.synthetic import synthetic_prompt_ids

    w = _wrapper(arch, torch.bfloat16, 9 if arch.startswith("tiny") else None)
    if not arch.startswith("tiny"):
        torch.manual_seed(4321)
    V = w.model.config.vocab_size
    c_ids, c_mask, _ = synthetic_prompt_ids(50, 24, 8, V, seed=2)
    b_ids, b_mask, _ = synthetic_prompt_ids(100, 24, 8, V, seed=3)
    _, vr, _ = w.engine.extract_vectors(c_ids, c_mask, b_ids, b_mask, layers, want_f32=False, want_norm=False)
    crows = w.engine.capture_rows(c_ids, c_mask, layers).cpu()
    for k, layer in enumerate(layers):
        cacts = ref_hooks.capture(w.model, c_ids, c_mask, layer)
        bacts = ref_hooks.capture(w.model, b_ids, b_mask, layer)
        want = ref_hooks.concept_vectors(cacts, bacts, "baseline")
        got = vr[k].cpu()
        rel = ((got.float() - want.float()).norm(dim=1) / want.float().norm(dim=1)).max().item()
        act_rel = ((crows[k].float() - cacts.float()).norm(dim=1) / cacts.float().norm(dim=1)).max().item()
        print(f"{arch} layer {layer}: vectors rel L2 {rel:.3g}, captured rows rel L2 {act_rel:.3g}, "
              f"identical vectors {torch.equal(got, want)}")
        assert rel <= 1e-3, (arch, layer, rel)
        assert torch.equal(crows[k], cacts), (arch, layer)
        assert torch.equal(got, want), (arch, layer)


def test_string_api_matches_hook_path_text():
    """The reference-facing string surface: same text as the hook path produces (with its A3
    slicing reproduced via compat_slice)."""
    from introspective_awareness_b200 import steering_utils as su
    from introspective_awareness_b200 import vector_utils as vu

    w = _wrapper("tiny-llama", torch.bfloat16, 7)
    concepts = ["Dust", "Trees", "Oceans"]
    vecs = vu.extract_concept_vectors_batch(w, concepts, vu.get_baseline_words(10), layer_idx=1)
    assert all(v.device.type == "cpu" and v.dtype == torch.bfloat16 for v in vecs.values())
    built = [su.build_trial(w, n) for n in (1, 2, 13)]
    prompts, t0 = [p for p, _ in built], [t for _, t in built]
    w.compat_slice = True
    texts = w.generate_batch_with_multi_steering(prompts, 1, [vecs[c] for c in concepts], strength=4.0,
                                                 max_new_tokens=12, temperature=0.0, steering_start_positions=t0)
    enc = w.tokenizer(prompts, return_tensors="pt", padding=True, truncation=True)
    ref = ref_hooks.steered_generate(w.model, enc["input_ids"], enc["attention_mask"], 1, [vecs[c] for c in concepts],
                                     4.0, t0, 12)
    lens = enc["attention_mask"].sum(1).tolist()
    full = torch.cat([enc["input_ids"], ref["new_ids"]], dim=1)
    want = [w.tokenizer.decode(full[i, lens[i]:], skip_special_tokens=True).strip() for i in range(3)]
    assert texts == want
    # single-prompt surface = batch of one with the same semantics
    one = w.generate_with_steering(prompts[0], 1, vecs["Dust"], strength=4.0, max_new_tokens=12, temperature=0.0,
                                   steering_start_pos=t0[0])
    w.compat_slice = False
    assert one == w.tokenizer.decode(ref_hooks.steered_generate(
        w.model, enc["input_ids"][:1, enc["input_ids"].shape[1] - lens[0]:], enc["attention_mask"][:1, enc["input_ids"].shape[1] - lens[0]:],
        1, [vecs["Dust"]], 4.0, [t0[0]], 12)["new_ids"][0], skip_special_tokens=True).strip()
    # a vector index beyond the table is refused up front (the kernels take no table length)
    ids_, mask_ = w._encode(prompts)
    with pytest.raises(IndexError, match="row_vec"):
        w.steered_generate_ids(ids_, mask_, 1, torch.stack([vecs[c] for c in concepts]), [0, 1, 3], [4.0] * 3, t0, 4)
    # same-vector batch entry point works (the reference's raises UnboundLocalError)
    outs = w.generate_batch_with_steering(prompts, 1, vecs["Dust"], strength=2.0, max_new_tokens=4, temperature=0.0,
                                          steering_start_pos=t0[0])
    assert len(outs) == 3
    # sampling path runs and is reproducible under a seeded generator
    ids, mask = w._encode(prompts)
    g1 = torch.Generator(device="cuda").manual_seed(3)
    a = w.steered_generate_ids(ids, mask, 1, torch.stack([vecs[c] for c in concepts]), [0, 1, 2], [4.0] * 3, t0, 8,
                               temperature=1.0, generator=g1).clone()
    g2 = torch.Generator(device="cuda").manual_seed(3)
    b = w.steered_generate_ids(ids, mask, 1, torch.stack([vecs[c] for c in concepts]), [0, 1, 2], [4.0] * 3, t0, 8,
                               temperature=1.0, generator=g2).clone()
    assert torch.equal(a, b)


def test_sweep_driver_end_to_end(tmp_path):
    """Multi-layer extraction + one batch per layer with per-row strength + reference file layout."""
    import json

    from introspective_awareness_b200 import sweep, vector_utils as vu

    w = _wrapper("tiny-llama", torch.bfloat16, 7)
    w.model_name = "tiny/llama"
    concepts = ["Dust", "Trees", "Oceans"]
    out = sweep.run_sweep(w, concepts, vu.get_baseline_words(10), [0.3, 0.8], [2.0, 8.0], n_trials=4, max_new_tokens=6,
                          temperature=0.0, batch_size=16, output_dir=str(tmp_path))
    assert set(out) == {(0.3, 2.0), (0.3, 8.0), (0.8, 2.0), (0.8, 8.0)}
    base = tmp_path / "tiny_llama"
    blob = json.load(open(base / "layer_0.80_strength_8.0" / "results.json"))
    assert blob["n_samples"] == 3 * (2 + 2 + 2) and blob["metrics"]["layer_idx"] == 3
    kinds = {r["trial_type"] for r in blob["results"]}
    assert kinds == {"injection", "control", "forced_injection"}
    assert [r["trial_type"] for r in blob["results"]] == ["injection"] * 6 + ["control"] * 6 + ["forced_injection"] * 6
    assert [r["trial"] for r in blob["results"]][:2] == [1, 2] and {r["trial"] for r in blob["results"][6:12]} == {3, 4}
    assert {r["trial"] for r in blob["results"][12:]} == {5, 6}
    assert all(r["strength"] == 8.0 for r in blob["results"])
    v = torch.load(base / "vectors" / "layer_0.30" / "Trees.pt")
    assert v.shape == (w.engine.hidden,) and v.dtype == torch.bfloat16
    # one batch per layer must equal the reference-style call for a single (layer, strength) config
    vecs = vu.extract_concept_vectors_batch(w, concepts, vu.get_baseline_words(10), layer_idx=3)
    assert torch.equal(vecs["Trees"], torch.load(base / "vectors" / "layer_0.80" / "Trees.pt"))
    from introspective_awareness_b200 import steering_utils as su

    inj = [r for r in blob["results"] if r["trial_type"] == "injection"]
    built = [su.build_trial(w, r["trial"]) for r in inj]
    texts = w.generate_batch_with_multi_steering([p for p, _ in built], 3, [vecs[r["concept"]] for r in inj], strength=8.0,
                                                 max_new_tokens=6, temperature=0.0,
                                                 steering_start_positions=[t for _, t in built])
    assert texts == [r["response"] for r in inj]


@pytest.mark.parametrize("arch", ["tiny-llama", "tiny-gemma2", "llama-8b-4layer", "gemma-2-9b-4layer"])
def test_runtime_optimisations_do_not_change_bits(arch):
    """K2 for every RMSNorm, the fused rotary kernel, the Gemma-2 tail variant, grouped-query SDPA, the
    CUDA graphs, the once-per-forward mask conversion, the engine-driven Llama block (residual add inside K2,
    rotary + KV append, SiLU-gate multiply) and the single-GEMM q/k/v and gate/up projections, each switched
    off in turn: same logits, bit for bit (Gemma-2's grouped-query mode is checked at its real head size above)."""
    from introspective_awareness_b200.synthetic import synthetic_prompt_ids

    w = _wrapper(arch, torch.bfloat16, 9 if arch.startswith("tiny") else None, attention="hf")
    V = w.model.config.vocab_size
    rows = 40
    ids, mask, lens = synthetic_prompt_ids(rows, 48, 32, V, seed=2)
    table = (torch.randn(8, w.engine.hidden, generator=torch.Generator().manual_seed(3)) * 0.5).to(torch.bfloat16)
    t0 = (lens - 20).tolist()
    row_vec = [i % 9 - 1 for i in range(rows)]  # includes control rows
    alphas = [[1.0, 3.0, 8.0][i % 3] for i in range(rows)]
    eng = w.engine

    def run():
        eng._session = None
        out = w.steered_generate_ids(ids, mask, 1, table, row_vec, alphas, t0, 12, keep_logits=True).cpu().clone()
        return out, torch.stack(w.last_session.step_logits).cpu()

    base_ids, base_logits = run()
    settings = [("fused_norms", False), ("fused_rope", False), ("graph_decode", False), ("additive_masks", False)]
    if eng.native_block:  # Llama / Qwen2: the block driven by the engine vs the HF layer forwards
        settings += [("native_block", False), ("fuse_projections", False)]
    if eng.is_gemma2:
        settings += [("fuse_gemma_tail", False), ("gemma_tail_everywhere", False), ("table_mlp", False), ("fuse_projections", False),
                     ("gemma_native_attention", False)]
    else:
        settings.append(("attention", "gqa"))
    for name, value in settings:
        if name == "graph_decode":
            orig_session = eng.session
            eng.session = lambda *a, **k: orig_session(*a, **{**k, "use_graphs": False})
        elif name == "attention":
            from introspective_awareness_b200 import engine as E

            E._register_attention()
            old = eng.attention
            eng.attention = value
        else:
            old = getattr(eng, name)
            setattr(eng, name, value)
        try:
            got_ids, got_logits = run()
        finally:
            if name == "graph_decode":
                eng.session = orig_session
            elif name == "attention":
                eng.attention = old
            else:
                setattr(eng, name, old)
        assert torch.equal(got_logits, base_logits), name
        assert torch.equal(got_ids, base_ids), name


def test_fp16_model_against_hook_path():
    """fp16 residual stream: K2 / rope fp16 paths inside the runtime."""
    from introspective_awareness_b200.model_utils import ModelWrapper
    from introspective_awareness_b200.synthetic import synthetic_prompt_ids

    w = ModelWrapper.from_arch("tiny-qwen2", device=_dev(), dtype=torch.float16, seed=4)
    V = w.model.config.vocab_size
    ids, mask, lens = synthetic_prompt_ids(24, 40, 30, V, seed=6)
    table = (torch.randn(6, w.engine.hidden, generator=torch.Generator().manual_seed(1)) * 0.5).to(torch.float16)
    t0 = (lens - 12).tolist()
    row_vec = [i % 6 for i in range(24)]
    ours = w.steered_generate_ids(ids, mask, 2, table, row_vec, [4.0] * 24, t0, 16, keep_logits=True).cpu()
    ol = torch.stack(w.last_session.step_logits).cpu()
    ref = ref_hooks.steered_generate(w.model, ids, mask, 2, [table[j] for j in row_vec], 4.0, t0, 16, want_logits=True)
    assert (ol - ref["logits"]).abs().max().item() <= 2e-2
    assert torch.equal(ours, ref["new_ids"])
