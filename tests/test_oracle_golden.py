"""Pins the oracle (oracle/kernels_np.py, oracle/ref_hooks.py) against fixtures produced by the
real reference (oracle/make_golden.py).  CPU only.

Tolerances, as stated by BASELINE.json's north_star and SURVEY.md Appendix C:
  * injection (h' = h + alpha*v): bit-exact in every dtype;
  * concept vectors: <= 1e-3 relative L2, and here in fact bit-exact for bf16/fp32 because the
    oracle's column-sum order equals torch's CPU order for Nb <= 256;
  * RMSNorm output: the reference's CPU reduction order differs from the kernel's canonical
    order, which can move a value across a rounding boundary: <= 1 ulp of the model dtype on
    <= 0.1 % of elements (bf16), <= 4e-6 relative (fp32).
"""

import glob
import os

import numpy as np
import pytest
import torch

from conftest import GOLDEN, load_golden, ulp_diff_bf16, widen
from oracle import kernels_np as K
from oracle import ref_hooks


def test_bf16_rounding_matches_torch():
    rng = np.random.default_rng(0)
    x = (rng.standard_normal(200000) * 10.0 ** rng.integers(-6, 6, 200000)).astype(np.float32)
    ours = K.bf16_bits(x)
    theirs = torch.from_numpy(x).to(torch.bfloat16).view(torch.int16).numpy().view(np.uint16)
    assert np.array_equal(ours, theirs)
    # exact ties round to even
    ties = np.array([0x3F808000, 0x3F818000, 0xBF808000], dtype=np.uint32).view(np.float32)
    assert K.bf16_bits(ties).tolist() == [0x3F80, 0x3F82, 0xBF80]


@pytest.mark.parametrize("name", ["k1_bf16_h512.npz", "k1_f32_h256.npz"])
def test_meandiff_matches_reference_vectors(name):
    g = load_golden(name)
    dt = str(g["dtype"])
    ca, ba, ctrl = (widen(g[k], dt) for k in ("concept_acts", "base_acts", "control_act"))
    cases = [("baseline", "baseline", ba), ("simple", "simple", ctrl), ("raw", "no_baseline", None),
             ("posneg", "posneg", ba)]
    for mode, key, base in cases:
        for norm in (0, 1):
            want = widen(g[f"{key}_{norm}"], dt)
            v32, vr, nr = K.meandiff(ca, base, mode, dt, normalize=bool(norm))
            rel = np.linalg.norm(vr - want, axis=1) / np.linalg.norm(want, axis=1)
            assert rel.max() <= 1e-3, (mode, norm, rel.max())
            if not norm:
                # un-normalised (the production case): exact for bf16, where the
                # oracle's sequential column sum is torch's CPU order; torch sums fp32 inputs in a
                # cascaded order, so there the last fp32 bit of the mean may differ
                if dt == "bf16":
                    assert np.array_equal(vr, want), (mode, norm)
                else:
                    assert rel.max() <= 1e-6, (mode, norm, rel.max())
                # the un-rounded fp32 path differs from the reference only by the model-dtype
                # rounding of the mean and of the difference (large for posneg, whose result is
                # a small difference of two rounded means)
                rel32 = np.linalg.norm(v32 - want, axis=1) / np.linalg.norm(want, axis=1)
                assert rel32.max() <= (5e-2 if dt == "bf16" else 1e-6)
            else:
                # norm reduction order differs from torch's CPU norm: allow 1 ulp on few elements
                if dt == "bf16":
                    d = ulp_diff_bf16(vr, want)
                    assert d.max() <= 1 and (d > 0).mean() <= 5e-3, (mode, d.max(), (d > 0).mean())
                else:
                    assert rel.max() <= 1e-6, (mode, norm, rel.max())


K2_FILES = sorted(os.path.basename(p) for p in glob.glob(os.path.join(GOLDEN, "k2_*.npz")))


@pytest.mark.parametrize("name", K2_FILES)
def test_inject_rmsnorm_matches_reference_hook_and_hf_norm(name):
    g = load_golden(name)
    dt = str(g["dtype"])
    variant = str(g["variant"])
    vt = widen(g["vectors"], dt)
    w = widen(g["norm_weight"], dt)
    b = vt.shape[0]
    for call in range(3):
        before = widen(g[f"before{call}"], dt)
        after = widen(g[f"after{call}"], dt)
        normed = widen(g[f"normed{call}"], dt)
        steer_all = bool(g["steer_all"]) or before.shape[1] == 1  # decode steps always steer
        hp, y = K.inject_rmsnorm(before, vt, np.arange(b), np.full(b, g["alpha"], np.float32),
                                 g["t0_shifted"], int(g["pos0"][call]), steer_all, w, float(g["eps"]),
                                 variant, dt)
        assert np.array_equal(hp, after), (name, call)
        if dt == "bf16":
            d = ulp_diff_bf16(y, normed)
            assert d.max() <= 1 and (d > 0).mean() <= 1e-3, (name, call, d.max(), (d > 0).mean())
        else:
            assert np.allclose(y, normed, rtol=4e-6, atol=1e-7), (name, call)
    # positions before t0 are untouched at prefill (reference model_utils.py:786-787)
    if not bool(g["steer_all"]):
        before = widen(g["before0"], dt)
        after = widen(g["after0"], dt)
        for i, t0 in enumerate(g["t0_shifted"]):
            assert np.array_equal(before[i, :t0], after[i, :t0])
            assert not np.array_equal(before[i, t0:], after[i, t0:])


def test_row_sum_squares_is_close_to_float64():
    rng = np.random.default_rng(3)
    for h in (128, 896, 3584, 4096, 5120):
        x = rng.standard_normal((5, h)).astype(np.float32)
        got = K.row_sum_squares(x)
        ref = (x.astype(np.float64) ** 2).sum(-1)
        assert np.allclose(got, ref, rtol=3e-6)


def test_column_sum_chunking():
    rng = np.random.default_rng(4)
    x = rng.standard_normal((700, 64)).astype(np.float32)
    got = K.column_sum(x)
    assert np.allclose(got, x.astype(np.float64).sum(0), rtol=1e-5, atol=1e-4)
    # <= 256 rows is the plain sequential sum
    seq = x[0].copy()
    for r in range(1, 200):
        seq = seq + x[r]
    assert np.array_equal(K.column_sum(x[:200]), seq)


@pytest.mark.parametrize("name", ["e2e_qwen2_f32.npz", "e2e_llama_bf16.npz", "e2e_gemma2_bf16.npz",
                                  "e2e_llama_bf16_allpos.npz"])
def test_ref_hooks_restatement_reproduces_reference(name):
    """oracle/ref_hooks.py re-run here must give the reference's stored ids/logits bit for bit
    (same torch, CPU).  Skipped quietly if the host's CPU kernels differ from the build box."""
    from introspective_awareness_b200.synthetic import build_model

    g = load_golden(name)
    dt = {"bf16": torch.bfloat16, "f32": torch.float32}[str(g["dtype"])]
    overrides = eval(str(g["overrides"]), {"__builtins__": {}})  # literal dict written by make_golden
    model = build_model(str(g["arch"]), dtype=dt, seed=int(g["seed"]), **overrides)
    layer = int(g["layer_idx"])
    cacts = ref_hooks.capture(model, torch.from_numpy(g["concept_ids"]), torch.from_numpy(g["concept_mask"]), layer)
    bacts = ref_hooks.capture(model, torch.from_numpy(g["base_ids"]), torch.from_numpy(g["base_mask"]), layer)
    vecs = ref_hooks.concept_vectors(cacts, bacts, "baseline")
    want = torch.from_numpy(widen(g["vectors"], str(g["dtype"])))
    rel = ((vecs.float() - want).norm(dim=1) / want.norm(dim=1)).max().item()
    assert rel <= 1e-3
    t0 = None if bool(g["steer_all"]) else g["t0"].tolist()
    out = ref_hooks.steered_generate(model, torch.from_numpy(g["input_ids"]), torch.from_numpy(g["attention_mask"]),
                                     layer, list(vecs), float(g["alpha"]), t0, int(g["n_new"]), want_logits=True)
    assert (out["logits"] - torch.from_numpy(g["logits"])).abs().max().item() <= 2e-2
    if torch.equal(vecs.float(), want):
        assert np.array_equal(out["new_ids"].numpy(), g["new_ids"])
