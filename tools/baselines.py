"""The two extra baselines BASELINE.md §4 asks for, measured on the GPU box:

  (2) path-only CPU micro-timing: the reference hook body (clone + per-row add, model_utils.py:767-788)
      followed by HF's RMSNorm on CPU tensors [R, 1, H] / [B, T, H], at the kernel ladder shapes, and the
      CPU mean-diff (vector_utils.py:541-547) — like-for-like GB/s next to K2 / K1;
  (3) the reference-on-GPU "before" number: the same random-init Llama-3.1-8B on the B200 driven by
      the reference's own mechanism (forward hook + `model.generate`, restated in oracle/ref_hooks.py),
      configs[1] = 50 concepts x 4 strengths: the reference needs one call per strength (scalar
      `strength`, model_utils.py:692), i.e. 4 calls of 50 rows.

    python tools/baselines.py [--skip-gpu] [--skip-cpu] > gpurun_out/baselines.jsonl
"""

from __future__ import annotations

import argparse
import json
import os
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import bench  # noqa: E402
from oracle import ref_hooks  # noqa: E402


def cpu_path_ladder():
    from transformers.models.llama.modeling_llama import LlamaRMSNorm

    torch.set_num_threads(os.cpu_count() or 1)
    H = 4096
    norm = LlamaRMSNorm(H, 1e-5).to(torch.bfloat16)
    for rows, T in ((200, 1), (3000, 1), (200, 128)):
        g = torch.Generator().manual_seed(0)
        h = torch.randn(rows, T, H, generator=g).to(torch.bfloat16)
        sv = (torch.randn(rows, H, generator=g) * 0.4).to(torch.bfloat16)
        starts = [0] * rows
        with torch.no_grad():
            ref_hooks.hook_body_and_norm(h, sv, starts, norm)
            reps = 3 if rows * T > 5000 else 10
            t = time.perf_counter()
            for _ in range(reps):
                ref_hooks.hook_body_and_norm(h, sv, starts, norm)
            dt = (time.perf_counter() - t) / reps
        algo = rows * T * H * 2 * 4 + H * 2
        print(json.dumps({"baseline": "cpu_path_hook_plus_rmsnorm", "rows": rows, "T": T, "H": H, "threads": torch.get_num_threads(),
                          "ms": dt * 1e3, "gbs_algorithmic": algo / dt / 1e9}), flush=True)
    for n_layers in (1, 32):
        g = torch.Generator().manual_seed(1)
        conc = [(torch.randn(50, H, generator=g) * 3 + 5).to(torch.bfloat16) for _ in range(n_layers)]
        base = [(torch.randn(100, H, generator=g) * 3 + 5).to(torch.bfloat16) for _ in range(n_layers)]
        t = time.perf_counter()
        for _ in range(5):
            for c, b in zip(conc, base):
                ref_hooks.concept_vectors(c, b, "baseline")
        dt = (time.perf_counter() - t) / 5
        algo = n_layers * (150 * H * 2 + 50 * H * 2)
        print(json.dumps({"baseline": "cpu_path_meandiff", "layers": n_layers, "ms": dt * 1e3, "gbs_algorithmic": algo / dt / 1e9}),
              flush=True)


def ref_on_gpu():
    from introspective_awareness_b200.model_utils import get_layer_at_fraction
    from introspective_awareness_b200.synthetic import build_model

    dev = "cuda:0"
    model = build_model(bench.ARCH, dtype=torch.bfloat16, device=dev, seed=None)

    class _W:
        n_layers = model.config.num_hidden_layers

    layer = get_layer_at_fraction(_W, bench.LAYER_FRACTION)
    wl = bench.make_workload(model.config.vocab_size, 0)

    def step():
        cacts = ref_hooks.capture(model, wl["c_ids"], wl["c_mask"], layer)      # hook + .cpu(), full-depth forward
        bacts = ref_hooks.capture(model, wl["b_ids"], wl["b_mask"], layer)
        vecs = ref_hooks.concept_vectors(cacts, bacts, "baseline")               # CPU tensor math
        outs = []
        for k, a in enumerate(bench.STRENGTHS):                                    # one generate call per strength
            sel = list(range(k * bench.N_CONCEPTS, (k + 1) * bench.N_CONCEPTS))
            ids, mask = wl["ids"][sel], wl["mask"][sel]
            keep = int(mask.sum(1).max())
            out = ref_hooks.steered_generate(model, ids[:, -keep:], mask[:, -keep:], layer, [vecs[i % bench.N_CONCEPTS] for i in sel],
                                             a, [wl["t0"][i] for i in sel], bench.N_NEW)
            outs.append(out["new_ids"])
        return outs

    for _ in range(2):
        step()
    torch.cuda.synchronize()
    t = time.perf_counter()
    reps = 3
    for _ in range(reps):
        step()
    torch.cuda.synchronize()
    dt = (time.perf_counter() - t) / reps
    print(json.dumps({"baseline": "reference_hook_path_on_gpu", "arch": bench.ARCH, "rows_per_step": wl["rows"],
                      "calls_per_step": len(bench.STRENGTHS), "new_tokens": bench.N_NEW, "ms_per_step": dt * 1e3,
                      "trials_per_s": wl["rows"] / dt,
                      "note": "torch-eager forward hooks + HF generate (DynamicCache, no CUDA graphs), stock SDPA"}), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--skip-gpu", action="store_true")
    ap.add_argument("--skip-cpu", action="store_true")
    args = ap.parse_args()
    if not args.skip_gpu:
        ref_on_gpu()
    if not args.skip_cpu:
        cpu_path_ladder()


if __name__ == "__main__":
    main()
