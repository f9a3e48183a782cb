"""One bench step split into phases (CUDA events), with cudaProfilerStart/Stop around a single
step so that `ncu --profile-from-start off` sees exactly one step's kernels.

    python tools/profile_step.py [--arch llama-3.1-8b] [--no-graphs]
    ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none --csv \
        --log-file gpurun_out/step_launches.csv python tools/profile_step.py
"""

from __future__ import annotations

import argparse
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import bench  # noqa: E402
from introspective_awareness_b200.engine import SteerSpec  # noqa: E402
from introspective_awareness_b200.model_utils import ModelWrapper, get_layer_at_fraction  # noqa: E402
from introspective_awareness_b200.stub_tokenizer import StubTokenizer  # noqa: E402
from introspective_awareness_b200.synthetic import build_model  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--arch", default=bench.ARCH)
    ap.add_argument("--no-graphs", action="store_true")
    ap.add_argument("--new-tokens", type=int, default=bench.N_NEW)
    ap.add_argument("--attention", default="auto", choices=["auto", "gqa", "hf"])
    ap.add_argument("--eager-norms", action="store_true")
    ap.add_argument("--profile-phase", default="all", choices=["all", "extract", "prefill_decode"])
    args = ap.parse_args()
    dev = torch.device("cuda:0")
    model = build_model(args.arch, dtype=torch.bfloat16, device="cuda:0", seed=None)
    w = ModelWrapper(args.arch, device="cuda:0", dtype=torch.bfloat16, model=model,
                     tokenizer=StubTokenizer(vocab_size=model.config.vocab_size), attention=args.attention)
    w.engine.fused_norms = not args.eager_norms
    eng = w.engine
    layer = get_layer_at_fraction(w, bench.LAYER_FRACTION)
    wl = bench.make_workload(model.config.vocab_size, 0)
    on = lambda t: t.to(dev)
    c_ids, c_mask, b_ids, b_mask = on(wl["c_ids"]), on(wl["c_mask"]), on(wl["b_ids"]), on(wl["b_mask"])
    ids, mask = on(wl["ids"]), on(wl["mask"])
    rows = wl["rows"]
    pads = bench.T_PROMPT - wl["lens"]
    t0 = (torch.tensor(wl["t0"]) + pads).to(torch.int32).to(dev)
    row_vec = torch.tensor(wl["row_vec"], dtype=torch.int32, device=dev)
    row_alpha = torch.tensor(wl["row_alpha"], dtype=torch.float32, device=dev)
    sess = eng.session(rows, bench.T_PROMPT, args.new_tokens, padded=True, use_graphs=not args.no_graphs)

    def ev():
        e = torch.cuda.Event(enable_timing=True)
        e.record()
        return e

    def step(profile=False):
        marks = [ev()]
        if profile and args.profile_phase in ("all", "extract"):
            torch.cuda.profiler.start()
        _, vr, _ = eng.extract_vectors(c_ids, c_mask, b_ids, b_mask, [layer], want_f32=False)
        if profile and args.profile_phase == "extract":
            torch.cuda.profiler.stop()
        marks.append(ev())
        if profile and args.profile_phase == "prefill_decode":
            torch.cuda.profiler.start()
        steer = SteerSpec(layer, vr[0], row_vec, row_alpha, t0)
        # split generate into its two halves by timing a 1-token session separately is intrusive;
        # instead time the whole call and a prefill-only call
        sess.generate(ids, mask, steer)
        if profile and args.profile_phase in ("all", "prefill_decode"):
            torch.cuda.profiler.stop()
        marks.append(ev())
        torch.cuda.synchronize()
        return [marks[i].elapsed_time(marks[i + 1]) for i in range(len(marks) - 1)]

    for _ in range(3):
        step()
    t = step(profile=True)
    # prefill-only timing: same session shape with 1 new token
    s1 = eng.session(rows, bench.T_PROMPT, 1, padded=True, use_graphs=False)
    steer = SteerSpec(layer, torch.zeros(bench.N_CONCEPTS, eng.hidden, dtype=torch.bfloat16, device=dev), row_vec, row_alpha, t0)
    for _ in range(2):
        s1.generate(ids, mask, steer)
    a = ev()
    s1.generate(ids, mask, steer)
    b = ev()
    torch.cuda.synchronize()
    prefill = a.elapsed_time(b)
    print(json.dumps({"arch": args.arch, "graphs": not args.no_graphs, "attention": args.attention, "fused_norms": not args.eager_norms, "extract_ms": t[0], "generate_ms": t[1],
                      "prefill_ms": prefill, "decode_ms": t[1] - prefill,
                      "decode_ms_per_step": (t[1] - prefill) / max(args.new_tokens - 1, 1)}))


if __name__ == "__main__":
    main()
