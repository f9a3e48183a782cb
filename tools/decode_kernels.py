"""Which kernels make up ONE decode step of an architecture?  torch.profiler over an un-graphed decode step
(device time per kernel name).   python tools/decode_kernels.py --arch gemma-2-9b [--attention hf|gqa]"""
import argparse
import collections
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402
from introspective_awareness_b200.model_utils import ModelWrapper  # noqa: E402
from introspective_awareness_b200.stub_tokenizer import StubTokenizer  # noqa: E402
from introspective_awareness_b200.synthetic import build_model  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--arch", default="gemma-2-9b")
ap.add_argument("--attention", default="auto")
ap.add_argument("--layers", type=int, default=6, help="build only this many layers (per-layer breakdown is what matters)")
a = ap.parse_args()
model = build_model(a.arch, dtype=torch.bfloat16, device="cuda:0", seed=None, num_hidden_layers=a.layers)
w = ModelWrapper(a.arch, device="cuda:0", dtype=torch.bfloat16, model=model, tokenizer=StubTokenizer(vocab_size=model.config.vocab_size),
                 attention=a.attention)
wl = bench.make_workload(model.config.vocab_size, 0)
H = w.engine.hidden
table = (torch.randn(bench.N_CONCEPTS, H) * 0.3).to(torch.bfloat16)
orig = w.engine.session
w.engine.session = lambda *x, **k: orig(*x, **{**k, "use_graphs": False})
args = (wl["ids"], wl["mask"], a.layers // 2, table, wl["row_vec"], wl["row_alpha"], wl["t0"], 4)
w.steered_generate_ids(*args)
torch.cuda.synchronize()
with torch.profiler.profile(activities=[torch.profiler.ProfilerActivity.CUDA]) as prof:
    w.steered_generate_ids(*args)
    torch.cuda.synchronize()
agg = collections.defaultdict(lambda: [0, 0.0])
for e in prof.events():
    if e.device_type == torch.autograd.DeviceType.CUDA:
        agg[e.name[:90]][0] += 1
        agg[e.name[:90]][1] += e.device_time
tot = sum(v[1] for v in agg.values())
print(json.dumps({"arch": a.arch, "layers": a.layers, "attention": w.engine.attention, "total_us_prefill_plus_3_decode": tot}))
for k, (n, t) in sorted(agg.items(), key=lambda kv: -kv[1][1])[:22]:
    print(f"{n:5d} {t:10.1f} us {100 * t / tot:5.1f}%  {k}")
