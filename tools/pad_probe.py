"""Does extra left padding change the bits?  (Would decide whether the padded width of a batch could be rounded up to
a bucket so that fewer distinct batch shapes — each costing max_new_tokens - 1 graph captures — occur in a sweep.)
Same rows, same t0 (the pad shift is applied by the runtime), width W against W + extra: compares the 32-step logits."""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from introspective_awareness_b200.model_utils import ModelWrapper  # noqa: E402
from introspective_awareness_b200.synthetic import synthetic_prompt_ids  # noqa: E402

for arch in ("llama-8b-4layer", "gemma-2-9b-4layer"):
    torch.manual_seed(7)
    w = ModelWrapper.from_arch(arch, device="cuda:0", dtype=torch.bfloat16, seed=None)
    V = w.model.config.vocab_size
    rows, T = 64, 121
    ids, mask, lens = synthetic_prompt_ids(rows, T, 100, V, seed=1)
    table = (torch.randn(16, w.engine.hidden, generator=torch.Generator().manual_seed(3)) * 0.5).to(torch.bfloat16)
    t0 = (lens - 20).tolist()
    row_vec = [i % 16 for i in range(rows)]
    alphas = [[1.0, 2.0, 4.0, 8.0][i % 4] for i in range(rows)]

    def run(extra):
        i2 = torch.cat([torch.zeros(rows, extra, dtype=ids.dtype), ids], 1)
        m2 = torch.cat([torch.zeros(rows, extra, dtype=mask.dtype), mask], 1)
        out = w.steered_generate_ids(i2, m2, 2, table, row_vec, alphas, t0, 32, keep_logits=True).cpu().clone()
        return out, torch.stack(w.last_session.step_logits).cpu()

    base_ids, base_lg = run(0)
    for extra in (1, 7, 8):
        o, lg = run(extra)
        print(json.dumps({"arch": arch, "width": T, "extra_left_pad": extra, "max_abs_logit_diff": (lg - base_lg).abs().max().item(),
                          "first_step_max_abs": (lg[0] - base_lg[0]).abs().max().item(),
                          "rows_with_identical_tokens": (o == base_ids).all(1).float().mean().item()}), flush=True)
    w.cleanup()
