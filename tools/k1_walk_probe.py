"""K1 column-walker probe: bits of the walker (n_layers * H >= 65536 and no norm outputs) against the staged kernel (the
same call WITH norm outputs), then cold-cache time per launch of the production call (no norm pass) on a ladder of layer
counts.  The variants this kernel was chosen from are recorded in profiles/r2_k1_walk_probe.jsonl."""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from introspective_awareness_b200 import _lib, kernels  # noqa: E402
from tools.kbench import time_launch  # noqa: E402

dev = torch.device("cuda:0")
g = torch.Generator(device=dev).manual_seed(0)
flush = torch.empty(512 << 20, dtype=torch.uint8, device=dev)
ok = True
for dtype in (torch.bfloat16, torch.float16, torch.float32):
    for (L, C, Nb, H) in ((80, 7, 100, 896), (16, 50, 256, 4096), (20, 51, 1, 3584), (4096, 1, 77, 16), (70, 120, 101, 1024)):
        ca = torch.randn(L, C, H, device=dev, generator=g).to(dtype)
        ba = torch.randn(L, Nb, H, device=dev, generator=g).to(dtype)
        for mode in ("baseline", "raw"):
            f_new, r_new, _ = kernels.meandiff_compact(ca, None if mode == "raw" else ba, mode, False, want_f32=True, want_norm=False)
            f_old, r_old, _ = kernels.meandiff_compact(ca, None if mode == "raw" else ba, mode, False, want_f32=True, want_norm=True)
            ok = ok and torch.equal(f_new.view(torch.int32), f_old.view(torch.int32)) and torch.equal(
                r_new.view(torch.uint8), r_old.view(torch.uint8))
print(json.dumps({"walker_bits_equal_to_staged": ok}), flush=True)
lib = _lib.load()
for (L, C, Nb, H) in ((1, 50, 100, 4096), (8, 50, 100, 4096), (16, 50, 100, 4096), (32, 50, 100, 4096), (64, 50, 100, 5120), (28, 50, 100, 3584)):
    ca = torch.randn(L, C, H, device=dev, generator=g).to(torch.bfloat16)
    ba = torch.randn(L, Nb, H, device=dev, generator=g).to(torch.bfloat16)
    ws = torch.empty(kernels.workspace_bytes(L, C, Nb, H) + 256, dtype=torch.uint8, device=dev)
    out_r = torch.empty(L, C, H, dtype=torch.bfloat16, device=dev)
    args = (ca.data_ptr(), C, ba.data_ptr(), Nb, L, H, _lib.IA_MODE_BASELINE, 0, 0, out_r.data_ptr(), 0, ws.data_ptr(), ws.numel(),
            kernels._dt(ca), torch.cuda.current_stream().cuda_stream)

    def fn():
        assert lib.ia_meandiff_compact(*args) == 0

    ms = time_launch(fn, 30, flush)
    algo = L * ((C + Nb) * H * 2 + C * H * 2)
    print(json.dumps({"kernel": "walker" if L * H >= 65536 else "staged", "L": L, "C": C, "Nb": Nb, "H": H, "us": round(ms * 1e3, 2),
                      "gbs": round(algo / ms / 1e6, 1)}), flush=True)
