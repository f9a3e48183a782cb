"""One cold launch of K1 at the BASELINE sweep shape (32 layers x (50 + 100) rows x 4096, bf16, compact rows, no norm pass) after
a few warm ones — the command `ncu --set full -k regex:k1_ -s 3 -c 1` is pointed at."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from introspective_awareness_b200 import _lib, kernels  # noqa: E402

L, C, Nb, H = (int(x) for x in (sys.argv[1:5] if len(sys.argv) >= 5 else (32, 50, 100, 4096)))
dev = torch.device("cuda:0")
g = torch.Generator(device=dev).manual_seed(0)
ca = torch.randn(L, C, H, device=dev, generator=g).to(torch.bfloat16)
ba = torch.randn(L, Nb, H, device=dev, generator=g).to(torch.bfloat16)
ws = torch.empty(kernels.workspace_bytes(L, C, Nb, H) + 256, dtype=torch.uint8, device=dev)
out_r = torch.empty(L, C, H, dtype=torch.bfloat16, device=dev)
flush = torch.empty(512 << 20, dtype=torch.uint8, device=dev)
lib = _lib.load()
for _ in range(5):
    flush.zero_()
    rc = lib.ia_meandiff_compact(ca.data_ptr(), C, ba.data_ptr(), Nb, L, H, _lib.IA_MODE_BASELINE, 0, 0, out_r.data_ptr(), 0,
                                 ws.data_ptr(), ws.numel(), kernels._dt(ca), torch.cuda.current_stream().cuda_stream)
    assert rc == 0, rc
torch.cuda.synchronize()
print("k1 once: ok")
