"""Small run of every kernel family for `compute-sanitizer --tool memcheck` (one tool per gpurun call)."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from introspective_awareness_b200 import _lib, kernels  # noqa: E402

dev = "cuda:0"
g = torch.Generator(device=dev).manual_seed(0)
for H in (128, 896, 4096, 5120, 132):
    for dtype in (torch.bfloat16, torch.float32):
        rows, T = 7, 5
        h = torch.randn(rows, T, H, device=dev, generator=g).to(dtype)
        vt = torch.randn(3, H, device=dev, generator=g).to(dtype)
        w = torch.ones(H, device=dev, dtype=dtype)
        rv = torch.tensor([0, 1, 2, -1, 0, 1, 2], dtype=torch.int32, device=dev)
        al = torch.full((rows,), 3.0, device=dev)
        t0 = torch.tensor([0, 1, 2, 3, 4, 5, 6], dtype=torch.int32, device=dev)
        for impl in ("regs", "auto") + (("tma",) if H % 128 == 0 else ()):
            kernels.inject_rmsnorm(h.clone(), w, 1e-5, "llama", vtable=vt, row_vec=rv, row_alpha=al, row_t0=t0, impl=impl)
            kernels.inject_rmsnorm(h.clone(), w, 1e-6, "gemma", impl=impl)
        if H % 128 == 0 and (dtype != torch.float32 or H <= 2048):
            kernels.gemma2_tail_inject_rmsnorm(h.clone(), h.clone(), w, w, 1e-6, vtable=vt, row_vec=rv, row_alpha=al, row_t0=t0)
for (L, C, Nb, H, T) in ((2, 5, 100, 896, 3), (1, 3, 700, 1024, 2), (3, 50, 100, 136, 2)):
    cl = [torch.randn(C, T, H, device=dev, generator=g).to(torch.bfloat16) for _ in range(L)]
    bl = [torch.randn(Nb, T, H, device=dev, generator=g).to(torch.bfloat16) for _ in range(L)]
    for mode in ("baseline", "posneg", "raw"):
        kernels.gather_meandiff(cl, None if mode == "raw" else bl, mode, True)
    rows, s = kernels.gather_segsum(bl)
    kernels.meandiff_finalize(torch.stack([c[:, -1] for c in cl]).contiguous(), s, Nb, "baseline", False)
    lib = _lib.load()
    cptr, bptr = kernels._ptr_table(cl, "c"), kernels._ptr_table(bl, "b")
    outr = torch.empty(L, C, H, dtype=torch.bfloat16, device=dev)
    rc = lib.ia_gather_meandiff_static(cptr.data_ptr(), T, 0, 0, C, bptr.data_ptr(), T, 0, 0, Nb, L, H, 0, 0, 0, outr.data_ptr(), 0, 0, 0)
    assert rc == 0
q = torch.randn(3, 9, 8 * 64, device=dev, generator=g).to(torch.bfloat16).view(3, 9, 8, 64).transpose(1, 2)
k = torch.randn(3, 9, 2 * 64, device=dev, generator=g).to(torch.bfloat16).view(3, 9, 2, 64).transpose(1, 2)
cs = torch.randn(3, 9, 64, device=dev, generator=g).to(torch.bfloat16)
assert kernels.rope(q, k, cs, cs) is not None
torch.cuda.synchronize()
print("sanitize smoke: all kernels ran")
