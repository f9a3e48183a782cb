"""Probe: how fast does stock cuBLAS run the decode-step projections (M = 200 rows) of Llama-3.1-8B, against the
time HBM needs to stream the weights once?  Weights rotate over 8 copies so that every call reads cold weights.
    python tools/gemm_probe.py [--lt]      (--lt: torch.backends.cuda.preferred_blas_library("cublaslt"))
"""
import json
import sys

import torch

if "--lt" in sys.argv:
    torch.backends.cuda.preferred_blas_library("cublaslt")
dev = "cuda:0"
M = 200
shapes = {"qkv": (6144, 4096), "o": (4096, 4096), "gate_up": (28672, 4096), "down": (4096, 14336), "gate": (14336, 4096)}
for name, (n, k) in shapes.items():
    ws = [torch.randn(n, k, device=dev, dtype=torch.bfloat16) * 0.02 for _ in range(8)]
    x = torch.randn(M, k, device=dev, dtype=torch.bfloat16)
    for w in ws:
        torch.nn.functional.linear(x, w)
    torch.cuda.synchronize()
    reps = 64
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for i in range(reps):
        torch.nn.functional.linear(x, ws[i % 8])
    b.record()
    torch.cuda.synchronize()
    us = a.elapsed_time(b) * 1e3 / reps
    byts = n * k * 2
    print(json.dumps({"gemm": name, "N": n, "K": k, "M": M, "us": round(us, 2), "weight_gbs": round(byts / us / 1e3, 1),
                      "lt": "--lt" in sys.argv}), flush=True)
    del ws
