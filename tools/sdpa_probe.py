"""Probe: decode-step attention (200 rows, 8 KV heads x 4 query heads, head_dim 128, ~145 cached positions, padding mask)
through torch SDPA in its grouped-query form and in the folded form (query heads of a group along the query-length axis)."""
import json

import torch
import torch.nn.functional as F

dev = "cuda:0"
B, nh, nkv, D, L, Lmax = 200, 32, 8, 128, 145, 160
g = nh // nkv
gen = torch.Generator(device=dev).manual_seed(0)
caches = [(torch.randn(B, nkv, Lmax, D, device=dev, generator=gen).to(torch.bfloat16),
           torch.randn(B, nkv, Lmax, D, device=dev, generator=gen).to(torch.bfloat16)) for _ in range(8)]
q = torch.randn(B, 1, nh, D, device=dev, generator=gen).to(torch.bfloat16).transpose(1, 2)
valid = torch.ones(B, 1, 1, L, dtype=torch.bool, device=dev)
valid[:, :, :, :7] = torch.rand(B, 1, 1, 7, device=dev, generator=gen) > 0.5
mask = torch.where(valid, torch.zeros((), dtype=torch.bfloat16, device=dev), torch.full((), float("-inf"), dtype=torch.bfloat16, device=dev))


def grouped(k, v):
    return F.scaled_dot_product_attention(q, k[:, :, :L], v[:, :, :L], attn_mask=mask, enable_gqa=True).transpose(1, 2).contiguous()


def folded(k, v):
    o = F.scaled_dot_product_attention(q.reshape(B, nkv, g, D), k[:, :, :L], v[:, :, :L], attn_mask=mask)
    return o.view(B, nkv, g, 1, D).permute(0, 3, 1, 2, 4).reshape(B, 1, nh, D).contiguous()


def repeat(k, v):
    kk = k[:, :, :L][:, :, None].expand(B, nkv, g, L, D).reshape(B, nh, L, D)
    vv = v[:, :, :L][:, :, None].expand(B, nkv, g, L, D).reshape(B, nh, L, D)
    return F.scaled_dot_product_attention(q, kk, vv, attn_mask=mask).transpose(1, 2).contiguous()


ref = repeat(*caches[0])
for name, fn in (("grouped", grouped), ("folded", folded), ("repeat_kv", repeat)):
    out = fn(*caches[0])
    for i in range(8):
        fn(*caches[i])
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for i in range(64):
        fn(*caches[i % 8])
    b.record()
    torch.cuda.synchronize()
    print(json.dumps({"form": name, "us": round(a.elapsed_time(b) * 1e3 / 64, 2), "equals_repeat_kv": bool(torch.equal(out, ref)),
                      "max_abs_diff": float((out.float() - ref.float()).abs().max())}), flush=True)
