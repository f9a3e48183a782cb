"""GPU checks of the runtime-support kernels that are not hot-path rows: the fused rotary
embedding must produce the bits of transformers' eager apply_rotary_pos_emb."""

import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16, torch.float32])
@pytest.mark.parametrize("B,T,nh,nkv,D", [(200, 1, 32, 8, 128), (3, 37, 14, 2, 64), (2, 16, 16, 8, 256), (1, 5, 4, 4, 32)])
def test_fused_rope_is_bit_identical_to_hf(dtype, B, T, nh, nkv, D):
    if not torch.cuda.is_available():
        pytest.skip("needs a GPU")
    from transformers.models.llama.modeling_llama import apply_rotary_pos_emb

    from introspective_awareness_b200 import kernels

    dev = "cuda:0"
    g = torch.Generator(device=dev).manual_seed(B * 1000 + T)
    q = torch.randn(B, T, nh * D, device=dev, generator=g).to(dtype).view(B, T, nh, D).transpose(1, 2)
    k = torch.randn(B, T, nkv * D, device=dev, generator=g).to(dtype).view(B, T, nkv, D).transpose(1, 2)
    ang = torch.rand(B, T, D // 2, device=dev, generator=g) * 6.28
    emb = torch.cat((ang, ang), dim=-1)
    cos, sin = emb.cos().to(dtype), emb.sin().to(dtype)
    want_q, want_k = apply_rotary_pos_emb(q, k, cos, sin)
    got = kernels.rope(q, k, cos, sin)
    assert got is not None
    torch.cuda.synchronize()
    assert got[0].stride() == q.stride()
    assert torch.equal(got[0], want_q) and torch.equal(got[1], want_k)
    # a layout the kernel does not cover is declined, not mis-read
    if T > 1 and nh > 1:
        assert kernels.rope(q.contiguous(), k, cos, sin) is None
