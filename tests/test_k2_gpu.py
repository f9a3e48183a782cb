"""GPU parity of K2 (inject + RMSNorm) through the C ABI.

Three anchors:
  1. the numpy oracle (oracle/kernels_np.py), with IA_RSQRT_IEEE so that both sides use the same
     correctly-rounded 1/sqrt: BIT-EXACT in bf16, fp16 and fp32;
  2. the eager torch chain on the same GPU (the reference hook's ops + HF RMSNorm ops), default
     flags: BIT-EXACT for rows >= 16 — the kernel sums the variance in ATen's order;
  3. the committed fixtures recorded from the real reference on CPU: h' bit-exact; y within
     1 ulp of the model dtype on <= 0.1 % of elements (bf16) / 4e-6 relative (fp32), because
     torch's CPU reduction order is not the GPU's.
"""

import glob
import os

import numpy as np
import pytest
import torch

from conftest import GOLDEN, load_golden, ulp_diff_bf16, widen
from oracle import kernels_np as K

pytestmark = pytest.mark.gpu

TORCH_DT = {"bf16": torch.bfloat16, "f16": torch.float16, "f32": torch.float32}


def _dev():
    if not torch.cuda.is_available():
        pytest.skip("needs a GPU")
    return torch.device("cuda:0")


def to_model(x32: np.ndarray, dt: str, dev) -> torch.Tensor:
    return torch.from_numpy(np.ascontiguousarray(x32, dtype=np.float32)).to(dev).to(TORCH_DT[dt])


def make_inputs(rows, T, H, dt, seed, n_vec=7, gemma=False):
    rng = np.random.default_rng(seed)
    h = rng.standard_normal((rows, T, H)).astype(np.float32)
    h[..., :4] *= 100.0  # massive channels
    vt = (rng.standard_normal((n_vec, H)) * 0.1).astype(np.float32)
    w = (rng.standard_normal(H) * 0.02 + (0.0 if gemma else 1.0)).astype(np.float32)
    h, vt_m, w = K.rd(h, dt), K.rd(vt, dt), K.rd(w, dt)
    row_vec = rng.integers(-1, n_vec, size=rows).astype(np.int32)  # -1 = control row
    alpha = rng.choice(np.array([1.0, 2.0, 3.0, 4.0, 8.0, 0.37], np.float32), size=rows)
    t0 = rng.integers(0, T + 2, size=rows).astype(np.int32)  # some rows start beyond the sequence
    return h, vt, vt_m, w, row_vec, alpha, t0


def run_kernel(h, vt, w, row_vec, alpha, t0, pos0, steer_all, eps, variant, dt, dev, vec_f32=False, ieee=True,
               impl="auto"):
    from introspective_awareness_b200 import kernels

    h_t = to_model(h, dt, dev)
    vt_t = torch.from_numpy(vt).to(dev) if vec_f32 else to_model(vt, dt, dev)
    y = kernels.inject_rmsnorm(
        h_t, to_model(w, dt, dev), eps, variant, vtable=vt_t,
        row_vec=torch.from_numpy(row_vec).to(dev), row_alpha=torch.from_numpy(alpha).to(dev),
        row_t0=torch.from_numpy(t0).to(dev), pos0=pos0, steer_all=steer_all, ieee_rsqrt=ieee, impl=impl)
    torch.cuda.synchronize()
    return h_t.float().cpu().numpy(), y.float().cpu().numpy()


SHAPES = [  # rows, T, H
    (5, 3, 128), (4, 9, 896), (3, 5, 1024), (6, 4, 2048), (4, 3, 3584), (9, 2, 4096), (3, 3, 5120),
    (2, 3, 8192), (2, 2, 12288), (33, 1, 4096), (1, 1, 132),
]


@pytest.mark.parametrize("dt", ["bf16", "f16", "f32"])
@pytest.mark.parametrize("variant", ["llama", "gemma"])
def test_bit_exact_against_oracle(dt, variant):
    dev = _dev()
    for k, (rows, T, H) in enumerate(SHAPES):
        if dt == "f32" and H > 5120:
            continue
        h, vt, vt_m, w, row_vec, alpha, t0 = make_inputs(rows, T, H, dt, 100 + k, gemma=variant == "gemma")
        for steer_all, pos0 in ((False, 0), (True, 7)):
            eps = 1e-5 if variant == "llama" else 1e-6
            want_h, want_y = K.inject_rmsnorm(h, vt_m, row_vec, alpha, t0, pos0, steer_all, w, eps, variant, dt)
            for impl in ("regs", "auto"):
                got_h, got_y = run_kernel(h, vt_m, w, row_vec, alpha, t0, pos0, steer_all, eps, variant, dt, dev, impl=impl)
                assert np.array_equal(got_h, want_h), (dt, variant, rows, T, H, steer_all, impl)
                assert np.array_equal(got_y, want_y), (dt, variant, rows, T, H, steer_all, impl)


@pytest.mark.parametrize("dt", ["bf16", "f16", "f32"])
@pytest.mark.parametrize("variant", ["llama", "gemma"])
def test_tma_staged_variant_bit_exact_against_oracle(dt, variant):
    """IA_IMPL_TMA: per-warp cp.async.bulk + mbarrier ring.  Same bits as the oracle, including long
    grid-stride runs (many ring wrap-arounds per warp) and mixed steered / control / late-t0 rows."""
    dev = _dev()
    shapes = [(5, 3, 128), (4, 9, 896), (7, 5, 1024), (4, 3, 3584), (9, 2, 4096), (3, 3, 5120), (700, 3, 256), (1, 1, 128)]
    for k, (rows, T, H) in enumerate(shapes):
        if dt == "f32" and H > 4096:
            continue
        h, vt, vt_m, w, row_vec, alpha, t0 = make_inputs(rows, T, H, dt, 500 + k, gemma=variant == "gemma")
        for steer_all, pos0 in ((False, 0), (True, 3)):
            eps = 1e-5 if variant == "llama" else 1e-6
            want_h, want_y = K.inject_rmsnorm(h, vt_m, row_vec, alpha, t0, pos0, steer_all, w, eps, variant, dt)
            got_h, got_y = run_kernel(h, vt_m, w, row_vec, alpha, t0, pos0, steer_all, eps, variant, dt, dev, impl="tma")
            assert np.array_equal(got_h, want_h), (dt, variant, rows, T, H, steer_all)
            assert np.array_equal(got_y, want_y), (dt, variant, rows, T, H, steer_all)


def test_tma_variant_equals_register_kernel_at_full_size():
    from introspective_awareness_b200 import kernels

    dev = _dev()
    rows, T, H = 200, 128, 4096
    g = torch.Generator(device=dev).manual_seed(1)
    h = torch.randn(rows, T, H, device=dev, generator=g).to(torch.bfloat16)
    vt = (torch.randn(50, H, device=dev, generator=g) * 0.1).to(torch.bfloat16)
    w = (1 + 0.02 * torch.randn(H, device=dev, generator=g)).to(torch.bfloat16)
    row_vec = (torch.arange(rows, device=dev) % 51 - 1).to(torch.int32)
    alpha = torch.tensor([1.0, 2.0, 4.0, 8.0], device=dev)[torch.arange(rows, device=dev) % 4].contiguous()
    t0 = (100 - torch.arange(rows, device=dev) % 16).to(torch.int32)
    outs = {}
    for impl in ("regs", "tma"):
        hh = h.clone()
        y = kernels.inject_rmsnorm(hh, w, 1e-5, "llama", vtable=vt, row_vec=row_vec, row_alpha=alpha, row_t0=t0, impl=impl)
        torch.cuda.synchronize()
        outs[impl] = (hh, y)
    assert torch.equal(outs["regs"][0], outs["tma"][0]) and torch.equal(outs["regs"][1], outs["tma"][1])


def test_fp32_vector_table_with_bf16_model():
    dev = _dev()
    h, vt, vt_m, w, row_vec, alpha, t0 = make_inputs(6, 4, 4096, "bf16", 7)
    want_h, want_y = K.inject_rmsnorm(h, vt, row_vec, alpha, t0, 0, False, w, 1e-5, "llama", "bf16")
    got_h, got_y = run_kernel(h, vt, w, row_vec, alpha, t0, 0, False, 1e-5, "llama", "bf16", dev, vec_f32=True)
    assert np.array_equal(got_h, want_h) and np.array_equal(got_y, want_y)


def test_plain_rmsnorm_when_no_rows_are_steered():
    from introspective_awareness_b200 import kernels

    dev = _dev()
    h, _, _, w, _, _, _ = make_inputs(8, 2, 3584, "bf16", 9, gemma=True)
    h_t = to_model(h, "bf16", dev)
    keep = h_t.clone()
    y = kernels.inject_rmsnorm(h_t, to_model(w, "bf16", dev), 1e-6, "gemma", ieee_rsqrt=True)
    assert torch.equal(h_t, keep)
    want = K.rmsnorm(h.reshape(-1, 3584), w, 1e-6, "gemma", "bf16").reshape(h.shape)
    assert np.array_equal(y.float().cpu().numpy(), want)


def torch_chain(h_t, sv, starts, pos0, steer_all, norm):
    """The reference hook's ops (model_utils.py:767-788 / :802) followed by the HF norm module."""
    new = h_t.clone()
    rows, T, _ = h_t.shape
    for i in range(rows):
        if sv[i] is None:
            continue
        if steer_all:
            new[i, :, :] += sv[i]
        else:
            s = max(int(starts[i]) - pos0, 0)
            if s < T:
                new[i, s:, :] += sv[i]
    return new, norm(new)


@pytest.mark.parametrize("arch,H", [("llama", 4096), ("llama", 896), ("llama", 5120), ("gemma", 3584)])
@pytest.mark.parametrize("rows,T", [(200, 1), (16, 37), (64, 128)])
def test_bit_identical_to_eager_torch_on_gpu(arch, H, rows, T):
    """Default flags (GPU rsqrtf): same bits as the eager chain the reference runs on a GPU."""
    from transformers.models.gemma2.modeling_gemma2 import Gemma2RMSNorm
    from transformers.models.llama.modeling_llama import LlamaRMSNorm

    from introspective_awareness_b200 import kernels

    dev = _dev()
    dt = "bf16"
    h, vt, vt_m, w, row_vec, alpha, t0 = make_inputs(rows, T, H, dt, 31 + rows + H, gemma=arch == "gemma")
    eps = 1e-5 if arch == "llama" else 1e-6
    norm = (LlamaRMSNorm(H, eps) if arch == "llama" else Gemma2RMSNorm(H, eps)).to(dev).to(torch.bfloat16)
    with torch.no_grad():
        norm.weight.copy_(to_model(w, dt, dev))
    h_t = to_model(h, dt, dev)
    vt_t = to_model(vt_m, dt, dev)
    sv = [None if row_vec[i] < 0 else (vt_t[row_vec[i]] * float(alpha[i])).to(torch.bfloat16) for i in range(rows)]
    steer_all = T == 1
    with torch.no_grad():
        want_h, want_y = torch_chain(h_t, sv, t0, 0, steer_all, norm)
    got_h = h_t.clone()
    got_y = kernels.inject_rmsnorm(got_h, norm.weight.data, eps, arch, vtable=vt_t,
                                   row_vec=torch.from_numpy(row_vec).to(dev),
                                   row_alpha=torch.from_numpy(alpha).to(dev),
                                   row_t0=torch.from_numpy(t0).to(dev), pos0=0, steer_all=steer_all)
    torch.cuda.synchronize()
    assert torch.equal(got_h, want_h)
    mism = (got_y != want_y).float().mean().item()
    assert mism == 0.0, f"{mism:.3e} of y elements differ from the eager chain"


K2_FILES = sorted(os.path.basename(p) for p in glob.glob(os.path.join(GOLDEN, "k2_*.npz")))


@pytest.mark.parametrize("name", K2_FILES)
def test_against_reference_fixtures(name):
    dev = _dev()
    g = load_golden(name)
    dt = str(g["dtype"])
    variant = str(g["variant"])
    vt = widen(g["vectors"], dt)
    w = widen(g["norm_weight"], dt)
    b = vt.shape[0]
    for call in range(3):
        before = widen(g[f"before{call}"], dt)
        after = widen(g[f"after{call}"], dt)
        normed = widen(g[f"normed{call}"], dt)
        steer_all = bool(g["steer_all"]) or before.shape[1] == 1
        got_h, got_y = run_kernel(before, vt, w, np.arange(b, dtype=np.int32), np.full(b, g["alpha"], np.float32),
                                  g["t0_shifted"].astype(np.int32), int(g["pos0"][call]), steer_all, float(g["eps"]),
                                  variant, dt, dev)
        assert np.array_equal(got_h, after), (name, call)
        if dt == "bf16":
            d = ulp_diff_bf16(got_y, normed)
            assert d.max() <= 1 and (d > 0).mean() <= 1e-3, (name, call, d.max(), (d > 0).mean())
        else:
            assert np.allclose(got_y, normed, rtol=4e-6, atol=1e-7), (name, call)


@pytest.mark.parametrize("dt", ["bf16", "f32"])
def test_gemma2_tail_variant_against_oracle(dt):
    """Both implementations: register-resident ("regs") and bulk-copy staged ("tma", the default where it fits:
    MLP-output row staged, residual row through registers), including a run long enough for many ring
    wrap-arounds per warp and, for fp32, a row wider than the register form covers."""
    from introspective_awareness_b200 import kernels

    dev = _dev()
    for k, (rows, T, H) in enumerate([(4, 3, 128), (5, 2, 3584), (3, 2, 1024), (18, 1, 3584), (900, 3, 256)]):
        h, vt, vt_m, w, row_vec, alpha, t0 = make_inputs(rows, T, H, dt, 300 + k, gemma=True)
        rng = np.random.default_rng(k)
        mlp = K.rd(rng.standard_normal((rows, T, H)).astype(np.float32) * 3, dt)
        w_post = K.rd((rng.standard_normal(H) * 0.02).astype(np.float32), dt)
        eps = 1e-6
        base = K.gemma2_block_tail(mlp, h, w_post, eps, dt)
        want_h, want_y = K.inject_rmsnorm(base, vt_m, row_vec, alpha, t0, 0, False, w, eps, "gemma", dt)
        for impl in ("auto", "tma", "regs"):
            if impl == "regs" and dt == "f32" and H > 2048:
                continue  # the register form keeps an fp32 row of at most 2048 elements
            res_t = to_model(h, dt, dev)
            mlp_t = to_model(mlp, dt, dev)
            y = kernels.gemma2_tail_inject_rmsnorm(
                mlp_t, res_t, to_model(w_post, dt, dev), to_model(w, dt, dev), eps,
                vtable=to_model(vt_m, dt, dev), row_vec=torch.from_numpy(row_vec).to(dev),
                row_alpha=torch.from_numpy(alpha).to(dev), row_t0=torch.from_numpy(t0).to(dev), ieee_rsqrt=True, impl=impl)
            torch.cuda.synchronize()
            assert np.array_equal(res_t.float().cpu().numpy(), want_h), (dt, rows, T, H, impl)
            assert np.array_equal(y.float().cpu().numpy(), want_y), (dt, rows, T, H, impl)
            assert torch.equal(mlp_t, to_model(mlp, dt, dev))  # the MLP output is only read


def test_full_size_properties():
    """BASELINE configs[1] prefill shape (200 rows x 128 positions x 4096): too big for the numpy
    oracle in seconds, so check size-independent properties on the GPU."""
    from introspective_awareness_b200 import kernels

    dev = _dev()
    rows, T, H = 200, 128, 4096
    g = torch.Generator(device=dev).manual_seed(0)
    h = torch.randn(rows, T, H, device=dev, generator=g).to(torch.bfloat16)
    vt = (torch.randn(50, H, device=dev, generator=g) * 0.1).to(torch.bfloat16)
    w = (1 + 0.02 * torch.randn(H, device=dev, generator=g)).to(torch.bfloat16)
    row_vec = (torch.arange(rows, device=dev) % 51 - 1).to(torch.int32)  # every 51st row is a control
    alpha = torch.tensor([1.0, 2.0, 4.0, 8.0], device=dev)[torch.arange(rows, device=dev) % 4].contiguous()
    t0 = (100 - torch.arange(rows, device=dev) % 16).to(torch.int32)
    h1 = h.clone()
    y1 = kernels.inject_rmsnorm(h1, w, 1e-5, "llama", vtable=vt, row_vec=row_vec, row_alpha=alpha, row_t0=t0)
    # (a) untouched where the predicate is false
    pos = torch.arange(T, device=dev)[None, :]
    steered = (row_vec[:, None] >= 0) & (pos >= t0[:, None])
    assert torch.equal(h1[~steered], h[~steered])
    # (b) steered positions equal the two-rounding formula
    sv = (vt[row_vec.clamp(min=0).long()] * alpha[:, None]).to(torch.bfloat16)
    want = (h + sv[:, None, :]).to(torch.bfloat16)
    assert torch.equal(h1[steered], want[steered])
    # (c) y is exactly the plain norm of the updated stream (second call injects nothing)
    h2 = h1.clone()
    y2 = kernels.inject_rmsnorm(h2, w, 1e-5, "llama")
    assert torch.equal(h2, h1) and torch.equal(y2, y1)
    # (d) row permutation invariance: rows are independent
    perm = torch.randperm(rows, device=dev, generator=g)
    h3 = h[perm].clone()
    y3 = kernels.inject_rmsnorm(h3, w, 1e-5, "llama", vtable=vt, row_vec=row_vec[perm].contiguous(),
                                row_alpha=alpha[perm].contiguous(), row_t0=t0[perm].contiguous())
    assert torch.equal(h3, h1[perm]) and torch.equal(y3, y1[perm])


@pytest.mark.parametrize("dt", ["bf16", "f16", "f32"])
@pytest.mark.parametrize("variant", ["llama", "gemma"])
def test_residual_add_form_bit_exact_against_oracle(dt, variant):
    """ia_add_inject_rmsnorm: h <- rd(h + add) first (the decoder layer's residual add), then K2 as
    before.  Oracle = the same numpy K2 applied to rd(h + add).  Covers the bulk-copy kernel (H % 128 == 0),
    the two-pass kernel (other H, fp32 table) and un-steered launches, where h must still be rewritten."""
    from introspective_awareness_b200 import kernels

    dev = _dev()
    # (3000, 3, 256) is a long launch (>= 8192 vectors): branch row through registers, one-row stages;
    # the shorter ones stage it by bulk copy; H = 132 / 1000 take the two-pass kernel
    shapes = [(5, 3, 128), (4, 9, 896), (4, 3, 3584), (9, 2, 4096), (3, 3, 5120), (700, 3, 256), (1, 1, 132), (3, 2, 1000),
              (3000, 3, 256), (2100, 4, 1024)]
    for k, (rows, T, H) in enumerate(shapes):
        if dt == "f32" and H > 4096:
            continue
        h, vt, vt_m, w, row_vec, alpha, t0 = make_inputs(rows, T, H, dt, 900 + k, gemma=variant == "gemma")
        add = K.rd(np.random.default_rng(77 + k).standard_normal((rows, T, H)).astype(np.float32) * 0.7, dt)
        summed = K.rd(h + add, dt)
        eps = 1e-5 if variant == "llama" else 1e-6
        for steer_all, pos0, steered in ((False, 0, True), (True, 3, True), (False, 0, False)):
            if steered:
                want_h, want_y = K.inject_rmsnorm(summed, vt_m, row_vec, alpha, t0, pos0, steer_all, w, eps, variant, dt)
            else:
                want_h = summed
                want_y = K.rmsnorm(summed.reshape(rows * T, H), w, eps, variant, dt).reshape(rows, T, H)
            for vec_f32 in ((False, True) if dt == "bf16" and steered else (False,)):
                h_t, add_t = to_model(h, dt, dev), to_model(add, dt, dev)
                kw = {}
                if steered:
                    kw = dict(vtable=torch.from_numpy(vt_m).to(dev) if vec_f32 else to_model(vt_m, dt, dev),
                              row_vec=torch.from_numpy(row_vec).to(dev), row_alpha=torch.from_numpy(alpha).to(dev),
                              row_t0=torch.from_numpy(t0).to(dev), pos0=pos0, steer_all=steer_all)
                for impl in (("auto", "add_staged") if rows * T >= 8192 and not vec_f32 else ("auto",)):
                    h_t = to_model(h, dt, dev)
                    y = kernels.inject_rmsnorm(h_t, to_model(w, dt, dev), eps, variant, ieee_rsqrt=True, add=add_t, impl=impl, **kw)
                    torch.cuda.synchronize()
                    tag = (dt, variant, rows, T, H, steer_all, steered, vec_f32, impl)
                    assert np.array_equal(h_t.float().cpu().numpy(), want_h), tag
                    assert np.array_equal(y.float().cpu().numpy(), want_y), tag
                assert torch.equal(add_t, to_model(add, dt, dev))  # the branch output is only read


@pytest.mark.parametrize("H,rows,T", [(4096, 200, 1), (4096, 64, 128), (896, 16, 37), (5120, 200, 1)])
def test_residual_add_form_bit_identical_to_eager_torch(H, rows, T):
    """Default flags (hardware rsqrt): residual + branch -> hook add -> LlamaRMSNorm, eager on the same GPU."""
    from transformers.models.llama.modeling_llama import LlamaRMSNorm

    from introspective_awareness_b200 import kernels

    dev = _dev()
    g = torch.Generator(device=dev).manual_seed(H + rows)
    resid = torch.randn(rows, T, H, device=dev, generator=g).to(torch.bfloat16)
    branch = (torch.randn(rows, T, H, device=dev, generator=g) * 0.5).to(torch.bfloat16)
    table = (torch.randn(5, H, device=dev, generator=g) * 0.3).to(torch.bfloat16)
    row_vec = torch.arange(rows, device=dev, dtype=torch.int32) % 6 - 1
    alpha = torch.tensor([1.0, 2.0, 4.0, 8.0], device=dev)[torch.arange(rows, device=dev) % 4].float()
    t0 = (torch.arange(rows, device=dev, dtype=torch.int32) * 7) % (T + 1)
    norm = LlamaRMSNorm(H, 1e-5).to(dev).to(torch.bfloat16)
    with torch.no_grad():
        norm.weight.copy_((torch.randn(H, device=dev, generator=g) * 0.02 + 1).to(torch.bfloat16))
        want_h = resid + branch
        for b in range(rows):
            if int(row_vec[b]) >= 0:
                sv = table[int(row_vec[b])] * alpha[b].to(torch.bfloat16)  # the reference scales in the model dtype
                want_h[b, int(t0[b]):] = want_h[b, int(t0[b]):] + sv
        want_y = norm(want_h)
    # alpha values above are exactly representable in bf16, as the reference's strengths are
    h = resid.clone()
    y = kernels.inject_rmsnorm(h, norm.weight.data, 1e-5, "llama", vtable=table, row_vec=row_vec, row_alpha=alpha,
                               row_t0=t0, pos0=0, add=branch)
    torch.cuda.synchronize()
    assert torch.equal(h, want_h)
    assert torch.equal(y, want_y)
