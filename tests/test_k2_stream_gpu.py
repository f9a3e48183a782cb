"""GPU parity of K2's streaming implementation (long launches: row in registers, quarter-row bulk-copy landing
slots) and of the capture tap of every K2 implementation, through the C ABI.

Anchors: the numpy oracle (bit-exact, IA_RSQRT_IEEE) at sizes it finishes in seconds — including runs long
enough for every warp to go round its ring many times — and, at BASELINE configs[1]'s prompt-pass size
(25 600 x 4096), bit equality with the staged and register kernels the other K2 tests pin to the oracle, the
eager torch chain and the reference fixtures.
"""

import numpy as np
import pytest
import torch

from oracle import kernels_np as K
from test_k2_gpu import _dev, make_inputs, to_model

pytestmark = pytest.mark.gpu


def _run(form, dt, dev, h, vt_m, w, row_vec, alpha, t0, pos0, steer_all, eps, variant, impl, other=None, w_post=None,
         steered=True, cap=None):
    from introspective_awareness_b200 import kernels

    h_t = to_model(h, dt, dev)
    kw = {}
    if steered:
        kw = dict(vtable=to_model(vt_m, dt, dev), row_vec=torch.from_numpy(row_vec).to(dev),
                  row_alpha=torch.from_numpy(alpha).to(dev), row_t0=torch.from_numpy(t0).to(dev), pos0=pos0, steer_all=steer_all)
    cap_out = None
    if cap is not None:
        cap_out = torch.full((h.shape[0], h.shape[2]), float("nan"), dtype=h_t.dtype, device=dev)
        kw.update(cap_out=cap_out)
        if isinstance(cap, int):
            kw.update(cap_t=cap)
        else:
            kw.update(cap_tok=torch.from_numpy(cap).to(dev))
    if form == "tail":
        y = kernels.gemma2_tail_inject_rmsnorm(to_model(other, dt, dev), h_t, to_model(w_post, dt, dev), to_model(w, dt, dev), eps,
                                               ieee_rsqrt=True, impl=impl, **kw)
    else:
        y = kernels.inject_rmsnorm(h_t, to_model(w, dt, dev), eps, variant, ieee_rsqrt=True, impl=impl,
                                   add=to_model(other, dt, dev) if form == "add" else None, **kw)
    torch.cuda.synchronize()
    return h_t.float().cpu().numpy(), y.float().cpu().numpy(), None if cap_out is None else cap_out.float().cpu().numpy()


def _oracle(form, dt, h, vt_m, w, row_vec, alpha, t0, pos0, steer_all, eps, variant, other, w_post, steered):
    base = h
    if form == "add":
        base = K.rd(h + other, dt)
    elif form == "tail":
        base = K.gemma2_block_tail(other, h, w_post, eps, dt)
    if steered:
        return K.inject_rmsnorm(base, vt_m, row_vec, alpha, t0, pos0, steer_all, w, eps, variant, dt)
    rows, T, H = h.shape
    return base, K.rmsnorm(base.reshape(rows * T, H), w, eps, variant, dt).reshape(rows, T, H)


# rows, T, H: every supported width; (9000, 4, 1024) = 36 000 vectors -> ~20 trips round each warp's ring
SHAPES = [(7, 3, 1024), (9000, 4, 1024), (40, 50, 2048), (300, 7, 3584), (250, 9, 4096), (200, 7, 5120), (1, 1, 4096)]


@pytest.mark.parametrize("form", ["plain", "add", "tail"])
@pytest.mark.parametrize("dt", ["bf16", "f16"])
def test_streaming_kernel_bit_exact_against_oracle(form, dt):
    dev = _dev()
    for k, (rows, T, H) in enumerate(SHAPES):
        for variant in (("gemma",) if form == "tail" else ("llama", "gemma")):
            if variant == "gemma" and form != "tail" and H not in (1024, 3584):
                continue
            h, vt, vt_m, w, row_vec, alpha, t0 = make_inputs(rows, T, H, dt, 4000 + k, gemma=variant == "gemma")
            rng = np.random.default_rng(50 + k)
            other = K.rd(rng.standard_normal((rows, T, H)).astype(np.float32) * (3.0 if form == "tail" else 0.7), dt) \
                if form != "plain" else None
            w_post = K.rd((rng.standard_normal(H) * 0.02).astype(np.float32), dt) if form == "tail" else None
            eps = 1e-5 if variant == "llama" else 1e-6
            cases = [(False, 0, True), (True, 3, True), (False, 0, False)] if rows * T < 20000 else [(False, 0, True)]
            for steer_all, pos0, steered in cases:
                want_h, want_y = _oracle(form, dt, h, vt_m, w, row_vec, alpha, t0, pos0, steer_all, eps, variant, other, w_post, steered)
                got_h, got_y, _ = _run(form, dt, dev, h, vt_m, w, row_vec, alpha, t0, pos0, steer_all, eps, variant, "stream",
                                       other, w_post, steered)
                tag = (form, dt, variant, rows, T, H, steer_all, steered)
                assert np.array_equal(got_h, want_h), tag
                assert np.array_equal(got_y, want_y), tag


@pytest.mark.parametrize("form", ["plain", "add", "tail"])
def test_streaming_kernel_equals_the_other_kernels_at_full_size(form):
    """BASELINE configs[1] prompt-pass shape (200 rows x 128 positions x 4096; Gemma-2-9B's 3584 for the tail):
    the automatic choice for this launch is the streaming kernel; its output must be the staged kernel's and,
    for the plain form, the register kernel's, bit for bit."""
    from introspective_awareness_b200 import kernels

    dev = _dev()
    rows, T = 200, 128
    H = 3584 if form == "tail" else 4096
    g = torch.Generator(device=dev).manual_seed(0)
    h = torch.randn(rows, T, H, device=dev, generator=g).to(torch.bfloat16)
    other = (torch.randn(rows, T, H, device=dev, generator=g) * (3.0 if form == "tail" else 0.5)).to(torch.bfloat16)
    vt = (torch.randn(50, H, device=dev, generator=g) * 0.1).to(torch.bfloat16)
    w = ((0.0 if form == "tail" else 1.0) + 0.02 * torch.randn(H, device=dev, generator=g)).to(torch.bfloat16)
    w_post = (0.02 * torch.randn(H, device=dev, generator=g)).to(torch.bfloat16)
    row_vec = (torch.arange(rows, device=dev) % 51 - 1).to(torch.int32)
    alpha = torch.tensor([1.0, 2.0, 4.0, 8.0], device=dev)[torch.arange(rows, device=dev) % 4].contiguous()
    t0 = (100 - torch.arange(rows, device=dev) % 16).to(torch.int32)
    outs = {}
    for impl in ("auto", "stream", "tma") + (("regs",) if form == "plain" else ()):
        hh = h.clone()
        cap = torch.empty(rows, H, dtype=torch.bfloat16, device=dev)
        kw = dict(vtable=vt, row_vec=row_vec, row_alpha=alpha, row_t0=t0, impl=impl, cap_out=cap, cap_t=-1)
        if form == "tail":
            y = kernels.gemma2_tail_inject_rmsnorm(other, hh, w_post, w, 1e-6, **kw)
        else:
            y = kernels.inject_rmsnorm(hh, w, 1e-5, "llama", add=other if form == "add" else None, **kw)
        torch.cuda.synchronize()
        assert torch.equal(cap, hh[:, -1]), impl  # the tap wrote the updated last-position rows
        outs[impl] = (hh, y)
    for impl in outs:
        assert torch.equal(outs[impl][0], outs["stream"][0]), impl
        assert torch.equal(outs[impl][1], outs["stream"][1]), impl


@pytest.mark.parametrize("impl", ["auto", "regs", "tma", "stream"])
@pytest.mark.parametrize("form", ["plain", "add", "tail"])
def test_capture_tap_of_every_implementation(impl, form):
    """cap_out[b] = the UPDATED row (b, cap_t) or (b, cap_tok[b]), whichever kernel serves the launch: register,
    two-pass (H = 132), staged and streaming; nothing else changes (h', y as without the tap)."""
    dev = _dev()
    dt = "bf16"
    for k, (rows, T, H) in enumerate([(6, 5, 1024), (5, 4, 4096), (3, 3, 132), (9, 6, 3584)]):
        if impl in ("tma", "stream") and H % 128:
            continue
        if impl == "stream" and H not in (1024, 2048, 3584, 4096, 5120):
            continue
        if form == "tail" and (H % 128 or impl == "regs" and H > 8192):
            continue
        if form == "add" and impl == "regs" and H % 128 == 0:
            pass  # falls through to the two-pass kernel
        variant = "gemma" if form == "tail" else "llama"
        h, vt, vt_m, w, row_vec, alpha, t0 = make_inputs(rows, T, H, dt, 7000 + k, gemma=variant == "gemma")
        rng = np.random.default_rng(9 + k)
        other = K.rd(rng.standard_normal((rows, T, H)).astype(np.float32), dt) if form != "plain" else None
        w_post = K.rd((rng.standard_normal(H) * 0.02).astype(np.float32), dt) if form == "tail" else None
        eps = 1e-6
        want_h, want_y = _oracle(form, dt, h, vt_m, w, row_vec, alpha, t0, 0, False, eps, variant, other, w_post, True)
        tok = rng.integers(0, T, size=rows).astype(np.int32)
        for cap in (-1, 0, T - 2 if T > 1 else 0, tok):
            got_h, got_y, got_cap = _run(form, dt, dev, h, vt_m, w, row_vec, alpha, t0, 0, False, eps, variant, impl, other, w_post,
                                         True, cap=cap)
            pos = tok if not isinstance(cap, int) else np.full(rows, cap % T)
            assert np.array_equal(got_h, want_h) and np.array_equal(got_y, want_y), (impl, form, rows, T, H)
            assert np.array_equal(got_cap, want_h[np.arange(rows), pos]), (impl, form, rows, T, H, cap if isinstance(cap, int) else "tok")


def test_stream_flag_rejects_unsupported_shapes():
    from introspective_awareness_b200 import _lib, kernels

    dev = _dev()
    h = torch.zeros(4, 2, 896, dtype=torch.bfloat16, device=dev)
    w = torch.ones(896, dtype=torch.bfloat16, device=dev)
    with pytest.raises(_lib.KernelLibraryError):
        kernels.inject_rmsnorm(h, w, 1e-5, "llama", impl="stream")  # H = 896 is not a streaming width
    h32 = torch.zeros(4, 2, 4096, dtype=torch.float32, device=dev)
    with pytest.raises(_lib.KernelLibraryError):
        kernels.inject_rmsnorm(h32, torch.ones(4096, device=dev), 1e-5, "llama", impl="stream")  # 16-bit dtypes only
