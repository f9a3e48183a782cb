"""GPU parity of K1 (gather + segmented mean-diff + L2 norm) through the C ABI.

Against the numpy oracle: BIT-EXACT (same stated summation orders) for every mode, dtype,
normalize setting, gathered or compact inputs, chunked (> 256 baseline rows) or not.
Against the reference fixtures: un-normalised bf16 vectors bit-exact; everything
<= 1e-3 relative L2 (north_star's bound).
"""

import numpy as np
import pytest
import torch

from conftest import load_golden, ulp_diff_bf16, widen
from oracle import kernels_np as K

pytestmark = pytest.mark.gpu

TORCH_DT = {"bf16": torch.bfloat16, "f16": torch.float16, "f32": torch.float32}


def _dev():
    if not torch.cuda.is_available():
        pytest.skip("needs a GPU")
    return torch.device("cuda:0")


def to_model(x32, dt, dev):
    return torch.from_numpy(np.ascontiguousarray(x32, dtype=np.float32)).to(dev).to(TORCH_DT[dt])


def residual_batch(n_layers, B, T, H, dt, seed):
    rng = np.random.default_rng(seed)
    x = (rng.standard_normal((n_layers, B, T, H)) * 3 + 5).astype(np.float32)
    x[..., :4] *= 40
    return K.rd(x, dt)


@pytest.mark.parametrize("dt", ["bf16", "f16", "f32"])
@pytest.mark.parametrize("mode", ["baseline", "simple", "raw", "posneg"])
@pytest.mark.parametrize("normalize", [False, True])
def test_gathered_bit_exact_against_oracle(dt, mode, normalize):
    from introspective_awareness_b200 import kernels

    dev = _dev()
    for k, (L, C, Nb, H, Tc, Tb) in enumerate([(3, 5, 100, 896, 6, 9), (2, 50, 100, 4096, 4, 5), (1, 3, 7, 136, 3, 3)]):
        if mode == "simple":
            Nb = 1
        if dt == "f16":
            H = min(H, 896)
        conc = residual_batch(L, C, Tc, H, dt, 10 + k)
        base = residual_batch(L, Nb, Tb, H, dt, 20 + k)
        rng = np.random.default_rng(k)
        ctok = rng.integers(0, Tc, size=C).astype(np.int32)
        btok = np.full(Nb, Tb - 1, np.int32)
        crow = rng.permutation(C).astype(np.int32)
        cl = [to_model(conc[l], dt, dev) for l in range(L)]
        bl = [to_model(base[l], dt, dev) for l in range(L)]
        v32, vr, nr = kernels.gather_meandiff(cl, None if mode == "raw" else bl, mode, normalize,
                                              concept_rows=crow, concept_tok=ctok, base_tok=btok)
        torch.cuda.synchronize()
        for l in range(L):
            ca = conc[l][crow, ctok]
            ba = base[l][np.arange(Nb), btok]
            w32, wr, wn = K.meandiff(ca, ba, {"raw": "raw"}.get(mode, mode), dt, normalize)
            assert np.array_equal(vr[l].float().cpu().numpy(), wr), (dt, mode, normalize, k, l)
            assert np.array_equal(v32[l].cpu().numpy(), w32), (dt, mode, normalize, k, l)
            assert np.array_equal(nr[l].cpu().numpy(), wn), (dt, mode, normalize, k, l)


@pytest.mark.parametrize("dt", ["bf16", "f32"])
def test_chunked_baseline_and_sharded_path(dt):
    """> 256 baseline rows go through the chunked kernels; the rank-local segsum + finalize pair
    (what prompt-sharded extraction uses) must give the same bits as the fused call."""
    from introspective_awareness_b200 import kernels

    dev = _dev()
    L, C, Nb, H, T = 2, 9, 700, 1024, 3
    conc = residual_batch(L, C, T, H, dt, 1)
    base = residual_batch(L, Nb, T, H, dt, 2)
    cl = [to_model(conc[l], dt, dev) for l in range(L)]
    bl = [to_model(base[l], dt, dev) for l in range(L)]
    v32, vr, nr = kernels.gather_meandiff(cl, bl, "baseline", False)
    rows_c, _ = kernels.gather_segsum(cl, want_rows=True)
    _, bsum = kernels.gather_segsum(bl, want_rows=False)
    f32, fr, fn = kernels.meandiff_finalize(rows_c, bsum, Nb, "baseline", False)
    torch.cuda.synchronize()
    for l in range(L):
        w32, wr, wn = K.meandiff(conc[l][:, -1], base[l][:, -1], "baseline", dt, False)
        assert np.array_equal(vr[l].float().cpu().numpy(), wr)
        assert np.array_equal(v32[l].cpu().numpy(), w32)
        assert np.array_equal(nr[l].cpu().numpy(), wn)
        assert np.array_equal(bsum[l].cpu().numpy(), K.column_sum(base[l][:, -1]))
    assert torch.equal(fr, vr) and torch.equal(f32, v32) and torch.equal(fn, nr)
    assert torch.equal(rows_c, torch.stack([c[:, -1] for c in cl]))
    # posneg through the chunked path
    p32, pr, pn = kernels.gather_meandiff(bl, cl, "posneg", True)
    torch.cuda.synchronize()
    w32, wr, wn = K.meandiff(base[0][:, -1], conc[0][:, -1], "posneg", dt, True)
    assert np.array_equal(pr[0].float().cpu().numpy(), wr) and np.array_equal(p32[0].cpu().numpy(), w32)


@pytest.mark.parametrize("name", ["k1_bf16_h512.npz", "k1_f32_h256.npz"])
def test_against_reference_fixtures(name):
    from introspective_awareness_b200 import kernels

    dev = _dev()
    g = load_golden(name)
    dt = str(g["dtype"])
    ca, ba, ctrl = (to_model(widen(g[k], dt), dt, dev).unsqueeze(1) for k in ("concept_acts", "base_acts", "control_act"))
    cases = [("baseline", "baseline", ba), ("simple", "simple", ctrl), ("raw", "no_baseline", None), ("posneg", "posneg", ba)]
    for mode, key, base in cases:
        for norm in (0, 1):
            want = widen(g[f"{key}_{norm}"], dt)
            _, vr, _ = kernels.gather_meandiff([ca], None if base is None else [base], mode, bool(norm))
            got = vr[0].float().cpu().numpy()
            rel = np.linalg.norm(got - want, axis=1) / np.linalg.norm(want, axis=1)
            assert rel.max() <= 1e-3, (mode, norm, rel.max())  # north_star tolerance
            if dt == "bf16" and not norm:
                assert np.array_equal(got, want), (mode, norm)
            elif dt == "bf16":
                d = ulp_diff_bf16(got, want)
                assert d.max() <= 1 and (d > 0).mean() <= 5e-3
            else:
                assert rel.max() <= 1e-6


def test_full_size_properties():
    """Llama-3.1-8B layer sweep shape: 32 layers x (50 + 100) prompts x 4096."""
    from introspective_awareness_b200 import kernels

    dev = _dev()
    L, C, Nb, H, T = 32, 50, 100, 4096, 12
    g = torch.Generator(device=dev).manual_seed(0)
    cl = [(torch.randn(C, T, H, device=dev, generator=g) * 3 + 5).to(torch.bfloat16) for _ in range(L)]
    bl = [(torch.randn(Nb, T, H, device=dev, generator=g) * 3 + 5).to(torch.bfloat16) for _ in range(L)]
    v32, vr, nr = kernels.gather_meandiff(cl, bl, "baseline", False)
    # torch restatement on the GPU: fp32 mean, round, subtract, round (same order: sequential rows)
    for l in (0, 13, 31):
        s = torch.zeros(H, device=dev)
        for r in range(Nb):
            s = s + bl[l][r, -1].float()
        mean = (s / Nb).to(torch.bfloat16)
        want = cl[l][:, -1] - mean
        assert torch.equal(vr[l], want)
        assert torch.allclose(nr[l], want.float().norm(dim=1).to(torch.bfloat16).float(), rtol=2 ** -7)
    # linearity in the concept rows: shifting every concept activation by a representable constant
    # shifts the fp32 vectors by exactly that constant
    cl2 = [c + 2.0 for c in cl]
    v32b, _, _ = kernels.gather_meandiff(cl2, bl, "baseline", False)
    exact = torch.stack([(c2[:, -1].float() - c[:, -1].float()) for c, c2 in zip(cl, cl2)])
    assert torch.allclose(v32b - v32, exact, atol=2e-5)
    # normalised rows have unit norm
    _, vn, _ = kernels.gather_meandiff(cl, bl, "baseline", True)
    assert torch.allclose(vn.float().norm(dim=-1), torch.ones(L, C, device=dev), atol=2e-2)


@pytest.mark.parametrize("dt", ["bf16", "f16", "f32"])
@pytest.mark.parametrize("mode", ["baseline", "simple", "raw", "posneg"])
def test_compact_entry_bit_exact_against_oracle_and_gathered(dt, mode):
    """ia_meandiff_compact (rows as K2's tap leaves them, [L, n, H]) against the numpy oracle and against
    ia_gather_meandiff on the same rows: same bits, for 1..8 shares of the concept rows per (layer, tile), for
    fewer rows than one cp.async group, for the chunked path (> 256 baseline rows) and for widths that are not
    a multiple of the 256-column tile."""
    from introspective_awareness_b200 import kernels

    dev = _dev()
    shapes = [(3, 5, 100, 896), (32, 50, 100, 1024), (1, 50, 100, 4096), (2, 3, 2, 136), (1, 1, 1, 4096), (2, 9, 300, 512),
              (1, 61, 37, 3584)]
    for k, (L, C, Nb, H) in enumerate(shapes):
        if mode == "simple":
            Nb = 1
        if dt == "f16":
            H = min(H, 896)
        conc = residual_batch(L, C, 1, H, dt, 100 + k)[:, :, 0]
        base = residual_batch(L, Nb, 1, H, dt, 200 + k)[:, :, 0]
        ct, bt = to_model(conc, dt, dev), to_model(base, dt, dev)
        for normalize in (False, True):
            v32, vr, nr = kernels.meandiff_compact(ct, None if mode == "raw" else bt, mode, normalize)
            g32, gr, gn = kernels.gather_meandiff([ct[l].unsqueeze(1) for l in range(L)],
                                                  None if mode == "raw" else [bt[l].unsqueeze(1) for l in range(L)], mode, normalize)
            torch.cuda.synchronize()
            assert torch.equal(vr, gr) and torch.equal(v32, g32) and torch.equal(nr, gn), (dt, mode, normalize, k)
            for l in range(L):
                w32, wr, wn = K.meandiff(conc[l], base[l], mode, dt, normalize)
                assert np.array_equal(vr[l].float().cpu().numpy(), wr), (dt, mode, normalize, k, l)
                assert np.array_equal(v32[l].cpu().numpy(), w32), (dt, mode, normalize, k, l)
                assert np.array_equal(nr[l].cpu().numpy(), wn), (dt, mode, normalize, k, l)
        if mode == "baseline":
            # without the norm outputs no norm pass runs and the vectors are the same
            _, vr2, none = kernels.meandiff_compact(ct, bt, mode, False, want_f32=False, want_norm=False)
            assert none is None and torch.equal(vr2, kernels.meandiff_compact(ct, bt, mode, False)[1])
            s = kernels.segsum_compact(bt)
            for l in range(L):
                assert np.array_equal(s[l].cpu().numpy(), K.column_sum(base[l])), (dt, k, l)


@pytest.mark.parametrize("dt", ["bf16", "f16", "f32"])
@pytest.mark.parametrize("mode", ["baseline", "simple", "raw"])
def test_column_walker_bit_exact_against_oracle_and_staged_kernel(dt, mode):
    """Calls with n_layers * H >= 65536 and no norm outputs run the column-walker kernel (a thread per 4 bytes of a
    row, 50 rows per request): the same bits as the oracle and as the staged kernel (the same call WITH norm outputs),
    compact and gathered, for row counts on both sides of a request (49 / 50 / 51 / 101 / 256 rows)."""
    from introspective_awareness_b200 import kernels

    dev = _dev()
    shapes = [(80, 7, 100, 896), (16, 50, 49, 4096), (20, 51, 256, 3584), (70, 101, 50, 1024)]
    for k, (L, C, Nb, H) in enumerate(shapes):
        assert L * H >= 65536
        if mode == "simple":
            Nb = 1
        conc = residual_batch(L, C, 1, H, dt, 300 + k)[:, :, 0]
        base = residual_batch(L, Nb, 1, H, dt, 400 + k)[:, :, 0]
        ct, bt = to_model(conc, dt, dev), to_model(base, dt, dev)
        bt_arg = None if mode == "raw" else bt
        w32, wr, none = kernels.meandiff_compact(ct, bt_arg, mode, False, want_f32=True, want_norm=False)
        s32, sr, _ = kernels.meandiff_compact(ct, bt_arg, mode, False, want_f32=True, want_norm=True)
        _, wr_only, _ = kernels.meandiff_compact(ct, bt_arg, mode, False, want_f32=False, want_norm=False)
        g32, gr, _ = kernels.gather_meandiff([ct[l].unsqueeze(1) for l in range(L)],
                                             None if mode == "raw" else [bt[l].unsqueeze(1) for l in range(L)], mode, False,
                                             want_norm=False)
        torch.cuda.synchronize()
        assert none is None
        assert torch.equal(wr, sr) and torch.equal(w32, s32), (dt, mode, k)
        assert torch.equal(wr_only, sr) and torch.equal(gr, sr) and torch.equal(g32, s32), (dt, mode, k)
        for l in (0, L // 2, L - 1):
            o32, orr, _ = K.meandiff(conc[l], base[l], mode, dt, False)
            assert np.array_equal(wr[l].float().cpu().numpy(), orr), (dt, mode, k, l)
            assert np.array_equal(w32[l].cpu().numpy(), o32), (dt, mode, k, l)
    # gathered, with row and token indices
    L, C, Nb, H, Tc, Tb = 16, 9, 60, 4096, 3, 4
    conc, base = residual_batch(L, C, Tc, H, dt, 500), residual_batch(L, Nb, Tb, H, dt, 501)
    rng = np.random.default_rng(7)
    ctok, btok = rng.integers(0, Tc, size=C).astype(np.int32), rng.integers(0, Tb, size=Nb).astype(np.int32)
    crow = rng.permutation(C).astype(np.int32)
    cl, bl = [to_model(conc[l], dt, dev) for l in range(L)], [to_model(base[l], dt, dev) for l in range(L)]
    if mode == "simple":
        return
    v32, vr, _ = kernels.gather_meandiff(cl, None if mode == "raw" else bl, mode, False, concept_rows=crow, concept_tok=ctok,
                                         base_tok=btok, want_norm=False)
    torch.cuda.synchronize()
    for l in (0, L - 1):
        o32, orr, _ = K.meandiff(conc[l][crow, ctok], base[l][np.arange(Nb), btok], mode, dt, False)
        assert np.array_equal(vr[l].float().cpu().numpy(), orr) and np.array_equal(v32[l].cpu().numpy(), o32), (dt, mode, l)
