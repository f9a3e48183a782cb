"""ORACLE — test infrastructure, not product code.

CPU (numpy) restatement of the arithmetic of the two hot-path kernels, written against the
reference's rounding points so that results can be compared bit for bit.  Only `tests/`,
`__graft_entry__.smoke()` and `bench.py`'s CPU-baseline leg may import this module; nothing
under `introspective_awareness_b200/` does.

Parity status: PINNED.  `tests/test_oracle_golden.py` checks every function here against
fixtures in `tests/golden/` that were produced by running the reference's own
`model_utils.py` / `vector_utils.py` (unmodified, imported from /root/reference) in the build
container; `oracle/make_golden.py` is the generating script.

What follows the reference, by function:

* `meandiff`           — reference vector_utils.py:541-548 (baseline), :570-572 (simple),
                         :586-588 (no_baseline), :103-109 (pos/neg mean difference).
* `inject`             — reference model_utils.py:737 (scale + cast) and :774-788 (the add).
* `rmsnorm`            — transformers `LlamaRMSNorm.forward` / `Qwen2RMSNorm.forward` /
                         `Gemma2RMSNorm.forward` (third-party, pinned 4.57.1 in the reference's
                         uv.lock:3088-3089; 5.5.0 installed here; the three forwards are
                         unchanged between them).  Reached implicitly through
                         reference model_utils.py:338 and :847.
* `inject_rmsnorm`     — the fusion of the two.

All floating-point work is done in numpy float32 with explicit rounding to the model dtype
(bf16 / fp16 / fp32), never in float64, so every intermediate matches what a GPU thread computes.
The summation ORDERS are part of the contract and are stated on each function; the CUDA
kernels implement the same orders so that kernel-vs-oracle comparisons are exact.
"""

from __future__ import annotations

from typing import Optional, Tuple

import numpy as np

F32 = np.float32

# ---------------------------------------------------------------------------
# dtype emulation
# ---------------------------------------------------------------------------


def bf16_bits(x: np.ndarray) -> np.ndarray:
    """float32 -> bf16 bit pattern (uint16), round-to-nearest-even; finite inputs."""
    u = np.ascontiguousarray(x, dtype=F32).view(np.uint32)
    bias = ((u >> np.uint32(16)) & np.uint32(1)) + np.uint32(0x7FFF)
    return ((u + bias) >> np.uint32(16)).astype(np.uint16)


def bf16_to_f32(bits: np.ndarray) -> np.ndarray:
    return (bits.astype(np.uint32) << np.uint32(16)).view(F32)


def rd(x: np.ndarray, dtype: str) -> np.ndarray:
    """Round a float32 array to `dtype` and return it widened back to float32."""
    x = np.asarray(x, dtype=F32)
    if dtype == "f32":
        return x
    if dtype == "bf16":
        return bf16_to_f32(bf16_bits(x))
    if dtype == "f16":
        return x.astype(np.float16).astype(F32)
    raise ValueError(dtype)


# ---------------------------------------------------------------------------
# K1: gather + segmented mean-difference + L2 norm
# ---------------------------------------------------------------------------

SUM_CHUNK = 256  # rows per sequential chunk in the canonical column-sum order
TILE_COLS = 256  # columns per norm partial


def column_sum(rows: np.ndarray) -> np.ndarray:
    """fp32 column sums of `rows` [N, H] (values already widened to float32).

    Canonical order: rows are cut into chunks of SUM_CHUNK consecutive rows; inside a chunk the
    rows are added one after another starting from the first row; chunk sums are then added in
    ascending chunk order.  For N <= 256 (the reference's 100 baseline words) this is the plain
    sequential sum, which is also what torch's CPU `sum(dim=0)` produces for this shape.
    """
    rows = np.asarray(rows, dtype=F32)
    n = rows.shape[0]
    total = None
    for c0 in range(0, n, SUM_CHUNK):
        part = rows[c0].copy()
        for r in range(c0 + 1, min(c0 + SUM_CHUNK, n)):
            part = part + rows[r]
        total = part if total is None else total + part
    if total is None:
        total = np.zeros(rows.shape[1:], dtype=F32)
    return total.astype(F32)


def _sq_norm(v: np.ndarray) -> np.ndarray:
    """Σ v² over the last axis in the kernel's order: per 256-column tile, each of 32 lanes owns
    8 consecutive columns (lane l -> columns 8l..8l+7 of the tile) and accumulates them left to
    right with fused multiply-adds; lanes combine by xor-butterfly (16, 8, 4, 2, 1); tiles are
    added in ascending order.  Short last tiles behave as if zero padded."""
    v = np.asarray(v, dtype=F32)
    h = v.shape[-1]
    pad = (-h) % TILE_COLS
    if pad:
        v = np.concatenate([v, np.zeros(v.shape[:-1] + (pad,), dtype=F32)], axis=-1)
    t = v.reshape(v.shape[:-1] + (-1, 32, 8))
    acc = np.zeros(t.shape[:-1], dtype=F32)
    for i in range(8):
        # fmaf(x, x, acc): one rounding.  float64 holds x*x + acc exactly enough: the product of
        # two float32 is exact in float64 and the sum is then rounded once to float32.
        acc = (t[..., i].astype(np.float64) * t[..., i].astype(np.float64) + acc.astype(np.float64)).astype(F32)
    for off in (16, 8, 4, 2, 1):
        idx = np.arange(32) ^ off
        acc = acc + acc[..., idx]
    tiles = acc[..., 0]
    total = tiles[..., 0].copy()
    for k in range(1, tiles.shape[-1]):
        total = total + tiles[..., k]
    return total.astype(F32)


def meandiff(concept: np.ndarray, base: Optional[np.ndarray], mode: str, dtype: str,
             normalize: bool = False, base_sum: Optional[np.ndarray] = None,
             n_base: Optional[int] = None):
    """Concept vectors from captured activations.

    concept [C, H], base [Nb, H]: activations already in the model dtype, widened to float32.
    mode: "baseline" | "simple" (same arithmetic, Nb == 1) | "raw" | "posneg".
    Returns (v_f32, v_rounded, norm_rounded):
      v_rounded  — the reference's result: mean accumulated in fp32 and rounded to the model
                   dtype, difference rounded to the model dtype (reference vector_utils.py:541-545).
      v_f32      — the same difference without either rounding (north_star's fp32 path).
      norm_rounded — ‖v_rounded‖ accumulated in fp32 and rounded to the model dtype, taken before
                   any normalisation (reference vector_utils.py:547 `vec.norm()`).
    With normalize, both vectors are divided by (norm + 1e-8) (reference vector_utils.py:547).
    `base_sum`/`n_base` let a caller supply an already reduced (e.g. all-reduced) column sum.
    """
    concept = np.asarray(concept, dtype=F32)
    if mode == "raw":
        v32 = concept.copy()
        vr = concept.copy()
    else:
        if base_sum is None:
            base_sum = column_sum(base)
            n_base = base.shape[0]
        mean32 = (np.asarray(base_sum, dtype=F32) / F32(n_base)).astype(F32)
        mean_r = rd(mean32, dtype)
        if mode == "posneg":
            pos32 = (column_sum(concept) / F32(concept.shape[0])).astype(F32)
            v32 = (pos32 - mean32)[None, :]
            vr = rd(rd(pos32, dtype) - mean_r, dtype)[None, :]
        elif mode in ("baseline", "simple"):
            v32 = (concept - mean32[None, :]).astype(F32)
            vr = rd(concept - mean_r[None, :], dtype)
        else:
            raise ValueError(mode)
    norm32 = np.sqrt(_sq_norm(v32)).astype(F32)
    norm_r = rd(np.sqrt(_sq_norm(vr)).astype(F32), dtype)
    if normalize:
        v32 = (v32 / (norm32 + F32(1e-8))[:, None]).astype(F32)
        den = rd(norm_r + F32(1e-8), dtype)
        vr = rd(vr / den[:, None], dtype)
    return v32.astype(F32), vr.astype(F32), norm_r.astype(F32)


# ---------------------------------------------------------------------------
# K2: per-row injection fused with the next RMSNorm
# ---------------------------------------------------------------------------


def inject(h: np.ndarray, vtable: np.ndarray, row_vec: np.ndarray, row_alpha: np.ndarray,
           row_t0: Optional[np.ndarray], pos0: int, steer_all: bool, dtype: str) -> np.ndarray:
    """h [B, T, H] (model dtype, widened).  Row b, position p receives
    RD(h + RD(v[row_vec[b]] * alpha[b])) when row_vec[b] >= 0 and (steer_all or
    p + pos0 >= row_t0[b]); otherwise it is left untouched.
    (reference model_utils.py:737 for the scaling, :774-788 for the predicate and the add.)"""
    h = np.asarray(h, dtype=F32)
    out = h.copy()
    b, t, _ = h.shape
    pos = np.arange(t) + int(pos0)
    for i in range(b):
        vi = int(row_vec[i])
        if vi < 0:
            continue
        sv = rd(np.asarray(vtable[vi], dtype=F32) * F32(row_alpha[i]), dtype)
        sel = np.ones(t, dtype=bool) if steer_all else pos >= int(row_t0[i])
        out[i, sel] = rd(h[i, sel] + sv[None, :], dtype)
    return out


def row_sum_squares(x: np.ndarray) -> np.ndarray:
    """Σ x² over the last axis of x [R, H] (float32), H % 4 == 0, in the kernel's order, which is
    the order ATen's CUDA `mean(-1)` uses for a contiguous fp32 [R >= 16, 128 <= H <= 8191] tensor
    (torch 2.11 ATen/native/cuda/Reduce.cuh: one warp per row, 4-wide vector loads): squares are
    rounded products; lane l owns the 4-element groups l, l+32, l+64, … and keeps one
    accumulator per element slot, adding groups in ascending order; the four accumulators are
    combined ((a0+a1)+a2)+a3; lanes combine by shuffle-down with offsets 16, 8, 4, 2, 1."""
    x = np.asarray(x, dtype=F32)
    r, h = x.shape
    assert h % 4 == 0
    sq = (x * x).astype(F32)
    g = h // 4
    per_lane = -(-g // 32)
    pad = per_lane * 32 - g
    sq = sq.reshape(r, g, 4)
    if pad:
        sq = np.concatenate([sq, np.zeros((r, pad, 4), dtype=F32)], axis=1)
    sq = sq.reshape(r, per_lane, 32, 4)
    acc = sq[:, 0].copy()
    for j in range(1, per_lane):
        acc = acc + sq[:, j]
    s = ((acc[..., 0] + acc[..., 1]) + acc[..., 2]) + acc[..., 3]  # [R, 32]
    for off in (16, 8, 4, 2, 1):
        shifted = np.concatenate([s[:, off:], s[:, -off:]], axis=1)  # lanes past the end keep junk
        s = s + shifted
    return s[:, 0].astype(F32)


def rmsnorm(x: np.ndarray, w: np.ndarray, eps: float, variant: str, dtype: str,
            rsqrt: str = "ieee") -> np.ndarray:
    """x [R, H] already in the model dtype (widened).  variant "llama" (also Qwen2):
    y = RD(w * RD(x * r)); variant "gemma": y = RD((x * r) * (1 + w)); r = rsqrt(mean(x²) + eps)
    in fp32.  rsqrt="ieee" is 1/sqrt with correctly rounded division and square root (what torch's
    CPU kernels and the CUDA kernel's IA_RSQRT_IEEE flag compute); the GPU's default `rsqrtf`
    differs from it by at most 2 ulp and cannot be reproduced on a CPU."""
    x = np.asarray(x, dtype=F32)
    w = np.asarray(w, dtype=F32)
    h = x.shape[-1]
    var = (row_sum_squares(x) * (F32(1.0) / F32(h))).astype(F32)
    assert rsqrt == "ieee"
    r = (F32(1.0) / np.sqrt((var + F32(eps)).astype(F32))).astype(F32)
    n = (x * r[:, None]).astype(F32)
    if variant == "llama":
        return rd(w[None, :] * rd(n, dtype), dtype)
    if variant == "gemma":
        return rd(n * (F32(1.0) + w)[None, :].astype(F32), dtype)
    raise ValueError(variant)


def inject_rmsnorm(h, vtable, row_vec, row_alpha, row_t0, pos0, steer_all, w, eps, variant, dtype
                   ) -> Tuple[np.ndarray, np.ndarray]:
    """Fused K2: returns (h', y) with h' as `inject` and y = rmsnorm(h')."""
    hp = inject(h, vtable, row_vec, row_alpha, row_t0, pos0, steer_all, dtype)
    b, t, hd = hp.shape
    y = rmsnorm(hp.reshape(b * t, hd), w, eps, variant, dtype).reshape(b, t, hd)
    return hp, y


def gemma2_block_tail(mlp_out, residual, w_post, eps, dtype):
    """Gemma-2 end of layer: h = RD(residual + gemma_norm(mlp_out; w_post))
    (transformers Gemma2DecoderLayer.forward: post_feedforward_layernorm then residual add)."""
    b, t, hd = mlp_out.shape
    n = rmsnorm(np.asarray(mlp_out, F32).reshape(b * t, hd), w_post, eps, "gemma", dtype).reshape(b, t, hd)
    return rd(np.asarray(residual, F32) + n, dtype)
