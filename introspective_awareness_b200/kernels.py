"""Torch-tensor front end of the C ABI: validates tensors, passes raw pointers + the current
CUDA stream.  PyTorch only owns the memory and the stream here; all arithmetic is in
libia_b200.so.  Every function raises if the inputs are not CUDA tensors — there is no CPU path.
"""

from __future__ import annotations

from typing import Optional, Sequence, Tuple

import torch

from . import _lib
from ._lib import (IA_BF16, IA_F16, IA_F32, IA_IMPL_ADD_STAGED, IA_IMPL_REGS, IA_IMPL_STREAM, IA_IMPL_TMA, IA_MODE_BASELINE,
                   IA_MODE_POSNEG, IA_MODE_RAW, IA_MODE_SIMPLE, IA_NORM_GEMMA, IA_NORM_LLAMA, IA_RSQRT_IEEE, IA_STEER_ALL, check)

_IMPL = {"auto": 0, "tma": IA_IMPL_TMA, "regs": IA_IMPL_REGS, "add_staged": IA_IMPL_ADD_STAGED, "stream": IA_IMPL_STREAM}

_DTYPES = {torch.bfloat16: IA_BF16, torch.float16: IA_F16, torch.float32: IA_F32}
MODES = {"baseline": IA_MODE_BASELINE, "simple": IA_MODE_SIMPLE, "no_baseline": IA_MODE_RAW, "raw": IA_MODE_RAW,
         "posneg": IA_MODE_POSNEG}
VARIANTS = {"llama": IA_NORM_LLAMA, "qwen2": IA_NORM_LLAMA, "gemma": IA_NORM_GEMMA, "gemma2": IA_NORM_GEMMA}


def _dt(t: torch.Tensor) -> int:
    try:
        return _DTYPES[t.dtype]
    except KeyError:
        raise TypeError(f"unsupported dtype {t.dtype}") from None


def _cuda(t: Optional[torch.Tensor], name: str, dtype=None) -> int:
    if t is None:
        return 0
    if not t.is_cuda:
        raise RuntimeError(f"{name} must be a CUDA tensor: the steering/extraction kernels have no CPU path")
    if not t.is_contiguous():
        raise ValueError(f"{name} must be contiguous")
    if dtype is not None and t.dtype != dtype:
        raise TypeError(f"{name} must be {dtype}, got {t.dtype}")
    return t.data_ptr()


def _stream() -> int:
    return torch.cuda.current_stream().cuda_stream


def _tap(cap_out: Optional[torch.Tensor], cap_tok: Optional[torch.Tensor], cap_t: int, rows: int, hid: int, dtype):
    if cap_out is None:
        return 0, 0, 0
    if tuple(cap_out.shape) != (rows, hid):
        raise ValueError("cap_out must be [rows, H]")
    return _cuda(cap_out, "cap_out", dtype), _cuda(cap_tok, "cap_tok", torch.int32), int(cap_t)


def inject_rmsnorm(h: torch.Tensor, weight: torch.Tensor, eps: float, variant: str,
                   vtable: Optional[torch.Tensor] = None, row_vec: Optional[torch.Tensor] = None,
                   row_alpha: Optional[torch.Tensor] = None, row_t0: Optional[torch.Tensor] = None,
                   pos0: int = 0, steer_all: bool = False, out: Optional[torch.Tensor] = None,
                   ieee_rsqrt: bool = False, impl: str = "auto", add: Optional[torch.Tensor] = None,
                   cap_out: Optional[torch.Tensor] = None, cap_tok: Optional[torch.Tensor] = None, cap_t: int = -1
                   ) -> torch.Tensor:
    """K2.  `h` [rows, T, H] is updated in place with the per-row injection; returns y = RMSNorm(h').
    With row_vec None this is a plain RMSNorm of `variant`.  impl: "auto" | "regs" | "tma" | "stream".
    add: a branch output [rows, T, H] that is added to h first (h <- h + add, rounded), the decoder
    layer's residual add folded into the same pass; h is then rewritten for every vector.
    cap_out [rows, H]: capture tap — also receives the UPDATED row of position cap_tok[b] (int32 [rows]) or,
    without cap_tok, position cap_t (-1 = last) of every row."""
    if h.dim() != 3:
        raise ValueError("h must be [rows, T, H]")
    rows, t_len, hid = h.shape
    y = torch.empty_like(h) if out is None else out
    flags = (IA_STEER_ALL if steer_all else 0) | (IA_RSQRT_IEEE if ieee_rsqrt else 0) | _IMPL[impl]
    tap = _tap(cap_out, cap_tok, cap_t, rows, hid, h.dtype)
    lib = _lib.load()
    if add is not None:
        if add.shape != h.shape:
            raise ValueError("add must have the shape of h")
        rc = lib.ia_add_inject_rmsnorm_tap(
            _cuda(h, "h"), _cuda(add, "add", h.dtype), _cuda(y, "out", h.dtype), _cuda(vtable, "vtable"),
            _cuda(row_vec, "row_vec", torch.int32), _cuda(row_alpha, "row_alpha", torch.float32),
            _cuda(row_t0, "row_t0", torch.int32), _cuda(weight, "weight", h.dtype), float(eps), rows, t_len, hid, int(pos0),
            _dt(h), _dt(vtable) if vtable is not None else _dt(h), VARIANTS[variant], flags, *tap, _stream())
        check(rc, "ia_add_inject_rmsnorm")
        return y
    rc = lib.ia_inject_rmsnorm_tap(
        _cuda(h, "h"), _cuda(y, "out", h.dtype), _cuda(vtable, "vtable"), _cuda(row_vec, "row_vec", torch.int32),
        _cuda(row_alpha, "row_alpha", torch.float32), _cuda(row_t0, "row_t0", torch.int32),
        _cuda(weight, "weight", h.dtype), float(eps), rows, t_len, hid, int(pos0), _dt(h),
        _dt(vtable) if vtable is not None else _dt(h), VARIANTS[variant], flags, *tap, _stream())
    check(rc, "ia_inject_rmsnorm")
    return y


def gemma2_tail_inject_rmsnorm(mlp_out: torch.Tensor, residual: torch.Tensor, w_post: torch.Tensor,
                               w_next: torch.Tensor, eps: float, vtable=None, row_vec=None, row_alpha=None,
                               row_t0=None, pos0: int = 0, steer_all: bool = False,
                               out: Optional[torch.Tensor] = None, ieee_rsqrt: bool = False, impl: str = "auto",
                               cap_out: Optional[torch.Tensor] = None, cap_tok: Optional[torch.Tensor] = None,
                               cap_t: int = -1) -> torch.Tensor:
    """K2 Gemma-2 block tail: residual <- inject(residual + gemma_norm(mlp_out; w_post)), returns
    gemma_norm(residual; w_next).  cap_out / cap_tok / cap_t: capture tap as in `inject_rmsnorm`."""
    rows, t_len, hid = residual.shape
    y = torch.empty_like(residual) if out is None else out
    flags = (IA_STEER_ALL if steer_all else 0) | (IA_RSQRT_IEEE if ieee_rsqrt else 0) | _IMPL[impl]
    tap = _tap(cap_out, cap_tok, cap_t, rows, hid, residual.dtype)
    lib = _lib.load()
    rc = lib.ia_gemma2_tail_inject_rmsnorm_tap(
        _cuda(mlp_out, "mlp_out", residual.dtype), _cuda(residual, "residual"), _cuda(y, "out", residual.dtype),
        _cuda(vtable, "vtable"), _cuda(row_vec, "row_vec", torch.int32), _cuda(row_alpha, "row_alpha", torch.float32),
        _cuda(row_t0, "row_t0", torch.int32), _cuda(w_post, "w_post", residual.dtype),
        _cuda(w_next, "w_next", residual.dtype), float(eps), rows, t_len, hid, int(pos0), _dt(residual),
        _dt(vtable) if vtable is not None else _dt(residual), flags, *tap, _stream())
    check(rc, "ia_gemma2_tail_inject_rmsnorm")
    return y


_PTR_TABLES: dict = {}


def _ptr_table(tensors: Sequence[torch.Tensor], name: str) -> torch.Tensor:
    """Device array of the tensors' base addresses.  Cached by value: graph-replayed forwards hand
    back the same buffers every step, so the (synchronous) upload happens once per shape, not per call."""
    ptrs = tuple(_cuda(t, name) for t in tensors)
    key = (tensors[0].device, ptrs)
    table = _PTR_TABLES.get(key)
    if table is None:
        if len(_PTR_TABLES) >= 64:
            _PTR_TABLES.clear()
        table = torch.tensor(ptrs, dtype=torch.int64).to(tensors[0].device, non_blocking=False)
        _PTR_TABLES[key] = table
    return table


def _idx(x, device) -> Optional[torch.Tensor]:
    if x is None:
        return None
    if isinstance(x, torch.Tensor):
        return x.to(device=device, dtype=torch.int32).contiguous()
    return torch.tensor(list(x), dtype=torch.int32, device=device)


def workspace_bytes(n_layers: int, n_concept: int, n_base: int, hidden: int, mode: str = "baseline") -> int:
    n = _lib.load().ia_meandiff_workspace_bytes(n_layers, n_concept, n_base, hidden, MODES[mode])
    if n < 0:
        check(int(n), "ia_meandiff_workspace_bytes")
    return int(n)


def gather_meandiff(concept_layers: Sequence[torch.Tensor], base_layers: Optional[Sequence[torch.Tensor]],
                    mode: str = "baseline", normalize: bool = False, concept_rows=None, concept_tok=None,
                    base_rows=None, base_tok=None, n_concept: Optional[int] = None, n_base: Optional[int] = None,
                    want_f32: bool = True, workspace: Optional[torch.Tensor] = None, want_norm: bool = True
                    ) -> Tuple[Optional[torch.Tensor], torch.Tensor, Optional[torch.Tensor]]:
    """K1.  concept_layers / base_layers: one [B, T, H] residual tensor per selected layer.
    Returns (v_f32 [L, n_out, H] or None, v_rounded [L, n_out, H] model dtype, norm [L, n_out] or None).
    Without want_norm and normalize the norm pass is not launched at all."""
    first = concept_layers[0]
    dev, dtype = first.device, first.dtype
    n_layers, hid = len(concept_layers), first.shape[-1]
    tc = first.shape[1]
    n_concept = first.shape[0] if n_concept is None else n_concept
    if concept_rows is not None:
        n_concept = len(concept_rows)
    imode = MODES[mode]
    if imode == IA_MODE_RAW:
        base_layers, n_base, tb = None, 0, 1
    else:
        if base_layers is None or len(base_layers) != n_layers:
            raise ValueError("base_layers must have one tensor per layer")
        tb = base_layers[0].shape[1]
        n_base = base_layers[0].shape[0] if n_base is None else n_base
        if base_rows is not None:
            n_base = len(base_rows)
    n_out = 1 if imode == IA_MODE_POSNEG else n_concept
    out_f32 = torch.empty((n_layers, n_out, hid), dtype=torch.float32, device=dev) if want_f32 else None
    out_r = torch.empty((n_layers, n_out, hid), dtype=dtype, device=dev)
    out_n = torch.empty((n_layers, n_out), dtype=torch.float32, device=dev) if want_norm else None
    need = workspace_bytes(n_layers, n_concept, n_base, hid, mode)
    if workspace is None or workspace.numel() * workspace.element_size() < need:
        workspace = torch.empty(max(need, 256), dtype=torch.uint8, device=dev)
    cptr = _ptr_table(concept_layers, "concept_layers")
    bptr = _ptr_table(base_layers, "base_layers") if base_layers is not None else None
    crow, ctok = _idx(concept_rows, dev), _idx(concept_tok, dev)
    brow, btok = _idx(base_rows, dev), _idx(base_tok, dev)
    lib = _lib.load()
    rc = lib.ia_gather_meandiff(
        cptr.data_ptr(), tc, _cuda(crow, "concept_rows"), _cuda(ctok, "concept_tok"), n_concept,
        bptr.data_ptr() if bptr is not None else 0, tb, _cuda(brow, "base_rows"), _cuda(btok, "base_tok"), n_base,
        n_layers, hid, imode, int(bool(normalize)), _cuda(out_f32, "out_f32"), out_r.data_ptr(), _cuda(out_n, "out_norm"),
        workspace.data_ptr(), workspace.numel() * workspace.element_size(), _dt(first), _stream())
    check(rc, "ia_gather_meandiff")
    # keep index/pointer tensors alive until the launch is enqueued (it is, the call is synchronous on the host)
    return out_f32, out_r, out_n


def meandiff_compact(concept_acts: torch.Tensor, base_acts: Optional[torch.Tensor], mode: str = "baseline",
                     normalize: bool = False, want_f32: bool = True, want_norm: bool = True,
                     workspace: Optional[torch.Tensor] = None
                     ) -> Tuple[Optional[torch.Tensor], torch.Tensor, Optional[torch.Tensor]]:
    """K1 over compact rows (what K2's tap leaves behind): concept_acts [L, C, H], base_acts [L, Nb, H] or None.
    Returns (v_f32 [L, n_out, H] or None, v_rounded [L, n_out, H] model dtype, norm [L, n_out] or None)."""
    if concept_acts.dim() != 3:
        raise ValueError("concept_acts must be [n_layers, n_concept, H]")
    n_layers, n_concept, hid = concept_acts.shape
    dev, dtype = concept_acts.device, concept_acts.dtype
    imode = MODES[mode]
    n_base = 0
    if imode != IA_MODE_RAW:
        if base_acts is None or base_acts.dim() != 3 or base_acts.shape[0] != n_layers or base_acts.shape[2] != hid:
            raise ValueError("base_acts must be [n_layers, n_base, H]")
        n_base = base_acts.shape[1]
    n_out = 1 if imode == IA_MODE_POSNEG else n_concept
    out_f32 = torch.empty((n_layers, n_out, hid), dtype=torch.float32, device=dev) if want_f32 else None
    out_r = torch.empty((n_layers, n_out, hid), dtype=dtype, device=dev)
    out_n = torch.empty((n_layers, n_out), dtype=torch.float32, device=dev) if want_norm else None
    need = workspace_bytes(n_layers, n_concept, n_base, hid, mode)
    if workspace is None or workspace.numel() * workspace.element_size() < need:
        workspace = torch.empty(max(need, 256), dtype=torch.uint8, device=dev)
    rc = _lib.load().ia_meandiff_compact(
        _cuda(concept_acts, "concept_acts"), n_concept, _cuda(base_acts if imode != IA_MODE_RAW else None, "base_acts", dtype),
        n_base, n_layers, hid, imode, int(bool(normalize)), _cuda(out_f32, "out_f32"), out_r.data_ptr(), _cuda(out_n, "out_norm"),
        workspace.data_ptr(), workspace.numel() * workspace.element_size(), _dt(concept_acts), _stream())
    check(rc, "ia_meandiff_compact")
    return out_f32, out_r, out_n


def segsum_compact(rows: torch.Tensor) -> torch.Tensor:
    """K1a over compact rows [L, n, H]: fp32 column sums [L, H] (the rank-local half of sharded extraction)."""
    n_layers, n_rows, hid = rows.shape
    if n_rows == 0:  # a rank that holds none of the rows contributes an exact zero
        return torch.zeros((n_layers, hid), dtype=torch.float32, device=rows.device)
    out_sum = torch.empty((n_layers, hid), dtype=torch.float32, device=rows.device)
    need = workspace_bytes(n_layers, 0, n_rows, hid, "baseline")
    ws = torch.empty(max(need, 256), dtype=torch.uint8, device=rows.device)
    rc = _lib.load().ia_segsum_compact(_cuda(rows, "rows"), n_rows, n_layers, hid, out_sum.data_ptr(), ws.data_ptr(), ws.numel(),
                                       _dt(rows), _stream())
    check(rc, "ia_segsum_compact")
    return out_sum


def gather_segsum(layers: Sequence[torch.Tensor], rows=None, tok=None, n_rows: Optional[int] = None,
                  want_rows: bool = True) -> Tuple[Optional[torch.Tensor], torch.Tensor]:
    """K1a (rank-local): returns (rows [L, n, H] model dtype or None, column sums [L, H] fp32)."""
    first = layers[0]
    dev, hid, n_layers = first.device, first.shape[-1], len(layers)
    n_rows = first.shape[0] if n_rows is None else n_rows
    if rows is not None:
        n_rows = len(rows)
    out_rows = torch.empty((n_layers, n_rows, hid), dtype=first.dtype, device=dev) if want_rows else None
    out_sum = torch.empty((n_layers, hid), dtype=torch.float32, device=dev)
    need = workspace_bytes(n_layers, 0, n_rows, hid, "baseline")
    ws = torch.empty(max(need, 256), dtype=torch.uint8, device=dev)
    ptrs = _ptr_table(layers, "layers")
    ridx, tidx = _idx(rows, dev), _idx(tok, dev)
    rc = _lib.load().ia_gather_segsum(ptrs.data_ptr(), first.shape[1], _cuda(ridx, "rows"), _cuda(tidx, "tok"), n_rows,
                                      n_layers, hid, _cuda(out_rows, "out_rows"), out_sum.data_ptr(), ws.data_ptr(),
                                      ws.numel(), _dt(first), _stream())
    check(rc, "ia_gather_segsum")
    return out_rows, out_sum


def meandiff_finalize(concept_acts: torch.Tensor, base_sum: Optional[torch.Tensor], n_base_total: int,
                      mode: str = "baseline", normalize: bool = False, want_f32: bool = True
                      ) -> Tuple[Optional[torch.Tensor], torch.Tensor, torch.Tensor]:
    """K1b: concept_acts [L, C, H] (model dtype) and all-reduced base_sum [L, H] fp32."""
    n_layers, n_concept, hid = concept_acts.shape
    dev = concept_acts.device
    out_f32 = torch.empty((n_layers, n_concept, hid), dtype=torch.float32, device=dev) if want_f32 else None
    out_r = torch.empty_like(concept_acts)
    out_n = torch.empty((n_layers, n_concept), dtype=torch.float32, device=dev)
    need = workspace_bytes(n_layers, n_concept, 0, hid, mode)
    ws = torch.empty(max(need, 256), dtype=torch.uint8, device=dev)
    rc = _lib.load().ia_meandiff_finalize(
        _cuda(concept_acts, "concept_acts"), n_concept, _cuda(base_sum, "base_sum", torch.float32), int(n_base_total),
        n_layers, hid, MODES[mode], int(bool(normalize)), _cuda(out_f32, "out_f32"), out_r.data_ptr(),
        out_n.data_ptr(), ws.data_ptr(), ws.numel(), _dt(concept_acts), _stream())
    check(rc, "ia_meandiff_finalize")
    return out_f32, out_r, out_n


def rope(q: torch.Tensor, k: torch.Tensor, cos: torch.Tensor, sin: torch.Tensor):
    """Fused rotary embedding.  q [B, n_heads, T, D] and k [B, n_kv, T, D] must be transposed views of
    contiguous [B, T, heads, D] tensors (what the HF projections produce); cos/sin [B, T, D] contiguous.
    Returns (q_out, k_out) with the same shapes and strides.  Returns None when the layout is different
    (the caller then uses the stock function)."""
    if not (q.is_cuda and q.dim() == 4 and k.dim() == 4 and q.dtype in _DTYPES and k.dtype == q.dtype
            and cos.dtype == q.dtype and sin.dtype == q.dtype):
        return None
    b, nh, t, d = q.shape
    nkv = k.shape[1]
    if d % 16 or k.shape != (b, nkv, t, d) or tuple(cos.shape) != (b, t, d) or tuple(sin.shape) != (b, t, d):
        return None
    if q.stride() != (t * nh * d, d, nh * d, 1) or k.stride() != (t * nkv * d, d, nkv * d, 1):
        return None
    if not (cos.is_contiguous() and sin.is_contiguous()):
        return None
    qo, ko = torch.empty_like(q), torch.empty_like(k)
    if qo.stride() != q.stride() or ko.stride() != k.stride():
        return None
    rc = _lib.load().ia_rope(q.data_ptr(), k.data_ptr(), cos.data_ptr(), sin.data_ptr(), qo.data_ptr(), ko.data_ptr(),
                             b * t, nh, nkv, d, _dt(q), _stream())
    check(rc, "ia_rope")
    return qo, ko


def _rows_view(t: torch.Tensor, width: int, name: str) -> Tuple[int, int]:
    """(pointer, row stride in elements) of a [..., width] tensor whose leading dims collapse to rows with
    one stride — a dense tensor or a column slice of one (e.g. part of a fused projection output)."""
    if not t.is_cuda:
        raise RuntimeError(f"{name} must be a CUDA tensor")
    if t.shape[-1] != width or t.stride(-1) != 1:
        raise ValueError(f"{name}: last dim must be {width} and dense")
    rs = t.stride(-2) if t.dim() >= 2 else width
    for d in range(t.dim() - 2):
        if t.shape[d] != 1 and t.stride(d) != t.stride(d + 1) * t.shape[d + 1]:
            raise ValueError(f"{name}: rows do not have a single stride")
    return t.data_ptr(), rs


def rope_append(q: torch.Tensor, k: torch.Tensor, v: torch.Tensor, cos: torch.Tensor, sin: torch.Tensor,
                k_dst: torch.Tensor, v_dst: torch.Tensor, pos0: int, n_heads: int, n_kv: int, head_dim: int) -> torch.Tensor:
    """Rotary embedding of q and k and the KV append in one launch.  q [B, T, n_heads*D], k / v
    [B, T, n_kv*D] (dense, or column slices of one fused projection output); cos / sin [B, T, D] dense;
    k_dst / v_dst [B, n_kv, L, D] (any batch/head/position strides, D dense): positions pos0 .. pos0+T-1
    are written.  Returns rotated q as a [B, n_heads, T, D] view of a dense [B, T, n_heads, D] tensor."""
    b, t = q.shape[0], q.shape[1]
    qp, qrs = _rows_view(q, n_heads * head_dim, "q")
    kp, krs = _rows_view(k, n_kv * head_dim, "k")
    vp, vrs = _rows_view(v, n_kv * head_dim, "v")
    if tuple(cos.shape) != (b, t, head_dim) or tuple(sin.shape) != (b, t, head_dim):
        raise ValueError("cos/sin must be [B, T, D]")
    if k_dst.shape != v_dst.shape or k_dst.stride() != v_dst.stride() or k_dst.stride(3) != 1:
        raise ValueError("k_dst / v_dst must share a layout with a dense head dimension")
    if k_dst.shape[0] != b or k_dst.shape[1] != n_kv or k_dst.shape[3] != head_dim or pos0 + t > k_dst.shape[2]:
        raise ValueError("destination does not fit")
    q_out = torch.empty((b, t, n_heads, head_dim), dtype=q.dtype, device=q.device)
    rc = _lib.load().ia_rope_append(qp, kp, vp, qrs, krs, vrs, _cuda(cos, "cos", q.dtype), _cuda(sin, "sin", q.dtype),
                                    q_out.data_ptr(), k_dst.data_ptr(), v_dst.data_ptr(), k_dst.stride(0), k_dst.stride(1),
                                    k_dst.stride(2), b * t, t, int(pos0), n_heads, n_kv, head_dim, _dt(q), _stream())
    check(rc, "ia_rope_append")
    return q_out.transpose(1, 2)


def silu_mul(gate: torch.Tensor, up: torch.Tensor) -> torch.Tensor:
    """round(round(silu(gate)) * up) in one pass; gate / up [..., I] dense or column slices of one tensor."""
    width = gate.shape[-1]
    gp, grs = _rows_view(gate, width, "gate")
    up_p, urs = _rows_view(up, width, "up")
    if up.shape != gate.shape or up.dtype != gate.dtype:
        raise ValueError("gate and up must match")
    out = torch.empty(gate.shape, dtype=gate.dtype, device=gate.device)
    rows = out.numel() // width
    rc = _lib.load().ia_silu_mul(gp, up_p, out.data_ptr(), rows, width, grs, urs, _dt(gate), _stream())
    check(rc, "ia_silu_mul")
    return out


def activation_table(act_fn, dtype: torch.dtype, device) -> torch.Tensor:
    """act_fn of every one of the 65 536 values of a 16-bit dtype, indexed by bit pattern: the table `act_mul_table`
    looks the gate activation up in.  Computed by the model's own activation module on the device, so its entries are
    that library's bits (NaN patterns included)."""
    if dtype not in (torch.bfloat16, torch.float16):
        raise TypeError("activation tables enumerate a 16-bit dtype")
    bits = torch.arange(65536, dtype=torch.int32, device=device).to(torch.int16)  # 0..65535 wraps to the same bit patterns
    with torch.no_grad():
        return act_fn(bits.view(dtype)).contiguous()


def act_mul_table(gate: torch.Tensor, up: torch.Tensor, table: torch.Tensor) -> torch.Tensor:
    """round(table[gate] * up) in one pass; gate / up [..., I] dense or column slices of one tensor, 16-bit dtype."""
    width = gate.shape[-1]
    gp, grs = _rows_view(gate, width, "gate")
    up_p, urs = _rows_view(up, width, "up")
    if up.shape != gate.shape or up.dtype != gate.dtype or table.dtype != gate.dtype or table.numel() != 65536:
        raise ValueError("gate, up and the 65536-entry table must share a 16-bit dtype")
    out = torch.empty(gate.shape, dtype=gate.dtype, device=gate.device)
    rows = out.numel() // width
    rc = _lib.load().ia_act_mul_table(gp, up_p, _cuda(table, "table"), out.data_ptr(), rows, width, grs, urs, _dt(gate), _stream())
    check(rc, "ia_act_mul_table")
    return out
