"""Decode runtime around the two kernels: one process per GPU, static KV cache, one CUDA graph per
decode step, the steering kernel (K2) sitting where the reference had a forward hook.

What is stock and what is ours (north_star: "model GEMMs left to stock cuBLAS"):
  * stock: every Hugging Face sub-module of the decoder layer — q/k/v/o/MLP `nn.Linear` (cuBLAS),
    rotary embedding, SDPA attention, the RMSNorms that are NOT fused — called exactly as
    `LlamaDecoderLayer.forward` / `Qwen2DecoderLayer.forward` / `Gemma2DecoderLayer.forward` call them,
    so the numerics outside the hot path are the reference's;
  * ours: the layer loop (no `model.generate`, no hooks, no per-row Python loops), the KV cache
    (preallocated, exact-length views, so attention never touches masked tail slots), per-step CUDA
    graphs, and at layer L the fused injection + next-RMSNorm kernel in place of
    reference model_utils.py:750-808 + the next layer's `input_layernorm` (or `model.norm`).

The reference drives `model.generate` with a hook (reference model_utils.py:811-847); the forward
hook fires once for the prompt and once per cached decode step, which is exactly
the call pattern of `Session.generate` below.

Like the reference's wrapper (hooks kept on the instance, model_utils.py:136) an engine is not
re-entrant: while one of its forwards runs it points the model's attention-function name and the
modelling module's `apply_rotary_pos_emb` at its own implementations and restores both afterwards,
so use one engine per model object and one thread per engine.
"""

from __future__ import annotations

import contextlib
import gc
from dataclasses import dataclass
from typing import Dict, List, Optional, Sequence, Tuple

import torch

from . import kernels


@contextlib.contextmanager
def _no_gc_during_capture():
    """Keep Python's cyclic collector out of a stream capture.  An older Session (graphs, pooled
    buffers) that is only reachable through a reference cycle is finalised whenever the collector
    happens to run; if that is inside a capture, destroying its CUDA graphs is an operation the
    capture forbids and the capture is invalidated (torch.cuda.graph no longer collects up front).
    So: collect once before capturing, and hold the collector off until the capture has ended."""
    was_enabled = gc.isenabled()
    gc.collect()
    gc.disable()
    try:
        yield
    finally:
        if was_enabled:
            gc.enable()


ATTN_GQA = "ia_b200_gqa_sdpa"


def _gqa_sdpa_forward(module, query, key, value, attention_mask, dropout: float = 0.0, scaling=None, **kwargs):
    """Attention function for HF's registry: torch SDPA with native grouped-query support.

    transformers' stock `sdpa_attention_forward` materialises K and V `num_key_value_groups` times
    (`repeat_kv`) whenever a mask is present, i.e. for every left-padded batch — at B=200 that is as
    many bytes per decode step as the model weights.  Same math, K/V read once."""
    b, nh, t, d = query.shape
    nkv = key.shape[1]
    if nkv != nh and d > 128:
        return _gqa_folded(query, key, value, attention_mask, scaling), None
    causal = t > 1 and attention_mask is None
    out = torch.nn.functional.scaled_dot_product_attention(
        query, key, value, attn_mask=attention_mask, dropout_p=0.0, scale=scaling, is_causal=causal,
        enable_gqa=nkv != nh)
    return out.transpose(1, 2).contiguous(), None


def _fold_mask(m: torch.Tensor, g: int) -> torch.Tensor:
    """[B, 1, T, L] -> [B, 1, g*T, L]: the same T rows once per query head of the group."""
    b, one, t, L = m.shape
    return m[:, :, None].expand(b, one, g, t, L).reshape(b, one, g * t, L)


def _gqa_folded(query, key, value, attention_mask, scaling):
    """Grouped-query attention for head sizes torch's grouped-query kernels do not take (Gemma-2: 256).
    The g query heads that share a KV head are laid side by side along the QUERY-LENGTH axis of that KV
    head — q [B, n_kv, g*T, D] against K/V [B, n_kv, L, D] — so the ordinary fused kernel runs with K/V read
    in place.  Every query row still sees the same keys in the same order through the same kernel, which is
    why the result equals the repeat_kv form; what disappears is the g-fold copy of K and V per layer.
    The engine hands over prompt-pass masks already folded (once per forward); an unfolded one is folded here."""
    b, nh, t, d = query.shape
    nkv, L = key.shape[1], key.shape[2]
    g = nh // nkv
    q = query.reshape(b, nkv, g * t, d)
    m = attention_mask
    if m is None and t > 1:  # causal without padding: spell the mask out, the folded rows are not in causal order
        m = torch.ones(t, L, dtype=torch.bool, device=query.device).tril(L - t)[None, None]
        m = torch.where(m, torch.zeros((), dtype=query.dtype, device=query.device),
                        torch.full((), float("-inf"), dtype=query.dtype, device=query.device))
    if m is not None and t > 1 and m.shape[2] == t:
        m = _fold_mask(m, g)
    out = torch.nn.functional.scaled_dot_product_attention(q, key, value, attn_mask=m, dropout_p=0.0, scale=scaling,
                                                           is_causal=False)
    # [B, n_kv, g, T, D] -> [B, T, n_kv, g, D] = [B, T, n_heads, D]
    return out.view(b, nkv, g, t, d).permute(0, 3, 1, 2, 4).reshape(b, t, nh, d).contiguous()


def _register_attention():
    from transformers import AttentionInterface

    AttentionInterface.register(ATTN_GQA, _gqa_sdpa_forward)


class _FusedLinear:
    """Several nn.Linear modules that read the same input, as one GEMM.  The weights are concatenated
    once and the modules' own parameters re-pointed at row slices of that buffer, so nothing is stored
    twice and the stock HF forward (the reference hook path in the parity tests) keeps working on the
    same values."""

    def __init__(self, mods):
        self.mods = list(mods)
        self.sizes = [m.weight.shape[0] for m in self.mods]
        with torch.no_grad():
            self.W = torch.cat([m.weight.data for m in self.mods], dim=0).contiguous()
            has_bias = self.mods[0].bias is not None
            self.b = torch.cat([m.bias.data for m in self.mods], dim=0).contiguous() if has_bias else None
            off = 0
            for m, n in zip(self.mods, self.sizes):
                m.weight.data = self.W[off:off + n]
                if has_bias:
                    m.bias.data = self.b[off:off + n]
                off += n

    def __call__(self, x: torch.Tensor):
        return torch.nn.functional.linear(x, self.W, self.b).split(self.sizes, dim=-1)


@dataclass
class SteerSpec:
    """Per-row steering for one batch: row b adds alpha[b] * vtable[row_vec[b]] at layer
    `layer_idx` from absolute position row_t0[b] on (prompt) and at every generated token.
    row_vec[b] < 0 marks a control row; row_t0 None steers every position
    (reference model_utils.py:798-802)."""

    layer_idx: Optional[int]          # None: `row_vec` is [n_layers, B], one row of vector indices per decoder layer
    vtable: torch.Tensor              # [n_vec, H] model dtype or fp32, on the device
    row_vec: torch.Tensor             # [B] int32 (or [n_layers, B], see layer_idx)
    row_alpha: torch.Tensor           # [B] float32
    row_t0: Optional[torch.Tensor]    # [B] int32 absolute (pad-shifted) start, or None

    def at(self, i: int) -> Optional["SteerSpec"]:
        """The steering that applies after decoder layer i, or None.  With per-layer index rows (the decode
        graphs: every layer's block-tail K2 launch reads its own row of indices, -1 = leave the row alone, so
        WHICH layer is steered is data, not part of the captured graph) every layer gets a spec."""
        if self.layer_idx is None:
            return SteerSpec(i, self.vtable, self.row_vec[i], self.row_alpha, self.row_t0)
        return self if self.layer_idx == i else None


class StaticKV:
    """Preallocated K/V with the `update(k, v, layer_idx)` method HF attention modules call.
    `pos` is a host integer, so the returned views have exact length — the same shapes a
    DynamicCache would hand to SDPA — and are static inside a per-step CUDA graph."""

    def __init__(self, n_layers: int, batch: int, n_kv: int, max_len: int, head_dim: int, dtype, device,
                 sliding_layers: Sequence[int] = (), sliding_window: Optional[int] = None):
        self.k = [torch.empty(batch, n_kv, max_len, head_dim, dtype=dtype, device=device) for _ in range(n_layers)]
        self.v = [torch.empty(batch, n_kv, max_len, head_dim, dtype=dtype, device=device) for _ in range(n_layers)]
        self.pos = 0
        self.max_len = max_len
        # Gemma-2's sliding-attention layers: on a cached step HF's sliding cache hands attention the last
        # `sliding_window` keys only (the new one included), not the whole history under a mask — same maths, but
        # the attention kernel then sums over another number of keys, i.e. in another order.  Same views here.
        self.sliding_layers = frozenset(sliding_layers)
        self.sliding_window = sliding_window

    def update(self, key_states, value_states, layer_idx, *args, **kwargs):
        t = key_states.shape[2]
        end = self.pos + t
        if end > self.max_len:
            raise RuntimeError("StaticKV overflow")
        self.k[layer_idx][:, :, self.pos:end].copy_(key_states)
        self.v[layer_idx][:, :, self.pos:end].copy_(value_states)
        lo = self.window_start(layer_idx, t, end)
        return self.k[layer_idx][:, :, lo:end], self.v[layer_idx][:, :, lo:end]

    def window_start(self, layer_idx: int, t: int, end: int) -> int:
        """First key a cached step (t == 1) of a sliding-attention layer sees; 0 everywhere else."""
        if t == 1 and self.sliding_window is not None and layer_idx in self.sliding_layers and end > self.sliding_window:
            return end - self.sliding_window
        return 0

    def bytes(self) -> int:
        return sum(t.numel() * t.element_size() for t in self.k + self.v)


class DecoderEngine:
    """Layer loop over a Hugging Face Llama / Qwen2 / Gemma-2 causal LM."""

    def __init__(self, model, fuse_gemma_tail: bool = True, attention: str = "auto", fused_norms: Optional[bool] = None):
        """attention: "hf" keeps the model's own attention function; "gqa" (= "auto" on a CUDA device; on the CPU
        "auto" means "hf") swaps in
        `_gqa_sdpa_forward`, which never materialises K/V per query head (HF's `repeat_kv`): torch's
        grouped-query kernel where head_dim <= 128 (measured 3.2x faster decode on Llama-3.1-8B), the folded form
        `_gqa_folded` above that (Gemma-2's 256: 1.9x faster bench step).  Both are bit-identical to the reference
        hook path on the same device (tests/test_e2e_gpu.py).
        fused_norms: run every RMSNorm of the stack through K2 (one launch instead of torch's ~7
        eager kernels, same bits); default = on CUDA."""
        self.model = model
        inner = model.model
        if not hasattr(inner, "layers") and hasattr(inner, "language_model"):
            inner = inner.language_model  # reference model_utils.py:287-289
        self.inner = inner
        self.layers = list(inner.layers)
        self.n_layers = len(self.layers)
        cfg = model.config
        self.cfg = cfg
        first = self.layers[0]
        self.is_gemma2 = hasattr(first, "pre_feedforward_layernorm")
        self.variant = "gemma" if "gemma" in type(model).__name__.lower() else "llama"
        self.fuse_gemma_tail = fuse_gemma_tail and self.is_gemma2
        norm0 = first.input_layernorm
        self.eps = float(getattr(norm0, "variance_epsilon", getattr(norm0, "eps", 1e-6)))
        p = next(model.parameters())
        self.device, self.dtype = p.device, p.dtype
        self.hidden = cfg.hidden_size
        # the block-tail variant keeps the row in registers: 16-bit dtypes up to H = 8192, fp32 up to 2048
        # (wider fp32 models take the unfused tail: HF post-FFN norm + add, then K2)
        if self.dtype == torch.float32 and self.hidden > 2048 or self.hidden > 8192 or self.hidden % 128:
            self.fuse_gemma_tail = False
        self.n_kv = cfg.num_key_value_heads
        self.head_dim = getattr(cfg, "head_dim", None) or cfg.hidden_size // cfg.num_attention_heads
        self.softcap = getattr(cfg, "final_logit_softcapping", None)
        self.sliding_window = getattr(cfg, "sliding_window", None) if self.is_gemma2 else None
        # launches of OUR kernels (K2 incl. plain-norm use, K1 calls, fused rope); graph replays add the
        # number of launches recorded when the graph was captured
        self.counters = {"k2_launches": 0, "k1_launches": 0, "rope_launches": 0, "act_launches": 0}
        self.fused_norms = (self.device.type == "cuda") if fused_norms is None else fused_norms
        self.graph_extraction = True
        # position ids of the extraction forwards: "arange" = the reference's (no position_ids passed, so a
        # left-padded row is numbered from its pad count); "mask" = pad-aware, as generation numbers them
        self.extraction_positions = "arange"
        self.additive_masks = True   # convert boolean attention masks once per forward, not once per layer
        # rotary embedding: one fused launch (same rounding points) instead of HF's ~8 eager kernels
        import sys as _sys
        self._rope_module = _sys.modules[type(first.self_attn).__module__]
        self.fused_rope = self.device.type == "cuda"
        self._extract_graphs: Dict[tuple, tuple] = {}
        self._sessions: Dict[tuple, "Session"] = {}   # insertion order = least recently used first
        self.max_sessions = 3
        if attention == "auto":
            attention = "gqa" if self.device.type == "cuda" else "hf"
        self.attention = attention
        if attention == "gqa":
            _register_attention()
        elif attention != "hf":
            raise ValueError(attention)
        # bench.py sets this to a list to get (start, end) CUDA events around each eager (prompt-pass)
        # K2 launch, recorded on the launching stream; None = no instrumentation.
        self.k2_events: Optional[list] = None
        # Llama / Qwen2 on CUDA: the decoder block is driven from here (`_run_layers_native`) instead of
        # through the HF layer/attention/MLP forwards — same cuBLAS projections, same SDPA call, but the
        # elementwise glue between them is three launches of ours (residual add folded into K2, rotary +
        # KV append, SiLU-gate multiply) and q/k/v and gate/up run as one GEMM each where that is
        # bit-identical to the separate ones on this device (checked per shape, see `_fused_ok`).
        self.native_block = (self.device.type == "cuda" and not self.is_gemma2 and self.hidden % 4 == 0
                             and self.head_dim % 16 == 0 and getattr(cfg, "intermediate_size", 0) % 8 == 0
                             and getattr(cfg, "hidden_act", None) == "silu"
                             and all(hasattr(first.self_attn, n) for n in ("q_proj", "k_proj", "v_proj", "o_proj"))
                             and all(hasattr(first.mlp, n) for n in ("gate_proj", "up_proj", "down_proj")))
        self.fuse_projections = True
        self.gemma_tail_everywhere = True   # Gemma-2: block-tail K2 at both residual joins of every layer
        # Gemma-2's gated MLP driven from here: gate/up as one GEMM where that is bit-identical, GELU(gate) * up in one
        # pass through a table of the module's own act_fn over all 65 536 values of the dtype (`_mlp_table`)
        self.table_mlp = (self.device.type == "cuda" and self.is_gemma2 and self.dtype in (torch.bfloat16, torch.float16)
                          and getattr(cfg, "intermediate_size", 0) % 8 == 0
                          and all(hasattr(first.mlp, n) for n in ("gate_proj", "up_proj", "down_proj", "act_fn")))
        self._act_table: Optional[torch.Tensor] = None
        # ... and its attention block: Gemma2Attention.forward is LlamaAttention.forward plus two keyword arguments
        # for the attention function, so `_attn_native` serves it (one q/k/v GEMM where bit-identical, rotary + KV
        # append in one launch, the sliding layers' key window on cached steps)
        self.gemma_native_attention = (self.device.type == "cuda" and self.is_gemma2 and self.head_dim % 16 == 0
                                       and all(hasattr(first.self_attn, n) for n in ("q_proj", "k_proj", "v_proj", "o_proj", "scaling")))
        self._qkv: Optional[list] = None   # per layer _FusedLinear, built on first use
        self._gu: Optional[list] = None
        self._fused_checked: Dict[tuple, bool] = {}

    # -- masks / positions --------------------------------------------------------------------
    @staticmethod
    def position_ids(attention_mask: torch.Tensor) -> torch.Tensor:
        pos = attention_mask.long().cumsum(-1) - 1
        return pos.masked_fill(attention_mask == 0, 1)

    def _layer_kind(self, i: int) -> str:
        # transformers Gemma2Model.forward picks the mask by config.layer_types[i]
        kinds = getattr(self.cfg, "layer_types", None)
        return kinds[i] if kinds else "full_attention"

    def prefill_masks(self, attention_mask: torch.Tensor, padded: Optional[bool] = None) -> Dict[str, Optional[torch.Tensor]]:
        """4-D boolean masks [B,1,T,T] (True = attend): causal AND key-not-padding, plus the window
        for Gemma-2 sliding layers.  None when nothing is padded (SDPA then runs is_causal).
        `padded` may be given by a caller that already knows (avoids a host sync, e.g. under capture)."""
        b, t = attention_mask.shape
        valid = attention_mask.bool()
        need_window = self.sliding_window is not None and t > self.sliding_window
        all_valid = bool(valid.all()) if padded is None else not padded
        if all_valid and not need_window:
            return {"full_attention": None, "sliding_attention": None}
        q = torch.arange(t, device=attention_mask.device)
        causal = q[None, :] <= q[:, None]
        full = (causal[None, None] & valid[:, None, None, :])
        out = {"full_attention": None if all_valid else full, "sliding_attention": None if all_valid else full}
        if need_window:
            win = (q[:, None] - q[None, :]) < self.sliding_window
            out["sliding_attention"] = full & win[None, None]
        return out

    def norm(self, module, x: torch.Tensor) -> torch.Tensor:
        """An un-steered RMSNorm of the stack: K2 with no rows steered (bit-identical to the eager
        module for >= 16 row vectors), or the HF module itself."""
        if self.fused_norms and x.is_cuda and x.dim() == 3:
            self.counters["k2_launches"] += 1
            return kernels.inject_rmsnorm(x if x.is_contiguous() else x.contiguous(), module.weight, self.eps, self.variant)
        return module(x)

    @staticmethod
    def _additive(masks, dtype):
        """Boolean masks -> the additive masks torch's fused SDPA backends consume, once per forward.
        Given a boolean mask, `scaled_dot_product_attention` itself runs `where(mask, 0, -inf)` in the
        query dtype (plus two scalar fills) inside EVERY layer's call — three launches per layer for a
        value that does not change across layers.  Same expression here, so the kernel inputs and the
        results are the same bits."""
        done, out = {}, {}
        for kind, m in masks.items():
            if m is None or m.dtype != torch.bool:
                out[kind] = m
                continue
            if id(m) not in done:
                done[id(m)] = torch.where(m, torch.zeros((), dtype=dtype, device=m.device),
                                          torch.full((), float("-inf"), dtype=dtype, device=m.device))
            out[kind] = done[id(m)]
        return out

    # -- the layer loop -----------------------------------------------------------------------
    def run_layers(self, *args, **kwargs):
        """`_run_layers` with this engine's attention function selected for the duration of the
        call only, so the HF model object keeps its stock behaviour for anyone else (the parity
        tests run the reference hook path on the very same model)."""
        stock_attn = self.cfg._attn_implementation
        stock_rope = getattr(self._rope_module, "apply_rotary_pos_emb", None)
        if self.attention == "gqa":
            self.cfg._attn_implementation = ATTN_GQA
        if self.fused_rope and stock_rope is not None:
            def fused(q, k, cos, sin, unsqueeze_dim=1):
                out = kernels.rope(q, k, cos, sin) if unsqueeze_dim == 1 else None
                if out is not None:
                    self.counters["rope_launches"] += 1
                    return out
                return stock_rope(q, k, cos, sin, unsqueeze_dim)

            self._rope_module.apply_rotary_pos_emb = fused
        if self.additive_masks and args[0].is_cuda:
            args = list(args)
            args[2] = self._additive(args[2], args[0].dtype)
        g = self.cfg.num_attention_heads // self.n_kv
        if self.attention == "gqa" and self.head_dim > 128 and g > 1 and args[0].shape[1] > 1:
            # folded grouped-query attention (`_gqa_folded`): fold the prompt-pass masks once, not per layer
            args = list(args)
            done = {}
            t = args[0].shape[1]
            for kind, m in list(args[2].items()):
                if m is not None and m.shape[2] == t:
                    if id(m) not in done:
                        done[id(m)] = _fold_mask(m, g)
                    args[2] = {**args[2], kind: done[id(m)]}
        try:
            return self._run_layers(*args, **kwargs)
        finally:
            self.cfg._attn_implementation = stock_attn
            if stock_rope is not None:
                self._rope_module.apply_rotary_pos_emb = stock_rope

    def _run_layers(self, h: torch.Tensor, pos_emb, masks, cache, steer: Optional[SteerSpec] = None,
                    pos0: int = 0, decode: bool = False, stop_after: Optional[int] = None,
                    capture: Optional[Dict[int, torch.Tensor]] = None, cap_t: int = -1):
        """Runs decoder layers on h [B,T,H].  Returns (final_normed_or_None, capture).
        With `stop_after` the loop ends after that layer and nothing is normed.
        capture: {layer index: [B, H] buffer}; the residual stream after that layer at position `cap_t` of every
        row is written into the buffer (reference model_utils.py:312-321: the hook's slice of the layer output) —
        by the K2 launch that finishes the layer where there is one (its tap output), by a row slice otherwise."""
        if self.native_block and h.is_cuda and self.fused_norms:
            return self._run_layers_native(h, pos_emb, masks, cache, steer, pos0, decode, stop_after, capture, cap_t)
        if self.is_gemma2 and self.fuse_gemma_tail and self.gemma_tail_everywhere and h.is_cuda and self.fused_norms:
            return self._run_layers_gemma(h, pos_emb, masks, cache, steer, pos0, decode, stop_after, capture, cap_t)
        captured = capture or {}
        want = set(captured)
        last = self.n_layers - 1 if stop_after is None else stop_after
        normed_in = None  # y from K2 = already-normed input of the next layer
        final = None
        for i in range(last + 1):
            layer = self.layers[i]
            residual = h
            x = normed_in if normed_in is not None else self.norm(layer.input_layernorm, h)
            normed_in = None
            attn, _ = layer.self_attn(hidden_states=x, position_embeddings=pos_emb,
                                      attention_mask=masks[self._layer_kind(i)], past_key_values=cache)
            st = steer.at(i) if steer is not None else None
            steer_here = st is not None
            if self.is_gemma2:
                h = residual + self.norm(layer.post_attention_layernorm, attn)
                residual = h
                m = layer.mlp(self.norm(layer.pre_feedforward_layernorm, h))
                if steer_here and self.fuse_gemma_tail:
                    nxt = self.layers[i + 1].input_layernorm if i + 1 < self.n_layers else self.inner.norm
                    ev = None
                    if self.k2_events is not None and not decode:
                        ev = (torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True))
                        ev[0].record()
                    y = kernels.gemma2_tail_inject_rmsnorm(
                        m.contiguous(), residual, layer.post_feedforward_layernorm.weight, nxt.weight, self.eps,
                        vtable=st.vtable, row_vec=st.row_vec, row_alpha=st.row_alpha,
                        row_t0=None if decode else st.row_t0, pos0=pos0,
                        steer_all=decode or st.row_t0 is None)
                    if ev is not None:
                        ev[1].record()
                        self.k2_events.append((ev[0], ev[1], residual.numel() * residual.element_size()))
                    self.counters["k2_launches"] += 1
                    h = residual  # updated in place by the kernel
                    if i + 1 < self.n_layers:
                        normed_in = y
                    else:
                        final = y
                    if i in want:
                        captured[i].copy_(h[:, cap_t])
                    continue
                h = residual + self.norm(layer.post_feedforward_layernorm, m)
            else:
                h = residual + attn
                residual = h
                h = residual + layer.mlp(self.norm(layer.post_attention_layernorm, h))
            if steer_here:
                nxt = self.layers[i + 1].input_layernorm if i + 1 < self.n_layers else self.inner.norm
                if not h.is_contiguous():
                    h = h.contiguous()
                ev = None
                if self.k2_events is not None and not decode:
                    ev = (torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True))
                    ev[0].record()
                y = kernels.inject_rmsnorm(h, nxt.weight, self.eps, self.variant, vtable=st.vtable,
                                           row_vec=st.row_vec, row_alpha=st.row_alpha,
                                           row_t0=None if decode else st.row_t0, pos0=pos0,
                                           steer_all=decode or st.row_t0 is None)
                if ev is not None:
                    ev[1].record()
                    self.k2_events.append((ev[0], ev[1], h.numel() * h.element_size()))
                self.counters["k2_launches"] += 1
                if i + 1 < self.n_layers:
                    normed_in = y
                else:
                    final = y
            if i in want:
                captured[i].copy_(h[:, cap_t])
        if stop_after is None and final is None:
            final = self.norm(self.inner.norm, h)
        return final, captured

    # -- the Llama / Qwen2 block driven from here ------------------------------------------------
    def _fused_sets(self):
        if self._qkv is None:
            with torch.inference_mode(False), torch.no_grad():  # ordinary tensors: the model outlives this call
                self._qkv = [_FusedLinear([l.self_attn.q_proj, l.self_attn.k_proj, l.self_attn.v_proj]) for l in self.layers]
                self._gu = [_FusedLinear([l.mlp.gate_proj, l.mlp.up_proj]) for l in self.layers]
        return self._qkv, self._gu

    def _fused_ok(self, kind: str, shape) -> bool:
        """May projections `kind` ("qkv" | "gu") run as ONE GEMM for inputs of `shape` [B, T, H]?  Only if
        cuBLAS then produces the very bits of the separate GEMMs on this device — decided by running
        both once on layer 0's weights (all layers share the shapes, hence the kernel choice).  The
        check needs a host sync, so a shape first met under stream capture stays on separate GEMMs."""
        if not self.fuse_projections:
            return False
        key = (kind, tuple(shape))
        ok = self._fused_checked.get(key)
        if ok is None:
            if torch.cuda.is_current_stream_capturing():
                return False
            qkv, gu = self._fused_sets()
            f = (qkv if kind == "qkv" else gu)[0]
            g = torch.Generator(device=self.device).manual_seed(1234)
            x = torch.randn(tuple(shape), device=self.device, dtype=torch.float32, generator=g).to(self.dtype)
            fused = f(x)
            ok = all(torch.equal(part, m(x)) for part, m in zip(fused, f.mods))
            self._fused_checked[key] = ok
        return ok

    def _attention_fn(self):
        if self.attention == "gqa":
            return _gqa_sdpa_forward
        mod = self._rope_module
        return mod.ALL_ATTENTION_FUNCTIONS.get_interface(self.cfg._attn_implementation, mod.eager_attention_forward)

    def _attn_native(self, layer, i: int, x: torch.Tensor, pos_emb, mask, cache) -> torch.Tensor:
        """transformers' LlamaAttention / Qwen2Attention.forward with the same library calls in the
        same order; rotary embedding, the rotated-k tensor and the two cache copies are one launch."""
        att = layer.self_attn
        b, t, _ = x.shape
        nh, nkv, d = self.cfg.num_attention_heads, self.n_kv, self.head_dim
        if self._fused_ok("qkv", x.shape):
            q, k, v = self._qkv[i](x)
        else:
            q, k, v = att.q_proj(x), att.k_proj(x), att.v_proj(x)
        cos, sin = pos_emb
        if cache is not None:
            end = cache.pos + t
            if end > cache.max_len:
                raise RuntimeError("StaticKV overflow")
            kd, vd, pos = cache.k[i], cache.v[i], cache.pos
        else:  # extraction forwards keep no cache: dense [B, T, n_kv, D], seen as HF's transposed views
            kd = torch.empty((b, t, nkv, d), dtype=x.dtype, device=x.device).transpose(1, 2)
            vd = torch.empty((b, t, nkv, d), dtype=x.dtype, device=x.device).transpose(1, 2)
            pos, end = 0, t
        qr = kernels.rope_append(q, k, v, cos.contiguous(), sin.contiguous(), kd, vd, pos, nh, nkv, d)
        self.counters["rope_launches"] += 1
        extra = {"sliding_window": att.sliding_window} if hasattr(att, "sliding_window") else {}
        if hasattr(att, "attn_logit_softcapping"):  # Gemma2Attention.forward hands it to the attention function
            extra["softcap"] = att.attn_logit_softcapping
        lo = cache.window_start(i, t, end) if cache is not None else 0  # Gemma-2 sliding layers on cached steps
        out, _ = self._attention_fn()(att, qr, kd[:, :, lo:end], vd[:, :, lo:end], mask, dropout=0.0, scaling=att.scaling,
                                      **extra)
        return att.o_proj(out.reshape(b, t, -1))

    def _mlp_native(self, layer, i: int, x: torch.Tensor) -> torch.Tensor:
        mlp = layer.mlp
        if self._fused_ok("gu", x.shape):
            g, u = self._gu[i](x)
        else:
            g, u = mlp.gate_proj(x), mlp.up_proj(x)
        self.counters["act_launches"] += 1
        return mlp.down_proj(kernels.silu_mul(g, u))

    def _mlp_table(self, layer, i: int, x: torch.Tensor) -> torch.Tensor:
        """transformers' Gemma2MLP.forward — down_proj(act_fn(gate_proj(x)) * up_proj(x)) — with the same GEMMs and the
        two eager elementwise kernels replaced by one table pass: the table is the module's own act_fn (ATen's
        tanh-GELU) evaluated on every value of the dtype, so both rounding points keep their bits."""
        mlp = layer.mlp
        if self._act_table is None:
            with torch.inference_mode(False):  # an ordinary tensor: it outlives this call
                self._act_table = kernels.activation_table(mlp.act_fn, self.dtype, self.device)
        if self._fused_ok("gu", x.shape):
            g, u = self._gu[i](x)
        else:
            g, u = mlp.gate_proj(x), mlp.up_proj(x)
        self.counters["act_launches"] += 1
        return mlp.down_proj(kernels.act_mul_table(g, u, self._act_table))

    def _k2(self, h, weight, add=None, steer: Optional[SteerSpec] = None, pos0: int = 0, decode: bool = False,
            timed: bool = False, cap_out: Optional[torch.Tensor] = None, cap_t: int = -1):
        ev = None
        if timed and self.k2_events is not None and not decode:
            ev = (torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True))
            ev[0].record()
        if steer is None:
            y = kernels.inject_rmsnorm(h, weight, self.eps, self.variant, add=add, cap_out=cap_out, cap_t=cap_t)
        else:
            y = kernels.inject_rmsnorm(h, weight, self.eps, self.variant, vtable=steer.vtable, row_vec=steer.row_vec,
                                       row_alpha=steer.row_alpha, row_t0=None if decode else steer.row_t0, pos0=pos0,
                                       steer_all=decode or steer.row_t0 is None, add=add, cap_out=cap_out, cap_t=cap_t)
        if ev is not None:
            ev[1].record()
            self.k2_events.append((ev[0], ev[1], h.numel() * h.element_size()))
        self.counters["k2_launches"] += 1
        return y

    def _gemma_tail(self, branch, h, w_post, w_next, steer=None, pos0=0, decode=False, timed=False, cap_out=None,
                    cap_t: int = -1):
        """h <- h + gemma_norm(branch; w_post) [+ alpha*v]; returns gemma_norm(h; w_next).  One launch for what the
        HF layer does as norm kernel + add kernel (+ hook) + the next norm."""
        ev = None
        if timed and self.k2_events is not None and not decode:
            ev = (torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True))
            ev[0].record()
        kw = {}
        if steer is not None:
            kw = dict(vtable=steer.vtable, row_vec=steer.row_vec, row_alpha=steer.row_alpha,
                      row_t0=None if decode else steer.row_t0, pos0=pos0, steer_all=decode or steer.row_t0 is None)
        y = kernels.gemma2_tail_inject_rmsnorm(branch if branch.is_contiguous() else branch.contiguous(), h, w_post, w_next,
                                               self.eps, cap_out=cap_out, cap_t=cap_t, **kw)
        if ev is not None:
            ev[1].record()
            self.k2_events.append((ev[0], ev[1], h.numel() * h.element_size()))
        self.counters["k2_launches"] += 1
        return y

    def _run_layers_gemma(self, h, pos_emb, masks, cache, steer, pos0, decode, stop_after, capture, cap_t=-1):
        """Gemma-2 with the block-tail form of K2 at BOTH residual joins of every layer (after attention:
        post_attention_layernorm + add + pre_feedforward_layernorm; after the MLP: post_feedforward_layernorm + add
        [+ injection] + the next layer's input_layernorm).  Attention and MLP stay the HF modules.  The residual
        stream is one buffer updated in place."""
        captured = capture or {}
        last = self.n_layers - 1 if stop_after is None else stop_after
        if not h.is_contiguous():
            h = h.contiguous()
        x = self.norm(self.layers[0].input_layernorm, h)
        final = None
        for i in range(last + 1):
            layer = self.layers[i]
            if self.gemma_native_attention:
                attn = self._attn_native(layer, i, x, pos_emb, masks[self._layer_kind(i)], cache)
            else:
                attn, _ = layer.self_attn(hidden_states=x, position_embeddings=pos_emb,
                                          attention_mask=masks[self._layer_kind(i)], past_key_values=cache)
            x = self._gemma_tail(attn, h, layer.post_attention_layernorm.weight, layer.pre_feedforward_layernorm.weight)
            m = self._mlp_table(layer, i, x) if self.table_mlp else layer.mlp(x)
            st = steer.at(i) if steer is not None else None
            nxt = self.layers[i + 1].input_layernorm if i + 1 < self.n_layers else self.inner.norm
            if i == last and stop_after is not None and st is None and i in captured:
                # nothing downstream: finish the captured rows only (B vectors instead of B*T), in the capture buffer
                rows = captured[i]
                rows.copy_(h[:, cap_t])
                self._gemma_tail(m[:, cap_t].unsqueeze(1), rows.unsqueeze(1), layer.post_feedforward_layernorm.weight, nxt.weight)
                break
            y = self._gemma_tail(m, h, layer.post_feedforward_layernorm.weight, nxt.weight, steer=st, pos0=pos0,
                                 decode=decode, timed=st is not None, cap_out=captured.get(i), cap_t=cap_t)
            if i < last:
                x = y
            elif stop_after is None:
                final = y
        return final, captured

    def _run_layers_native(self, h, pos_emb, masks, cache, steer, pos0, decode, stop_after, capture, cap_t=-1):
        """Same contract as `_run_layers`.  The residual stream `h` is ONE buffer updated in place by K2
        (h <- h + branch [+ alpha*v]); K2's second output is the next sub-block's normed input."""
        captured = capture or {}
        last = self.n_layers - 1 if stop_after is None else stop_after
        if not h.is_contiguous():
            h = h.contiguous()
        x = self._k2(h, self.layers[0].input_layernorm.weight)
        final = None
        for i in range(last + 1):
            layer = self.layers[i]
            attn = self._attn_native(layer, i, x, pos_emb, masks[self._layer_kind(i)], cache)
            x = self._k2(h, layer.post_attention_layernorm.weight, add=attn)
            m = self._mlp_native(layer, i, x)
            st = steer.at(i) if steer is not None else None
            if i == last and stop_after is not None and st is None:
                # nothing reads a normed output after the last requested layer: finish the captured rows only
                if i in captured:
                    torch.add(h[:, cap_t], m[:, cap_t], out=captured[i])
                else:
                    h.add_(m)
            else:
                nxt = self.layers[i + 1].input_layernorm if i + 1 < self.n_layers else self.inner.norm
                y = self._k2(h, nxt.weight, add=m, steer=st, pos0=pos0, decode=decode, timed=st is not None,
                             cap_out=captured.get(i), cap_t=cap_t)
                if i < last:
                    x = y
                elif stop_after is None:
                    final = y
        return final, captured

    def logits(self, normed_last: torch.Tensor) -> torch.Tensor:
        """lm_head (+ Gemma-2 final soft-capping) on [B,1,H] -> [B,V]."""
        lg = self.model.lm_head(normed_last)
        if self.softcap is not None:
            lg = torch.tanh(lg / self.softcap) * self.softcap
        return lg[:, -1, :]

    # -- extraction ---------------------------------------------------------------------------
    @torch.inference_mode()
    def capture_rows(self, input_ids: torch.Tensor, attention_mask: torch.Tensor, layer_indices: Sequence[int],
                     token_idx: int = -1, slot: int = 0) -> torch.Tensor:
        """One forward, stopped after the deepest requested layer; returns the residual stream after each requested
        layer at position `token_idx` of every row, compactly: [n_layers, B, H] in the model dtype, on the device
        (reference model_utils.py:312-338 runs the whole model once per layer, slices the hooked layer output and
        ships it to the host).  On the kernel paths the rows are written by the K2 launch that finishes the layer
        (its tap output): no copy of the residual stream is kept and nothing re-reads it."""
        layer_indices = [int(i) for i in layer_indices]
        for i in layer_indices:
            if not 0 <= i < self.n_layers:
                raise IndexError(f"layer {i} out of range")
        t_len = input_ids.shape[1]
        cap_t = token_idx if token_idx >= 0 else t_len + token_idx  # the reference slices the padded tensor
        if not 0 <= cap_t < t_len:
            raise IndexError(f"token_idx {token_idx} out of range for width {t_len}")
        padded = not bool(attention_mask.all())
        # small (launch-bound) forwards replay as a CUDA graph; large ones are GPU-bound anyway and a
        # graph would pin their activations in a private pool
        if self.device.type == "cuda" and self.graph_extraction and input_ids.numel() <= 16384:
            return self._capture_rows_graphed(input_ids, attention_mask, layer_indices, padded, cap_t, slot)
        return self._capture_rows_eager(input_ids.to(self.device), attention_mask.to(self.device), layer_indices, padded, cap_t)

    def _capture_rows_eager(self, ids, mask, layer_indices, padded, cap_t):
        h = self.inner.embed_tokens(ids)
        if self.extraction_positions == "arange":
            # what the reference's forward sees: `model(**inputs)` passes no position_ids (model_utils.py:338), so
            # Hugging Face numbers the PADDED sequence 0..T-1 and a left-padded row's positions start at its pad count
            pos = torch.arange(ids.shape[1], device=ids.device)[None].expand(ids.shape[0], -1)
        else:
            pos = self.position_ids(mask)
        pos_emb = self.inner.rotary_emb(h, pos)
        uniq = sorted(set(layer_indices))
        buf = torch.empty((len(uniq), ids.shape[0], self.hidden), dtype=h.dtype, device=h.device)
        self.run_layers(h, pos_emb, self.prefill_masks(mask, padded), None, stop_after=uniq[-1],
                        capture={layer: buf[k] for k, layer in enumerate(uniq)}, cap_t=cap_t)
        if uniq == layer_indices:
            return buf
        return buf[[uniq.index(i) for i in layer_indices]]

    def _capture_rows_graphed(self, input_ids, attention_mask, layer_indices, padded, cap_t, slot):
        """The extraction forward is ~1500 small launches (short prompts), i.e. CPU-launch-bound when
        run eagerly: replay it as one CUDA graph per (batch, width, layers) shape.  The returned
        tensor lives in the graph's pool and is overwritten by the next replay of the same shape and
        `slot` (callers that need two captures alive at once, e.g. concept and baseline batches, use
        different slots)."""
        key = (tuple(input_ids.shape), tuple(layer_indices), padded, cap_t, slot, self.extraction_positions)
        entry = self._extract_graphs.get(key)
        if entry is None:
            if len(self._extract_graphs) >= 8:
                self._extract_graphs.clear()
            ids_s = torch.zeros(input_ids.shape, dtype=torch.long, device=self.device)
            mask_s = torch.ones(attention_mask.shape, dtype=torch.long, device=self.device)
            ids_s.copy_(input_ids)
            mask_s.copy_(attention_mask)
            side = torch.cuda.Stream()
            side.wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(side):
                for _ in range(2):
                    self._capture_rows_eager(ids_s, mask_s, layer_indices, padded, cap_t)
            torch.cuda.current_stream().wait_stream(side)
            torch.cuda.synchronize()
            g = torch.cuda.CUDAGraph()
            before = dict(self.counters)
            with _no_gc_during_capture(), torch.cuda.graph(g):
                out = self._capture_rows_eager(ids_s, mask_s, layer_indices, padded, cap_t)
            g.ia_launches = {k: self.counters[k] - before[k] for k in before}
            for k in before:
                self.counters[k] = before[k]
            entry = (g, ids_s, mask_s, out)
            self._extract_graphs[key] = entry
        g, ids_s, mask_s, out = entry
        ids_s.copy_(input_ids, non_blocking=True)
        mask_s.copy_(attention_mask, non_blocking=True)
        g.replay()
        for k, v in g.ia_launches.items():
            self.counters[k] += v
        return out

    @torch.inference_mode()
    def extract_vectors(self, concept_ids, concept_mask, base_ids, base_mask, layer_indices: Sequence[int],
                        mode: str = "baseline", normalize: bool = False, token_idx: int = -1, want_f32: bool = True,
                        want_norm: bool = True):
        """K1 over any number of layers: two forwards (concept batch, baseline batch — the
        reference also runs them separately, vector_utils.py:524-540), each leaving its captured rows
        compactly in HBM through K2's tap, and one mean-diff (+ norm) launch over them.  Returns
        (v_f32, v_rounded, norm) on the device, each [n_layers, n_out, H] / [n_layers, n_out]."""
        if not self.device.type == "cuda":
            raise RuntimeError("concept-vector extraction needs the CUDA kernels; there is no CPU path")
        crows = self.capture_rows(concept_ids, concept_mask, layer_indices, token_idx, slot=0)
        brows = None
        if mode not in ("no_baseline", "raw"):
            brows = self.capture_rows(base_ids, base_mask, layer_indices, token_idx, slot=1)
        out = kernels.meandiff_compact(crows, brows, mode, normalize, want_f32=want_f32, want_norm=want_norm)
        self.counters["k1_launches"] += 2 if (want_norm or normalize) else 1  # mean-diff kernel (+ norm pass)
        return out

    # -- generation ---------------------------------------------------------------------------
    def session(self, batch: int, prompt_len: int, max_new_tokens: int, padded: bool, n_vec_cap: int = 64,
                vec_dtype: Optional[torch.dtype] = None, use_graphs: Optional[bool] = None) -> "Session":
        """The static buffers and step graphs for one batch shape.  A sweep alternates between a few shapes (the
        padded width moves by a token when trial numbers gain a digit; the last batch of a layer is shorter), and a
        new shape costs max_new_tokens - 1 graph captures, so the last `max_sessions` shapes are kept, least
        recently used first out — fewer when their KV caches would not fit next to a new one."""
        key = (batch, prompt_len, max_new_tokens, padded, n_vec_cap, vec_dtype)
        sess = self._sessions.pop(key, None)
        if sess is None:
            need = 2 * self.n_layers * batch * self.n_kv * (prompt_len + max_new_tokens) * self.head_dim * self.model.dtype.itemsize
            while self._sessions and (len(self._sessions) >= self.max_sessions or not self._fits(need)):
                self._sessions.pop(next(iter(self._sessions)))  # oldest first
                if self.device.type == "cuda":
                    gc.collect()
                    torch.cuda.empty_cache()
            sess = Session(self, key, use_graphs)
        self._sessions[key] = sess  # most recently used last
        return sess

    def _fits(self, kv_bytes: int) -> bool:
        if self.device.type != "cuda":
            return True
        free, _ = torch.cuda.mem_get_info(self.device)
        return free + torch.cuda.memory_reserved(self.device) - torch.cuda.memory_allocated(self.device) > 2 * kv_bytes

    @property
    def _session(self):
        """The most recently used session (None when there is none); assigning None drops every kept session."""
        return next(reversed(self._sessions.values())) if self._sessions else None

    @_session.setter
    def _session(self, value):
        if value is not None:
            raise ValueError("sessions are created by DecoderEngine.session()")
        self._sessions.clear()
        if self.device.type == "cuda":
            gc.collect()
            torch.cuda.empty_cache()


class Session:
    """Static buffers + per-step CUDA graphs for one (batch, prompt_len, max_new_tokens) shape."""

    def __init__(self, eng: DecoderEngine, key, use_graphs: Optional[bool]):
        self.eng, self.key = eng, key
        b, t, n, padded, n_vec_cap, vec_dtype = key
        dev = eng.device
        self.B, self.T, self.N, self.padded = b, t, n, padded
        sliding = [i for i in range(eng.n_layers) if eng._layer_kind(i) == "sliding_attention"] if eng.sliding_window else []
        self.cache = StaticKV(eng.n_layers, b, eng.n_kv, t + max(n - 1, 0), eng.head_dim, eng.dtype, dev,
                              sliding_layers=sliding, sliding_window=eng.sliding_window)
        self.key_valid = torch.ones(b, t + n, dtype=torch.bool, device=dev)
        self.tok = torch.zeros(b, 1, dtype=torch.long, device=dev)
        self.posid = torch.zeros(b, 1, dtype=torch.long, device=dev)
        self.out_ids = torch.zeros(b, n, dtype=torch.long, device=dev)
        self.logits_buf: Optional[torch.Tensor] = None
        self.unfinished = torch.ones(b, dtype=torch.bool, device=dev)
        self.vtable = torch.zeros(n_vec_cap, eng.hidden, dtype=vec_dtype or eng.dtype, device=dev)
        self.row_vec = torch.full((b,), -1, dtype=torch.int32, device=dev)
        # decode graphs: one row of vector indices per decoder layer, -1 everywhere except the steered layer
        self.row_vec_layers = torch.full((eng.n_layers, b), -1, dtype=torch.int32, device=dev)
        self.row_alpha = torch.zeros(b, dtype=torch.float32, device=dev)
        self.row_t0 = torch.zeros(b, dtype=torch.int32, device=dev)
        self.use_graphs = (dev.type == "cuda") if use_graphs is None else use_graphs
        self.graphs: Dict[Tuple, List] = {}
        self.pool = None
        self.step_logits: List[torch.Tensor] = []
        # every token a row has seen (prompt incl. pads, then generated), for the repetition penalty; slots not yet
        # written repeat column 0, which is part of the history anyway, so penalising over the whole buffer is exact
        self.hist: Optional[torch.Tensor] = None
        self._hist_buf: Optional[torch.Tensor] = None
        self._eos_ids: Dict[tuple, torch.Tensor] = {}

    def _eos_tensor(self, eos: Optional[tuple]) -> Optional[torch.Tensor]:
        if not eos:
            return None
        t = self._eos_ids.get(eos)
        if t is None:
            t = torch.tensor(list(eos), dtype=torch.long, device=self.eng.device)
            self._eos_ids[eos] = t
        return t

    def _penalise(self, lg: torch.Tensor, penalty: float) -> torch.Tensor:
        """HF RepetitionPenaltyLogitsProcessor over the row's whole history: a seen token's logit is divided by the
        penalty when positive and multiplied by it when negative (applies to greedy decoding too)."""
        if penalty == 1.0:
            return lg
        lg = lg.float()  # HF runs its logits processors on an fp32 copy of the last-position logits
        seen = lg.gather(1, self.hist)
        seen = torch.where(seen < 0, seen * penalty, seen / penalty)
        return lg.scatter(1, self.hist, seen)

    # one cached decode step: feeds self.tok at cache position T+s-1, leaves next token in self.tok
    def _step(self, s: int, steer: Optional[SteerSpec], greedy: bool, eos: Optional[tuple], pad: int,
              penalty: float = 1.0):
        eng = self.eng
        eng_cache = self.cache
        eng_cache.pos = self.T + s - 1
        length = self.T + s
        h = eng.inner.embed_tokens(self.tok)
        pos_emb = eng.inner.rotary_emb(h, self.posid)
        full = self.key_valid[:, None, None, :length] if self.padded else None
        masks = {"full_attention": full, "sliding_attention": full}
        if eng.sliding_window is not None and length > eng.sliding_window:
            # sliding layers get the last `sliding_window` keys (StaticKV.update): every one of them is inside the
            # window, so only the padding mask is left, cut to the same keys
            masks["sliding_attention"] = None if full is None else full[..., length - eng.sliding_window:]
        final, _ = eng.run_layers(h, pos_emb, masks, eng_cache, steer=steer, pos0=self.T + s - 1, decode=True)
        lg = eng.logits(final)
        if self.logits_buf is not None:
            self.logits_buf.copy_(lg)
        if greedy:
            nxt = self._penalise(lg, penalty).argmax(dim=-1)
            self._commit(nxt, s, eos, pad)

    def _commit(self, nxt: torch.Tensor, s: int, eos: Optional[tuple], pad: int):
        """A finished row emits the pad token from then on; a row finishes on ANY id of the model's EOS set, as HF
        `generate` does (Llama-3.x-Instruct lists three ids and ends a turn with the last of them)."""
        eos_t = self._eos_tensor(eos)
        if eos_t is not None:
            nxt = torch.where(self.unfinished, nxt, torch.full_like(nxt, pad))
            self.unfinished &= ~torch.isin(nxt, eos_t)
        self.out_ids[:, s].copy_(nxt)
        if self.hist is not None:
            self.hist[:, self.T + s].copy_(nxt)
        self.tok.copy_(nxt[:, None])
        self.posid.add_(1)

    def _ensure_graphs(self, gkey, steer, greedy, eos, pad, penalty=1.0):
        if gkey in self.graphs:
            return self.graphs[gkey]
        # warm-up on a side stream (lazy cuBLAS/SDPA init must not happen under capture)
        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            for _ in range(2):
                self._step(1, steer, greedy, eos, pad, penalty)
        torch.cuda.current_stream().wait_stream(side)
        torch.cuda.synchronize()
        graphs = []
        with _no_gc_during_capture():
            for s in range(1, self.N):
                g = torch.cuda.CUDAGraph()
                before = dict(self.eng.counters)
                with torch.cuda.graph(g, pool=self.pool):
                    self._step(s, steer, greedy, eos, pad, penalty)
                # launches recorded into this graph (they did not run): credited on every replay instead
                g.ia_launches = {k: self.eng.counters[k] - before[k] for k in before}
                for k in before:
                    self.eng.counters[k] = before[k]
                if self.pool is None:
                    self.pool = g.pool()
                graphs.append(g)
        self.graphs[gkey] = graphs
        return graphs

    @torch.inference_mode()
    def generate(self, input_ids: torch.Tensor, attention_mask: torch.Tensor, steer: Optional[SteerSpec] = None,
                 temperature: float = 0.0, top_k: Optional[int] = 50, eos_token_id=None,
                 pad_token_id: int = 0, generator: Optional[torch.Generator] = None,
                 keep_logits: bool = False, top_p: Optional[float] = None, repetition_penalty: float = 1.0,
                 do_sample: Optional[bool] = None) -> torch.Tensor:
        """Prefill + N-1 cached decode steps.  Returns new token ids [B, N] (device tensor owned by
        the session; clone to keep).  With keep_logits, self.step_logits holds [N] x [B,V] float32 (raw logits,
        HF's `output_logits`).  eos_token_id: one id or a sequence of ids (HF stops a row on any of them).
        do_sample None = sample iff temperature > 0; the warpers follow HF's order: repetition penalty,
        temperature, top-k, top-p."""
        eng = self.eng
        ids = input_ids.to(eng.device, non_blocking=True)
        mask = attention_mask.to(eng.device, non_blocking=True)
        assert ids.shape == (self.B, self.T), (ids.shape, self.B, self.T)
        greedy = (temperature <= 0) if do_sample is None else not do_sample
        if not greedy and temperature <= 0:
            raise ValueError("sampling needs a positive temperature")
        eos_token_id = _eos_tuple(eos_token_id)
        penalty = float(repetition_penalty or 1.0)
        if penalty != 1.0:
            if self._hist_buf is None:  # allocated once: decode graphs captured with a penalty point at it
                self._hist_buf = torch.zeros(self.B, self.T + self.N, dtype=torch.long, device=eng.device)
            self.hist = self._hist_buf
            self.hist[:, : self.T].copy_(ids)
            self.hist[:, self.T:] = ids[:, :1]
        else:
            self.hist = None
        self.step_logits = []
        st = None
        if steer is not None:
            if steer.vtable.device.type != "cuda":
                raise RuntimeError("steering needs the CUDA kernels; there is no CPU path")
            n_vec = steer.vtable.shape[0]
            self.vtable[:n_vec].copy_(steer.vtable, non_blocking=True)
            self.row_vec.copy_(steer.row_vec, non_blocking=True)
            self.row_alpha.copy_(steer.row_alpha, non_blocking=True)
            if steer.row_t0 is not None:
                self.row_t0.copy_(steer.row_t0, non_blocking=True)
            st = SteerSpec(steer.layer_idx, self.vtable, self.row_vec, self.row_alpha,
                           self.row_t0 if steer.row_t0 is not None else None)
            self.row_vec_layers.fill_(-1)
            self.row_vec_layers[steer.layer_idx].copy_(self.row_vec)
        self.key_valid[:, : self.T].copy_(mask.bool())
        self.unfinished.fill_(True)

        # ---- prefill (eager: one pass of large GEMMs) ----
        self.cache.pos = 0
        h = eng.inner.embed_tokens(ids)
        pos = eng.position_ids(mask)
        pos_emb = eng.inner.rotary_emb(h, pos)
        final, _ = eng.run_layers(h, pos_emb, eng.prefill_masks(mask), self.cache, steer=st, pos0=0, decode=False)
        lg = eng.logits(final[:, -1:, :])
        if keep_logits:
            self.step_logits.append(lg.float().clone())
        self.posid.copy_((mask.long().sum(-1, keepdim=True)) - 1)  # position of the last prompt token
        lgp = self._penalise(lg, penalty)
        nxt = lgp.argmax(dim=-1) if greedy else self._sample(lgp, temperature, top_k, generator, top_p)
        self._commit(nxt, 0, eos_token_id, pad_token_id)  # also advances posid to the new token's position

        # ---- decode ----
        if self.N > 1:
            want_buf = keep_logits or not greedy
            if want_buf and self.logits_buf is None:
                self.logits_buf = torch.empty(self.B, lg.shape[-1], dtype=lg.dtype, device=eng.device)
                self.graphs.clear()
            graphs = None
            # decode steps steer every position of a steered row, whatever t0 was; the steered LAYER is data (the
            # per-layer index rows), so one set of step graphs serves every layer of a sweep
            st_dec = None if st is None else SteerSpec(None, self.vtable, self.row_vec_layers, self.row_alpha, None)
            if self.use_graphs:
                gkey = (st is not None, greedy, eos_token_id, pad_token_id, self.logits_buf is not None,
                        penalty if greedy else 1.0)
                if gkey not in self.graphs:
                    # capturing runs warm-up steps that scribble on the state: save and restore it
                    state = [self.tok, self.posid, self.out_ids, self.unfinished] + ([self.hist] if self.hist is not None else [])
                    keep = [t.clone() for t in state]
                    graphs = self._ensure_graphs(gkey, st_dec, greedy, eos_token_id, pad_token_id, penalty)
                    for dst, src in zip(state, keep):
                        dst.copy_(src)
                graphs = self.graphs[gkey]
            for s in range(1, self.N):
                if graphs is not None:
                    graphs[s - 1].replay()
                    for k, v in graphs[s - 1].ia_launches.items():
                        eng.counters[k] += v
                else:
                    self._step(s, st_dec, greedy, eos_token_id, pad_token_id, penalty)
                if keep_logits:
                    self.step_logits.append(self.logits_buf.float().clone())
                if not greedy:
                    nxt = self._sample(self._penalise(self.logits_buf, penalty), temperature, top_k, generator, top_p)
                    self._commit(nxt, s, eos_token_id, pad_token_id)
                if eos_token_id is not None and s % 16 == 0 and not bool(self.unfinished.any()):
                    self.out_ids[:, s + 1:].fill_(pad_token_id)
                    break
        return self.out_ids

    @staticmethod
    def sampling_probs(logits: torch.Tensor, temperature: float, top_k: Optional[int], top_p: Optional[float] = None
                       ) -> torch.Tensor:
        """The distribution HF `generate(do_sample=True, temperature=...)` samples from: temperature scaling, then
        top-k, then top-p (TemperatureLogitsWarper, TopKLogitsWarper, TopPLogitsWarper in that order), then softmax.
        reference model_utils.py:833-837 passes only the temperature; top_k / top_p come from the checkpoint's
        generation_config (top_k 50 by default; Llama-3.1 ships top_p 0.9, Qwen2.5 top_k 20 / top_p 0.8)."""
        lg = logits.float() / temperature
        if top_k is not None and 0 < top_k < lg.shape[-1]:
            kth = lg.topk(top_k, dim=-1).values[..., -1:]
            lg = lg.masked_fill(lg < kth, float("-inf"))
        if top_p is not None and 0.0 < top_p < 1.0:
            # drop the low-probability tail whose cumulative mass stays within 1 - top_p; the most likely token
            # always survives
            asc, order = torch.sort(lg, dim=-1, descending=False)
            tail = asc.softmax(dim=-1).cumsum(dim=-1) <= (1.0 - top_p)
            tail[..., -1] = False
            lg = lg.masked_fill(tail.scatter(-1, order, tail), float("-inf"))
        return torch.softmax(lg, dim=-1)

    @staticmethod
    def _sample(logits: torch.Tensor, temperature: float, top_k: Optional[int], generator,
                top_p: Optional[float] = None) -> torch.Tensor:
        probs = Session.sampling_probs(logits, temperature, top_k, top_p)
        return torch.multinomial(probs, 1, generator=generator).squeeze(-1)


def _eos_tuple(eos) -> Optional[tuple]:
    """None, one id or any sequence of ids -> a sorted tuple (hashable: part of the decode-graph key) or None."""
    if eos is None:
        return None
    if isinstance(eos, int):
        return (int(eos),)
    if isinstance(eos, torch.Tensor):
        eos = eos.flatten().tolist()
    ids = tuple(sorted({int(e) for e in eos}))
    return ids or None
