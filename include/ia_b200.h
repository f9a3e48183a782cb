/*
 * ia_b200.h — C ABI of libia_b200.so: the sm_100a kernels behind the concept-vector
 * extraction / residual-stream injection hot path of tim-hua-01/introspective-awareness.
 *
 * The reference has no FFI of its own (it is pure Python).  Each entry point
 * below replaces a stretch of reference Python that runs inside a forward hook or right
 * after one; the citation on each says which.  Conventions, all entry points:
 *
 *   - plain pointers and sizes only; every pointer except `stream` is a DEVICE pointer;
 *   - nothing is allocated, freed or synchronised inside; the caller owns every buffer;
 *   - work is enqueued on `stream` (a cudaStream_t passed as void*, NULL = default stream)
 *     and the call returns immediately;
 *   - return value: 0 = enqueued; negative = argument error (IA_ERR_*), nothing was launched;
 *     positive = a cudaError_t raised by the launch;
 *   - stateless and thread-safe: calls on different streams may overlap as long as their
 *     buffers (including the workspace) do not — except ia_gather_meandiff_static, which keeps
 *     process-wide state (one built-in scratch array per GPU; see its comment below).
 *
 * Element types.  `dtype` is the MODEL dtype (the residual stream and the norm weight);
 * all arithmetic is fp32 with results rounded to `dtype` exactly where the reference's torch
 * ops round.
 */
#ifndef IA_B200_H
#define IA_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#if defined(__GNUC__)
#define IA_API __attribute__((visibility("default")))
#else
#define IA_API
#endif

#define IA_ABI_VERSION 1

/* element types */
#define IA_BF16 0
#define IA_F16 1
#define IA_F32 2

/* RMSNorm flavours (ia_inject_rmsnorm `variant`) */
#define IA_NORM_LLAMA 0 /* y = w * round(x * r)        LlamaRMSNorm / Qwen2RMSNorm */
#define IA_NORM_GEMMA 1 /* y = round(x * r * (1 + w))  Gemma2RMSNorm               */

/* ia_inject_rmsnorm `flags` */
#define IA_STEER_ALL 1   /* ignore row_t0: steer every position (decode steps; positions=None) */
#define IA_RSQRT_IEEE 2  /* r = 1/sqrt(.) with IEEE div+sqrt (torch CPU) instead of rsqrtf (torch CUDA) */
#define IA_IMPL_TMA 4    /* force the bulk-copy (TMA, cp.async.bulk + mbarrier) staged implementation */
#define IA_IMPL_REGS 8   /* force the register-resident implementation */
#define IA_IMPL_ADD_STAGED 16 /* ia_add_inject_rmsnorm: bulk-copy the branch row into the stage even in long launches */
#define IA_IMPL_STREAM 32 /* force the streaming implementation (row in registers, quarter-row bulk-copy landing slots) */
/* no IMPL bit: automatic, by measurement — streaming for launches of >= 3000 vectors of the two-operand forms
 * (residual add, Gemma-2 block tail; plain form too at H = 5120) where it applies (16-bit dtypes, table in the model
 * dtype, H in 1024/2048/3584/4096/5120); TMA staging otherwise when a row is >= 4 KiB (always for the two-operand
 * forms) and the table is in the model dtype; else registers. */

/* extraction methods (ia_gather_meandiff / ia_meandiff_finalize `mode`) */
#define IA_MODE_BASELINE 0 /* v_c = act_c - mean_b(act_b)      vector_utils.py:541-545           */
#define IA_MODE_SIMPLE 1   /* v_c = act_c - act_control (same arithmetic, n_base == 1) :570        */
#define IA_MODE_RAW 2      /* v_c = act_c  (also: plain gather of captured rows)      :586        */
#define IA_MODE_POSNEG 3   /* v   = mean_c(act_c) - mean_b(act_b), one vector          :103-105    */

/* argument errors */
#define IA_ERR_NULL -1      /* a required pointer is NULL */
#define IA_ERR_SHAPE -2     /* a size is out of range (H % 4 != 0, negative count, ...) */
#define IA_ERR_DTYPE -3     /* unknown dtype / variant / mode */
#define IA_ERR_ALIGN -4     /* a pointer is not 16-byte aligned */
#define IA_ERR_WORKSPACE -5 /* workspace too small */
#define IA_ERR_TOO_LARGE -6 /* size beyond a documented maximum */

/* capacity of the built-in scratch used by ia_gather_meandiff_static */
#define IA_STATIC_SCRATCH_BYTES (4 * 1024 * 1024)

/* Library version (== IA_ABI_VERSION) and a static description of an error code. */
IA_API int ia_version(void);
IA_API const char *ia_error_string(int code);

/*
 * K2 — per-row injection fused into the next RMSNorm.
 *
 * Replaces: the steering hook body of ModelWrapper.generate_batch_with_multi_steering
 * (reference model_utils.py:737 scale+cast, :767 clone, :774-777 decode adds, :784-788 prefill
 * adds, :798-802 all-position add) AND the HF RMSNorm that consumes the hooked tensor next
 * (transformers LlamaRMSNorm / Qwen2RMSNorm / Gemma2RMSNorm .forward, reached through
 * reference model_utils.py:847).  Also covers the single-vector hooks at :377-405 and :592-629
 * (all rows pointing at the same table entry).
 *
 * h [rows, T, H] (dtype, contiguous) is the output of decoder layer L, updated IN PLACE:
 *   for row b, position p:  if row_vec[b] >= 0 and (flags & IA_STEER_ALL or p + pos0 >= row_t0[b])
 *       h[b,p,:] = round(h[b,p,:] + round(vtable[row_vec[b],:] * row_alpha[b]))
 * y [rows, T, H] (dtype) receives RMSNorm(h[b,p,:]; w, eps) of the UPDATED row, in `variant`.
 *
 * vtable [n_vec, H] holds un-scaled concept vectors in `vec_dtype` (IA_F32 or == dtype);
 * row_vec/row_alpha/row_t0 are [rows]; row_t0 is the absolute start position in the padded
 * sequence (the reference's t0 + left-pad count, :822-825) and may be NULL with IA_STEER_ALL.
 * Requirements: H % 4 == 0, 4 <= H <= 65536, pointers 16-byte aligned, rows*T < 2^31.
 * The variance is summed in the order ATen's CUDA mean(-1) uses for these shapes, so with
 * default flags the outputs are bit-identical to the eager torch chain on the same GPU.
 */
IA_API int ia_inject_rmsnorm(void *h, void *y, const void *vtable, const int32_t *row_vec,
                      const float *row_alpha, const int32_t *row_t0, const void *w, float eps,
                      int64_t rows, int64_t T, int64_t H, int32_t pos0, int32_t dtype,
                      int32_t vec_dtype, int32_t variant, int32_t flags, void *stream);

/*
 * K2, Gemma-2 block-tail variant: also absorbs the end of decoder layer L.
 *
 * Replaces, in addition to the above: Gemma2DecoderLayer.forward's
 * `post_feedforward_layernorm(mlp_out)` and `residual + hidden_states` (transformers
 * modeling_gemma2.py), i.e. the last two ops that produce the tensor the reference hook sees.
 *   h[b,p,:] = round(residual[b,p,:] + gemma_norm(mlp_out[b,p,:]; w_post, eps))   then as above.
 * `residual` [rows,T,H] is updated in place to the steered h; `mlp_out` is read only;
 * y = gemma_norm(h; w_next, eps).
 */
IA_API int ia_gemma2_tail_inject_rmsnorm(const void *mlp_out, void *residual, void *y, const void *vtable,
                                  const int32_t *row_vec, const float *row_alpha,
                                  const int32_t *row_t0, const void *w_post, const void *w_next,
                                  float eps, int64_t rows, int64_t T, int64_t H, int32_t pos0,
                                  int32_t dtype, int32_t vec_dtype, int32_t flags, void *stream);

/*
 * K1 — fused residual gather + segmented fp32 mean-difference + L2 norm.
 *
 * Replaces: the capture hook's slice + `.cpu()` (reference model_utils.py:312-321) and the CPU
 * vector arithmetic of extract_concept_vectors_batch (reference vector_utils.py:541-548, :570-573,
 * :586-589) and extract_concept_vector (:103-109), for any number of layers in one launch
 * sequence (the reference re-runs the whole forward per layer, detect_injected_thoughts.py:1551).
 *
 * concept_ptrs / base_ptrs: DEVICE arrays of n_layers pointers, each to a residual-stream
 * tensor [*, Tc, H] / [*, Tb, H] (dtype, contiguous) — the outputs of the selected decoder
 * layers for the concept-prompt batch and the baseline-prompt batch (they may be the same
 * tensors).  Row r of a segment is read at (row_idx[r] * T + tok_idx[r]) * H.
 * Outputs, for n_out = (mode == IA_MODE_POSNEG ? 1 : n_concept):
 *   out_f32     [n_layers, n_out, H] float, may be NULL: the difference with no intermediate
 *               rounding (fp32 accumulate, fp32 subtract);
 *   out_rounded [n_layers, n_out, H] dtype: the reference's result — mean rounded to dtype,
 *               difference rounded to dtype;
 *   out_norm    [n_layers, n_out] float, may be NULL: ||out_rounded|| (fp32 accumulate, rounded
 *               to dtype), taken before normalisation.
 * normalize != 0 divides both outputs by (norm + 1e-8) as the reference does; with normalize == 0 and
 * out_norm == NULL the norm pass is not launched at all (the reference's driver never normalises).
 * Baseline rows are summed sequentially in chunks of 256 (chunk sums added in order), so for
 * n_base <= 256 the mean equals torch's CPU result bit for bit.
 * workspace: at least ia_meandiff_workspace_bytes(...) bytes, 16-byte aligned; contents are
 * scratch (no need to clear).  Requirements: H % 8 == 0.
 */
IA_API int64_t ia_meandiff_workspace_bytes(int64_t n_layers, int64_t n_concept, int64_t n_base, int64_t H,
                                    int32_t mode);

IA_API int ia_gather_meandiff(const void *const *concept_ptrs, int64_t Tc, const int32_t *concept_row,
                       const int32_t *concept_tok, int64_t n_concept, const void *const *base_ptrs,
                       int64_t Tb, const int32_t *base_row, const int32_t *base_tok, int64_t n_base,
                       int64_t n_layers, int64_t H, int32_t mode, int32_t normalize, float *out_f32,
                       void *out_rounded, float *out_norm, void *workspace, int64_t workspace_bytes,
                       int32_t dtype, void *stream);

/*
 * K1 without a caller-supplied workspace.  No counterpart in the reference (which never manages
 * device memory itself); it exists for hosts that cannot allocate.
 *
 * Some callers of this library may use only the device buffers they were given: code embedded in a
 * process whose GPU memory is partitioned up front by someone else, callers inside latency-critical
 * or capture-bound regions where cudaMalloc's device-wide synchronisation is not acceptable, and
 * minimal FFI bindings that bring no allocator.  They can still pass inputs and outputs, but they
 * cannot produce K1's scratch.  This variant takes the same arguments as ia_gather_meandiff minus
 * `workspace` / `workspace_bytes` and uses a fixed IA_STATIC_SCRATCH_BYTES (4 MiB) array that is part
 * of the library image — each GPU that loads the library carries it until the process exits.
 *
 *   - If ia_meandiff_workspace_bytes(...) for the request exceeds IA_STATIC_SCRATCH_BYTES the call
 *     returns IA_ERR_TOO_LARGE and launches nothing (4 MiB covers 64 layers x 50 concepts x H = 5120).
 *   - The first call on a device resolves the array's device address; issue it before any stream
 *     capture begins.
 *   - All calls on a device share the one array: calls that may overlap on the GPU race.  Nothing in
 *     it needs clearing between calls; each call overwrites what it reads.
 *   - No cudaMalloc, cudaFree or allocating library routine is called.
 * Results are bit-identical to ia_gather_meandiff.
 */
IA_API int ia_gather_meandiff_static(const void *const *concept_ptrs, int64_t Tc, const int32_t *concept_row,
                              const int32_t *concept_tok, int64_t n_concept, const void *const *base_ptrs,
                              int64_t Tb, const int32_t *base_row, const int32_t *base_tok, int64_t n_base,
                              int64_t n_layers, int64_t H, int32_t mode, int32_t normalize, float *out_f32,
                              void *out_rounded, float *out_norm, int32_t dtype, void *stream);

/*
 * K1a — rank-local half of K1 for prompt-sharded extraction: gathers `n_rows` rows per layer
 * into out_rows [n_layers, n_rows, H] (dtype, may be NULL) and writes their fp32 column sums
 * (same chunked order) to out_sum [n_layers, H].  The caller all-reduces out_sum across ranks
 * (NCCL) and all-gathers out_rows, then calls ia_meandiff_finalize.  No reference counterpart
 * beyond the lines cited for K1: the reference has no data parallelism.
 */
IA_API int ia_gather_segsum(const void *const *layer_ptrs, int64_t T, const int32_t *row_idx,
                     const int32_t *tok_idx, int64_t n_rows, int64_t n_layers, int64_t H,
                     void *out_rows, float *out_sum, void *workspace, int64_t workspace_bytes,
                     int32_t dtype, void *stream);

/*
 * K1b — finishes K1 from compact inputs: concept_acts [n_layers, n_concept, H] (dtype) and
 * base_sum [n_layers, H] fp32 column sums over n_base_total rows (ignored for IA_MODE_RAW).
 * Same outputs and rounding as ia_gather_meandiff (reference vector_utils.py:541-548).
 */
IA_API int ia_meandiff_finalize(const void *concept_acts, int64_t n_concept, const float *base_sum,
                         int64_t n_base_total, int64_t n_layers, int64_t H, int32_t mode,
                         int32_t normalize, float *out_f32, void *out_rounded, float *out_norm,
                         void *workspace, int64_t workspace_bytes, int32_t dtype, void *stream);

/*
 * K1 from compact inputs — what the K2 tap entry points write: concept_acts [n_layers, n_concept, H] and
 * base_acts [n_layers, n_base, H] (dtype, contiguous; base_acts ignored for IA_MODE_RAW).
 *
 * Replaces the same reference lines as ia_gather_meandiff — the capture hook's slice + `.cpu()` (reference
 * model_utils.py:312-321, done by the tap) and the vector arithmetic of vector_utils.py:541-548, :570-573, :586-589,
 * :103-109 — with no pointer table and no pass over the full residual tensors.  Same outputs, rounding points,
 * summation order and workspace (ia_meandiff_workspace_bytes) as ia_gather_meandiff; bit-identical to it on the
 * same rows.  For n_base <= 256 (the reference's 100 baseline words) this is ONE launch (+ the norm pass when
 * normalize or out_norm asks for it).
 */
IA_API int ia_meandiff_compact(const void *concept_acts, int64_t n_concept, const void *base_acts, int64_t n_base,
                        int64_t n_layers, int64_t H, int32_t mode, int32_t normalize, float *out_f32,
                        void *out_rounded, float *out_norm, void *workspace, int64_t workspace_bytes,
                        int32_t dtype, void *stream);

/*
 * K1a from compact rows: fp32 column sums out_sum [n_layers, H] of rows [n_layers, n_rows, H] (dtype), in the
 * chunked sequential order of ia_gather_segsum — the rank-local half of prompt-sharded extraction when the rows
 * were captured by the K2 tap.  No reference counterpart beyond the lines cited for K1 (no data parallelism there).
 */
IA_API int ia_segsum_compact(const void *rows, int64_t n_rows, int64_t n_layers, int64_t H, float *out_sum,
                      void *workspace, int64_t workspace_bytes, int32_t dtype, void *stream);

/*
 * Runtime support (not a hot-path row of SURVEY.md §8): rotary position embedding of q and k in one
 * launch, with the rounding points of transformers' apply_rotary_pos_emb
 * (modeling_llama.py; reached through reference model_utils.py:338 and :847 like every other model op):
 *     out = round(round(x*cos) + round(rotate_half(x)*sin)).
 * q/q_out are [tokens, n_heads, head_dim], k/k_out [tokens, n_kv_heads, head_dim], cos/sin
 * [tokens, head_dim], all contiguous in `dtype`; head_dim % 16 == 0.  Bit-identical to the eager ops.
 */
IA_API int ia_rope(const void *q, const void *k, const void *cosv, const void *sinv, void *q_out, void *k_out,
            int64_t tokens, int64_t n_heads, int64_t n_kv_heads, int64_t head_dim, int32_t dtype, void *stream);

/*
 * K2 with the decoder layer's residual add folded in:
 *     h <- round(h + add);  h <- round(h + round(v*alpha)) where steered;  y <- rmsnorm(h; w)
 * `add` [rows, T, H] is the branch output (attention or MLP) the layer adds to the residual stream
 * just before the norm — transformers' `hidden_states = residual + hidden_states`
 * (modeling_llama.py LlamaDecoderLayer.forward), which the reference runs between its hook
 * (model_utils.py:767-788) and the next RMSNorm.  Everything else as ia_inject_rmsnorm (same
 * rounding points, same summation order, same flags); h is rewritten for every vector.
 */
IA_API int ia_add_inject_rmsnorm(void *h, const void *add, void *y, const void *vtable, const int32_t *row_vec,
                          const float *row_alpha, const int32_t *row_t0, const void *w, float eps,
                          int64_t rows, int64_t T, int64_t H, int32_t pos0, int32_t dtype,
                          int32_t vec_dtype, int32_t variant, int32_t flags, void *stream);

/*
 * K2 with a capture tap — the three entry points above with one more output.
 *
 * Replaces, in addition to what the un-tapped entry point replaces: the capture hook of
 * ModelWrapper.extract_activations (reference model_utils.py:312-321: slice the hooked layer output at
 * token_idx, detach, `.cpu()`), for the layer whose output this launch finishes.  The kernel already has the
 * row in hand when it writes it back, so the captured row costs one extra H-element store and no extra read:
 *     cap_out[b, :] = h[b, p, :]  (the UPDATED row)   for p = (cap_tok ? cap_tok[b] : cap_t)
 * cap_out [rows, H] (dtype, 16-byte aligned) is compact — what K1 (ia_meandiff_compact) consumes; cap_tok
 * [rows] int32 or NULL; cap_t may be negative (-1 = last position, the reference's token_idx default).
 * cap_out == NULL makes these identical to the un-tapped calls.
 */
IA_API int ia_inject_rmsnorm_tap(void *h, void *y, const void *vtable, const int32_t *row_vec, const float *row_alpha,
                          const int32_t *row_t0, const void *w, float eps, int64_t rows, int64_t T, int64_t H,
                          int32_t pos0, int32_t dtype, int32_t vec_dtype, int32_t variant, int32_t flags,
                          void *cap_out, const int32_t *cap_tok, int32_t cap_t, void *stream);
IA_API int ia_add_inject_rmsnorm_tap(void *h, const void *add, void *y, const void *vtable, const int32_t *row_vec,
                              const float *row_alpha, const int32_t *row_t0, const void *w, float eps,
                              int64_t rows, int64_t T, int64_t H, int32_t pos0, int32_t dtype, int32_t vec_dtype,
                              int32_t variant, int32_t flags, void *cap_out, const int32_t *cap_tok,
                              int32_t cap_t, void *stream);
IA_API int ia_gemma2_tail_inject_rmsnorm_tap(const void *mlp_out, void *residual, void *y, const void *vtable,
                                      const int32_t *row_vec, const float *row_alpha, const int32_t *row_t0,
                                      const void *w_post, const void *w_next, float eps, int64_t rows, int64_t T,
                                      int64_t H, int32_t pos0, int32_t dtype, int32_t vec_dtype, int32_t flags,
                                      void *cap_out, const int32_t *cap_tok, int32_t cap_t, void *stream);

/*
 * Runtime support: ia_rope plus the KV-cache append in the same launch.  q/k/v are [tokens, heads, head_dim]
 * with their own token-row strides in elements (they may be slices of one fused projection output);
 * q_out is dense [tokens, n_heads, head_dim]; rotated k and plain v are written to
 * dst + b*dst_batch_stride + head*dst_head_stride + (pos0 + t)*dst_pos_stride, b = token / tokens_per_row,
 * t = token % tokens_per_row — what transformers does with `past_key_values.update(...)` after
 * apply_rotary_pos_emb (modeling_llama.py LlamaAttention.forward).  Bit-identical to those ops.
 */
IA_API int ia_rope_append(const void *q, const void *k, const void *v, int64_t q_row_stride, int64_t k_row_stride,
                   int64_t v_row_stride, const void *cosv, const void *sinv, void *q_out, void *k_dst, void *v_dst,
                   int64_t dst_batch_stride, int64_t dst_head_stride, int64_t dst_pos_stride, int64_t tokens,
                   int64_t tokens_per_row, int64_t pos0, int64_t n_heads, int64_t n_kv_heads, int64_t head_dim,
                   int32_t dtype, void *stream);

/*
 * Runtime support: out = round(round(silu(gate)) * up) over [rows, width], the two eager kernels of
 * transformers' LlamaMLP.forward (`act_fn(gate_proj(x)) * up_proj(x)`) in one pass with both rounding
 * points kept; gate/up have their own row strides in elements, out is dense.  Bit-identical.
 */
IA_API int ia_silu_mul(const void *gate, const void *up, void *out, int64_t rows, int64_t width, int64_t gate_row_stride,
                int64_t up_row_stride, int32_t dtype, void *stream);

/*
 * Runtime support: out = round(table[gate] * up) over [rows, width] for a 16-bit dtype — `act_fn(gate_proj(x)) * up_proj(x)`
 * of a gated MLP (transformers Gemma2MLP / LlamaMLP .forward) in one pass for ANY activation function: `table` holds
 * act_fn of each of the 65 536 bit patterns of the dtype (index = the bit pattern; the caller fills it once by running
 * its model's own act_fn on all of them), so the first rounding point is the library's own result and the second is
 * the packed single-rounding multiply: bit-identical to the two eager kernels.  gate/up have their own row strides in
 * elements, out is dense; width % 8 == 0; pointers 16-byte aligned.
 */
IA_API int ia_act_mul_table(const void *gate, const void *up, const void *table, void *out, int64_t rows, int64_t width,
                     int64_t gate_row_stride, int64_t up_row_stride, int32_t dtype, void *stream);

#ifdef __cplusplus
}
#endif
#endif /* IA_B200_H */
