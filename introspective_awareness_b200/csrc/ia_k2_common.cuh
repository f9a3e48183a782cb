// Shared pieces of the K2 kernels (register-resident and bulk-copy staged): parameters and the
// arithmetic with the reference's rounding points (see ia_inject_rmsnorm.cu for the contract).
#pragma once

#include "ia_common.cuh"

namespace ia {

struct K2Params {
    void *h;            // [nvec, H] in/out (TAIL: residual in/out)
    void *y;            // [nvec, H] out
    const void *mlp;    // TAIL only: [nvec, H] MLP output
    const void *add;    // optional [nvec, H]: branch output added to h first (h = rd(h + add)), else NULL
    const void *vtable; // [n_vec, H]
    const int32_t *row_vec;
    const float *row_alpha;
    const int32_t *row_t0;
    const void *w;      // next norm weight [H]
    const void *w_post; // TAIL only: post-feedforward norm weight [H]
    float eps;
    float inv_h;
    int nvec;  // rows * T
    int T;
    int H;
    int pos0;
    int steer_all;
    // optional tap (the extraction capture folded into this pass): the UPDATED row of position cap_pos(row) is
    // also written to cap_out[row, :] — cap_pos(row) = cap_tok[row] if cap_tok else cap_t
    void *cap_out;           // [rows, H] or NULL
    const int32_t *cap_tok;  // [rows] or NULL
    int cap_t;
};

__device__ __forceinline__ bool tap_here(const K2Params &p, int row, int pos) {
    return p.cap_out != nullptr && pos == (p.cap_tok ? __ldg(p.cap_tok + row) : p.cap_t);
}

template <bool IEEE> __device__ __forceinline__ float rinv(float var, float eps) {
    float t = __fadd_rn(var, eps);
    if (IEEE) return __fdiv_rn(1.0f, __fsqrt_rn(t));
    return rsqrtf(t);
}

__device__ __forceinline__ float warp_total(float a0, float a1, float a2, float a3) {
    float s = __fadd_rn(__fadd_rn(__fadd_rn(a0, a1), a2), a3);
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) s = __fadd_rn(s, __shfl_down_sync(FULL, s, off));
    return __shfl_sync(FULL, s, 0);
}

// cp.async (LDGSTS) of one 4-element group: asynchronous, no register staging.
template <int BYTES> __device__ __forceinline__ void cp_async_group(void *smem_dst, const void *gsrc) {
    const unsigned d = (unsigned)__cvta_generic_to_shared(smem_dst);
    asm volatile("cp.async.ca.shared.global [%0], [%1], %2;\n" ::"r"(d), "l"(gsrc), "n"(BYTES));
}
__device__ __forceinline__ void cp_async_wait_all() {
    asm volatile("cp.async.commit_group;\n" ::);
    asm volatile("cp.async.wait_group 0;\n" ::: "memory");
}

template <typename T>
__device__ __forceinline__ void sumsq4(const Pack4<T> &x, float &a0, float &a1, float &a2, float &a3) {
    float f[4];
    Elt<T>::unpack(x, f);
    a0 = __fadd_rn(a0, __fmul_rn(f[0], f[0]));
    a1 = __fadd_rn(a1, __fmul_rn(f[1], f[1]));
    a2 = __fadd_rn(a2, __fmul_rn(f[2], f[2]));
    a3 = __fadd_rn(a3, __fmul_rn(f[3], f[3]));
}

// y group = norm(x group): LLAMA rounds x*r to the dtype before the weight multiply.
template <typename T, int VARIANT>
__device__ __forceinline__ void norm4(const Pack4<T> &x, float r, const Pack4<T> &w, Pack4<T> &out) {
    float f[4], wf[4];
    Elt<T>::unpack(x, f);
    Elt<T>::unpack(w, wf);
#pragma unroll
    for (int i = 0; i < 4; ++i) f[i] = __fmul_rn(f[i], r);
    if (VARIANT == IA_NORM_LLAMA) {
        Pack4<T> tmp;
        Elt<T>::rd4(f, tmp);
#pragma unroll
        for (int i = 0; i < 4; ++i) f[i] = __fmul_rn(wf[i], f[i]);
    } else {
#pragma unroll
        for (int i = 0; i < 4; ++i) f[i] = __fmul_rn(f[i], __fadd_rn(1.0f, wf[i]));
    }
    Elt<T>::rd4(f, out);
}

// x group += round(v group * alpha), rounded to the dtype.  `Injector` is built once per row.
template <typename T, typename VT> struct Injector {
    float alpha;
    __device__ __forceinline__ explicit Injector(float a) : alpha(a) {}
    __device__ __forceinline__ void apply(Pack4<T> &x, const Pack4<VT> &v) const {
        float f[4], vf[4];
        Elt<T>::unpack(x, f);
        Elt<VT>::unpack(v, vf);
#pragma unroll
        for (int i = 0; i < 4; ++i) vf[i] = __fmul_rn(vf[i], alpha);
        Pack4<T> tmp;
        Elt<T>::rd4(vf, tmp);
#pragma unroll
        for (int i = 0; i < 4; ++i) f[i] = __fadd_rn(f[i], vf[i]);
        Elt<T>::rd4(f, x);
    }
};

// x group = round(x group + a group): the decoder layer's `residual + branch_output`.
template <typename T> __device__ __forceinline__ void add4(Pack4<T> &x, const Pack4<T> &a) {
    float f[4], af[4];
    Elt<T>::unpack(x, f);
    Elt<T>::unpack(a, af);
#pragma unroll
    for (int i = 0; i < 4; ++i) f[i] = __fadd_rn(f[i], af[i]);
    Elt<T>::rd4(f, x);
}

// ---- bf16 fast paths: two elements per instruction, same bits -----------------------------------
// torch computes a bf16 op as: widen to fp32, one fp32 op (rounded), round to bf16.  For bf16
// operands that double rounding is provably the single rounding `mul.rn.bf16x2` / `add.rn.bf16x2`
// perform on the exact result:
//   * a product of two bf16 values has <= 16 significant bits, so the fp32 product is exact;
//   * a sum of two bf16 values is inexact in fp32 only when the operands' exponents differ by >= 17,
//     and then the small operand is < 2^-16 of the large one — far from any bf16 rounding boundary
//     of the result (half an ulp is 2^-9), so rounding the fp32 value or the exact value to bf16
//     gives the same number.
// The `_rn` intrinsic forms are never contracted into an fma.  (tests: bit-exact against the numpy
// oracle and against the eager torch chain on the GPU.)
__device__ __forceinline__ __nv_bfloat162 as_bf162(unsigned u) { return *reinterpret_cast<__nv_bfloat162 *>(&u); }
__device__ __forceinline__ unsigned as_u32(__nv_bfloat162 v) { return *reinterpret_cast<unsigned *>(&v); }

template <> struct Injector<__nv_bfloat16, __nv_bfloat16> {
    float alpha;
    __device__ __forceinline__ explicit Injector(float a) : alpha(a) {}
    __device__ __forceinline__ void apply(Pack4<__nv_bfloat16> &x, const Pack4<__nv_bfloat16> &v) const {
        // v*alpha with an arbitrary fp32 alpha: fp32 product (rounded), then to bf16 — as torch does;
        // the add is the packed single-rounding form (see above).
        float vf[4];
        Elt<__nv_bfloat16>::unpack(v, vf);
        const __nv_bfloat162 s0 = __floats2bfloat162_rn(__fmul_rn(vf[0], alpha), __fmul_rn(vf[1], alpha));
        const __nv_bfloat162 s1 = __floats2bfloat162_rn(__fmul_rn(vf[2], alpha), __fmul_rn(vf[3], alpha));
        x.raw.x = as_u32(__hadd2_rn(as_bf162(x.raw.x), s0));
        x.raw.y = as_u32(__hadd2_rn(as_bf162(x.raw.y), s1));
    }
};

template <> __device__ __forceinline__ void add4<__nv_bfloat16>(Pack4<__nv_bfloat16> &x, const Pack4<__nv_bfloat16> &a) {
    x.raw.x = as_u32(__hadd2_rn(as_bf162(x.raw.x), as_bf162(a.raw.x)));
    x.raw.y = as_u32(__hadd2_rn(as_bf162(x.raw.y), as_bf162(a.raw.y)));
}

template <>
__device__ __forceinline__ void norm4<__nv_bfloat16, IA_NORM_LLAMA>(const Pack4<__nv_bfloat16> &x, float r,
                                                                    const Pack4<__nv_bfloat16> &w,
                                                                    Pack4<__nv_bfloat16> &out) {
    float f[4];
    Elt<__nv_bfloat16>::unpack(x, f);
    const __nv_bfloat162 n0 = __floats2bfloat162_rn(__fmul_rn(f[0], r), __fmul_rn(f[1], r));
    const __nv_bfloat162 n1 = __floats2bfloat162_rn(__fmul_rn(f[2], r), __fmul_rn(f[3], r));
    out.raw.x = as_u32(__hmul2_rn(as_bf162(w.raw.x), n0));  // weight * round(x*r): bf16 x bf16, exact in fp32
    out.raw.y = as_u32(__hmul2_rn(as_bf162(w.raw.y), n1));
}

// ---- mbarrier / bulk-copy (TMA) helpers shared by the staged and the streaming kernels ----------
__device__ __forceinline__ unsigned smem_u32(const void *p) { return (unsigned)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(void *bar, unsigned count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;\n" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(void *bar, unsigned bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;\n" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(void *bar, unsigned parity) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "WAIT_%=:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra DONE_%=;\n"
        "bra WAIT_%=;\n"
        "DONE_%=:\n"
        "}\n" ::"r"(smem_u32(bar)),
        "r"(parity)
        : "memory");
}
__device__ __forceinline__ void bulk_g2s(void *smem_dst, const void *gsrc, unsigned bytes, void *bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];\n" ::"r"(
                     smem_u32(smem_dst)),
                 "l"(gsrc), "r"(bytes), "r"(smem_u32(bar))
                 : "memory");
}
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;\n" ::: "memory"); }

struct RowMeta {
    int vi;
    float alpha;
    bool steer;
    bool tap;  // this vector's updated row is also written to cap_out[row]
    int row;
};

template <typename T> __device__ __forceinline__ RowMeta load_meta(const K2Params &p, int vec) {
    RowMeta m;
    const int row = vec / p.T;
    const int pos = vec - row * p.T;
    m.vi = p.row_vec ? __ldg(p.row_vec + row) : -1;
    const int t0 = (p.row_vec && !p.steer_all) ? __ldg(p.row_t0 + row) : 0;
    m.alpha = p.row_vec ? __ldg(p.row_alpha + row) : 0.0f;
    m.steer = m.vi >= 0 && (p.steer_all || (pos + p.pos0 >= t0));
    m.row = row;
    m.tap = tap_here(p, row, pos);
    return m;
}


// Bulk-copy (TMA) staged implementation, ia_inject_rmsnorm_tma.cu.  Returns IA_ERR_DTYPE / IA_ERR_SHAPE for
// combinations it does not cover (callers then use the register kernel).
int launch_k2_tma(const K2Params &p, int dtype, int vec_dtype, int variant, int flags, cudaStream_t s);
// Streaming implementation (row in registers, quarter-row landing slots), ia_inject_rmsnorm_stream.cu.
// k2_stream_supported: does it cover this shape / dtype combination at all?
bool k2_stream_supported(const K2Params &p, int dtype, int vec_dtype);
int launch_k2_stream(const K2Params &p, int dtype, int vec_dtype, int variant, int flags, cudaStream_t s);
int sm_count();

}  // namespace ia
