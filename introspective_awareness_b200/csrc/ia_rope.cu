// Fused rotary position embedding for q and k in one launch (runtime support, not a §8 hot-path row).
//
// transformers' apply_rotary_pos_emb is, per tensor, mul + slice + neg + cat + mul + add: six
// eager kernels over a strided [B, heads, T, D] view, each rounding to the model dtype.  This kernel
// keeps exactly those rounding points —
//     out = RD( RD(x * cos) + RD(rotate_half(x) * sin) ),   rotate_half(x) = cat(-x[D/2:], x[:D/2])
// — so its output is bit-identical, and touches each element once.  Memory layout is the one the
// projections produce: [B, T, heads, D] (what HF hands over as a transposed view); cos/sin are [B, T, D].

#include "ia_common.cuh"

namespace ia {

template <typename T> struct Rope8;  // 8 consecutive elements: load, rotate-multiply-add, store

template <> struct Rope8<__nv_bfloat16> {
    // packed two-at-a-time bf16 math; single rounding == torch's fp32-then-round for bf16 operands
    // (see ia_k2_common.cuh)
    __device__ static __forceinline__ void run(const __nv_bfloat16 *x, const __nv_bfloat16 *c, const __nv_bfloat16 *s, int half,
                                               __nv_bfloat16 *o) {
        const uint4 x1 = *reinterpret_cast<const uint4 *>(x), x2 = *reinterpret_cast<const uint4 *>(x + half);
        const uint4 c1 = __ldg(reinterpret_cast<const uint4 *>(c)), c2 = __ldg(reinterpret_cast<const uint4 *>(c + half));
        const uint4 s1 = __ldg(reinterpret_cast<const uint4 *>(s)), s2 = __ldg(reinterpret_cast<const uint4 *>(s + half));
        uint4 o1, o2;
        const unsigned *px1 = &x1.x, *px2 = &x2.x, *pc1 = &c1.x, *pc2 = &c2.x, *ps1 = &s1.x, *ps2 = &s2.x;
        unsigned *po1 = &o1.x, *po2 = &o2.x;
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const __nv_bfloat162 a = *reinterpret_cast<const __nv_bfloat162 *>(px1 + i);
            const __nv_bfloat162 b = *reinterpret_cast<const __nv_bfloat162 *>(px2 + i);
            const __nv_bfloat162 nb = __hneg2(b);
            const __nv_bfloat162 lo = __hadd2_rn(__hmul2_rn(a, *reinterpret_cast<const __nv_bfloat162 *>(pc1 + i)),
                                                 __hmul2_rn(nb, *reinterpret_cast<const __nv_bfloat162 *>(ps1 + i)));
            const __nv_bfloat162 hi = __hadd2_rn(__hmul2_rn(b, *reinterpret_cast<const __nv_bfloat162 *>(pc2 + i)),
                                                 __hmul2_rn(a, *reinterpret_cast<const __nv_bfloat162 *>(ps2 + i)));
            po1[i] = *reinterpret_cast<const unsigned *>(&lo);
            po2[i] = *reinterpret_cast<const unsigned *>(&hi);
        }
        *reinterpret_cast<uint4 *>(o) = o1;
        *reinterpret_cast<uint4 *>(o + half) = o2;
    }
};

template <typename T> struct Rope8 {
    __device__ static __forceinline__ void run(const T *x, const T *c, const T *s, int half, T *o) {
#pragma unroll
        for (int q = 0; q < 2; ++q) {  // two groups of four
            float a[4], b[4], c1[4], c2[4], s1[4], s2[4], lo[4], hi[4];
            Elt<T>::unpack(ld4<T>(x + 4 * q), a);
            Elt<T>::unpack(ld4<T>(x + half + 4 * q), b);
            Elt<T>::unpack(ld4_ro<T>(c + 4 * q), c1);
            Elt<T>::unpack(ld4_ro<T>(c + half + 4 * q), c2);
            Elt<T>::unpack(ld4_ro<T>(s + 4 * q), s1);
            Elt<T>::unpack(ld4_ro<T>(s + half + 4 * q), s2);
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                lo[i] = Elt<T>::rd(__fadd_rn(Elt<T>::rd(__fmul_rn(a[i], c1[i])), Elt<T>::rd(__fmul_rn(-b[i], s1[i]))));
                hi[i] = Elt<T>::rd(__fadd_rn(Elt<T>::rd(__fmul_rn(b[i], c2[i])), Elt<T>::rd(__fmul_rn(a[i], s2[i]))));
            }
            Pack4<T> p;
            Elt<T>::pack_exact(lo, p);
            st4<T>(o + 4 * q, p);
            Elt<T>::pack_exact(hi, p);
            st4<T>(o + half + 4 * q, p);
        }
    }
};

// One thread per 8-element chunk of the first half of a head; q heads first, then k heads.
template <typename T>
__global__ void __launch_bounds__(256) rope_kernel(const T *q, const T *k, const T *cosv, const T *sinv, T *qo, T *ko,
                                                   long tokens, int nh, int nkv, int D) {
    const int half = D >> 1;
    const int chunks = half >> 3;
    const long total = tokens * (nh + nkv) * chunks;
    for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long)gridDim.x * blockDim.x) {
        const int ch = (int)(i % chunks);
        const long r = i / chunks;
        const int head = (int)(r % (nh + nkv));
        const long tok = r / (nh + nkv);
        const T *c = cosv + tok * D + ch * 8;
        const T *s = sinv + tok * D + ch * 8;
        if (head < nh) {
            const long off = (tok * nh + head) * (long)D + ch * 8;
            Rope8<T>::run(q + off, c, s, half, qo + off);
        } else {
            const long off = (tok * nkv + (head - nh)) * (long)D + ch * 8;
            Rope8<T>::run(k + off, c, s, half, ko + off);
        }
    }
}

template <typename T>
static int launch_rope(const void *q, const void *k, const void *c, const void *s, void *qo, void *ko, long tokens, int nh,
                       int nkv, int D, cudaStream_t st) {
    const long total = tokens * (nh + nkv) * (D >> 4);
    long blocks = (total + 255) / 256;
    if (blocks > 148L * 16) blocks = 148L * 16;
    if (blocks < 1) blocks = 1;
    rope_kernel<T><<<(unsigned)blocks, 256, 0, st>>>(static_cast<const T *>(q), static_cast<const T *>(k),
                                                     static_cast<const T *>(c), static_cast<const T *>(s),
                                                     static_cast<T *>(qo), static_cast<T *>(ko), tokens, nh, nkv, D);
    return (int)cudaGetLastError();
}


// Rotary embedding of q and k plus the KV-cache append, one launch: q goes to a dense
// [tokens, n_heads, D] buffer, rotated k and the untouched v go straight to their cache slots
// dst[b][head][pos0 + t][:] (element strides sb / sh / sp), b = token / T, t = token % T.  Same
// arithmetic as rope_kernel; what disappears is the separate rotated-k tensor and the two strided
// copies that moved k and v into the cache.  Inputs may be slices of one fused projection output:
// each has its own token-row stride.
struct RopeAppend {
    const void *q, *k, *v;
    long q_rs, k_rs, v_rs;  // elements between consecutive tokens
    const void *cosv, *sinv;
    void *q_out, *k_dst, *v_dst;
    long sb, sh, sp;
    long tokens;
    int T, pos0, nh, nkv, D;
};

template <typename T> __device__ __forceinline__ void copy8x2(const T *x, int half, T *o) {
    if (sizeof(T) == 2) {
        *reinterpret_cast<uint4 *>(o) = *reinterpret_cast<const uint4 *>(x);
        *reinterpret_cast<uint4 *>(o + half) = *reinterpret_cast<const uint4 *>(x + half);
    } else {
#pragma unroll
        for (int q = 0; q < 2; ++q) {
            *reinterpret_cast<uint4 *>(o + 4 * q) = *reinterpret_cast<const uint4 *>(x + 4 * q);
            *reinterpret_cast<uint4 *>(o + half + 4 * q) = *reinterpret_cast<const uint4 *>(x + half + 4 * q);
        }
    }
}

template <typename T> __global__ void __launch_bounds__(256) rope_append_kernel(const RopeAppend a) {
    const int half = a.D >> 1;
    const int chunks = half >> 3;
    const int heads = a.nh + 2 * a.nkv;
    const long total = a.tokens * heads * chunks;
    const T *cosv = static_cast<const T *>(a.cosv), *sinv = static_cast<const T *>(a.sinv);
    for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long)gridDim.x * blockDim.x) {
        const int ch = (int)(i % chunks);
        const long r = i / chunks;
        const int head = (int)(r % heads);
        const long tok = r / heads;
        const T *c = cosv + tok * a.D + ch * 8;
        const T *s = sinv + tok * a.D + ch * 8;
        if (head < a.nh) {
            const T *src = static_cast<const T *>(a.q) + tok * a.q_rs + (long)head * a.D + ch * 8;
            Rope8<T>::run(src, c, s, half, static_cast<T *>(a.q_out) + (tok * a.nh + head) * (long)a.D + ch * 8);
        } else {
            const long b = tok / a.T;
            const long t = tok - b * a.T;
            const bool is_k = head < a.nh + a.nkv;
            const int hh = head - a.nh - (is_k ? 0 : a.nkv);
            const long off = b * a.sb + hh * a.sh + (a.pos0 + t) * a.sp + ch * 8;
            if (is_k)
                Rope8<T>::run(static_cast<const T *>(a.k) + tok * a.k_rs + (long)hh * a.D + ch * 8, c, s, half,
                              static_cast<T *>(a.k_dst) + off);
            else
                copy8x2<T>(static_cast<const T *>(a.v) + tok * a.v_rs + (long)hh * a.D + ch * 8, half,
                           static_cast<T *>(a.v_dst) + off);
        }
    }
}

template <typename T> static int launch_rope_append(const RopeAppend &a, cudaStream_t st) {
    const long total = a.tokens * (a.nh + 2 * a.nkv) * (a.D >> 4);
    long blocks = (total + 255) / 256;
    if (blocks > 148L * 16) blocks = 148L * 16;
    if (blocks < 1) blocks = 1;
    rope_append_kernel<T><<<(unsigned)blocks, 256, 0, st>>>(a);
    return (int)cudaGetLastError();
}

}  // namespace ia

extern "C" int ia_rope(const void *q, const void *k, const void *cosv, const void *sinv, void *q_out, void *k_out,
                       int64_t tokens, int64_t n_heads, int64_t n_kv_heads, int64_t head_dim, int32_t dtype, void *stream) {
    using namespace ia;
    if (!q || !k || !cosv || !sinv || !q_out || !k_out) return IA_ERR_NULL;
    if (tokens < 0 || n_heads < 1 || n_kv_heads < 1 || head_dim < 16 || (head_dim & 15) || head_dim > 1024) return IA_ERR_SHAPE;
    if (!aligned16(q) || !aligned16(k) || !aligned16(cosv) || !aligned16(sinv) || !aligned16(q_out) || !aligned16(k_out))
        return IA_ERR_ALIGN;
    if (tokens == 0) return 0;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    switch (dtype) {
    case IA_BF16: return launch_rope<__nv_bfloat16>(q, k, cosv, sinv, q_out, k_out, tokens, (int)n_heads, (int)n_kv_heads, (int)head_dim, st);
    case IA_F16: return launch_rope<__half>(q, k, cosv, sinv, q_out, k_out, tokens, (int)n_heads, (int)n_kv_heads, (int)head_dim, st);
    case IA_F32: return launch_rope<float>(q, k, cosv, sinv, q_out, k_out, tokens, (int)n_heads, (int)n_kv_heads, (int)head_dim, st);
    default: return IA_ERR_DTYPE;
    }
}

extern "C" int ia_rope_append(const void *q, const void *k, const void *v, int64_t q_row_stride, int64_t k_row_stride,
                              int64_t v_row_stride, const void *cosv, const void *sinv, void *q_out, void *k_dst, void *v_dst,
                              int64_t dst_batch_stride, int64_t dst_head_stride, int64_t dst_pos_stride, int64_t tokens,
                              int64_t tokens_per_row, int64_t pos0, int64_t n_heads, int64_t n_kv_heads, int64_t head_dim,
                              int32_t dtype, void *stream) {
    using namespace ia;
    if (!q || !k || !v || !cosv || !sinv || !q_out || !k_dst || !v_dst) return IA_ERR_NULL;
    if (tokens < 0 || tokens_per_row < 1 || pos0 < 0 || n_heads < 1 || n_kv_heads < 1 || head_dim < 16 || (head_dim & 15) ||
        head_dim > 1024)
        return IA_ERR_SHAPE;
    if (tokens % tokens_per_row) return IA_ERR_SHAPE;
    const int64_t vec = dtype == IA_F32 ? 4 : 8;  // elements per 16 bytes: every stride keeps 16-byte alignment
    if (q_row_stride < n_heads * head_dim || k_row_stride < n_kv_heads * head_dim || v_row_stride < n_kv_heads * head_dim)
        return IA_ERR_SHAPE;
    if (q_row_stride % vec || k_row_stride % vec || v_row_stride % vec || dst_batch_stride % vec || dst_head_stride % vec ||
        dst_pos_stride % vec || dst_pos_stride < head_dim)
        return IA_ERR_ALIGN;
    if (!aligned16(q) || !aligned16(k) || !aligned16(v) || !aligned16(cosv) || !aligned16(sinv) || !aligned16(q_out) ||
        !aligned16(k_dst) || !aligned16(v_dst))
        return IA_ERR_ALIGN;
    if (tokens == 0) return 0;
    RopeAppend a;
    a.q = q; a.k = k; a.v = v; a.q_rs = q_row_stride; a.k_rs = k_row_stride; a.v_rs = v_row_stride;
    a.cosv = cosv; a.sinv = sinv; a.q_out = q_out; a.k_dst = k_dst; a.v_dst = v_dst;
    a.sb = dst_batch_stride; a.sh = dst_head_stride; a.sp = dst_pos_stride; a.tokens = tokens;
    a.T = (int)tokens_per_row; a.pos0 = (int)pos0; a.nh = (int)n_heads; a.nkv = (int)n_kv_heads; a.D = (int)head_dim;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    switch (dtype) {
    case IA_BF16: return launch_rope_append<__nv_bfloat16>(a, st);
    case IA_F16: return launch_rope_append<__half>(a, st);
    case IA_F32: return launch_rope_append<float>(a, st);
    default: return IA_ERR_DTYPE;
    }
}
