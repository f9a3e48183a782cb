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
