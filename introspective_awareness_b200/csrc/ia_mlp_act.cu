// SiLU-gate multiply of the Llama / Qwen2 MLP in one pass (runtime support, not a §8 hot-path row).
//
// transformers computes `act_fn(gate_proj(x)) * up_proj(x)` as two eager kernels: silu (rounds to the
// model dtype) and mul (rounds again).  This kernel keeps both rounding points —
//     out = RD( RD( g / (1 + exp(-g)) ) * u ),   fp32 arithmetic inside, as ATen's silu kernel does
// — so the result is bit-identical, and the [tokens, I] intermediate is written once instead of twice
// and read twice instead of three times.  gate and up may be the two halves of one fused projection
// output: each has its own row stride.

#include "ia_common.cuh"

namespace ia {

template <typename T> __device__ __forceinline__ void silu_mul4(const Pack4<T> &g, const Pack4<T> &u, Pack4<T> &o) {
    float gf[4], uf[4];
    Elt<T>::unpack(g, gf);
    Elt<T>::unpack(u, uf);
#pragma unroll
    for (int i = 0; i < 4; ++i) gf[i] = __fdiv_rn(gf[i], __fadd_rn(1.0f, expf(-gf[i])));
    Pack4<T> tmp;
    Elt<T>::rd4(gf, tmp);  // silu output in the model dtype
#pragma unroll
    for (int i = 0; i < 4; ++i) gf[i] = __fmul_rn(gf[i], uf[i]);
    Elt<T>::rd4(gf, o);
}

// One thread per 16 bytes of a row (8 x 16-bit or 4 x fp32 elements).
template <typename T>
__global__ void __launch_bounds__(256) silu_mul_kernel(const T *gate, const T *up, T *out, long rows, int I, long g_rs, long u_rs) {
    constexpr int E = 16 / sizeof(T);
    const int per_row = I / E;
    const long total = rows * per_row;
    for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long)gridDim.x * blockDim.x) {
        const long r = i / per_row;
        const int c = (int)(i - r * per_row) * E;
        const T *g = gate + r * g_rs + c;
        const T *u = up + r * u_rs + c;
        T *o = out + r * (long)I + c;
        if (sizeof(T) == 2) {
            const uint4 gv = *reinterpret_cast<const uint4 *>(g), uv = *reinterpret_cast<const uint4 *>(u);
            Pack4<T> g0, g1, u0, u1, o0, o1;
            g0.raw = *reinterpret_cast<const decltype(g0.raw) *>(&gv.x);
            g1.raw = *reinterpret_cast<const decltype(g0.raw) *>(&gv.z);
            u0.raw = *reinterpret_cast<const decltype(g0.raw) *>(&uv.x);
            u1.raw = *reinterpret_cast<const decltype(g0.raw) *>(&uv.z);
            silu_mul4<T>(g0, u0, o0);
            silu_mul4<T>(g1, u1, o1);
            uint4 ov;
            *reinterpret_cast<decltype(g0.raw) *>(&ov.x) = o0.raw;
            *reinterpret_cast<decltype(g0.raw) *>(&ov.z) = o1.raw;
            *reinterpret_cast<uint4 *>(o) = ov;
        } else {
            Pack4<T> ov;
            silu_mul4<T>(ld4<T>(g), ld4<T>(u), ov);
            st4<T>(o, ov);
        }
    }
}

template <typename T>
static int launch_silu_mul(const void *gate, const void *up, void *out, long rows, int I, long g_rs, long u_rs, cudaStream_t st) {
    constexpr int E = 16 / sizeof(T);
    const long total = rows * (I / E);
    long blocks = (total + 255) / 256;
    if (blocks > 148L * 32) blocks = 148L * 32;
    if (blocks < 1) blocks = 1;
    silu_mul_kernel<T><<<(unsigned)blocks, 256, 0, st>>>(static_cast<const T *>(gate), static_cast<const T *>(up),
                                                         static_cast<T *>(out), rows, I, g_rs, u_rs);
    return (int)cudaGetLastError();
}

}  // namespace ia

extern "C" int ia_silu_mul(const void *gate, const void *up, void *out, int64_t rows, int64_t width, int64_t gate_row_stride,
                           int64_t up_row_stride, int32_t dtype, void *stream) {
    using namespace ia;
    if (!gate || !up || !out) return IA_ERR_NULL;
    const int64_t vec = dtype == IA_F32 ? 4 : 8;
    if (rows < 0 || width < vec || width % vec || width >= ((int64_t)1 << 31)) return IA_ERR_SHAPE;
    if (gate_row_stride < width || up_row_stride < width) return IA_ERR_SHAPE;
    if (gate_row_stride % vec || up_row_stride % vec || !aligned16(gate) || !aligned16(up) || !aligned16(out)) return IA_ERR_ALIGN;
    if (rows == 0) return 0;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    switch (dtype) {
    case IA_BF16: return launch_silu_mul<__nv_bfloat16>(gate, up, out, rows, (int)width, gate_row_stride, up_row_stride, st);
    case IA_F16: return launch_silu_mul<__half>(gate, up, out, rows, (int)width, gate_row_stride, up_row_stride, st);
    case IA_F32: return launch_silu_mul<float>(gate, up, out, rows, (int)width, gate_row_stride, up_row_stride, st);
    default: return IA_ERR_DTYPE;
    }
}
