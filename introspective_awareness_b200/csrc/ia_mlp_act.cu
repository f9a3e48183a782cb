// SiLU-gate multiply of the Llama / Qwen2 MLP in one pass (runtime support, not a §8 hot-path row).
//
// transformers computes `act_fn(gate_proj(x)) * up_proj(x)` as two eager kernels: silu (rounds to the
// model dtype) and mul (rounds again).  This kernel keeps both rounding points —
//     out = RD( RD( g / (1 + exp(-g)) ) * u ),   fp32 arithmetic inside, as ATen's silu kernel does
// — so the result is bit-identical, and the [tokens, I] intermediate is written once instead of twice
// and read twice instead of three times.  gate and up may be the two halves of one fused projection
// output: each has its own row stride.

#include <atomic>

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

// Large launches of 16-bit dtypes (prompt passes: hundreds of millions of elements).  The direct kernel above is
// ALU-bound there — an exact `expf` and an IEEE division per element cost ~25 instructions where the memory system
// needs the time of ~10 — but RD(silu(x)) of a 16-bit x is a pure function of 65 536 bit patterns.  Each CTA first
// fills a 128 KB table in shared memory with that function, evaluated by the very expression of the direct kernel
// (so the bits cannot differ), then streams: per element one 16-bit table read and the packed multiply by `up`
// (the product of two 16-bit values is exact in fp32, so `mul.rn` on the pair rounds once, where torch's
// widen -> multiply -> round does).  Measured on B200, 25 600 x 14 336 bf16 (a Llama-3.1-8B prompt pass): 386 us cold =
// 5.7 TB/s, 0.87 of the measured copy bandwidth, against ~550 us for the direct kernel inside the bench step (ALU-bound,
// so it also slows with the power-capped SM clock; this one does not).  The table costs ~15 us per launch, so only
// launches of >= 64 M elements take this path.
template <typename T> __device__ __forceinline__ unsigned short silu_bits(unsigned short xb);
template <> __device__ __forceinline__ unsigned short silu_bits<__nv_bfloat16>(unsigned short xb) {
    const float x = __uint_as_float((unsigned)xb << 16);
    const __nv_bfloat16 r = __float2bfloat16_rn(__fdiv_rn(x, __fadd_rn(1.0f, expf(-x))));
    return *reinterpret_cast<const unsigned short *>(&r);
}
template <> __device__ __forceinline__ unsigned short silu_bits<__half>(unsigned short xb) {
    const float x = __half2float(*reinterpret_cast<const __half *>(&xb));
    const __half r = __float2half_rn(__fdiv_rn(x, __fadd_rn(1.0f, expf(-x))));
    return *reinterpret_cast<const unsigned short *>(&r);
}
template <typename T> __device__ __forceinline__ unsigned mul2(unsigned a, unsigned b);
template <> __device__ __forceinline__ unsigned mul2<__nv_bfloat16>(unsigned a, unsigned b) {
    const __nv_bfloat162 r = __hmul2_rn(*reinterpret_cast<const __nv_bfloat162 *>(&a), *reinterpret_cast<const __nv_bfloat162 *>(&b));
    return *reinterpret_cast<const unsigned *>(&r);
}
template <> __device__ __forceinline__ unsigned mul2<__half>(unsigned a, unsigned b) {
    const __half2 r = __hmul2_rn(*reinterpret_cast<const __half2 *>(&a), *reinterpret_cast<const __half2 *>(&b));
    return *reinterpret_cast<const unsigned *>(&r);
}

constexpr int kSiluTableBytes = 65536 * 2;

template <typename T>
__global__ void __launch_bounds__(1024, 1) silu_mul_table_kernel(const T *gate, const T *up, T *out, long rows, int I, long g_rs, long u_rs) {
    extern __shared__ __align__(16) unsigned short silu_tab[];
    for (unsigned i = threadIdx.x; i < 65536u; i += blockDim.x) silu_tab[i] = silu_bits<T>((unsigned short)i);
    __syncthreads();
    const int per_row = I / 8;
    const long total = rows * per_row;
    for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long)gridDim.x * blockDim.x) {
        const long r = i / per_row;
        const int c = (int)(i - r * per_row) * 8;
        const uint4 gv = *reinterpret_cast<const uint4 *>(gate + r * g_rs + c);
        const uint4 uv = *reinterpret_cast<const uint4 *>(up + r * u_rs + c);
        const unsigned gw[4] = {gv.x, gv.y, gv.z, gv.w}, uw[4] = {uv.x, uv.y, uv.z, uv.w};
        unsigned ow[4];
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const unsigned s = (unsigned)silu_tab[gw[k] & 0xffffu] | ((unsigned)silu_tab[gw[k] >> 16] << 16);
            ow[k] = mul2<T>(s, uw[k]);
        }
        *reinterpret_cast<uint4 *>(out + r * (long)I + c) = make_uint4(ow[0], ow[1], ow[2], ow[3]);
    }
}

constexpr long kSiluTableMinElements = 64L * 1024 * 1024;

template <typename T> static int launch_silu_mul_table(const void *gate, const void *up, void *out, long rows, int I, long g_rs, long u_rs,
                                                       cudaStream_t st) {
    static std::atomic<bool> done[64];  // opt in to 128 KB of dynamic shared memory once per device
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) dev = 0;
    if (!done[dev].load(std::memory_order_acquire)) {
        cudaError_t e = cudaFuncSetAttribute(silu_mul_table_kernel<T>, cudaFuncAttributeMaxDynamicSharedMemorySize, kSiluTableBytes);
        if (e != cudaSuccess) return (int)e;
        done[dev].store(true, std::memory_order_release);
    }
    int sms = 148;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    silu_mul_table_kernel<T><<<(unsigned)sms, 1024, kSiluTableBytes, st>>>(static_cast<const T *>(gate), static_cast<const T *>(up),
                                                                          static_cast<T *>(out), rows, I, g_rs, u_rs);
    return (int)cudaGetLastError();
}

// out = RD(table[gate] * up): the gate activation looked up in a caller-supplied table of all 65 536 values of the
// 16-bit dtype (128 KB, copied to shared memory by every CTA), then the packed single-rounding multiply.  The caller
// fills the table with whatever its model's activation function returns for each bit pattern — e.g. ATen's tanh-GELU
// for Gemma-2, whose fused-multiply-add contractions a hand-written GELU would have to mirror to reproduce its bits —
// so the result is bit-identical to `act_fn(gate) * up` by construction.
template <typename T>
__global__ void __launch_bounds__(1024, 1) act_mul_table_kernel(const T *gate, const T *up, const unsigned short *table, T *out,
                                                                long rows, int I, long g_rs, long u_rs) {
    extern __shared__ __align__(16) unsigned short act_tab[];
    for (unsigned i = threadIdx.x; i < 65536u / 8; i += blockDim.x)
        reinterpret_cast<uint4 *>(act_tab)[i] = __ldg(reinterpret_cast<const uint4 *>(table) + i);
    __syncthreads();
    const int per_row = I / 8;
    const long total = rows * per_row;
    for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long)gridDim.x * blockDim.x) {
        const long r = i / per_row;
        const int c = (int)(i - r * per_row) * 8;
        const uint4 gv = *reinterpret_cast<const uint4 *>(gate + r * g_rs + c);
        const uint4 uv = *reinterpret_cast<const uint4 *>(up + r * u_rs + c);
        const unsigned gw[4] = {gv.x, gv.y, gv.z, gv.w}, uw[4] = {uv.x, uv.y, uv.z, uv.w};
        unsigned ow[4];
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const unsigned s = (unsigned)act_tab[gw[k] & 0xffffu] | ((unsigned)act_tab[gw[k] >> 16] << 16);
            ow[k] = mul2<T>(s, uw[k]);
        }
        *reinterpret_cast<uint4 *>(out + r * (long)I + c) = make_uint4(ow[0], ow[1], ow[2], ow[3]);
    }
}

template <typename T> static int launch_act_mul_table(const void *gate, const void *up, const void *table, void *out, long rows, int I,
                                                      long g_rs, long u_rs, cudaStream_t st) {
    static std::atomic<bool> done[64];
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) dev = 0;
    if (!done[dev].load(std::memory_order_acquire)) {
        cudaError_t e = cudaFuncSetAttribute(act_mul_table_kernel<T>, cudaFuncAttributeMaxDynamicSharedMemorySize, kSiluTableBytes);
        if (e != cudaSuccess) return (int)e;
        done[dev].store(true, std::memory_order_release);
    }
    int sms = 148;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    // every CTA pays for a 128 KB table copy (an L2 hit after the first): small launches use fewer CTAs
    const long per_cta = 1024L * 8 * 16;
    long ctas = (rows * (long)I + per_cta - 1) / per_cta;
    if (ctas > sms) ctas = sms;
    if (ctas < 1) ctas = 1;
    act_mul_table_kernel<T><<<(unsigned)ctas, 1024, kSiluTableBytes, st>>>(static_cast<const T *>(gate), static_cast<const T *>(up),
                                                                          static_cast<const unsigned short *>(table),
                                                                          static_cast<T *>(out), rows, I, g_rs, u_rs);
    return (int)cudaGetLastError();
}

template <typename T>
static int launch_silu_mul(const void *gate, const void *up, void *out, long rows, int I, long g_rs, long u_rs, cudaStream_t st) {
    constexpr int E = 16 / sizeof(T);
    const long total = rows * (I / E);
    if constexpr (sizeof(T) == 2) {
        if (rows * (long)I >= kSiluTableMinElements) return launch_silu_mul_table<T>(gate, up, out, rows, I, g_rs, u_rs, st);
    }
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

extern "C" int ia_act_mul_table(const void *gate, const void *up, const void *table, void *out, int64_t rows, int64_t width,
                                int64_t gate_row_stride, int64_t up_row_stride, int32_t dtype, void *stream) {
    using namespace ia;
    if (!gate || !up || !out || !table) return IA_ERR_NULL;
    if (dtype != IA_BF16 && dtype != IA_F16) return IA_ERR_DTYPE;  // the table enumerates a 16-bit dtype
    if (rows < 0 || width < 8 || width % 8 || width >= ((int64_t)1 << 31)) return IA_ERR_SHAPE;
    if (gate_row_stride < width || up_row_stride < width) return IA_ERR_SHAPE;
    if (gate_row_stride % 8 || up_row_stride % 8 || !aligned16(gate) || !aligned16(up) || !aligned16(out) || !aligned16(table))
        return IA_ERR_ALIGN;
    if (rows == 0) return 0;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    if (dtype == IA_BF16)
        return launch_act_mul_table<__nv_bfloat16>(gate, up, table, out, rows, (int)width, gate_row_stride, up_row_stride, st);
    return launch_act_mul_table<__half>(gate, up, table, out, rows, (int)width, gate_row_stride, up_row_stride, st);
}
