// Shared device helpers: element traits with the reference's rounding points made explicit.
#pragma once

#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/ia_b200.h"

namespace ia {

constexpr unsigned FULL = 0xffffffffu;

// Four consecutive elements of the model dtype, as they sit in memory.
template <typename T> struct Pack4;
template <> struct Pack4<__nv_bfloat16> { uint2 raw; };
template <> struct Pack4<__half> { uint2 raw; };
template <> struct Pack4<float> { float4 raw; };

template <typename T> struct Elt;

template <> struct Elt<__nv_bfloat16> {
    using type = __nv_bfloat16;
    static constexpr int kBytes = 2;
    __device__ static __forceinline__ float rd(float x) { return __bfloat162float(__float2bfloat16_rn(x)); }
    __device__ static __forceinline__ void unpack(const Pack4<type> &p, float (&f)[4]) {
        f[0] = __uint_as_float(p.raw.x << 16);
        f[1] = __uint_as_float(p.raw.x & 0xffff0000u);
        f[2] = __uint_as_float(p.raw.y << 16);
        f[3] = __uint_as_float(p.raw.y & 0xffff0000u);
    }
    // f[] must already hold bf16-representable values (outputs of rd()).
    __device__ static __forceinline__ void pack_exact(const float (&f)[4], Pack4<type> &p) {
        p.raw.x = (__float_as_uint(f[0]) >> 16) | (__float_as_uint(f[1]) & 0xffff0000u);
        p.raw.y = (__float_as_uint(f[2]) >> 16) | (__float_as_uint(f[3]) & 0xffff0000u);
    }
    // Round four floats to bf16 (RNE) with the packed two-at-a-time conversion (F2FP, ALU pipe)
    // instead of four scalar F2F; leaves the rounded values widened in f[] and packed in p.
    __device__ static __forceinline__ void rd4(float (&f)[4], Pack4<type> &p) {
        __nv_bfloat162 a = __floats2bfloat162_rn(f[0], f[1]);
        __nv_bfloat162 b = __floats2bfloat162_rn(f[2], f[3]);
        p.raw.x = *reinterpret_cast<unsigned *>(&a);
        p.raw.y = *reinterpret_cast<unsigned *>(&b);
        unpack(p, f);
    }
};

template <> struct Elt<__half> {
    using type = __half;
    static constexpr int kBytes = 2;
    __device__ static __forceinline__ float rd(float x) { return __half2float(__float2half_rn(x)); }
    __device__ static __forceinline__ void unpack(const Pack4<type> &p, float (&f)[4]) {
        __half2 a = *reinterpret_cast<const __half2 *>(&p.raw.x);
        __half2 b = *reinterpret_cast<const __half2 *>(&p.raw.y);
        f[0] = __low2float(a); f[1] = __high2float(a);
        f[2] = __low2float(b); f[3] = __high2float(b);
    }
    __device__ static __forceinline__ void pack_exact(const float (&f)[4], Pack4<type> &p) {
        __half2 a = __floats2half2_rn(f[0], f[1]);
        __half2 b = __floats2half2_rn(f[2], f[3]);
        p.raw.x = *reinterpret_cast<unsigned *>(&a);
        p.raw.y = *reinterpret_cast<unsigned *>(&b);
    }
    __device__ static __forceinline__ void rd4(float (&f)[4], Pack4<type> &p) {
        pack_exact(f, p);
        unpack(p, f);
    }
};

template <> struct Elt<float> {
    using type = float;
    static constexpr int kBytes = 4;
    __device__ static __forceinline__ float rd(float x) { return x; }
    __device__ static __forceinline__ void unpack(const Pack4<type> &p, float (&f)[4]) {
        f[0] = p.raw.x; f[1] = p.raw.y; f[2] = p.raw.z; f[3] = p.raw.w;
    }
    __device__ static __forceinline__ void pack_exact(const float (&f)[4], Pack4<type> &p) {
        p.raw = make_float4(f[0], f[1], f[2], f[3]);
    }
    __device__ static __forceinline__ void rd4(float (&f)[4], Pack4<type> &p) { pack_exact(f, p); }
};

template <typename T> __device__ __forceinline__ Pack4<T> ld4(const T *p) {
    Pack4<T> r;
    r.raw = *reinterpret_cast<const decltype(r.raw) *>(p);
    return r;
}
template <typename T> __device__ __forceinline__ Pack4<T> ld4_ro(const T *p) {
    Pack4<T> r;
    r.raw = __ldg(reinterpret_cast<const decltype(r.raw) *>(p));
    return r;
}
template <typename T> __device__ __forceinline__ void st4(T *p, const Pack4<T> &v) {
    *reinterpret_cast<decltype(v.raw) *>(p) = v.raw;
}

// 16-byte vector of the model dtype: 8 x 16-bit or 4 x fp32 elements.
template <typename T> struct Vec16 { static constexpr int kElts = 16 / sizeof(T); };

inline bool aligned16(const void *p) { return (reinterpret_cast<uintptr_t>(p) & 15u) == 0; }

}  // namespace ia
