// K1: fused residual gather + segmented fp32 mean-difference + L2 norm (sm_100a).
//
// A lane owns 8 consecutive columns (one 128-bit access of a 16-bit dtype, two of fp32); a CTA of
// 8 warps owns a 256-column tile of one layer.  Rows are gathered straight out of the residual-
// stream tensors the forward pass produced (pointer table + row/token index), so the captured
// activations never make a separate trip through HBM or the host.  The rows that feed a column sum
// are staged into shared memory with cp.async by all 8 warps at once (up to 128 rows = 64 KB in
// flight per CTA, one memory round trip), then ONE warp adds them in the contractual sequential
// order out of shared memory; the concept rows are then split over the 8 warps, 4 loads deep.
//
// Numeric contract (include/ia_b200.h, oracle/kernels_np.py:meandiff):
//   column sums: sequential over rows inside chunks of 256 rows, chunk sums added in order;
//   mean = sum / n (IEEE division), rounded to the model dtype for the reference-rounded output;
//   squared norms: per lane fmaf over its 8 columns left to right, xor-butterfly 16..1 across the
//   warp, tiles added in ascending order by one thread.
// HBM-bound streaming: algorithmic bytes per layer (n_concept + n_base)*H*e_in + n_out*H*e_out.

#include <algorithm>
#include <atomic>
#include <type_traits>
#include <mutex>

#include "ia_common.cuh"

namespace ia {

constexpr int kChunkRows = 256;
constexpr int kTileCols = 256;

template <typename T> __device__ __forceinline__ void load8(const T *p, float (&f)[8]);
template <> __device__ __forceinline__ void load8<__nv_bfloat16>(const __nv_bfloat16 *p, float (&f)[8]) {
    const uint4 r = __ldg(reinterpret_cast<const uint4 *>(p));
    f[0] = __uint_as_float(r.x << 16); f[1] = __uint_as_float(r.x & 0xffff0000u);
    f[2] = __uint_as_float(r.y << 16); f[3] = __uint_as_float(r.y & 0xffff0000u);
    f[4] = __uint_as_float(r.z << 16); f[5] = __uint_as_float(r.z & 0xffff0000u);
    f[6] = __uint_as_float(r.w << 16); f[7] = __uint_as_float(r.w & 0xffff0000u);
}
template <> __device__ __forceinline__ void load8<__half>(const __half *p, float (&f)[8]) {
    const uint4 r = __ldg(reinterpret_cast<const uint4 *>(p));
    const __half2 *h = reinterpret_cast<const __half2 *>(&r);
#pragma unroll
    for (int i = 0; i < 4; ++i) { f[2 * i] = __low2float(h[i]); f[2 * i + 1] = __high2float(h[i]); }
}
template <> __device__ __forceinline__ void load8<float>(const float *p, float (&f)[8]) {
    const float4 a = __ldg(reinterpret_cast<const float4 *>(p));
    const float4 b = __ldg(reinterpret_cast<const float4 *>(p) + 1);
    f[0] = a.x; f[1] = a.y; f[2] = a.z; f[3] = a.w; f[4] = b.x; f[5] = b.y; f[6] = b.z; f[7] = b.w;
}

// v[] must hold values representable in T.
template <typename T> __device__ __forceinline__ void store8(T *p, const float (&v)[8]);
template <> __device__ __forceinline__ void store8<__nv_bfloat16>(__nv_bfloat16 *p, const float (&v)[8]) {
    uint4 r;
    r.x = (__float_as_uint(v[0]) >> 16) | (__float_as_uint(v[1]) & 0xffff0000u);
    r.y = (__float_as_uint(v[2]) >> 16) | (__float_as_uint(v[3]) & 0xffff0000u);
    r.z = (__float_as_uint(v[4]) >> 16) | (__float_as_uint(v[5]) & 0xffff0000u);
    r.w = (__float_as_uint(v[6]) >> 16) | (__float_as_uint(v[7]) & 0xffff0000u);
    *reinterpret_cast<uint4 *>(p) = r;
}
template <> __device__ __forceinline__ void store8<__half>(__half *p, const float (&v)[8]) {
    uint4 r;
    __half2 *h = reinterpret_cast<__half2 *>(&r);
#pragma unroll
    for (int i = 0; i < 4; ++i) h[i] = __halves2half2(__float2half_rn(v[2 * i]), __float2half_rn(v[2 * i + 1]));
    *reinterpret_cast<uint4 *>(p) = r;
}
template <> __device__ __forceinline__ void store8<float>(float *p, const float (&v)[8]) {
    reinterpret_cast<float4 *>(p)[0] = make_float4(v[0], v[1], v[2], v[3]);
    reinterpret_cast<float4 *>(p)[1] = make_float4(v[4], v[5], v[6], v[7]);
}

__device__ __forceinline__ float sq8_warp(const float (&v)[8]) {
    float a = 0.f;
#pragma unroll
    for (int i = 0; i < 8; ++i) a = __fmaf_rn(v[i], v[i], a);
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) a = __fadd_rn(a, __shfl_xor_sync(FULL, a, off));
    return a;
}

// Where a segment's rows live: gathered from per-layer residual tensors, or compact.
template <typename T> struct RowSource {
    const void *const *ptrs;  // gathered: [n_layers] tensor base pointers; compact: nullptr
    const T *compact;         // compact: [n_layers, n, H]
    const int32_t *row;
    const int32_t *tok;
    int64_t T_len;
    int64_t n;
    __device__ __forceinline__ const T *at(int layer, int64_t r, int H) const {
        if (ptrs) {
            const T *base = static_cast<const T *>(ptrs[layer]);
            const int64_t b = row ? row[r] : r;
            const int64_t t = tok ? tok[r] : T_len - 1;
            return base + (b * T_len + t) * (int64_t)H;
        }
        return compact + ((int64_t)layer * n + r) * (int64_t)H;
    }
};

struct K1Out {
    float *out_f32;     // [n_layers, n_out, H] or null
    void *out_rounded;  // [n_layers, n_out, H]
    float *part32;      // [n_layers, n_out, ntiles] squared-norm partials of out_f32
    float *partr;       // [n_layers, n_out, ntiles] squared-norm partials of out_rounded
    int ntiles;
};

// Eight consecutive elements as they sit in memory (kept packed while in flight).
template <typename T> struct Raw8 { uint4 v; };
template <> struct Raw8<float> { float4 a, b; };

template <typename T> __device__ __forceinline__ Raw8<T> ldg_raw8(const T *p) {
    Raw8<T> r;
    r.v = __ldg(reinterpret_cast<const uint4 *>(p));
    return r;
}
template <> __device__ __forceinline__ Raw8<float> ldg_raw8<float>(const float *p) {
    Raw8<float> r;
    r.a = __ldg(reinterpret_cast<const float4 *>(p));
    r.b = __ldg(reinterpret_cast<const float4 *>(p) + 1);
    return r;
}
template <typename T> __device__ __forceinline__ void widen8(const Raw8<T> &r, float (&f)[8]);
template <> __device__ __forceinline__ void widen8<__nv_bfloat16>(const Raw8<__nv_bfloat16> &r, float (&f)[8]) {
    f[0] = __uint_as_float(r.v.x << 16); f[1] = __uint_as_float(r.v.x & 0xffff0000u);
    f[2] = __uint_as_float(r.v.y << 16); f[3] = __uint_as_float(r.v.y & 0xffff0000u);
    f[4] = __uint_as_float(r.v.z << 16); f[5] = __uint_as_float(r.v.z & 0xffff0000u);
    f[6] = __uint_as_float(r.v.w << 16); f[7] = __uint_as_float(r.v.w & 0xffff0000u);
}
template <> __device__ __forceinline__ void widen8<__half>(const Raw8<__half> &r, float (&f)[8]) {
    const __half2 *h = reinterpret_cast<const __half2 *>(&r.v);
#pragma unroll
    for (int i = 0; i < 4; ++i) { f[2 * i] = __low2float(h[i]); f[2 * i + 1] = __high2float(h[i]); }
}
template <> __device__ __forceinline__ void widen8<float>(const Raw8<float> &r, float (&f)[8]) {
    f[0] = r.a.x; f[1] = r.a.y; f[2] = r.a.z; f[3] = r.a.w; f[4] = r.b.x; f[5] = r.b.y; f[6] = r.b.z; f[7] = r.b.w;
}

// Rows in flight per warp while emitting concept rows: 8 x 16 B for 16-bit dtypes (all of a warp's
// share of 50 concepts in one round trip), 4 x 32 B for fp32.
template <typename T> struct RowBatch { static constexpr int kRows = sizeof(T) == 4 ? 4 : 8; };

// Issue the loads of concept rows cb, cb+cstep, ... (one batch).  Nothing here depends on the mean.
template <typename T>
__device__ __forceinline__ void load_batch(const RowSource<T> &csrc, int layer, int H, int col, bool active, int64_t cb,
                                           int64_t cstep, Raw8<T> (&raw)[RowBatch<T>::kRows]) {
#pragma unroll
    for (int u = 0; u < RowBatch<T>::kRows; ++u) {
        const int64_t c = cb + u * cstep;
        if (active && c < csrc.n) raw[u] = ldg_raw8<T>(csrc.at(layer, c, H) + col);
    }
}

// Difference against the mean, both roundings, stores and norm partials for one loaded batch.
template <typename T>
__device__ __forceinline__ void emit_batch(const RowSource<T> &csrc, int layer, int tile, int lane, int H, int col, bool active,
                                           bool raw_mode, const float (&mean32)[8], const float (&meanr)[8], int64_t cb,
                                           int64_t cstep, const Raw8<T> (&raw)[RowBatch<T>::kRows], const K1Out &o) {
    const int64_t n_out = csrc.n;
#pragma unroll
    for (int u = 0; u < RowBatch<T>::kRows; ++u) {
        const int64_t c = cb + u * cstep;
        if (c >= n_out) break;  // warp-uniform
        float a[8], v32[8], vr[8];
        if (active) {
            widen8<T>(raw[u], a);
        } else {
#pragma unroll
            for (int i = 0; i < 8; ++i) a[i] = 0.f;
        }
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            v32[i] = raw_mode ? a[i] : __fsub_rn(a[i], mean32[i]);
            vr[i] = raw_mode ? a[i] : Elt<T>::rd(__fsub_rn(a[i], meanr[i]));
        }
        const int64_t orow = (int64_t)layer * n_out + c;
        if (active) {
            if (o.out_f32) store8<float>(o.out_f32 + orow * H + col, v32);
            store8<T>(static_cast<T *>(o.out_rounded) + orow * H + col, vr);
        }
        const float p32 = sq8_warp(v32);
        const float pr = sq8_warp(vr);
        if (lane == 0) {
            o.part32[orow * o.ntiles + tile] = p32;
            o.partr[orow * o.ntiles + tile] = pr;
        }
    }
}

// Concept rows c0, c0+cstep, ... of one (layer, tile), a batch of loads in flight at a time (the
// compiler cannot hoist loads above the output stores on its own).
template <typename T>
__device__ __forceinline__ void emit_rows(const RowSource<T> &csrc, int layer, int tile, int lane, int H, int col,
                                          bool active, bool raw_mode, const float (&mean32)[8], const float (&meanr)[8],
                                          int64_t c0, int64_t cstep, const K1Out &o) {
    for (int64_t cb = c0; cb < csrc.n; cb += cstep * RowBatch<T>::kRows) {
        Raw8<T> raw[RowBatch<T>::kRows];
        load_batch<T>(csrc, layer, H, col, active, cb, cstep, raw);
        emit_batch<T>(csrc, layer, tile, lane, H, col, active, raw_mode, mean32, meanr, cb, cstep, raw, o);
    }
}

// ---- CTA-cooperative column sum ---------------------------------------------------------------
constexpr int kStageBytes = 64 * 1024;  // shared-memory staging buffer for baseline rows
constexpr int kMeanBytes = 2 * kTileCols * (int)sizeof(float);

__device__ __forceinline__ void cp_async16(void *smem_dst, const void *gsrc) {
    const unsigned d = (unsigned)__cvta_generic_to_shared(smem_dst);
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;\n" ::"r"(d), "l"(gsrc));
}
__device__ __forceinline__ void cp_async_drain() {
    asm volatile("cp.async.commit_group;\n" ::);
    asm volatile("cp.async.wait_group 0;\n" ::: "memory");
}

template <typename T> __device__ __forceinline__ void lds8(const unsigned char *p, float (&f)[8]);
template <> __device__ __forceinline__ void lds8<__nv_bfloat16>(const unsigned char *p, float (&f)[8]) {
    const uint4 r = *reinterpret_cast<const uint4 *>(p);
    f[0] = __uint_as_float(r.x << 16); f[1] = __uint_as_float(r.x & 0xffff0000u);
    f[2] = __uint_as_float(r.y << 16); f[3] = __uint_as_float(r.y & 0xffff0000u);
    f[4] = __uint_as_float(r.z << 16); f[5] = __uint_as_float(r.z & 0xffff0000u);
    f[6] = __uint_as_float(r.w << 16); f[7] = __uint_as_float(r.w & 0xffff0000u);
}
template <> __device__ __forceinline__ void lds8<__half>(const unsigned char *p, float (&f)[8]) {
    const uint4 r = *reinterpret_cast<const uint4 *>(p);
    const __half2 *h = reinterpret_cast<const __half2 *>(&r);
#pragma unroll
    for (int i = 0; i < 4; ++i) { f[2 * i] = __low2float(h[i]); f[2 * i + 1] = __high2float(h[i]); }
}
template <> __device__ __forceinline__ void lds8<float>(const unsigned char *p, float (&f)[8]) {
    const float4 a = reinterpret_cast<const float4 *>(p)[0], b = reinterpret_cast<const float4 *>(p)[1];
    f[0] = a.x; f[1] = a.y; f[2] = a.z; f[3] = a.w; f[4] = b.x; f[5] = b.y; f[6] = b.z; f[7] = b.w;
}

// Sequential fp32 column sum of rows [r0, r1) for this CTA's tile.  All warps stage rows into
// shared memory with cp.async, in passes of `pass_rows` rows that alternate between the two halves
// of `stage`; while warp 0 adds pass p in row order, the loads of pass p+1 are already in flight.
// On return warp 0 holds the sums for columns col..col+7 in acc[]; other warps hold zeros.
// With copy_to, the staged rows are also written out compactly (all warps, 16-byte stores).
template <typename T>
__device__ __forceinline__ void cta_column_sum(const RowSource<T> &src, int layer, int64_t r0, int64_t r1, int H, int col,
                                               bool active, unsigned char *stage, int pass_rows, float (&acc)[8], T *copy_to) {
    constexpr int kRowBytes = kTileCols * (int)sizeof(T);
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarps = blockDim.x >> 5;
#pragma unroll
    for (int i = 0; i < 8; ++i) acc[i] = 0.f;
    const int npass = (int)((r1 - r0 + pass_rows - 1) / pass_rows);
    auto issue = [&](int p) {
        const int64_t p0 = r0 + (int64_t)p * pass_rows;
        const int n = (int)((r1 - p0) < pass_rows ? (r1 - p0) : pass_rows);
        unsigned char *buf = stage + (size_t)(p & 1) * pass_rows * kRowBytes;
        if (active) {
            for (int r = warp; r < n; r += nwarps) {
                const T *g = src.at(layer, p0 + r, H) + col;
                unsigned char *d = buf + (size_t)r * kRowBytes + lane * 8 * sizeof(T);
                cp_async16(d, g);
                if (sizeof(T) == 4) cp_async16(d + 16, g + 4);
            }
        }
        asm volatile("cp.async.commit_group;\n" ::);
    };
    if (npass > 0) issue(0);
    for (int p = 0; p < npass; ++p) {
        if (p + 1 < npass) {
            issue(p + 1);
            asm volatile("cp.async.wait_group 1;\n" ::: "memory");
        } else {
            asm volatile("cp.async.wait_group 0;\n" ::: "memory");
        }
        __syncthreads();
        const int64_t p0 = r0 + (int64_t)p * pass_rows;
        const int n = (int)((r1 - p0) < pass_rows ? (r1 - p0) : pass_rows);
        const unsigned char *buf = stage + (size_t)(p & 1) * pass_rows * kRowBytes;
        if (warp == 0 && active) {
            const unsigned char *base = buf + lane * 8 * sizeof(T);
            int r = 0;
            for (; r + 8 <= n; r += 8) {
                float f[8][8];
#pragma unroll
                for (int u = 0; u < 8; ++u) lds8<T>(base + (size_t)(r + u) * kRowBytes, f[u]);
#pragma unroll
                for (int u = 0; u < 8; ++u) {
#pragma unroll
                    for (int i = 0; i < 8; ++i) acc[i] = __fadd_rn(acc[i], f[u][i]);
                }
            }
            for (; r < n; ++r) {
                float f[8];
                lds8<T>(base + (size_t)r * kRowBytes, f);
#pragma unroll
                for (int i = 0; i < 8; ++i) acc[i] = __fadd_rn(acc[i], f[i]);
            }
        }
        if (copy_to && active) {
            for (int r = warp; r < n; r += nwarps) {
                const unsigned char *sp = buf + (size_t)r * kRowBytes + lane * 8 * sizeof(T);
                unsigned char *dp = reinterpret_cast<unsigned char *>(copy_to + (p0 + r) * (int64_t)H + col);
                *reinterpret_cast<uint4 *>(dp) = *reinterpret_cast<const uint4 *>(sp);
                if (sizeof(T) == 4) *reinterpret_cast<uint4 *>(dp + 16) = *reinterpret_cast<const uint4 *>(sp + 16);
            }
        }
        __syncthreads();  // this half of the stage is refilled by pass p+2
    }
}

// Fused path: n_base <= 256 (or RAW).  One CTA (4 warps, 4 CTAs per SM) per (layer, tile, share of the concept
// rows).  Each warp's first batch of concept rows is requested BEFORE the baseline rows are staged and summed:
// only the use of a concept row depends on the mean, not its load, so both sets of loads share one memory round
// trip.  With `nsplit` > 1 (small problems only: fewer units than one wave) the concept rows of a (layer, tile) are
// dealt over nsplit CTAs, each of which stages and sums the baseline rows itself (the same rows in the same order:
// the same mean; the repeats are L2 hits), so that a few-layer extraction still fills the machine.
template <typename T>
__global__ void __launch_bounds__(128, 4) k1_fused_kernel(RowSource<T> csrc, RowSource<T> bsrc, int n_layers, int H, int raw,
                                                          int stage_rows, int nsplit, K1Out o) {
    extern __shared__ __align__(16) unsigned char smem[];
    float *mean_s = reinterpret_cast<float *>(smem);  // [2][kTileCols]: fp32 mean, rounded mean
    unsigned char *stage = smem + kMeanBytes;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarps = blockDim.x >> 5;
    const int split = blockIdx.x % nsplit;
    const int unit = blockIdx.x / nsplit;
    const int layer = unit / o.ntiles;
    const int tile = unit - layer * o.ntiles;
    const int col = tile * kTileCols + lane * 8;
    const bool active = col < H;
    const int64_t c0 = warp + (int64_t)nwarps * split, cstep = (int64_t)nwarps * nsplit;  // this warp's concept rows
    Raw8<T> first[RowBatch<T>::kRows];
    load_batch<T>(csrc, layer, H, col, active, c0, cstep, first);
    float mean32[8], meanr[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) { mean32[i] = 0.f; meanr[i] = 0.f; }
    if (!raw) {
        float acc[8];
        cta_column_sum<T>(bsrc, layer, 0, bsrc.n, H, col, active, stage, stage_rows, acc, nullptr);
        if (warp == 0) {
            const float nb = (float)bsrc.n;
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                const float m = active ? __fdiv_rn(acc[i], nb) : 0.f;
                mean_s[lane * 8 + i] = m;
                mean_s[kTileCols + lane * 8 + i] = Elt<T>::rd(m);
            }
        }
        __syncthreads();
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            mean32[i] = mean_s[lane * 8 + i];
            meanr[i] = mean_s[kTileCols + lane * 8 + i];
        }
    }
    emit_batch<T>(csrc, layer, tile, lane, H, col, active, raw != 0, mean32, meanr, c0, cstep, first, o);
    emit_rows<T>(csrc, layer, tile, lane, H, col, active, raw != 0, mean32, meanr, c0 + cstep * RowBatch<T>::kRows, cstep, o);
}

// ---- column-walker path -------------------------------------------------------------------------
// The same arithmetic with no shared memory, no barrier and no idle warp, for calls with many columns.  A THREAD owns
// four bytes of a row (two 16-bit columns or one fp32 column: a warp reads one full 128-byte line per load) of one
// layer, requests kWalkRows baseline rows at a time into registers, adds them in row order, and then walks the concept
// rows the same way.  The summation order is a per-column contract, so columns are the only parallelism on offer —
// n_layers * H * sizeof(T) / 4 threads.  For a whole sweep (32 x 4096 bf16: 2048 warps, ~14 per SM, each with 50
// line-sized loads outstanding) that keeps HBM busier than the staged kernel above, in which one warp in four adds
// while the others wait at a barrier; a few-layer call stays on the staged kernel, which splits the concept rows of a
// column over CTAs.  Used when nobody reads the squared-norm partials (the production call: SURVEY.md A7, the driver
// never normalises).  Measured on B200, cold L2, 50 concepts x 100 baseline rows, bf16 (tools/k1_walk_probe.py):
//      layers x H      staged     walker
//       1 x 4096      10.4 us    12.4 us
//       8 x 4096      14.1       14.3
//      28 x 3584      20.4       16.4 - 18.1
//      32 x 4096      22.6       19.4 - 20.9
//      64 x 5120      49.3       39.0 - 41.3
// Deeper variants measured and dropped: 25 / 32 / 64 rows per request (20.2 / 22.5 / 25.0 us at 32 x 4096), the next
// request issued before the current rows are used (20.6), eight bytes per thread (20.7), and all 150 rows of a thread
// requested at once through cp.async into a private shared-memory column (22.2).
constexpr int kWalkRows = 50;
constexpr int64_t kWalkMinColumns = 65536;  // n_layers * H from which the walker is used (16 layers x 4096)

template <typename T> __device__ __forceinline__ void walk_widen(uint32_t r, float (&f)[4 / sizeof(T)]) {
    if constexpr (std::is_same<T, __nv_bfloat16>::value) {
        f[0] = __uint_as_float(r << 16);
        f[1] = __uint_as_float(r & 0xffff0000u);
    } else if constexpr (std::is_same<T, __half>::value) {
        const __half2 h = *reinterpret_cast<const __half2 *>(&r);
        f[0] = __low2float(h);
        f[1] = __high2float(h);
    } else {
        f[0] = __uint_as_float(r);
    }
}
// f[] already rounded to T (outputs of Elt<T>::rd)
template <typename T> __device__ __forceinline__ uint32_t walk_pack(const float (&f)[4 / sizeof(T)]) {
    if constexpr (std::is_same<T, __nv_bfloat16>::value) {
        return (__float_as_uint(f[0]) >> 16) | (__float_as_uint(f[1]) & 0xffff0000u);
    } else if constexpr (std::is_same<T, __half>::value) {
        const __half2 h = __floats2half2_rn(f[0], f[1]);
        return *reinterpret_cast<const uint32_t *>(&h);
    } else {
        return __float_as_uint(f[0]);
    }
}

// One thread's view of a segment's rows in one layer.  Compact rows and rows gathered WITHOUT index arrays (row r of a
// [n, T, H] tensor at its last position) are a fixed stride apart; calls with row / token index arrays stay on the
// staged kernel, where a warp, not every thread, reads the indices.
template <typename T> struct WalkRows {
    const T *first;  // this thread's column group of row 0
    int64_t stride;  // elements between consecutive rows
    int64_t n;
    __device__ __forceinline__ WalkRows(const RowSource<T> &src, int layer, int H, int col) : n(src.n) {
        if (src.ptrs) {
            first = static_cast<const T *>(src.ptrs[layer]) + (src.T_len - 1) * (int64_t)H + col;
            stride = src.T_len * (int64_t)H;
        } else {
            first = src.compact + (int64_t)layer * src.n * H + col;
            stride = H;
        }
    }
    __device__ __forceinline__ const uint32_t *at(int64_t r) const { return reinterpret_cast<const uint32_t *>(first + r * stride); }
};

// kWalkRows rows r0.. of one column group, requested together; rows past the end are left untouched (and unused).
template <typename T> __device__ __forceinline__ void walk_request(const WalkRows<T> &rows, int64_t r0, uint32_t (&buf)[kWalkRows]) {
    if (r0 + kWalkRows <= rows.n) {
#pragma unroll
        for (int u = 0; u < kWalkRows; ++u) buf[u] = __ldg(rows.at(r0 + u));
    } else {
#pragma unroll
        for (int u = 0; u < kWalkRows; ++u)
            if (r0 + u < rows.n) buf[u] = __ldg(rows.at(r0 + u));
    }
}

template <typename T>
__global__ void __launch_bounds__(128) k1_walk_kernel(RowSource<T> csrc, RowSource<T> bsrc, int n_layers, int H, int raw,
                                                      float *out_f32, T *out_rounded) {
    constexpr int kN = 4 / (int)sizeof(T);  // columns per thread
    const int per_layer = H / kN;
    const int64_t unit = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (unit >= (int64_t)n_layers * per_layer) return;  // no barrier and no warp-wide operation below
    const int layer = (int)(unit / per_layer);
    const int col = (int)(unit - (int64_t)layer * per_layer) * kN;
    float mean32[kN], meanr[kN];
#pragma unroll
    for (int i = 0; i < kN; ++i) { mean32[i] = 0.f; meanr[i] = 0.f; }
    uint32_t buf[kWalkRows];
    if (!raw) {
        const int64_t nb = bsrc.n;
        const WalkRows<T> rows(bsrc, layer, H, col);
        float sum[kN];
#pragma unroll
        for (int i = 0; i < kN; ++i) sum[i] = 0.f;
#pragma unroll 1
        for (int64_t r0 = 0; r0 < nb; r0 += kWalkRows) {
            walk_request<T>(rows, r0, buf);
#pragma unroll
            for (int u = 0; u < kWalkRows; ++u) {
                if (r0 + u < nb) {
                    float a[kN];
                    walk_widen<T>(buf[u], a);
#pragma unroll
                    for (int i = 0; i < kN; ++i) sum[i] = __fadd_rn(sum[i], a[i]);
                }
            }
        }
        const float n = (float)nb;
#pragma unroll
        for (int i = 0; i < kN; ++i) {
            mean32[i] = __fdiv_rn(sum[i], n);
            meanr[i] = Elt<T>::rd(mean32[i]);
        }
    }
    const int64_t nc = csrc.n;
    const WalkRows<T> rows(csrc, layer, H, col);
    const int64_t obase = (int64_t)layer * nc * H + col;
#pragma unroll 1
    for (int64_t c0 = 0; c0 < nc; c0 += kWalkRows) {
        walk_request<T>(rows, c0, buf);
#pragma unroll
        for (int u = 0; u < kWalkRows; ++u) {
            if (c0 + u < nc) {
                float a[kN], vr[kN];
                walk_widen<T>(buf[u], a);
#pragma unroll
                for (int i = 0; i < kN; ++i) vr[i] = raw ? a[i] : Elt<T>::rd(__fsub_rn(a[i], meanr[i]));
                const int64_t off = obase + (c0 + u) * (int64_t)H;
                if (out_f32) {
#pragma unroll
                    for (int i = 0; i < kN; ++i) out_f32[off + i] = raw ? a[i] : __fsub_rn(a[i], mean32[i]);
                }
                *reinterpret_cast<uint32_t *>(out_rounded + off) = walk_pack<T>(vr);
            }
        }
    }
}

// Chunked column sums: one CTA per (chunk, layer, tile); also the rank-local gather.  Small CTAs
// (4 warps, 32 KB of staging) so that ~7 are resident per SM: a CTA's start-up and drain (pointer
// table -> first rows -> last adds -> store), during which it has no loads in flight, is then
// covered by the others.
constexpr int kSegsumStageBytes = 32 * 1024;
template <typename T>
__global__ void __launch_bounds__(128) k1_segsum_kernel(RowSource<T> src, int n_layers, int H, int ntiles, int nchunks,
                                                        float *part /* [nchunks, n_layers, H] */, T *copy_to) {
    extern __shared__ __align__(16) unsigned char smem[];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int64_t w = blockIdx.x;
    const int tile = (int)(w % ntiles);
    const int layer = (int)((w / ntiles) % n_layers);
    const int chunk = (int)(w / ((int64_t)ntiles * n_layers));
    const int col = tile * kTileCols + lane * 8;
    const bool active = col < H;
    const int64_t r0 = (int64_t)chunk * kChunkRows;
    const int64_t r1 = r0 + kChunkRows < src.n ? r0 + kChunkRows : src.n;
    float acc[8];
    cta_column_sum<T>(src, layer, r0, r1, H, col, active, smem, kSegsumStageBytes / 2 / (kTileCols * (int)sizeof(T)), acc,
                      copy_to ? copy_to + (int64_t)layer * src.n * H : nullptr);
    if (warp == 0 && active) store8<float>(part + ((int64_t)chunk * n_layers + layer) * H + col, acc);
}

// out[l, :] = part[0, l, :] + part[1, l, :] + ... in ascending chunk order.  The adds of a column are a
// strict chain, but the loads need not be: a CTA owns 32 columns, all its threads stage up to 1024
// chunk rows of those columns (128 B each) into shared memory at once with cp.async, then 32 threads
// add down their column out of shared memory.
constexpr int kCrCols = 32;
constexpr int kCrPass = 1024;
__global__ void __launch_bounds__(256) k1_chunk_reduce_kernel(const float *part, float *out, int64_t per_chunk, int nchunks) {
    extern __shared__ __align__(16) unsigned char smem[];
    float *sm = reinterpret_cast<float *>(smem);  // [kCrPass][kCrCols]
    const int t = threadIdx.x;
    const int64_t col0 = (int64_t)blockIdx.x * kCrCols;
    const int ncol = (int)((per_chunk - col0) < kCrCols ? (per_chunk - col0) : kCrCols);  // multiple of 8
    const int q = (t & 7) * 4;  // this thread's 4 columns inside the 32
    float s = 0.f;              // 0 + x is exact, so starting from zero equals starting from chunk 0
    for (int c0 = 0; c0 < nchunks; c0 += kCrPass) {
        const int n = (nchunks - c0) < kCrPass ? (nchunks - c0) : kCrPass;
        if (q < ncol) {
            for (int r = t >> 3; r < n; r += 32)
                cp_async16(sm + r * kCrCols + q, part + (int64_t)(c0 + r) * per_chunk + col0 + q);
        }
        cp_async_drain();
        __syncthreads();
        if (t < ncol) {
#pragma unroll 8
            for (int r = 0; r < n; ++r) s = __fadd_rn(s, sm[r * kCrCols + t]);
        }
        __syncthreads();
    }
    if (t < ncol) out[col0 + t] = s;
}

// Finalize from given column sums.  grid.y splits the concept rows.
template <typename T>
__global__ void __launch_bounds__(128) k1_finalize_kernel(RowSource<T> csrc, const float *base_sum, float n_base,
                                                          const float *pos_sum, float n_pos, int n_layers, int H, int mode,
                                                          K1Out o) {
    const int lane = threadIdx.x & 31;
    const int w = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (w >= n_layers * o.ntiles) return;
    const int layer = w / o.ntiles;
    const int tile = w - layer * o.ntiles;
    const int col = tile * kTileCols + lane * 8;
    const bool active = col < H;
    float mean32[8], meanr[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) { mean32[i] = 0.f; meanr[i] = 0.f; }
    if (mode != IA_MODE_RAW && active) {
        float s[8];
        load8<float>(base_sum + (int64_t)layer * H + col, s);
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            mean32[i] = __fdiv_rn(s[i], n_base);
            meanr[i] = Elt<T>::rd(mean32[i]);
        }
    }
    if (mode == IA_MODE_POSNEG) {
        if (blockIdx.y != 0) return;
        float v32[8], vr[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) { v32[i] = 0.f; vr[i] = 0.f; }
        if (active) {
            float s[8];
            load8<float>(pos_sum + (int64_t)layer * H + col, s);
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                const float p32 = __fdiv_rn(s[i], n_pos);
                v32[i] = __fsub_rn(p32, mean32[i]);
                vr[i] = Elt<T>::rd(__fsub_rn(Elt<T>::rd(p32), meanr[i]));
            }
            if (o.out_f32) store8<float>(o.out_f32 + (int64_t)layer * H + col, v32);
            store8<T>(static_cast<T *>(o.out_rounded) + (int64_t)layer * H + col, vr);
        }
        const float p32 = sq8_warp(v32);
        const float pr = sq8_warp(vr);
        if (lane == 0) {
            o.part32[(int64_t)layer * o.ntiles + tile] = p32;
            o.partr[(int64_t)layer * o.ntiles + tile] = pr;
        }
        return;
    }
    emit_rows<T>(csrc, layer, tile, lane, H, col, active, mode == IA_MODE_RAW, mean32, meanr, blockIdx.y, gridDim.y, o);
}

// One warp per output row: tile partials -> norms; optional normalisation.
template <typename T>
__global__ void __launch_bounds__(128) k1_norm_kernel(int64_t n_rows, int H, int normalize, float *out_norm, K1Out o) {
    const int lane = threadIdx.x & 31;
    const int64_t row = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (row >= n_rows) return;
    float t32 = 0.f, tr = 0.f;
    if (lane == 0) {
        t32 = o.part32[row * o.ntiles];
        tr = o.partr[row * o.ntiles];
        for (int k = 1; k < o.ntiles; ++k) {
            t32 = __fadd_rn(t32, o.part32[row * o.ntiles + k]);
            tr = __fadd_rn(tr, o.partr[row * o.ntiles + k]);
        }
    }
    t32 = __shfl_sync(FULL, t32, 0);
    tr = __shfl_sync(FULL, tr, 0);
    const float norm32 = __fsqrt_rn(t32);
    const float normr = Elt<T>::rd(__fsqrt_rn(tr));
    if (lane == 0 && out_norm) out_norm[row] = normr;
    if (!normalize) return;
    const float den32 = __fadd_rn(norm32, 1e-8f);
    const float denr = Elt<T>::rd(__fadd_rn(normr, 1e-8f));
    T *vr = static_cast<T *>(o.out_rounded) + row * H;
    for (int col = lane * 8; col < H; col += 256) {
        float f[8];
        // plain loads: this kernel writes the same buffer
        Pack4<T> a = ld4<T>(vr + col), b = ld4<T>(vr + col + 4);
        float fa[4], fb[4];
        Elt<T>::unpack(a, fa); Elt<T>::unpack(b, fb);
#pragma unroll
        for (int i = 0; i < 4; ++i) { f[i] = Elt<T>::rd(__fdiv_rn(fa[i], denr)); f[4 + i] = Elt<T>::rd(__fdiv_rn(fb[i], denr)); }
        store8<T>(vr + col, f);
        if (o.out_f32) {
            float *v = o.out_f32 + row * H + col;
            float4 x = reinterpret_cast<float4 *>(v)[0], y = reinterpret_cast<float4 *>(v)[1];
            x.x = __fdiv_rn(x.x, den32); x.y = __fdiv_rn(x.y, den32); x.z = __fdiv_rn(x.z, den32); x.w = __fdiv_rn(x.w, den32);
            y.x = __fdiv_rn(y.x, den32); y.y = __fdiv_rn(y.y, den32); y.z = __fdiv_rn(y.z, den32); y.w = __fdiv_rn(y.w, den32);
            reinterpret_cast<float4 *>(v)[0] = x; reinterpret_cast<float4 *>(v)[1] = y;
        }
    }
}

// ---------------------------------------------------------------------------------------------
// host side
// ---------------------------------------------------------------------------------------------

// cudaFuncSetAttribute once per kernel and device (the first call on a device is a set-up call: make it
// outside any stream capture).
template <typename K> static int opt_in_smem(K kern, std::atomic<bool> (&done)[64], int bytes) {
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) dev = 0;
    if (!done[dev].load(std::memory_order_acquire)) {
        cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
        if (e != cudaSuccess) return (int)e;
        done[dev].store(true, std::memory_order_release);
    }
    return 0;
}

template <typename T>
static int launch_walk(const RowSource<T> &csrc, const RowSource<T> &bsrc, int L, int H, int raw, float *out_f32, T *out_rounded,
                       cudaStream_t s) {
    const int64_t units = (int64_t)L * (H / (4 / (int)sizeof(T)));
    const unsigned grid = (unsigned)((units + 127) / 128);
    k1_walk_kernel<T><<<grid, 128, 0, s>>>(csrc, bsrc, L, H, raw, out_f32, out_rounded);
    return (int)cudaGetLastError();
}

static inline int64_t up64(int64_t x) { return (x + 63) / 64 * 64; }  // floats -> 256-byte multiples
static inline int nchunks_of(int64_t n) { return (int)((n + kChunkRows - 1) / kChunkRows); }

struct WsLayout {
    int64_t part32, partr, base_part, base_sum, pos_part, pos_sum, total;  // offsets in floats
};

static WsLayout ws_layout(int64_t L, int64_t n_concept, int64_t n_base, int64_t H, int mode) {
    const int64_t ntiles = (H + kTileCols - 1) / kTileCols;
    const int64_t n_out = mode == IA_MODE_POSNEG ? 1 : n_concept;
    WsLayout w{};
    int64_t off = 0;
    w.part32 = off; off += up64(L * n_out * ntiles);
    w.partr = off; off += up64(L * n_out * ntiles);
    w.base_part = off; off += up64((int64_t)nchunks_of(n_base) * L * H);
    w.base_sum = off; off += up64(L * H);
    w.pos_part = off; off += mode == IA_MODE_POSNEG ? up64((int64_t)nchunks_of(n_concept) * L * H) : 0;
    w.pos_sum = off; off += mode == IA_MODE_POSNEG ? up64(L * H) : 0;
    w.total = off;
    return w;
}

template <typename T>
static int run_segsum(const RowSource<T> &src, int L, int H, float *part, float *sum_out, T *copy_to, cudaStream_t s) {
    const int ntiles = (H + kTileCols - 1) / kTileCols;
    const int nch = nchunks_of(src.n);
    if (src.n == 0) return (int)cudaMemsetAsync(sum_out, 0, sizeof(float) * (size_t)L * H, s);
    const int64_t ctas = (int64_t)nch * L * ntiles;
    float *dst = nch == 1 ? sum_out : part;
    k1_segsum_kernel<T><<<(unsigned)ctas, 128, kSegsumStageBytes, s>>>(src, L, H, ntiles, nch, dst, copy_to);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return (int)e;
    if (nch > 1) {
        const int64_t per = (int64_t)L * H;
        const int cr_smem = kCrPass * kCrCols * (int)sizeof(float);
        static std::atomic<bool> cr_done[64];
        const int rc = opt_in_smem(k1_chunk_reduce_kernel, cr_done, cr_smem);
        if (rc) return rc;
        k1_chunk_reduce_kernel<<<(unsigned)((per + kCrCols - 1) / kCrCols), 256, cr_smem, s>>>(part, sum_out, per, nch);
        e = cudaGetLastError();
    }
    return (int)e;
}

template <typename T>
static int run_finalize(const RowSource<T> &csrc, const float *base_sum, int64_t n_base, const float *pos_sum,
                        int64_t n_pos, int L, int H, int mode, int normalize, float *out_norm, K1Out o, cudaStream_t s) {
    const int64_t n_out = mode == IA_MODE_POSNEG ? 1 : csrc.n;
    const int warps = L * o.ntiles;
    int split = 1;
    if (mode != IA_MODE_POSNEG) {
        // enough warps to fill the machine: ~8 per SM
        split = (int)std::min<int64_t>(std::max<int64_t>(1, (148 * 8 + warps - 1) / warps), std::max<int64_t>(1, n_out));
    }
    dim3 grid((unsigned)((warps + 3) / 4), (unsigned)split);
    k1_finalize_kernel<T><<<grid, 128, 0, s>>>(csrc, base_sum, (float)n_base, pos_sum, (float)n_pos, L, H, mode, o);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return (int)e;
    if (!normalize && !out_norm) return 0;
    const int64_t rows = (int64_t)L * n_out;
    k1_norm_kernel<T><<<(unsigned)((rows + 3) / 4), 128, 0, s>>>(rows, H, normalize, out_norm, o);
    return (int)cudaGetLastError();
}

template <typename T>
static int meandiff_t(const RowSource<T> &csrc, const RowSource<T> &bsrc, int L, int H, int mode, int normalize, float *out_f32,
                      void *out_rounded, float *out_norm, float *ws, cudaStream_t s) {
    const int64_t nc = csrc.n, nb = bsrc.n;
    const WsLayout wl = ws_layout(L, nc, nb, H, mode);
    const int ntiles = (H + kTileCols - 1) / kTileCols;
    K1Out o{out_f32, out_rounded, ws + wl.part32, ws + wl.partr, ntiles};
    const int64_t n_out = mode == IA_MODE_POSNEG ? 1 : nc;
    if (mode != IA_MODE_POSNEG && (mode == IA_MODE_RAW || nb <= kChunkRows)) {
        const bool want_norm = normalize || out_norm;
        const bool indexed = csrc.row || csrc.tok || (mode != IA_MODE_RAW && (bsrc.row || bsrc.tok));
        if (!want_norm && !indexed && (int64_t)L * H >= kWalkMinColumns)
            return launch_walk<T>(csrc, bsrc, L, H, mode == IA_MODE_RAW, out_f32, static_cast<T *>(out_rounded), s);
        // all baseline rows in one pass when they fit half the staging budget each way (two halves of
        // `stage_rows` rows), otherwise two alternating half-size passes
        const int row_bytes = kTileCols * (int)sizeof(T);
        const int half_rows = kStageBytes / 2 / row_bytes;
        const int stage_rows = nb <= 2 * half_rows ? (int)std::max<int64_t>(nb, 1) : half_rows;
        const int smem = kMeanBytes + (mode == IA_MODE_RAW ? 0 : (nb <= 2 * half_rows ? stage_rows : 2 * stage_rows) * row_bytes);
        static std::atomic<bool> done[64];
        const int rc = opt_in_smem(k1_fused_kernel<T>, done, kMeanBytes + kStageBytes);
        if (rc) return rc;
        // shares of the concept rows per (layer, tile): as many as still fit ONE wave of 4 CTAs per SM — measured on
        // B200: one layer x 16 tiles 15.9 -> 13.1 us with 13 shares; the 32-layer sweep is already a full wave, and
        // splitting it (3 shares, three times the baseline staging) costs 22.8 -> 30.0 us
        const int64_t units = (int64_t)L * ntiles;
        const int nsplit = (int)std::max<int64_t>(1, std::min<int64_t>((4 * 148) / units, std::max<int64_t>(1, (nc + 3) / 4)));
        k1_fused_kernel<T><<<(unsigned)(units * nsplit), 128, smem, s>>>(csrc, bsrc, L, H, mode == IA_MODE_RAW, stage_rows, nsplit, o);
        cudaError_t e = cudaGetLastError();
        if (e != cudaSuccess) return (int)e;
        if (!want_norm) return 0;  // nobody reads the norms
        const int64_t rows = (int64_t)L * n_out;
        k1_norm_kernel<T><<<(unsigned)((rows + 3) / 4), 128, 0, s>>>(rows, H, normalize, out_norm, o);
        return (int)cudaGetLastError();
    }
    int rc = run_segsum<T>(bsrc, L, H, ws + wl.base_part, ws + wl.base_sum, nullptr, s);
    if (rc) return rc;
    if (mode == IA_MODE_POSNEG) {
        rc = run_segsum<T>(csrc, L, H, ws + wl.pos_part, ws + wl.pos_sum, nullptr, s);
        if (rc) return rc;
    }
    return run_finalize<T>(csrc, ws + wl.base_sum, nb, ws + wl.pos_sum, nc, L, H, mode, normalize, out_norm, o, s);
}

template <typename T>
static int gather_meandiff_t(const void *const *cptrs, int64_t Tc, const int32_t *crow, const int32_t *ctok, int64_t nc,
                             const void *const *bptrs, int64_t Tb, const int32_t *brow, const int32_t *btok, int64_t nb,
                             int L, int H, int mode, int normalize, float *out_f32, void *out_rounded, float *out_norm,
                             float *ws, cudaStream_t s) {
    RowSource<T> csrc{cptrs, nullptr, crow, ctok, Tc, nc};
    RowSource<T> bsrc{bptrs, nullptr, brow, btok, Tb, nb};
    return meandiff_t<T>(csrc, bsrc, L, H, mode, normalize, out_f32, out_rounded, out_norm, ws, s);
}

template <typename T>
static int compact_meandiff_t(const void *cacts, int64_t nc, const void *bacts, int64_t nb, int L, int H, int mode, int normalize,
                              float *out_f32, void *out_rounded, float *out_norm, float *ws, cudaStream_t s) {
    RowSource<T> csrc{nullptr, static_cast<const T *>(cacts), nullptr, nullptr, 1, nc};
    RowSource<T> bsrc{nullptr, static_cast<const T *>(bacts), nullptr, nullptr, 1, nb};
    return meandiff_t<T>(csrc, bsrc, L, H, mode, normalize, out_f32, out_rounded, out_norm, ws, s);
}

static int check_sizes(int64_t L, int64_t H, int64_t n) {
    if (L < 1 || L > 4096 || H < 8 || H > 65536 || (H & 7) || n < 0 || n > ((int64_t)1 << 30)) return IA_ERR_SHAPE;
    return 0;
}

}  // namespace ia

extern "C" int64_t ia_meandiff_workspace_bytes(int64_t n_layers, int64_t n_concept, int64_t n_base, int64_t H, int32_t mode) {
    if (ia::check_sizes(n_layers, H, n_concept) || n_base < 0) return IA_ERR_SHAPE;
    return ia::ws_layout(n_layers, n_concept, n_base, H, mode).total * (int64_t)sizeof(float);
}

extern "C" int ia_gather_meandiff(const void *const *concept_ptrs, int64_t Tc, const int32_t *concept_row,
                                  const int32_t *concept_tok, int64_t n_concept, const void *const *base_ptrs,
                                  int64_t Tb, const int32_t *base_row, const int32_t *base_tok, int64_t n_base,
                                  int64_t n_layers, int64_t H, int32_t mode, int32_t normalize, float *out_f32,
                                  void *out_rounded, float *out_norm, void *workspace, int64_t workspace_bytes,
                                  int32_t dtype, void *stream) {
    using namespace ia;
    if (mode < IA_MODE_BASELINE || mode > IA_MODE_POSNEG) return IA_ERR_DTYPE;
    if (!concept_ptrs || !out_rounded || !workspace) return IA_ERR_NULL;
    if (mode != IA_MODE_RAW && !base_ptrs) return IA_ERR_NULL;
    int rc = check_sizes(n_layers, H, n_concept);
    if (rc) return rc;
    if (Tc < 1 || (mode != IA_MODE_RAW && (Tb < 1 || n_base < 1)) || n_base < 0) return IA_ERR_SHAPE;
    if (!aligned16(out_rounded) || !aligned16(workspace) || (out_f32 && !aligned16(out_f32))) return IA_ERR_ALIGN;
    if (mode == IA_MODE_RAW) n_base = 0;
    if (workspace_bytes < ia_meandiff_workspace_bytes(n_layers, n_concept, n_base, H, mode)) return IA_ERR_WORKSPACE;
    if (n_concept == 0) return 0;
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    float *ws = static_cast<float *>(workspace);
    switch (dtype) {
    case IA_BF16:
        return gather_meandiff_t<__nv_bfloat16>(concept_ptrs, Tc, concept_row, concept_tok, n_concept, base_ptrs, Tb, base_row,
                                                base_tok, n_base, (int)n_layers, (int)H, mode, normalize, out_f32, out_rounded,
                                                out_norm, ws, s);
    case IA_F16:
        return gather_meandiff_t<__half>(concept_ptrs, Tc, concept_row, concept_tok, n_concept, base_ptrs, Tb, base_row, base_tok,
                                         n_base, (int)n_layers, (int)H, mode, normalize, out_f32, out_rounded, out_norm, ws, s);
    case IA_F32:
        return gather_meandiff_t<float>(concept_ptrs, Tc, concept_row, concept_tok, n_concept, base_ptrs, Tb, base_row, base_tok,
                                        n_base, (int)n_layers, (int)H, mode, normalize, out_f32, out_rounded, out_norm, ws, s);
    default:
        return IA_ERR_DTYPE;
    }
}

extern "C" int ia_meandiff_compact(const void *concept_acts, int64_t n_concept, const void *base_acts, int64_t n_base,
                                   int64_t n_layers, int64_t H, int32_t mode, int32_t normalize, float *out_f32,
                                   void *out_rounded, float *out_norm, void *workspace, int64_t workspace_bytes,
                                   int32_t dtype, void *stream) {
    using namespace ia;
    if (mode < IA_MODE_BASELINE || mode > IA_MODE_POSNEG) return IA_ERR_DTYPE;
    if (!concept_acts || !out_rounded || !workspace) return IA_ERR_NULL;
    if (mode != IA_MODE_RAW && !base_acts) return IA_ERR_NULL;
    int rc = check_sizes(n_layers, H, n_concept);
    if (rc) return rc;
    if ((mode != IA_MODE_RAW && n_base < 1) || n_base < 0) return IA_ERR_SHAPE;
    if (!aligned16(concept_acts) || (base_acts && !aligned16(base_acts)) || !aligned16(out_rounded) || !aligned16(workspace) ||
        (out_f32 && !aligned16(out_f32)))
        return IA_ERR_ALIGN;
    if (mode == IA_MODE_RAW) n_base = 0;
    if (workspace_bytes < ia_meandiff_workspace_bytes(n_layers, n_concept, n_base, H, mode)) return IA_ERR_WORKSPACE;
    if (n_concept == 0) return 0;
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    float *ws = static_cast<float *>(workspace);
    switch (dtype) {
    case IA_BF16:
        return compact_meandiff_t<__nv_bfloat16>(concept_acts, n_concept, base_acts, n_base, (int)n_layers, (int)H, mode, normalize,
                                                 out_f32, out_rounded, out_norm, ws, s);
    case IA_F16:
        return compact_meandiff_t<__half>(concept_acts, n_concept, base_acts, n_base, (int)n_layers, (int)H, mode, normalize, out_f32,
                                          out_rounded, out_norm, ws, s);
    case IA_F32:
        return compact_meandiff_t<float>(concept_acts, n_concept, base_acts, n_base, (int)n_layers, (int)H, mode, normalize, out_f32,
                                         out_rounded, out_norm, ws, s);
    default:
        return IA_ERR_DTYPE;
    }
}

extern "C" int ia_gather_segsum(const void *const *layer_ptrs, int64_t T, const int32_t *row_idx, const int32_t *tok_idx,
                                int64_t n_rows, int64_t n_layers, int64_t H, void *out_rows, float *out_sum, void *workspace,
                                int64_t workspace_bytes, int32_t dtype, void *stream) {
    using namespace ia;
    if (!layer_ptrs || !out_sum || !workspace) return IA_ERR_NULL;
    int rc = check_sizes(n_layers, H, n_rows);
    if (rc) return rc;
    if (T < 1) return IA_ERR_SHAPE;
    if (!aligned16(out_sum) || !aligned16(workspace) || (out_rows && !aligned16(out_rows))) return IA_ERR_ALIGN;
    if (workspace_bytes < ia_meandiff_workspace_bytes(n_layers, 0, n_rows, H, IA_MODE_BASELINE)) return IA_ERR_WORKSPACE;
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    const WsLayout wl = ws_layout(n_layers, 0, n_rows, H, IA_MODE_BASELINE);
    float *part = static_cast<float *>(workspace) + wl.base_part;
    switch (dtype) {
    case IA_BF16: {
        RowSource<__nv_bfloat16> src{layer_ptrs, nullptr, row_idx, tok_idx, T, n_rows};
        return run_segsum<__nv_bfloat16>(src, (int)n_layers, (int)H, part, out_sum, static_cast<__nv_bfloat16 *>(out_rows), s);
    }
    case IA_F16: {
        RowSource<__half> src{layer_ptrs, nullptr, row_idx, tok_idx, T, n_rows};
        return run_segsum<__half>(src, (int)n_layers, (int)H, part, out_sum, static_cast<__half *>(out_rows), s);
    }
    case IA_F32: {
        RowSource<float> src{layer_ptrs, nullptr, row_idx, tok_idx, T, n_rows};
        return run_segsum<float>(src, (int)n_layers, (int)H, part, out_sum, static_cast<float *>(out_rows), s);
    }
    default:
        return IA_ERR_DTYPE;
    }
}

extern "C" int ia_segsum_compact(const void *rows, int64_t n_rows, int64_t n_layers, int64_t H, float *out_sum, void *workspace,
                                 int64_t workspace_bytes, int32_t dtype, void *stream) {
    using namespace ia;
    if (!rows || !out_sum || !workspace) return IA_ERR_NULL;
    int rc = check_sizes(n_layers, H, n_rows);
    if (rc) return rc;
    if (!aligned16(rows) || !aligned16(out_sum) || !aligned16(workspace)) return IA_ERR_ALIGN;
    if (workspace_bytes < ia_meandiff_workspace_bytes(n_layers, 0, n_rows, H, IA_MODE_BASELINE)) return IA_ERR_WORKSPACE;
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    const WsLayout wl = ws_layout(n_layers, 0, n_rows, H, IA_MODE_BASELINE);
    float *part = static_cast<float *>(workspace) + wl.base_part;
    switch (dtype) {
    case IA_BF16: {
        RowSource<__nv_bfloat16> src{nullptr, static_cast<const __nv_bfloat16 *>(rows), nullptr, nullptr, 1, n_rows};
        return run_segsum<__nv_bfloat16>(src, (int)n_layers, (int)H, part, out_sum, nullptr, s);
    }
    case IA_F16: {
        RowSource<__half> src{nullptr, static_cast<const __half *>(rows), nullptr, nullptr, 1, n_rows};
        return run_segsum<__half>(src, (int)n_layers, (int)H, part, out_sum, nullptr, s);
    }
    case IA_F32: {
        RowSource<float> src{nullptr, static_cast<const float *>(rows), nullptr, nullptr, 1, n_rows};
        return run_segsum<float>(src, (int)n_layers, (int)H, part, out_sum, nullptr, s);
    }
    default:
        return IA_ERR_DTYPE;
    }
}

extern "C" int ia_meandiff_finalize(const void *concept_acts, int64_t n_concept, const float *base_sum, int64_t n_base_total,
                                    int64_t n_layers, int64_t H, int32_t mode, int32_t normalize, float *out_f32,
                                    void *out_rounded, float *out_norm, void *workspace, int64_t workspace_bytes,
                                    int32_t dtype, void *stream) {
    using namespace ia;
    if (mode < IA_MODE_BASELINE || mode > IA_MODE_RAW) return IA_ERR_DTYPE;  // POSNEG needs both sums: use ia_gather_meandiff
    if (!concept_acts || !out_rounded || !workspace) return IA_ERR_NULL;
    if (mode != IA_MODE_RAW && (!base_sum || n_base_total < 1)) return IA_ERR_NULL;
    int rc = check_sizes(n_layers, H, n_concept);
    if (rc) return rc;
    if (!aligned16(concept_acts) || !aligned16(out_rounded) || !aligned16(workspace) || (out_f32 && !aligned16(out_f32)) ||
        (base_sum && !aligned16(base_sum)))
        return IA_ERR_ALIGN;
    if (workspace_bytes < ia_meandiff_workspace_bytes(n_layers, n_concept, 0, H, mode)) return IA_ERR_WORKSPACE;
    if (n_concept == 0) return 0;
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    const WsLayout wl = ws_layout(n_layers, n_concept, 0, H, mode);
    float *ws = static_cast<float *>(workspace);
    const int ntiles = (int)((H + kTileCols - 1) / kTileCols);
    K1Out o{out_f32, out_rounded, ws + wl.part32, ws + wl.partr, ntiles};
    switch (dtype) {
    case IA_BF16: {
        RowSource<__nv_bfloat16> c{nullptr, static_cast<const __nv_bfloat16 *>(concept_acts), nullptr, nullptr, 1, n_concept};
        return run_finalize<__nv_bfloat16>(c, base_sum, n_base_total, nullptr, 1, (int)n_layers, (int)H, mode, normalize, out_norm, o, s);
    }
    case IA_F16: {
        RowSource<__half> c{nullptr, static_cast<const __half *>(concept_acts), nullptr, nullptr, 1, n_concept};
        return run_finalize<__half>(c, base_sum, n_base_total, nullptr, 1, (int)n_layers, (int)H, mode, normalize, out_norm, o, s);
    }
    case IA_F32: {
        RowSource<float> c{nullptr, static_cast<const float *>(concept_acts), nullptr, nullptr, 1, n_concept};
        return run_finalize<float>(c, base_sum, n_base_total, nullptr, 1, (int)n_layers, (int)H, mode, normalize, out_norm, o, s);
    }
    default:
        return IA_ERR_DTYPE;
    }
}

// ---------------------------------------------------------------------------------------------
// ia_gather_meandiff_static: K1 for callers that are not allowed to allocate device memory.
//
// Who needs it: a host that only ever holds the buffers it was handed — a plugin loaded into
// somebody else's process whose device memory is carved up front by the embedding runtime, a
// real-time or capture-mode caller where cudaMalloc (a device-wide synchronising call) is
// off limits, or a thin FFI binding (cgo / JNI / ctypes) with no allocator of its own.  Such a
// caller can pass the inputs and outputs it already owns, but has nowhere to get K1's scratch
// (per-tile norm partials and chunked column sums) from.  Here the scratch lives in the library
// image instead: one fixed array per GPU that loads the library, held until the process exits.
//
// What a caller must know:
//   * the first call on each device looks up the array's device address (the only set-up there
//     is); make that first call outside any stream capture;
//   * requests whose scratch need exceeds IA_STATIC_SCRATCH_BYTES are refused with
//     IA_ERR_TOO_LARGE before anything is launched;
//   * every call on a device uses the same array, so two calls whose kernels can be in flight at
//     the same time (different streams, different host threads) race with each other — keep them
//     ordered; contents left by one call are simply overwritten by the next;
//   * the call itself launches kernels and nothing else: no cudaMalloc / cudaFree, no library
//     routine that allocates.
// ---------------------------------------------------------------------------------------------

namespace ia {

constexpr int64_t kStaticScratchBytes = IA_STATIC_SCRATCH_BYTES;  // 4 MiB: the whole Qwen2.5-32B sweep (64 layers x 5120) needs 3.1 MiB

// File-scope device array.  In host code this name is only a handle for the runtime's symbol
// table, never an address a kernel may be given; the address comes from cudaGetSymbolAddress below.
static __device__ __align__(256) unsigned char g_static_scratch[kStaticScratchBytes];

struct StaticScratchSlot {
    void *ptr = nullptr;
    bool ready = false;
};
static StaticScratchSlot g_scratch_slot[64];
static std::mutex g_scratch_mutex;

// Device address of this GPU's copy of the array; resolved once per device, with that device current.
static int static_scratch_ptr(void **out) {
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return (int)e;
    if (dev < 0 || dev >= 64) return IA_ERR_SHAPE;
    std::lock_guard<std::mutex> lock(g_scratch_mutex);
    StaticScratchSlot &slot = g_scratch_slot[dev];
    if (!slot.ready) {
        void *p = nullptr;
        e = cudaGetSymbolAddress(&p, g_static_scratch);
        if (e != cudaSuccess) return (int)e;
        if (p == nullptr || (reinterpret_cast<uintptr_t>(p) & 255u) != 0) return IA_ERR_ALIGN;
        slot.ptr = p;
        slot.ready = true;
    }
    *out = slot.ptr;
    return 0;
}

}  // namespace ia

extern "C" int ia_gather_meandiff_static(const void *const *concept_ptrs, int64_t Tc, const int32_t *concept_row,
                                         const int32_t *concept_tok, int64_t n_concept, const void *const *base_ptrs,
                                         int64_t Tb, const int32_t *base_row, const int32_t *base_tok, int64_t n_base,
                                         int64_t n_layers, int64_t H, int32_t mode, int32_t normalize, float *out_f32,
                                         void *out_rounded, float *out_norm, int32_t dtype, void *stream) {
    // Size check first, so that an oversized request fails the same way whether or not this
    // device has been set up yet.
    const int64_t need = ia_meandiff_workspace_bytes(n_layers, n_concept, mode == IA_MODE_RAW ? 0 : n_base, H, mode);
    if (need < 0) return (int)need;
    if (need > ia::kStaticScratchBytes) return IA_ERR_TOO_LARGE;
    void *scratch = nullptr;
    int rc = ia::static_scratch_ptr(&scratch);
    if (rc) return rc;
    // Same kernels and launch sequence as ia_gather_meandiff; only the scratch differs.  Every
    // scratch word the kernels read in a call was written earlier in the same call (column sums
    // before the finalize pass, norm partials before the norm pass), so nothing has to be cleared
    // and whatever the previous call left behind is never observed.
    return ia_gather_meandiff(concept_ptrs, Tc, concept_row, concept_tok, n_concept, base_ptrs, Tb, base_row, base_tok,
                              n_base, n_layers, H, mode, normalize, out_f32, out_rounded, out_norm, scratch,
                              ia::kStaticScratchBytes, dtype, stream);
}
