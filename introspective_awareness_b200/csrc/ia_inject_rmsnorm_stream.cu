// K2, streaming implementation for long launches on sm_100a (prompt passes: thousands of vectors).
//
// Same arithmetic, same summation order, same bits as the register and the staged kernels.  What is
// different is where a row lives while it is worked on, and therefore how many bytes a warp can keep
// in flight:
//   * the row under work is held in REGISTERS (packed, H/128 x 64 bits per lane), not in shared memory;
//   * shared memory is only a landing zone: each warp owns a ring of NS quarter-row slots, filled by
//     `cp.async.bulk` (TMA; SASS UBLKCP) and completed on one mbarrier per slot.  A slot is free again
//     as soon as its quarter row has been read into registers, so the next quarter rows — of this
//     vector's other operand, or of the warp's next vector — are requested while the current row is
//     still being reduced, normed and written out.  Every input operand (residual row, branch row of
//     the residual-add form, MLP-output row of the Gemma-2 block tail) arrives this way; nothing that
//     comes from DRAM waits on a register load.  Table rows (L2-resident, <= 0.5 MB) are fetched into
//     registers one quarter row ahead of their use.
// With the row out of shared memory a warp needs 8-20 KB of it instead of 16-32 KB, so 12-16 warps fit
// per SM in every form (the staged kernel: 6 with two staged operands), each with a continuous stream
// of bulk copies in flight instead of one row per iteration plus exposed register loads.
//
// Covers 16-bit model dtypes with the table in the model dtype and H in {3584, 4096, 5120} (Gemma-2-9B,
// Llama-3.1-8B, Qwen2.5-32B: the BASELINE.json architectures) plus {1024, 2048} for tests; forced with
// IA_IMPL_STREAM.  Chosen automatically for launches of >= 3000 vectors where it measured faster than the
// staged kernel on B200 (25 600 vectors, cold L2, every vector steered; staged kernel in brackets):
//   residual-add form   H = 4096: 147 us (167)   H = 5120: 181 us (213)   H = 3584: 129 us (144)
//   Gemma-2 block tail  H = 3584: 127 us (158)
//   plain form          H = 5120: 139 us (150); at 4096 / 3584 the two are level (116 / 111, 101 / 99): staged kept
// i.e. 0.87-0.89 of the measured copy bandwidth on the bytes DRAM must move for the two-operand forms (0.94 at
// 262 144 vectors), where a torch copy of the same byte count reaches 0.96.

#include <atomic>

#include "ia_k2_common.cuh"

namespace ia {

constexpr int kFormPlain = 0, kFormAdd = 1, kFormTail = 2;

// sum of squares of a 16-bit group: x*x is exact in fp32 for bf16 / fp16 inputs, so the fused form
// rounds once, exactly where `a + (x*x)` with separate roundings does — same bits, one instruction.
template <typename T> __device__ __forceinline__ void sumsq4_fma(const Pack4<T> &x, float &a0, float &a1, float &a2, float &a3) {
    float f[4];
    Elt<T>::unpack(x, f);
    a0 = __fmaf_rn(f[0], f[0], a0);
    a1 = __fmaf_rn(f[1], f[1], a1);
    a2 = __fmaf_rn(f[2], f[2], a2);
    a3 = __fmaf_rn(f[3], f[3], a3);
}

// A row is cut into kChunks pieces ("quarter rows" in the comments; 4, 5 or 8 of them); GPC = 4-element groups per
// lane per piece (H = 128 * kChunks * GPC); NS = landing slots per warp.
template <typename T, int VARIANT, bool IEEE, int FORM, int kChunks, int GPC, int NS, int MAXT>
__global__ void __launch_bounds__(MAXT, 1) k2_stream_kernel(const K2Params p) {
    extern __shared__ __align__(128) unsigned char smem[];
    constexpr int GPL = kChunks * GPC;                                // groups per lane per row
    constexpr int R = FORM == kFormPlain ? kChunks : 2 * kChunks;     // quarter-row loads per vector
    // NS divides R (slot index and parity static per load), or NS = M * R (the ring holds M whole vectors)
    static_assert(R % NS == 0 || NS % R == 0, "ring size must divide, or be a multiple of, the loads per vector");
    constexpr int M = NS >= R ? NS / R : 1;  // vectors ahead that are requested in full
    constexpr unsigned kChunkBytes = GPC * 32u * (unsigned)sizeof(Pack4<T>);
    constexpr int kChunkElts = GPC * 128;
    constexpr int H = GPL * 128;
    constexpr unsigned kRowBytes = (unsigned)H * sizeof(T);
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarps = blockDim.x >> 5;
    // layout: [weight row][post weight row (tail)][mbarriers nwarps*NS, padded to 128][slots nwarps*NS]
    T *w_s = reinterpret_cast<T *>(smem);
    T *wpost_s = w_s + H;
    constexpr unsigned kBarsOff = (FORM == kFormTail ? 2u : 1u) * kRowBytes;
    unsigned long long *bars = reinterpret_cast<unsigned long long *>(smem + kBarsOff) + warp * NS;
    const unsigned slots_off = kBarsOff + ((unsigned)(nwarps * NS * 8 + 127) & ~127u);
    unsigned char *my_slots = smem + slots_off + (size_t)warp * NS * kChunkBytes;

    for (int g = threadIdx.x; g < H / 4; g += blockDim.x) {
        cp_async_group<sizeof(Pack4<T>)>(w_s + 4 * g, static_cast<const T *>(p.w) + 4 * g);
        if (FORM == kFormTail) cp_async_group<sizeof(Pack4<T>)>(wpost_s + 4 * g, static_cast<const T *>(p.w_post) + 4 * g);
    }
    if (lane == 0) {
#pragma unroll
        for (int s = 0; s < NS; ++s) mbar_init(&bars[s], 1);
        fence_proxy_async();
    }
    __syncwarp();

    T *hbase = static_cast<T *>(p.h);
    T *ybase = static_cast<T *>(p.y);
    const T *vtab = static_cast<const T *>(p.vtable);
    const T *obase = static_cast<const T *>(FORM == kFormTail ? p.mlp : p.add);  // the second operand
    const int stride = gridDim.x * nwarps;
    int vec = blockIdx.x * nwarps + warp;

    // load number j of a vector into ring slot `sl`: which operand, which piece
    auto issue = [&](int v, int j, int sl) {  // lane 0 only
        const T *src;
        int c;
        if (FORM == kFormPlain) { src = hbase; c = j; }
        else if (FORM == kFormAdd) { src = (j & 1) ? obase : hbase; c = j >> 1; }
        else { src = j < kChunks ? obase : hbase; c = j % kChunks; }
        void *bar = &bars[sl];
        mbar_expect_tx(bar, kChunkBytes);
        bulk_g2s(my_slots + (size_t)sl * kChunkBytes, src + (size_t)v * H + (size_t)c * kChunkElts, kChunkBytes, bar);
    };
    // slot of load j of the vector handled in iteration `it`
    auto slot_of = [&](unsigned it, int j) { return NS >= R ? (int)(it % (unsigned)M) * R + j : j % NS; };
    // load j of vector v (iteration it) has been read into registers by every lane: request the load that
    // takes over its slot — NS loads further down the sequence
    auto refill = [&](unsigned it, int v, int j) {
        if (lane == 0) {
            fence_proxy_async();
            if (NS >= R) {
                const long vn = (long)v + (long)M * stride;
                if (vn < p.nvec) issue((int)vn, j, slot_of(it, j));
            } else {
                const int jn = j + NS;
                if (jn < R) issue(v, jn, jn % NS);
                else if (v + stride < p.nvec) issue(v + stride, jn - R, (jn - R) % NS);
            }
        }
    };

    RowMeta cur{};
    if (vec < p.nvec) {
        cur = load_meta<T>(p, vec);
        if (lane == 0) {
            if (NS >= R) {
#pragma unroll
                for (int m = 0; m < M; ++m) {
                    const long v = (long)vec + (long)m * stride;
                    if (v < p.nvec) {
#pragma unroll
                        for (int j = 0; j < R; ++j) issue((int)v, j, m * R + j);
                    }
                }
            } else {
#pragma unroll
                for (int j = 0; j < NS; ++j) issue(vec, j, j);
            }
        }
    }
    Pack4<T> vv[GPC];  // table groups of the quarter row that is processed next
    auto load_table = [&](int vi, int c) {
        const T *vp = vtab + (size_t)vi * H + (size_t)c * kChunkElts + 4 * lane;
#pragma unroll
        for (int g = 0; g < GPC; ++g) vv[g] = ld4_ro<T>(vp + 128 * g);
    };
    if (cur.steer) load_table(cur.vi, 0);
    cp_async_wait_all();
    __syncthreads();  // weight rows staged

    for (unsigned it = 0; vec < p.nvec; ++it) {
        const int vnext = vec + stride;
        RowMeta nxt{};
        if (vnext < p.nvec) nxt = load_meta<T>(p, vnext);
        T *hp = hbase + (size_t)vec * H + 4 * lane;
        T *yp = ybase + (size_t)vec * H + 4 * lane;
        T *cp = cur.tap ? static_cast<T *>(p.cap_out) + (size_t)cur.row * H + 4 * lane : nullptr;
        const Injector<T, T> inj(cur.alpha);
        auto parity = [&](int j) { return NS >= R ? (it / (unsigned)M) & 1u : (it * (unsigned)(R / NS) + (unsigned)(j / NS)) & 1u; };
        auto slot = [&](int j) { return reinterpret_cast<const T *>(my_slots + (size_t)slot_of(it, j) * kChunkBytes) + 4 * lane; };

        Pack4<T> xr[GPL];
        float a0 = 0.f, a1 = 0.f, a2 = 0.f, a3 = 0.f;
        float rm = 0.f;
        if (FORM == kFormTail) {
            // Gemma-2 block tail, first pass: the MLP-output row into registers, its 1/rms
#pragma unroll
            for (int c = 0; c < kChunks; ++c) {
                mbar_wait(&bars[slot_of(it, c)], parity(c));
                const T *s = slot(c);
#pragma unroll
                for (int g = 0; g < GPC; ++g) {
                    xr[c * GPC + g] = ld4<T>(s + 128 * g);
                    sumsq4_fma<T>(xr[c * GPC + g], a0, a1, a2, a3);
                }
                __syncwarp();
                refill(it, vec, c);
            }
            rm = rinv<IEEE>(__fmul_rn(warp_total(a0, a1, a2, a3), p.inv_h), p.eps);
            a0 = a1 = a2 = a3 = 0.f;
        }
#pragma unroll
        for (int c = 0; c < kChunks; ++c) {
            const int jh = FORM == kFormAdd ? 2 * c : (FORM == kFormTail ? kChunks + c : c);
            const int ja = 2 * c + 1;  // kFormAdd only
            mbar_wait(&bars[slot_of(it, jh)], parity(jh));
            if (FORM == kFormAdd) mbar_wait(&bars[slot_of(it, ja)], parity(ja));
            const T *hs = slot(jh);
            const T *as = slot(ja);
#pragma unroll
            for (int g = 0; g < GPC; ++g) {
                const int k = c * GPC + g;
                Pack4<T> x = ld4<T>(hs + 128 * g);
                if (FORM == kFormAdd) add4<T>(x, ld4<T>(as + 128 * g));
                if (FORM == kFormTail) {
                    Pack4<T> n;
                    norm4<T, IA_NORM_GEMMA>(xr[k], rm, ld4<T>(wpost_s + 4 * lane + 128 * k), n);
                    add4<T>(x, n);
                }
                if (cur.steer) inj.apply(x, vv[g]);
                xr[k] = x;
                if (FORM != kFormPlain || cur.steer) st4<T>(hp + 128 * k, x);
                if (cp) st4<T>(cp + 128 * k, x);
                sumsq4_fma<T>(x, a0, a1, a2, a3);
                // keep the compiler from hoisting every shared-memory load of the piece to its top: four groups
                // in flight per lane are enough to cover the LDS latency, more only costs registers
                if ((g & 3) == 3) asm volatile("" ::: "memory");
            }
            // table groups for the next quarter row (this vector's, or the next vector's first)
            if (c + 1 < kChunks) {
                if (cur.steer) load_table(cur.vi, c + 1);
            } else if (nxt.steer) {
                load_table(nxt.vi, 0);
            }
            __syncwarp();
            refill(it, vec, jh);
            if (FORM == kFormAdd) refill(it, vec, ja);
        }
        const float r = rinv<IEEE>(__fmul_rn(warp_total(a0, a1, a2, a3), p.inv_h), p.eps);
#pragma unroll
        for (int k = 0; k < GPL; ++k) {
            Pack4<T> out;
            norm4<T, VARIANT>(xr[k], r, ld4<T>(w_s + 4 * lane + 128 * k), out);
            st4<T>(yp + 128 * k, out);
            if ((k & 3) == 3) asm volatile("" ::: "memory");  // as above: do not hoist all the weight loads
        }
        vec = vnext;
        cur = nxt;
    }
}

template <typename T, int VARIANT, bool IEEE, int FORM, int kChunks, int GPC, int NS, int MAXT> static int launch_stream(const K2Params &p, cudaStream_t s) {
    constexpr size_t kChunkBytes = (size_t)GPC * 32 * sizeof(Pack4<T>);
    constexpr size_t kRowBytes = (size_t)kChunks * kChunkBytes;  // = H * sizeof(T)
    constexpr size_t wrows = FORM == kFormTail ? 2 : 1;
    const size_t budget = 220 * 1024;
    int nwarps = (int)((budget - wrows * kRowBytes - 1024) / (NS * kChunkBytes));
    if (nwarps > MAXT / 32) nwarps = MAXT / 32;
    const int sms = sm_count();
    const int per_sm = (p.nvec + sms - 1) / sms;
    if (nwarps > per_sm) nwarps = per_sm < 1 ? 1 : per_sm;
    const size_t smem = wrows * kRowBytes + (((size_t)nwarps * NS * 8 + 127) & ~(size_t)127) + (size_t)nwarps * NS * kChunkBytes;
    auto kern = k2_stream_kernel<T, VARIANT, IEEE, FORM, kChunks, GPC, NS, MAXT>;
    static std::atomic<bool> configured[64];  // large dynamic shared memory: opted in once per instantiation and device
    int dev = 0;
    cudaGetDevice(&dev);
    if (dev < 0 || dev >= 64) dev = 0;
    if (!configured[dev].load(std::memory_order_acquire)) {
        cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
        if (e != cudaSuccess) return (int)e;
        configured[dev].store(true, std::memory_order_release);
    }
    long ctas = ((long)p.nvec + nwarps - 1) / nwarps;
    if (ctas > sms) ctas = sms;
    kern<<<(unsigned)ctas, 32 * nwarps, smem, s>>>(p);
    return (int)cudaGetLastError();
}

template <typename T, int FORM, int kChunks, int GPC, int NS, int MAXT> static int stream_flags(const K2Params &p, int variant, int flags, cudaStream_t s) {
    const bool ieee = (flags & IA_RSQRT_IEEE) != 0;
    if (variant == IA_NORM_GEMMA)
        return ieee ? launch_stream<T, IA_NORM_GEMMA, true, FORM, kChunks, GPC, NS, MAXT>(p, s)
                    : launch_stream<T, IA_NORM_GEMMA, false, FORM, kChunks, GPC, NS, MAXT>(p, s);
    if (FORM == kFormTail) return IA_ERR_DTYPE;
    return ieee ? launch_stream<T, IA_NORM_LLAMA, true, FORM, kChunks, GPC, NS, MAXT>(p, s)
                : launch_stream<T, IA_NORM_LLAMA, false, FORM, kChunks, GPC, NS, MAXT>(p, s);
}

template <typename T, int kChunks, int GPC, int MAXT> static int stream_form(const K2Params &p, int variant, int flags, cudaStream_t s) {
    // ring = the next vector's loads in full for the two-operand forms, the next two vectors' for the plain form:
    // 2 * kChunks slots, i.e. two rows' worth of bytes in flight per warp either way
    if (p.mlp) return stream_flags<T, kFormTail, kChunks, GPC, 2 * kChunks, MAXT>(p, variant, flags, s);
    if (p.add) return stream_flags<T, kFormAdd, kChunks, GPC, 2 * kChunks, MAXT>(p, variant, flags, s);
    return stream_flags<T, kFormPlain, kChunks, GPC, 2 * kChunks, MAXT>(p, variant, flags, s);
}

template <typename T> static int stream_h(const K2Params &p, int variant, int flags, cudaStream_t s) {
    switch (p.H) {
    case 1024: return stream_form<T, 4, 2, 512>(p, variant, flags, s);
    case 2048: return stream_form<T, 4, 4, 512>(p, variant, flags, s);
    // pieces per row x groups per lane per piece x threads per CTA, as measured on B200 (25 600 vectors, cold):
    // 1 KiB pieces at H = 4096 (2 KiB pieces: residual-add form 181 us against 147); H = 5120 keeps 80 registers of
    // row per lane, so 8 warps (no spills: 181 us against 234 with 10 warps and spills)
    case 3584: return stream_form<T, 4, 7, 384>(p, variant, flags, s);
    case 4096: return stream_form<T, 8, 4, 384>(p, variant, flags, s);
    case 5120: return stream_form<T, 5, 8, 256>(p, variant, flags, s);
    default: return IA_ERR_SHAPE;
    }
}

bool k2_stream_supported(const K2Params &p, int dtype, int vec_dtype) {
    if (vec_dtype != dtype || (dtype != IA_BF16 && dtype != IA_F16)) return false;
    return p.H == 1024 || p.H == 2048 || p.H == 3584 || p.H == 4096 || p.H == 5120;
}

int launch_k2_stream(const K2Params &p, int dtype, int vec_dtype, int variant, int flags, cudaStream_t s) {
    if (!k2_stream_supported(p, dtype, vec_dtype)) return vec_dtype != dtype ? IA_ERR_DTYPE : IA_ERR_SHAPE;
    return dtype == IA_BF16 ? stream_h<__nv_bfloat16>(p, variant, flags, s) : stream_h<__half>(p, variant, flags, s);
}

}  // namespace ia
