// K2, bulk-copy (TMA) staged implementation for sm_100a — selected with IA_IMPL_TMA.
//
// Same arithmetic, same summation order, same bits as the register kernel; what changes is how
// the bytes arrive.  Each warp owns a two-stage ring in shared memory.  One lane issues
// `cp.async.bulk` (the TMA engine; SASS UBLKCP) for the NEXT vector's residual row and, when that
// vector is steered, its concept-table row, completing on a per-stage mbarrier, while the whole
// warp works on the CURRENT vector out of shared memory.  DRAM and L2 latency are therefore off
// the critical path and no registers are tied up by loads in flight; the row lives in shared memory
// (h' is written back in place for the second pass), so the kernel is not templated on H.
//
// One persistent CTA per SM; warps per CTA = what two stages per warp fit in 220 KB.  A stage is the
// residual row plus, for short launches, the table row; long launches read the table row from L2
// with batched loads instead, which halves the stage and doubles the warps (chosen per launch).
// Outputs leave through ordinary 64-bit stores (fire-and-forget; they never stall the warp).

#include <atomic>

#include "ia_k2_common.cuh"

namespace ia {

// S = stages per warp: one being consumed, S-1 being filled (prefetch distance S-1 vectors).
// STAGE_V: the concept-table row is bulk-copied into the stage too (16 KB per stage at H = 4096);
// otherwise it is read from L2 with batched 64-bit loads and a stage holds only the residual row.
// ADD: a second [nvec, H] input (the branch output of the decoder layer) is added first,
// h = rd(h + add), and h' is then written for every vector.  kAddStaged: the branch row is
// bulk-copied into the stage next to the residual row (short launches: everything arrives behind one
// barrier).  kAddLoaded: it comes through registers with batched loads, the stage stays one row and
// twice as many warps fit (long launches).
// kTail (Gemma-2 block tail): the staged row is the MLP output; it is normed with w_post first, the residual
// row comes through registers like kAddLoaded, h = rd(residual + gemma_norm(mlp_out)), then as before.
constexpr int kAddNone = 0, kAddStaged = 1, kAddLoaded = 2, kTail = 3;

template <typename T, int VARIANT, bool IEEE, int S, bool STAGE_V, int ADD>
__global__ void __launch_bounds__(512, 1) k2_tma_kernel(const K2Params p) {
    extern __shared__ __align__(128) unsigned char smem[];
    const int H = p.H;
    const int G = H >> 2;
    const unsigned row_bytes = (unsigned)H * sizeof(T);
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarps = blockDim.x >> 5;
    // layout: [weight row][mbarriers: nwarps*S x 8 B, padded to 128][stages: nwarps x S x (h row + v row)]
    T *w_s = reinterpret_cast<T *>(smem);
    T *wpost_s = w_s + H;  // kTail only: second weight row
    const unsigned bars_off = ((ADD == kTail ? 2u : 1u) * row_bytes + 127u) & ~127u;
    unsigned long long *bars = reinterpret_cast<unsigned long long *>(smem + bars_off) + warp * S;
    const unsigned stage_off = bars_off + ((unsigned)(nwarps * S * 8 + 127) & ~127u);
    constexpr int kRowsPerStage = 1 + (ADD == kAddStaged ? 1 : 0) + (STAGE_V ? 1 : 0);
    unsigned char *my_stage = smem + stage_off + (size_t)warp * S * kRowsPerStage * row_bytes;

    for (int g = threadIdx.x; g < G; g += blockDim.x) {
        cp_async_group<sizeof(Pack4<T>)>(w_s + 4 * g, static_cast<const T *>(p.w) + 4 * g);
        if (ADD == kTail) cp_async_group<sizeof(Pack4<T>)>(wpost_s + 4 * g, static_cast<const T *>(p.w_post) + 4 * g);
    }
    if (lane == 0) {
        for (int s = 0; s < S; ++s) mbar_init(&bars[s], 1);
        fence_proxy_async();  // make the initialised barriers visible to the async (TMA) proxy
    }
    __syncwarp();

    T *hbase = static_cast<T *>(p.h);
    T *ybase = static_cast<T *>(p.y);
    const T *vtab = static_cast<const T *>(p.vtable);
    const T *abase = static_cast<const T *>(p.add);
    const int stride = gridDim.x * nwarps;
    int vec = blockIdx.x * nwarps + warp;

    auto issue = [&](int stage, int v, const RowMeta &m) {
        // one lane: arm the barrier with the byte count, then start the bulk copies
        if (lane == 0) {
            unsigned char *dst = my_stage + (size_t)stage * kRowsPerStage * row_bytes;
            fence_proxy_async();  // earlier generic-proxy accesses to this stage are done (warp synced)
            const bool with_v = STAGE_V && m.steer;
            mbar_expect_tx(&bars[stage], (1u + (ADD == kAddStaged ? 1u : 0u) + (with_v ? 1u : 0u)) * row_bytes);
            bulk_g2s(dst, (ADD == kTail ? static_cast<const T *>(p.mlp) : hbase) + (size_t)v * H, row_bytes, &bars[stage]);
            if (ADD == kAddStaged) bulk_g2s(dst + row_bytes, abase + (size_t)v * H, row_bytes, &bars[stage]);
            // (an L2 bulk prefetch of the register-loaded branch row here was measured slower: 177 vs 168 us at 25 600 x 4096)
            if (with_v) bulk_g2s(dst + (ADD == kAddStaged ? 2 : 1) * row_bytes, vtab + (size_t)m.vi * H, row_bytes, &bars[stage]);
        }
    };

    // prologue: the first S-1 vectors of this warp are requested before anything else happens
    RowMeta meta[S];  // meta[k] belongs to the vector that sits k iterations ahead (ring, rotated below)
#pragma unroll
    for (int k = 0; k < S - 1; ++k) {
        const int v = vec + k * stride;
        meta[k] = RowMeta{};
        if (v < p.nvec) {
            meta[k] = load_meta<T>(p, v);
            issue(k, v, meta[k]);
        }
    }
    cp_async_wait_all();
    __syncthreads();  // weight row staged

    for (int it = 0; vec < p.nvec; ++it) {
        const int ahead = vec + (S - 1) * stride;
        meta[S - 1] = RowMeta{};
        if (ahead < p.nvec) {
            meta[S - 1] = load_meta<T>(p, ahead);
            issue((it + S - 1) % S, ahead, meta[S - 1]);
        }
        const RowMeta cur = meta[0];
        const int stage = it % S;
        // Operands that are not staged — the table row (L2) in the long-launch layouts, the branch row
        // (kAddLoaded) or the residual row (kTail) from HBM — come through registers in batches of 8 groups per
        // lane, software-pipelined: batch 0 is requested here, BEFORE the wait on the staged row (nothing in it
        // depends on that row), and batch b+1 while batch b is being worked on (two register buffers).
        const bool via_regs = !(ADD != kTail && STAGE_V && cur.steer) && (cur.steer || ADD != kAddNone);
        const T *vp = vtab + (size_t)(cur.steer ? cur.vi : 0) * H;
        const T *ap = (ADD == kTail ? hbase : abase) + (size_t)vec * H;
        Pack4<T> vvA[8], aaA[8], vvB[8], aaB[8];
        auto load_batch = [&](int g0, Pack4<T>(&vv)[8], Pack4<T>(&aa)[8]) {
            if (ADD == kAddLoaded) {
#pragma unroll
                for (int u = 0; u < 8; ++u)
                    if (g0 + 32 * u < G) aa[u] = ld4_ro<T>(ap + 4 * (g0 + 32 * u));
            }
            if (ADD == kTail) {  // the residual row is rewritten below: ordinary loads, not the read-only path
#pragma unroll
                for (int u = 0; u < 8; ++u)
                    if (g0 + 32 * u < G) aa[u] = ld4<T>(ap + 4 * (g0 + 32 * u));
            }
            if (cur.steer) {
#pragma unroll
                for (int u = 0; u < 8; ++u)
                    if (g0 + 32 * u < G) vv[u] = ld4_ro<T>(vp + 4 * (g0 + 32 * u));
            }
        };
        if (via_regs) load_batch(lane, vvA, aaA);
        mbar_wait(&bars[stage], (unsigned)(it / S) & 1u);
        T *h_s = reinterpret_cast<T *>(my_stage + (size_t)stage * kRowsPerStage * row_bytes);
        const T *a_s = reinterpret_cast<const T *>(reinterpret_cast<unsigned char *>(h_s) + row_bytes);  // kAddStaged only
        const T *v_s = reinterpret_cast<const T *>(reinterpret_cast<unsigned char *>(h_s) + (ADD == kAddStaged ? 2 : 1) * row_bytes);
        T *hp = hbase + (size_t)vec * H;
        T *yp = ybase + (size_t)vec * H;
        const Injector<T, T> inj(cur.alpha);

        float a0 = 0.f, a1 = 0.f, a2 = 0.f, a3 = 0.f;
        float rm = 0.f;  // kTail: 1/rms of the MLP output row
        if (ADD == kTail) {
#pragma unroll 4
            for (int g = lane; g < G; g += 32) sumsq4<T>(ld4<T>(h_s + 4 * g), a0, a1, a2, a3);
            rm = rinv<IEEE>(__fmul_rn(warp_total(a0, a1, a2, a3), p.inv_h), p.eps);
            a0 = a1 = a2 = a3 = 0.f;
        }
        if (ADD != kTail && STAGE_V && cur.steer) {
#pragma unroll 4
            for (int g = lane; g < G; g += 32) {
                Pack4<T> x = ld4<T>(h_s + 4 * g);
                if (ADD == kAddStaged) add4<T>(x, ld4<T>(a_s + 4 * g));
                if (ADD == kAddLoaded) add4<T>(x, ld4_ro<T>(abase + (size_t)vec * H + 4 * g));
                inj.apply(x, ld4<T>(v_s + 4 * g));
                st4<T>(hp + 4 * g, x);   // h' to HBM
                st4<T>(h_s + 4 * g, x);  // and kept for the second pass
                sumsq4<T>(x, a0, a1, a2, a3);
            }
        } else if (via_regs) {
            auto work_batch = [&](int g0, const Pack4<T>(&vv)[8], const Pack4<T>(&aa)[8]) {
#pragma unroll
                for (int u = 0; u < 8; ++u) {
                    const int g = g0 + 32 * u;
                    if (g < G) {
                        Pack4<T> x = ld4<T>(h_s + 4 * g);
                        if (ADD == kAddStaged) add4<T>(x, ld4<T>(a_s + 4 * g));
                        if (ADD == kAddLoaded) add4<T>(x, aa[u]);
                        if (ADD == kTail) {
                            Pack4<T> n;
                            norm4<T, IA_NORM_GEMMA>(x, rm, ld4<T>(wpost_s + 4 * g), n);
                            x = aa[u];
                            add4<T>(x, n);
                        }
                        if (cur.steer) inj.apply(x, vv[u]);
                        st4<T>(hp + 4 * g, x);
                        st4<T>(h_s + 4 * g, x);
                        sumsq4<T>(x, a0, a1, a2, a3);
                    }
                }
            };
            constexpr int kBatch = 32 * 8;
            if (sizeof(T) == 2) {
                for (int g0 = lane; g0 < G; g0 += 2 * kBatch) {
                    if (g0 + kBatch < G) load_batch(g0 + kBatch, vvB, aaB);
                    work_batch(g0, vvA, aaA);
                    if (g0 + kBatch < G) {
                        if (g0 + 2 * kBatch < G) load_batch(g0 + 2 * kBatch, vvA, aaA);
                        work_batch(g0 + kBatch, vvB, aaB);
                    }
                }
            } else {  // 4-byte elements: a second register buffer would spill; one batch at a time
                for (int g0 = lane; g0 < G; g0 += kBatch) {
                    if (g0 != lane) load_batch(g0, vvA, aaA);
                    work_batch(g0, vvA, aaA);
                }
            }
        } else {
#pragma unroll 4
            for (int g = lane; g < G; g += 32) sumsq4<T>(ld4<T>(h_s + 4 * g), a0, a1, a2, a3);
        }
        const float r = rinv<IEEE>(__fmul_rn(warp_total(a0, a1, a2, a3), p.inv_h), p.eps);
        if (cur.tap) {  // extraction capture: the updated row (it sits in the stage), compactly
            T *cp = static_cast<T *>(p.cap_out) + (size_t)cur.row * H;
#pragma unroll 4
            for (int g = lane; g < G; g += 32) st4<T>(cp + 4 * g, ld4<T>(h_s + 4 * g));
        }
#pragma unroll 4
        for (int g = lane; g < G; g += 32) {
            Pack4<T> out;
            norm4<T, VARIANT>(ld4<T>(h_s + 4 * g), r, ld4<T>(w_s + 4 * g), out);
            st4<T>(yp + 4 * g, out);
        }
        fence_proxy_async();  // this lane's shared-memory accesses are ordered before later bulk copies
        __syncwarp();         // every lane is done with this stage before it is refilled S-1 iterations on
        vec += stride;
#pragma unroll
        for (int k = 0; k + 1 < S; ++k) meta[k] = meta[k + 1];
    }
}

template <typename T, int VARIANT, bool IEEE, int S, bool STAGE_V, int ADD> static int launch_tma(const K2Params &p, cudaStream_t s) {
    const size_t row_bytes = (size_t)p.H * sizeof(T);
    const size_t budget = 220 * 1024;
    constexpr int kRowsPerStage = 1 + (ADD == kAddStaged ? 1 : 0) + (STAGE_V ? 1 : 0);
    const size_t wrows = ADD == kTail ? 2 : 1;
    if (wrows * row_bytes + 1024 + kRowsPerStage * S * row_bytes > budget) return IA_ERR_TOO_LARGE;
    int nwarps = (int)((budget - wrows * row_bytes - 1024) / (kRowsPerStage * S * row_bytes));
    if (nwarps > 16) nwarps = 16;
    const int sms = sm_count();
    // small problems: do not start more warps than there are vectors per SM
    const int per_sm = (p.nvec + sms - 1) / sms;
    if (nwarps > per_sm) nwarps = per_sm < 1 ? 1 : per_sm;
    const size_t bars_off = (wrows * row_bytes + 127) & ~(size_t)127;
    const size_t smem = bars_off + (((size_t)nwarps * S * 8 + 127) & ~(size_t)127) + (size_t)nwarps * kRowsPerStage * S * row_bytes;
    auto kern = k2_tma_kernel<T, VARIANT, IEEE, S, STAGE_V, ADD>;
    // opt in to large dynamic shared memory once per instantiation and device (first call = a warm-up
    // call, i.e. not under stream capture)
    static std::atomic<bool> configured[64];  // any host thread may be first on a device
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

template <typename T, int S, bool STAGE_V, int ADD> static int tma_flags_s(const K2Params &p, int variant, int flags, cudaStream_t s) {
    const bool ieee = (flags & IA_RSQRT_IEEE) != 0;
    if (variant == IA_NORM_GEMMA)
        return ieee ? launch_tma<T, IA_NORM_GEMMA, true, S, STAGE_V, ADD>(p, s)
                    : launch_tma<T, IA_NORM_GEMMA, false, S, STAGE_V, ADD>(p, s);
    return ieee ? launch_tma<T, IA_NORM_LLAMA, true, S, STAGE_V, ADD>(p, s) : launch_tma<T, IA_NORM_LLAMA, false, S, STAGE_V, ADD>(p, s);
}

template <typename T> static int tma_flags(const K2Params &p, int variant, int flags, cudaStream_t s) {
    // Long launches: table rows come from L2 and a stage holds only the residual row, which doubles the
    // warps per SM (12 at H = 4096) — measured 1.29 vs 1.18 of the copy-bandwidth peak at 262 144 vectors.
    // Short launches are latency-bound: there the table row rides in the stage (10.4 vs 12.9 us at 200).
    if (p.mlp) {  // Gemma-2 block tail
        if (variant != IA_NORM_GEMMA) return IA_ERR_DTYPE;
        return tma_flags_s<T, 2, false, kTail>(p, variant, flags, s);
    }
    if (p.add) {
        // long launches: one-row stages (12 warps at H = 4096), branch and table rows through registers;
        // short launches: all three rows ride in the stage behind one barrier
        const size_t rb = (size_t)p.H * sizeof(T);
        const bool three_fit = 7 * rb + 1024 <= 220 * 1024, two_fit = 5 * rb + 1024 <= 220 * 1024;
        if ((flags & IA_IMPL_ADD_STAGED) && two_fit) return tma_flags_s<T, 2, false, kAddStaged>(p, variant, flags, s);
        if (p.nvec >= 8192 || !three_fit) return tma_flags_s<T, 2, false, kAddLoaded>(p, variant, flags, s);
        return tma_flags_s<T, 2, true, kAddStaged>(p, variant, flags, s);
    }
    if (p.nvec >= 8192) return tma_flags_s<T, 2, false, kAddNone>(p, variant, flags, s);
    return tma_flags_s<T, 2, true, kAddNone>(p, variant, flags, s);
}

int launch_k2_tma(const K2Params &p, int dtype, int vec_dtype, int variant, int flags, cudaStream_t s) {
    if (vec_dtype != dtype) return IA_ERR_DTYPE;         // table rows are bulk-copied as model-dtype rows
    if ((p.H & 127) != 0) return IA_ERR_SHAPE;           // every lane owns the same number of groups
    if (((size_t)p.H * (dtype == IA_F32 ? 4 : 2)) % 16 != 0) return IA_ERR_SHAPE;  // bulk copies move multiples of 16 bytes
    switch (dtype) {
    case IA_BF16: return tma_flags<__nv_bfloat16>(p, variant, flags, s);
    case IA_F16: return tma_flags<__half>(p, variant, flags, s);
    case IA_F32: return tma_flags<float>(p, variant, flags, s);
    default: return IA_ERR_DTYPE;
    }
}

}  // namespace ia
