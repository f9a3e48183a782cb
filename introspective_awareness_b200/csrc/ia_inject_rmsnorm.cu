// K2: per-row concept injection fused into the next RMSNorm (sm_100a).
//
// One warp owns one (row, position) vector of the residual stream.  The vector is read once with
// coalesced vector loads, kept packed in registers across the reduction, and written back once
// (h', only if steered) plus once (y).  HBM traffic is the algorithmic minimum
// rows*T*H*e*(2 reads + 2 writes) + H*e; the concept table (<= 0.5 MB) and the norm weight live in
// L2 / shared memory.  No tensor-core work: < 2 flop/byte.
//
// Summation order of the variance (part of the numeric contract, see include/ia_b200.h and
// oracle/kernels_np.py:row_sum_squares): lane l owns the 4-element groups l, l+32, l+64, ...;
// one fp32 accumulator per element slot, groups added in ascending order; accumulators combined
// ((a0+a1)+a2)+a3; lanes combined by shuffle-down 16, 8, 4, 2, 1.  This is ATen's CUDA order for a
// contiguous fp32 [R>=16, H] mean(-1), which makes the kernel bit-identical to the eager chain
// hook-add -> .float() -> pow(2) -> mean -> +eps -> rsqrt -> mul -> .to(dtype) -> * weight.
// Every product and sum that torch rounds separately is issued as __fmul_rn/__fadd_rn so the
// compiler cannot contract them into an FMA.

#include "ia_k2_common.cuh"

namespace ia {

// Injects one table row into the register-resident vector and stores h'.  The h' stores may alias
// the table as far as the compiler knows, so it will not hoist table loads above them: VB loads
// are issued explicitly before the stores they precede, keeping VB of them in flight.
template <typename T, typename VT, int MAXJ>
__device__ __forceinline__ void inject_row_regs(Pack4<T> (&x)[MAXJ], const Injector<T, VT> &inj, const VT *vp, T *hp,
                                                int lane, int nj) {
    constexpr int VB = sizeof(VT) == 4 ? 4 : 8;  // 16 registers of table data in flight per lane
#pragma unroll
    for (int j0 = 0; j0 < MAXJ; j0 += VB) {
        Pack4<VT> vv[VB];
#pragma unroll
        for (int jj = 0; jj < VB; ++jj)
            if (j0 + jj < MAXJ && j0 + jj < nj) vv[jj] = ld4_ro<VT>(vp + 4 * (lane + 32 * (j0 + jj)));
#pragma unroll
        for (int jj = 0; jj < VB; ++jj) {
            if (j0 + jj < MAXJ && j0 + jj < nj) {
                inj.apply(x[j0 + jj], vv[jj]);
                st4<T>(hp + 4 * (lane + 32 * (j0 + jj)), x[j0 + jj]);
            }
        }
    }
}

// Register-resident implementation: MAXJ = groups of 4 elements per lane; requires H % 128 == 0
// (every lane owns the same number of groups, so all predicates are warp-uniform) and
// H <= 128 * MAXJ.  Grid-stride over row vectors, one warp per vector.
template <typename T, typename VT, int MAXJ, int VARIANT, bool IEEE, bool TAIL, int MINB>
__global__ void __launch_bounds__(128, MINB) k2_regs_kernel(const K2Params p) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    T *w_s = reinterpret_cast<T *>(smem_raw);
    T *wpost_s = w_s + p.H;  // TAIL only
    const int H = p.H;
    const int G = H >> 2;
    const int nj = G >> 5;  // groups per lane (uniform)
    for (int g = threadIdx.x; g < G; g += blockDim.x) {
        cp_async_group<sizeof(Pack4<T>)>(w_s + 4 * g, static_cast<const T *>(p.w) + 4 * g);
        if (TAIL) cp_async_group<sizeof(Pack4<T>)>(wpost_s + 4 * g, static_cast<const T *>(p.w_post) + 4 * g);
    }

    const int lane = threadIdx.x & 31;
    const int warps_per_cta = blockDim.x >> 5;
    const int stride = gridDim.x * warps_per_cta;
    T *hbase = static_cast<T *>(p.h);
    T *ybase = static_cast<T *>(p.y);
    const T *src = TAIL ? static_cast<const T *>(p.mlp) : hbase;

    int vec = blockIdx.x * warps_per_cta + (threadIdx.x >> 5);
    Pack4<T> x[MAXJ];
    // The first row's loads are in flight while the weights land in shared memory.
    if (vec < p.nvec) {
#pragma unroll
        for (int j = 0; j < MAXJ; ++j)
            if (j < nj) x[j] = ld4<T>(src + (size_t)vec * H + 4 * (lane + 32 * j));
    }
    cp_async_wait_all();
    __syncthreads();

    while (vec < p.nvec) {
        const int row = vec / p.T;
        const int pos = vec - row * p.T;
        // three independent metadata loads (no dependent chain)
        const int vi = p.row_vec ? __ldg(p.row_vec + row) : -1;
        const int t0 = (p.row_vec && !p.steer_all) ? __ldg(p.row_t0 + row) : 0;
        const float alpha = p.row_vec ? __ldg(p.row_alpha + row) : 0.0f;
        const bool steer = vi >= 0 && (p.steer_all || (pos + p.pos0 >= t0));
        T *hp = hbase + (size_t)vec * H;

        if (TAIL) {
            // Gemma-2 block tail: h = rd(residual + gemma_norm(mlp_out; w_post)); x holds mlp_out.
            float a0 = 0.f, a1 = 0.f, a2 = 0.f, a3 = 0.f;
#pragma unroll
            for (int j = 0; j < MAXJ; ++j)
                if (j < nj) sumsq4<T>(x[j], a0, a1, a2, a3);
            const float rm = rinv<IEEE>(__fmul_rn(warp_total(a0, a1, a2, a3), p.inv_h), p.eps);
#pragma unroll
            for (int j = 0; j < MAXJ; ++j) {
                if (j < nj) {
                    const int g = lane + 32 * j;
                    Pack4<T> n;
                    norm4<T, IA_NORM_GEMMA>(x[j], rm, ld4<T>(wpost_s + 4 * g), n);
                    float f[4], rf[4];
                    Elt<T>::unpack(n, f);
                    Elt<T>::unpack(ld4<T>(hp + 4 * g), rf);
#pragma unroll
                    for (int i = 0; i < 4; ++i) f[i] = __fadd_rn(rf[i], f[i]);
                    Elt<T>::rd4(f, x[j]);
                }
            }
        }

        if (steer) {
            const VT *vp = static_cast<const VT *>(p.vtable) + (size_t)vi * H;
            const Injector<T, VT> inj(alpha);
            inject_row_regs<T, VT, MAXJ>(x, inj, vp, hp, lane, nj);
        } else if (TAIL) {
#pragma unroll
            for (int j = 0; j < MAXJ; ++j)
                if (j < nj) st4<T>(hp + 4 * (lane + 32 * j), x[j]);
        }

        if (tap_here(p, row, pos)) {  // extraction capture: the updated row, compactly
            T *cp = static_cast<T *>(p.cap_out) + (size_t)row * H;
#pragma unroll
            for (int j = 0; j < MAXJ; ++j)
                if (j < nj) st4<T>(cp + 4 * (lane + 32 * j), x[j]);
        }

        float a0 = 0.f, a1 = 0.f, a2 = 0.f, a3 = 0.f;
#pragma unroll
        for (int j = 0; j < MAXJ; ++j)
            if (j < nj) sumsq4<T>(x[j], a0, a1, a2, a3);
        const float r = rinv<IEEE>(__fmul_rn(warp_total(a0, a1, a2, a3), p.inv_h), p.eps);

        T *yp = ybase + (size_t)vec * H;
#pragma unroll
        for (int j = 0; j < MAXJ; ++j) {
            if (j < nj) {
                const int g = lane + 32 * j;
                Pack4<T> out;
                norm4<T, VARIANT>(x[j], r, ld4<T>(w_s + 4 * g), out);
                st4<T>(yp + 4 * g, out);
            }
        }

        vec += stride;
        if (vec < p.nvec) {
#pragma unroll
            for (int j = 0; j < MAXJ; ++j)
                if (j < nj) x[j] = ld4<T>(src + (size_t)vec * H + 4 * (lane + 32 * j));
        }
    }
}

// Any H % 4 == 0: two passes over the row; the second pass re-reads h' (L1/L2 hit).
template <typename T, typename VT, int VARIANT, bool IEEE>
__global__ void __launch_bounds__(128) k2_twopass_kernel(const K2Params p) {
    const int lane = threadIdx.x & 31;
    const int warps_per_cta = blockDim.x >> 5;
    const int H = p.H, G = H >> 2;
    T *hbase = static_cast<T *>(p.h);
    T *ybase = static_cast<T *>(p.y);
    const T *w = static_cast<const T *>(p.w);
    for (int vec = blockIdx.x * warps_per_cta + (threadIdx.x >> 5); vec < p.nvec; vec += gridDim.x * warps_per_cta) {
        const int row = vec / p.T;
        const int pos = vec - row * p.T;
        const int vi = p.row_vec ? __ldg(p.row_vec + row) : -1;
        const int t0 = (p.row_vec && !p.steer_all) ? __ldg(p.row_t0 + row) : 0;
        const float alpha = p.row_vec ? __ldg(p.row_alpha + row) : 0.0f;
        const bool steer = vi >= 0 && (p.steer_all || (pos + p.pos0 >= t0));
        T *hp = hbase + (size_t)vec * H;
        const VT *vp = static_cast<const VT *>(p.vtable) + (size_t)(steer ? vi : 0) * H;
        const Injector<T, VT> inj(alpha);
        float a0 = 0.f, a1 = 0.f, a2 = 0.f, a3 = 0.f;
        const T *ap = p.add ? static_cast<const T *>(p.add) + (size_t)vec * H : nullptr;
        T *cp = tap_here(p, row, pos) ? static_cast<T *>(p.cap_out) + (size_t)row * H : nullptr;
        for (int g = lane; g < G; g += 32) {
            Pack4<T> x = ld4<T>(hp + 4 * g);
            if (ap) add4<T>(x, ld4_ro<T>(ap + 4 * g));
            if (steer) inj.apply(x, ld4_ro<VT>(vp + 4 * g));
            if (steer || ap) st4<T>(hp + 4 * g, x);
            if (cp) st4<T>(cp + 4 * g, x);
            sumsq4<T>(x, a0, a1, a2, a3);
        }
        const float r = rinv<IEEE>(__fmul_rn(warp_total(a0, a1, a2, a3), p.inv_h), p.eps);
        T *yp = ybase + (size_t)vec * H;
        for (int g = lane; g < G; g += 32) {
            Pack4<T> out;
            norm4<T, VARIANT>(ld4<T>(hp + 4 * g) /* own earlier store, or untouched input */, r, ld4_ro<T>(w + 4 * g), out);
            st4<T>(yp + 4 * g, out);
        }
    }
}

constexpr int kStreamMinVectors = 3000;  // measured level or ahead from 3000 vectors on; shorter launches are latency-bound: staged kernel

int sm_count() {
    static int cached[64] = {0};
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 148;
    if (cached[dev] == 0) {
        int n = 0;
        if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) n = 148;
        cached[dev] = n;
    }
    return cached[dev];
}

template <typename T, typename VT, int MAXJ, int VARIANT, bool IEEE, bool TAIL, int MINB>
static int launch_regs(const K2Params &p, cudaStream_t s) {
    auto kern = k2_regs_kernel<T, VT, MAXJ, VARIANT, IEEE, TAIL, MINB>;
    const size_t smem = (size_t)p.H * sizeof(T) * (TAIL ? 2 : 1);
    if (smem > 48 * 1024) {
        cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) return (int)e;
    }
    // small problems: fewer warps per CTA so that the rows spread over more SMs
    const int sms = sm_count();
    int warps = 4;
    if ((long)p.nvec < 4L * sms) warps = p.nvec >= 2L * sms ? 2 : 1;
    long ctas = ((long)p.nvec + warps - 1) / warps;
    const long cap = (long)sms * MINB;
    if (ctas > cap) ctas = cap;
    kern<<<(unsigned)ctas, 32 * warps, smem, s>>>(p);
    return (int)cudaGetLastError();
}

template <typename T, typename VT, int VARIANT, bool IEEE, bool TAIL>
static int dispatch_h(const K2Params &p, cudaStream_t s) {
    const int per_lane = ((p.H >> 2) + 31) / 32;
    constexpr bool wide = sizeof(T) == 4;
    const bool even = (p.H & 127) == 0;  // the register kernel wants the same group count on every lane
    if (p.add) {
        // residual-add form: bulk-copy kernel or this two-pass kernel only
        if (TAIL) return IA_ERR_SHAPE;
    } else if (!even) {
        if (TAIL) return IA_ERR_SHAPE;
    } else if (per_lane <= 8) return launch_regs<T, VT, 8, VARIANT, IEEE, TAIL, wide ? 4 : 8>(p, s);
    else if (per_lane <= 16) return launch_regs<T, VT, 16, VARIANT, IEEE, TAIL, wide ? 3 : 6>(p, s);
    if constexpr (!wide) if (even && !p.add) {
        // one CTA fewer per SM when extra registers are live (fp32 table groups, Gemma-2 tail)
        constexpr int D = (TAIL || sizeof(VT) == 4) ? 1 : 0;
        if (per_lane <= 28) return launch_regs<T, VT, 28, VARIANT, IEEE, TAIL, 5 - D>(p, s);
        if (per_lane <= 32) return launch_regs<T, VT, 32, VARIANT, IEEE, TAIL, 5 - D>(p, s);
        if (per_lane <= 40) return launch_regs<T, VT, 40, VARIANT, IEEE, TAIL, 4 - D>(p, s);
        if (per_lane <= 64) return launch_regs<T, VT, 64, VARIANT, IEEE, TAIL, 3 - D>(p, s);
    }
    if (TAIL) return IA_ERR_TOO_LARGE;
    long ctas = ((long)p.nvec + 3) / 4;
    const long cap = (long)sm_count() * 8;
    if (ctas > cap) ctas = cap;
    k2_twopass_kernel<T, VT, VARIANT, IEEE><<<(unsigned)ctas, 128, 0, s>>>(p);
    return (int)cudaGetLastError();
}

template <typename T, typename VT, bool TAIL> static int dispatch_flags(const K2Params &p, int variant, int flags, cudaStream_t s) {
    const bool ieee = (flags & IA_RSQRT_IEEE) != 0;
    if (TAIL || variant == IA_NORM_GEMMA)
        return ieee ? dispatch_h<T, VT, IA_NORM_GEMMA, true, TAIL>(p, s) : dispatch_h<T, VT, IA_NORM_GEMMA, false, TAIL>(p, s);
    return ieee ? dispatch_h<T, VT, IA_NORM_LLAMA, true, false>(p, s) : dispatch_h<T, VT, IA_NORM_LLAMA, false, false>(p, s);
}

template <bool TAIL> static int dispatch_dtype(const K2Params &p, int dtype, int vec_dtype, int variant, int flags, cudaStream_t s) {
    if (vec_dtype != dtype && vec_dtype != IA_F32) return IA_ERR_DTYPE;
    switch (dtype) {
    case IA_BF16:
        return vec_dtype == IA_F32 ? dispatch_flags<__nv_bfloat16, float, TAIL>(p, variant, flags, s)
                                   : dispatch_flags<__nv_bfloat16, __nv_bfloat16, TAIL>(p, variant, flags, s);
    case IA_F16:
        if (vec_dtype != IA_F16 || TAIL) return IA_ERR_DTYPE;  // fp16 models: table in fp16, no Gemma tail
        return dispatch_flags<__half, __half, false>(p, variant, flags, s);
    case IA_F32:
        return dispatch_flags<float, float, TAIL>(p, variant, flags, s);
    default:
        return IA_ERR_DTYPE;
    }
}

static int check_common(const void *h, const void *y, const void *vtable, const int32_t *row_vec,
                        const float *row_alpha, const int32_t *row_t0, const void *w, int64_t rows,
                        int64_t T, int64_t H, int32_t flags) {
    if (!h || !y || !w) return IA_ERR_NULL;
    if (row_vec && (!vtable || !row_alpha)) return IA_ERR_NULL;
    if (row_vec && !(flags & IA_STEER_ALL) && !row_t0) return IA_ERR_NULL;
    if (rows < 0 || T < 1 || H < 4 || H > 65536 || (H & 3)) return IA_ERR_SHAPE;
    if (rows * T >= (int64_t)1 << 31) return IA_ERR_SHAPE;
    if (!aligned16(h) || !aligned16(y) || !aligned16(w) || (vtable && !aligned16(vtable))) return IA_ERR_ALIGN;
    return 0;
}

// One body for the three K2 entry points and their tap forms.  `other` is the branch output (residual-add
// form), the MLP output (Gemma-2 block tail, with w_post) or NULL (plain form).
enum K2Form { kPlain = 0, kAddForm = 1, kTailForm = 2 };

static int k2_entry(K2Form form, void *h, const void *other, void *y, const void *vtable, const int32_t *row_vec,
                    const float *row_alpha, const int32_t *row_t0, const void *w_post, const void *w, float eps,
                    int64_t rows, int64_t T, int64_t H, int32_t pos0, int32_t dtype, int32_t vec_dtype, int32_t variant,
                    int32_t flags, void *cap_out, const int32_t *cap_tok, int32_t cap_t, void *stream) {
    if (form != kPlain && !other) return IA_ERR_NULL;
    if (form == kTailForm && !w_post) return IA_ERR_NULL;
    int rc = check_common(h, y, vtable, row_vec, row_alpha, row_t0, w, rows, T, H, flags);
    if (rc) return rc;
    if ((other && !aligned16(other)) || (w_post && !aligned16(w_post)) || (cap_out && !aligned16(cap_out))) return IA_ERR_ALIGN;
    if (form == kTailForm) variant = IA_NORM_GEMMA;
    if (variant != IA_NORM_LLAMA && variant != IA_NORM_GEMMA) return IA_ERR_DTYPE;
    if (cap_out && !cap_tok) {
        if (cap_t < 0) cap_t += (int32_t)T;  // -1 = the last position, as the reference's token_idx=-1
        if (cap_t < 0 || cap_t >= T) return IA_ERR_SHAPE;
    }
    if (rows == 0) return 0;
    K2Params p{};
    p.h = h; p.y = y; p.vtable = vtable; p.row_vec = row_vec; p.row_alpha = row_alpha; p.row_t0 = row_t0;
    p.mlp = form == kTailForm ? other : nullptr;
    p.add = form == kAddForm ? other : nullptr;
    p.w = w; p.w_post = form == kTailForm ? w_post : nullptr;
    p.eps = eps; p.inv_h = 1.0f / (float)H;
    p.nvec = (int)(rows * T); p.T = (int)T; p.H = (int)H; p.pos0 = pos0; p.steer_all = (flags & IA_STEER_ALL) ? 1 : 0;
    p.cap_out = cap_out; p.cap_tok = cap_tok; p.cap_t = cap_t;
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    // Implementation choice (same bits every way), as measured on B200 at 25 600 vectors, cold, every vector
    // steered (tools/kbench.py --r2, profiles/r2_kbench_ladder.jsonl; table in ia_inject_rmsnorm_stream.cu):
    //   * long launches of the two-operand forms (residual add, Gemma-2 block tail), and of the plain form at
    //     H = 5120: the streaming kernel — row in registers, quarter-row landing slots, every DRAM operand by
    //     bulk copy (add form at H = 4096: 147 us against 167);
    //   * everything else with rows of a few KiB, and every short launch: the bulk-copy staged kernel (lowest
    //     latency; plain form 111 us at 25 600 x 4096, the streaming kernel 116);
    //   * short rows, fp32 tables, other H: the register / two-pass kernels.
    const bool stream_ok = k2_stream_supported(p, dtype, vec_dtype);
    if (flags & IA_IMPL_STREAM) {
        if (!stream_ok) return vec_dtype != dtype ? IA_ERR_DTYPE : IA_ERR_SHAPE;
        return launch_k2_stream(p, dtype, vec_dtype, variant, flags, s);
    }
    const int64_t row_bytes = H * (dtype == IA_F32 ? 4 : 2);
    // staged kernel: the weight row(s) and two stages per warp must fit in shared memory
    const int64_t stage_rows = form == kPlain ? 5 : (form == kAddForm ? 3 : 4);
    const bool tma_ok = vec_dtype == dtype && (H & 127) == 0 && row_bytes * stage_rows + 1024 <= 220 * 1024;
    if (flags & IA_IMPL_TMA) {
        if (!tma_ok) return vec_dtype != dtype ? IA_ERR_DTYPE : IA_ERR_SHAPE;
        return launch_k2_tma(p, dtype, vec_dtype, variant, flags, s);
    }
    if (!(flags & IA_IMPL_REGS)) {
        if (stream_ok && p.nvec >= kStreamMinVectors && (form != kPlain || H == 5120))
            return launch_k2_stream(p, dtype, vec_dtype, variant, flags, s);
        if (tma_ok && (form != kPlain || row_bytes >= 4096)) return launch_k2_tma(p, dtype, vec_dtype, variant, flags, s);
    }
    return form == kTailForm ? dispatch_dtype<true>(p, dtype, vec_dtype, variant, flags, s)
                             : dispatch_dtype<false>(p, dtype, vec_dtype, variant, flags, s);
}

}  // namespace ia

extern "C" int ia_inject_rmsnorm_tap(void *h, void *y, const void *vtable, const int32_t *row_vec, const float *row_alpha,
                                     const int32_t *row_t0, const void *w, float eps, int64_t rows, int64_t T, int64_t H,
                                     int32_t pos0, int32_t dtype, int32_t vec_dtype, int32_t variant, int32_t flags,
                                     void *cap_out, const int32_t *cap_tok, int32_t cap_t, void *stream) {
    return ia::k2_entry(ia::kPlain, h, nullptr, y, vtable, row_vec, row_alpha, row_t0, nullptr, w, eps, rows, T, H, pos0, dtype,
                        vec_dtype, variant, flags, cap_out, cap_tok, cap_t, stream);
}

extern "C" int ia_add_inject_rmsnorm_tap(void *h, const void *add, void *y, const void *vtable, const int32_t *row_vec,
                                         const float *row_alpha, const int32_t *row_t0, const void *w, float eps,
                                         int64_t rows, int64_t T, int64_t H, int32_t pos0, int32_t dtype, int32_t vec_dtype,
                                         int32_t variant, int32_t flags, void *cap_out, const int32_t *cap_tok,
                                         int32_t cap_t, void *stream) {
    return ia::k2_entry(ia::kAddForm, h, add, y, vtable, row_vec, row_alpha, row_t0, nullptr, w, eps, rows, T, H, pos0, dtype,
                        vec_dtype, variant, flags, cap_out, cap_tok, cap_t, stream);
}

extern "C" int ia_gemma2_tail_inject_rmsnorm_tap(const void *mlp_out, void *residual, void *y, const void *vtable,
                                                 const int32_t *row_vec, const float *row_alpha, const int32_t *row_t0,
                                                 const void *w_post, const void *w_next, float eps, int64_t rows, int64_t T,
                                                 int64_t H, int32_t pos0, int32_t dtype, int32_t vec_dtype, int32_t flags,
                                                 void *cap_out, const int32_t *cap_tok, int32_t cap_t, void *stream) {
    return ia::k2_entry(ia::kTailForm, residual, mlp_out, y, vtable, row_vec, row_alpha, row_t0, w_post, w_next, eps, rows, T, H,
                        pos0, dtype, vec_dtype, IA_NORM_GEMMA, flags, cap_out, cap_tok, cap_t, stream);
}

extern "C" int ia_inject_rmsnorm(void *h, void *y, const void *vtable, const int32_t *row_vec,
                                 const float *row_alpha, const int32_t *row_t0, const void *w, float eps,
                                 int64_t rows, int64_t T, int64_t H, int32_t pos0, int32_t dtype,
                                 int32_t vec_dtype, int32_t variant, int32_t flags, void *stream) {
    return ia_inject_rmsnorm_tap(h, y, vtable, row_vec, row_alpha, row_t0, w, eps, rows, T, H, pos0, dtype, vec_dtype, variant,
                                 flags, nullptr, nullptr, 0, stream);
}

extern "C" int ia_gemma2_tail_inject_rmsnorm(const void *mlp_out, void *residual, void *y, const void *vtable,
                                             const int32_t *row_vec, const float *row_alpha,
                                             const int32_t *row_t0, const void *w_post, const void *w_next,
                                             float eps, int64_t rows, int64_t T, int64_t H, int32_t pos0,
                                             int32_t dtype, int32_t vec_dtype, int32_t flags, void *stream) {
    return ia_gemma2_tail_inject_rmsnorm_tap(mlp_out, residual, y, vtable, row_vec, row_alpha, row_t0, w_post, w_next, eps, rows,
                                             T, H, pos0, dtype, vec_dtype, flags, nullptr, nullptr, 0, stream);
}

extern "C" int ia_add_inject_rmsnorm(void *h, const void *add, void *y, const void *vtable, const int32_t *row_vec,
                                     const float *row_alpha, const int32_t *row_t0, const void *w, float eps,
                                     int64_t rows, int64_t T, int64_t H, int32_t pos0, int32_t dtype,
                                     int32_t vec_dtype, int32_t variant, int32_t flags, void *stream) {
    return ia_add_inject_rmsnorm_tap(h, add, y, vtable, row_vec, row_alpha, row_t0, w, eps, rows, T, H, pos0, dtype, vec_dtype,
                                     variant, flags, nullptr, nullptr, 0, stream);
}
