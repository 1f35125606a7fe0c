// Fused group asymmetric integer fake-quant forward and STE backward (sm_100a).
//
// One pass over HBM: every lane moves 32 B with LDG.E.256 / STG.E.256, a quantisation group of G elements is
// held by G*sizeof(T)/32 consecutive lanes, min/max are reduced with xor-shuffles inside that lane subset, the
// per-group scale / zero are derived op-by-op in T (see bdq_common.cuh:int_group_params) and every element is
// divided, rounded, clamped and dequantised in registers before a single 256-bit store.
//
// Reference semantics: quantization/quantizer.py:25-69 (pseudo_quantize_tensor), :119-182 (Ste*AsymQuantizer),
// backward per SURVEY.md App. B.
#include "bdq_common.cuh"
#include "fq_packed.cuh"
#include "fq_int_elem.cuh"

namespace bdq {

// threshold-formulated kernels for 16-bit dtypes, bits 2/3 (fakequant_thr.cu); return false if not applicable
bool launch_int_thr_fwd(const void* w, void* dq, const void* cmin, const void* cmax, int64_t n, int G, int bits, int dtype,
                        int round_mode, int sem, cudaStream_t st);
bool launch_int_thr_bwd(const void* w, const void* gy, void* gw, int64_t n, int G, int bits, int dtype, int sem,
                        cudaStream_t st, int n_extra, const void* const* extra);

constexpr int kThreads = 256;
constexpr int kUnroll = 2;  // 32-byte vectors in flight per lane

// Process one 32-byte vector of a 16-bit dtype with packed math (see fq_packed.cuh for the exactness argument).
// raw: input values (already clamped if clipping); returns packed dequantised values; cz[] (optional) = c - z.
template <int DT, int RM, bool WANT_CZ>
__device__ __forceinline__ void int_vec16_fast(const U8& raw, const IntQ& p, float qmax, U8& out, float* cz) {
    using P = P2<DT>;
    const typename P::h2 lo2 = P::dup(-p.z), hi2 = P::dup(qmax - p.z), s2 = P::dup(p.s);
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        const typename P::h2 x2 = P::from_bits(raw.r[i]);
        float q0a, q0b;
        const float ua = div_fast(P::lo(x2), p.s, p.y, q0a);
        const float ub = div_fast(P::hi(x2), p.s, p.y, q0b);
        const typename P::h2 u2 = P::pack(ua, ub);                      // rT(x / s)
        const float ra = round_T<DT, RM>(P::lo(u2)), rb = round_T<DT, RM>(P::hi(u2));
        const typename P::h2 cz2 = __hmin2(__hmax2(P::pack(ra, rb), lo2), hi2);  // clamp(r + z, 0, qmax) - z
        out.r[i] = P::bits(__hmul2(cz2, s2));                           // rT((c - z) * s)
        if constexpr (WANT_CZ) { cz[2 * i] = P::lo(cz2); cz[2 * i + 1] = P::hi(cz2); }
    }
}

template <int DT, int G, int RM, bool FAST, bool AUX, bool CLIP>
__global__ void __launch_bounds__(kThreads) int_fq_fwd_kernel(
    const uint8_t* __restrict__ w, uint8_t* __restrict__ dq, void* __restrict__ scales, void* __restrict__ zeros,
    uint8_t* __restrict__ codes, const void* __restrict__ cmin, const void* __restrict__ cmax, int64_t nvec,
    float qmax, int sem) {
    constexpr int VEC = VecN<DT>::v;
    constexpr int LPG = G / VEC;
    static_assert(LPG >= 1 && LPG <= 32, "group must span 1..32 lanes");
    const int lane = threadIdx.x & 31;
    const int64_t warp = (int64_t)blockIdx.x * (kThreads / 32) + (threadIdx.x >> 5);
    const int64_t nwarps = (int64_t)gridDim.x * (kThreads / 32);
    const int64_t nchunks = (nvec + 31) / 32;

    for (int64_t c0 = warp * kUnroll; c0 < nchunks; c0 += nwarps * kUnroll) {
        U8 raw[kUnroll];
        bool valid[kUnroll];
#pragma unroll
        for (int u = 0; u < kUnroll; ++u) {
            int64_t v = (c0 + u) * 32 + lane;
            valid[u] = v < nvec;
            if (valid[u]) raw[u] = ldg256(w + v * 32);
            else {
#pragma unroll
                for (int i = 0; i < 8; ++i) raw[u].r[i] = 0;
            }
        }
#pragma unroll
        for (int u = 0; u < kUnroll; ++u) {
            const int64_t v = (c0 + u) * 32 + lane;
            const int64_t g = v / LPG;
            float mx, mn;
            U8 out;
            float czs[VEC];
            IntQ p;
            if constexpr (DT != kF32) {
                using P = P2<DT>;
                if constexpr (CLIP) {
                    if (valid[u]) {
                        const typename P::h2 l2 = P::dup(load_scalar<DT>(cmin, g)), h2v = P::dup(load_scalar<DT>(cmax, g));
#pragma unroll
                        for (int i = 0; i < 8; ++i) raw[u].r[i] = P::bits(__hmin2(__hmax2(P::from_bits(raw[u].r[i]), l2), h2v));
                    }
                }
                vec_minmax16<DT>(raw[u], mx, mn);
                mx = group_max<LPG>(mx);
                mn = group_min<LPG>(mn);
                p = int_group_params<DT>(mx, mn, qmax, sem);
                const bool safe = FAST && (fmaxf(fabsf(mx), fabsf(mn)) * p.y < 1e30f);
                if (safe) {
                    int_vec16_fast<DT, RM, AUX>(raw[u], p, qmax, out, czs);
                } else {
                    int_vec_slow<DT, RM>(raw[u], p, qmax, out, czs);
                }
            } else {
                float f[VEC], o[VEC];
                unpack<DT>(raw[u], f);
                if constexpr (CLIP) {
                    if (valid[u]) {
                        float lo = load_scalar<DT>(cmin, g), hi = load_scalar<DT>(cmax, g);
#pragma unroll
                        for (int i = 0; i < VEC; ++i) f[i] = fminf(fmaxf(f[i], lo), hi);
                    }
                }
                mx = f[0];
                mn = f[0];
#pragma unroll
                for (int i = 1; i < VEC; ++i) { mx = fmaxf(mx, f[i]); mn = fminf(mn, f[i]); }
                mx = group_max<LPG>(mx);
                mn = group_min<LPG>(mn);
                p = int_group_params<DT>(mx, mn, qmax, sem);
#pragma unroll
                for (int i = 0; i < VEC; ++i) {
                    float qu;
                    float c = quant_code<DT, RM>(f[i], p, qmax, qu);
                    czs[i] = c - p.z;
                    o[i] = czs[i] * p.s;
                }
                pack<DT>(o, out);
            }
            if (valid[u]) {
                stg256(dq + v * 32, out);
                if constexpr (AUX) {
                    if (codes != nullptr) {
                        uint32_t* cp = reinterpret_cast<uint32_t*>(codes + v * VEC);
#pragma unroll
                        for (int i = 0; i < VEC / 4; ++i) {
                            uint32_t cw = 0;
#pragma unroll
                            for (int j = 0; j < 4; ++j) cw |= ((uint32_t)(czs[4 * i + j] + p.z) & 0xffu) << (8 * j);
                            cp[i] = cw;
                        }
                    }
                    if ((lane % LPG) == 0) {
                        if (scales != nullptr) store_scalar<DT>(scales, g, p.s);
                        if (zeros != nullptr) store_scalar<DT>(zeros, g, p.z);
                    }
                }
            }
        }
    }
}

template <int DT, int G, bool FAST>
__global__ void __launch_bounds__(kThreads) int_ste_bwd_kernel(
    const uint8_t* __restrict__ w, const uint8_t* __restrict__ gy, uint8_t* __restrict__ gw, int64_t nvec,
    float qmax, int sem) {
    constexpr int VEC = VecN<DT>::v;
    constexpr int LPG = G / VEC;
    const int lane = threadIdx.x & 31;
    const int64_t warp = (int64_t)blockIdx.x * (kThreads / 32) + (threadIdx.x >> 5);
    const int64_t nwarps = (int64_t)gridDim.x * (kThreads / 32);
    const int64_t nchunks = (nvec + 31) / 32;
    const float inv_qmax = 1.0f / qmax;

    for (int64_t c0 = warp * kUnroll; c0 < nchunks; c0 += nwarps * kUnroll) {
        U8 rw[kUnroll], rg[kUnroll];
        bool valid[kUnroll];
#pragma unroll
        for (int u = 0; u < kUnroll; ++u) {
            int64_t v = (c0 + u) * 32 + lane;
            valid[u] = v < nvec;
            if (valid[u]) {
                rw[u] = ldg256(w + v * 32);
                rg[u] = ldg256(gy + v * 32);
            } else {
#pragma unroll
                for (int i = 0; i < 8; ++i) { rw[u].r[i] = 0; rg[u].r[i] = 0; }
            }
        }
#pragma unroll
        for (int u = 0; u < kUnroll; ++u) {
            const int64_t v = (c0 + u) * 32 + lane;
            float f[VEC], g[VEC];
            unpack<DT>(rw[u], f);
            unpack<DT>(rg[u], g);
            float mx, mn;
            if constexpr (DT != kF32) {
                vec_minmax16<DT>(rw[u], mx, mn);
            } else {
                mx = f[0];
                mn = f[0];
#pragma unroll
                for (int i = 1; i < VEC; ++i) { mx = fmaxf(mx, f[i]); mn = fminf(mn, f[i]); }
            }
            mx = group_max<LPG>(mx);
            mn = group_min<LPG>(mn);
            const IntQ p = int_group_params<DT>(mx, mn, qmax, sem);
            const float lo = -p.z, hi = qmax - p.z;
            float Gs = 0.f;
            int cnt = 0;  // low 16 bits: #(x == max), high 16 bits: #(x == min)
            uint32_t mbits = 0;
            bool safe = false;
            if constexpr (DT != kF32) safe = FAST && (fmaxf(fabsf(mx), fabsf(mn)) * p.y < 1e30f);
            if (safe) {
                if constexpr (DT != kF32) {
                    using P = P2<DT>;
#pragma unroll
                    for (int i = 0; i < VEC; i += 2) {
                        float q0a, q0b;
                        const float ua = div_fast(f[i], p.s, p.y, q0a);
                        const float ub = div_fast(f[i + 1], p.s, p.y, q0b);
                        const typename P::h2 u2 = P::pack(ua, ub);
                        const float ra = round_T<DT, BDQ_ROUND_HALF_AWAY>(P::lo(u2));
                        const float rb = round_T<DT, BDQ_ROUND_HALF_AWAY>(P::hi(u2));
                        const bool ma = (ra >= lo) && (ra <= hi), mb = (rb >= lo) && (rb <= hi);
                        mbits |= (ma ? 1u : 0u) << i;
                        mbits |= (mb ? 1u : 0u) << (i + 1);
                        Gs += g[i] * (fminf(fmaxf(ra, lo), hi) - (ma ? q0a : 0.f));
                        Gs += g[i + 1] * (fminf(fmaxf(rb, lo), hi) - (mb ? q0b : 0.f));
                    }
                }
            } else {
#pragma unroll
                for (int i = 0; i < VEC; ++i) {
                    float qu;
                    float c = quant_code<DT, BDQ_ROUND_HALF_AWAY>(f[i], p, qmax, qu);
                    bool m = (qu >= 0.f) && (qu <= qmax);  // clamp backward: closed interval
                    mbits |= (m ? 1u : 0u) << i;
                    Gs += g[i] * ((c - p.z) - (m ? f[i] * p.y : 0.f));
                }
            }
#pragma unroll
            for (int i = 0; i < VEC; ++i) cnt += (f[i] == mx ? 1 : 0) + (f[i] == mn ? 65536 : 0);
            Gs = group_sum<LPG>(Gs);
            cnt = group_sum_i<LPG>(cnt);
            const float gs = (p.t1 >= 1e-5f) ? Gs * inv_qmax : 0.f;  // clamp(min=1e-5) cuts the scale path
            const float gmax = gs / (float)(cnt & 0xffff);
            const float gmin = gs / (float)(cnt >> 16);
            float o[VEC];
#pragma unroll
            for (int i = 0; i < VEC; ++i) {
                float r = ((mbits >> i) & 1u) ? g[i] : 0.f;
                if (f[i] == mx) r += gmax;
                if (f[i] == mn) r -= gmin;
                o[i] = r;
            }
            if (valid[u]) {
                U8 out;
                pack<DT>(o, out);
                stg256(gw + v * 32, out);
            }
        }
    }
}

// ---- generic group size (any G, e.g. q_group_size = -1 -> whole row): one warp per group, scalar strided loops
template <int DT, int RM, bool BWD>
__global__ void __launch_bounds__(kThreads) int_generic_kernel(
    const void* __restrict__ w, const void* __restrict__ gy, void* __restrict__ out, void* __restrict__ scales,
    void* __restrict__ zeros, uint8_t* __restrict__ codes, const void* __restrict__ cmin,
    const void* __restrict__ cmax, int64_t ngroups, int G, float qmax, int sem) {
    const int lane = threadIdx.x & 31;
    const int64_t warp = (int64_t)blockIdx.x * (kThreads / 32) + (threadIdx.x >> 5);
    const int64_t nwarps = (int64_t)gridDim.x * (kThreads / 32);
    for (int64_t g = warp; g < ngroups; g += nwarps) {
        const int64_t base = g * G;
        float lo = -INFINITY, hi = INFINITY;
        if (cmin != nullptr) { lo = load_scalar<DT>(cmin, g); hi = load_scalar<DT>(cmax, g); }
        float mx = -INFINITY, mn = INFINITY;
        for (int k = lane; k < G; k += 32) {
            float x = fminf(fmaxf(load_scalar<DT>(w, base + k), lo), hi);
            mx = fmaxf(mx, x);
            mn = fminf(mn, x);
        }
        mx = group_max<32>(mx);
        mn = group_min<32>(mn);
        const IntQ p = int_group_params<DT>(mx, mn, qmax, sem);
        if constexpr (!BWD) {
            for (int k = lane; k < G; k += 32) {
                float x = fminf(fmaxf(load_scalar<DT>(w, base + k), lo), hi);
                float qu;
                float c = quant_code<DT, RM>(x, p, qmax, qu);
                store_scalar<DT>(out, base + k, (c - p.z) * p.s);
                if (codes != nullptr) codes[base + k] = (uint8_t)c;
            }
            if (lane == 0) {
                if (scales != nullptr) store_scalar<DT>(scales, g, p.s);
                if (zeros != nullptr) store_scalar<DT>(zeros, g, p.z);
            }
        } else {
            float Gs = 0.f;
            int nmax = 0, nmin = 0;
            for (int k = lane; k < G; k += 32) {
                float x = load_scalar<DT>(w, base + k);
                float gk = load_scalar<DT>(gy, base + k);
                float qu;
                float c = quant_code<DT, BDQ_ROUND_HALF_AWAY>(x, p, qmax, qu);
                bool m = (qu >= 0.f) && (qu <= qmax);
                Gs += gk * ((c - p.z) - (m ? x * p.y : 0.f));
                nmax += (x == mx);
                nmin += (x == mn);
            }
            Gs = group_sum<32>(Gs);
            nmax = group_sum_i<32>(nmax);
            nmin = group_sum_i<32>(nmin);
            const float gs = (p.t1 >= 1e-5f) ? Gs / qmax : 0.f;
            for (int k = lane; k < G; k += 32) {
                float x = load_scalar<DT>(w, base + k);
                float gk = load_scalar<DT>(gy, base + k);
                float qu;
                quant_code<DT, BDQ_ROUND_HALF_AWAY>(x, p, qmax, qu);
                float r = ((qu >= 0.f) && (qu <= qmax)) ? gk : 0.f;
                if (x == mx) r += gs / (float)nmax;
                if (x == mn) r -= gs / (float)nmin;
                store_scalar<DT>(out, base + k, r);
            }
        }
    }
}

static int grid_for(int64_t work_warps, int unroll) {
    int64_t blocks = (work_warps + (kThreads / 32) * unroll - 1) / ((kThreads / 32) * unroll);
    int64_t cap = (int64_t)sm_count() * 60;   // whole waves for 3, 4, 5 or 6 resident CTAs per SM (see wave_grid)
    if (blocks > cap) blocks = cap;
    if (blocks < 1) blocks = 1;
    return (int)blocks;
}

template <int DT, int RM, bool FAST, bool AUX, bool CLIP>
static void launch_fwd_g(const uint8_t* wp, uint8_t* op, void* scales, void* zeros, uint8_t* codes, const void* cmin,
                         const void* cmax, int64_t nvec, int G, float qmax, int sem, int grid, cudaStream_t st) {
    if (G == 128) int_fq_fwd_kernel<DT, 128, RM, FAST, AUX, CLIP><<<grid, kThreads, 0, st>>>(wp, op, scales, zeros, codes, cmin, cmax, nvec, qmax, sem);
    else if (G == 64) int_fq_fwd_kernel<DT, 64, RM, FAST, AUX, CLIP><<<grid, kThreads, 0, st>>>(wp, op, scales, zeros, codes, cmin, cmax, nvec, qmax, sem);
    else int_fq_fwd_kernel<DT, 32, RM, FAST, AUX, CLIP><<<grid, kThreads, 0, st>>>(wp, op, scales, zeros, codes, cmin, cmax, nvec, qmax, sem);
}

template <int DT, int RM>
static int launch_fwd(const void* w, void* dq, void* scales, void* zeros, uint8_t* codes, const void* cmin,
                      const void* cmax, int64_t n, int G, float qmax, int sem, cudaStream_t st) {
    constexpr int VEC = VecN<DT>::v;
    const bool aligned = ((uintptr_t)w % 32 == 0) && ((uintptr_t)dq % 32 == 0) &&
                         (codes == nullptr || (uintptr_t)codes % 4 == 0);
    if (aligned && (G == 32 || G == 64 || G == 128) && G >= VEC) {
        int64_t nvec = n / VEC;
        int grid = grid_for((nvec + 31) / 32, kUnroll);
        const uint8_t* wp = (const uint8_t*)w;
        uint8_t* op = (uint8_t*)dq;
        const bool aux = scales || zeros || codes;
        const bool clip = cmin != nullptr;
        const bool fast = qmax <= 63.f;  // packed rounding identity holds for bits <= 6 (fq_packed.cuh)
        if (!aux && !clip) {
            if (fast) launch_fwd_g<DT, RM, true, false, false>(wp, op, scales, zeros, codes, cmin, cmax, nvec, G, qmax, sem, grid, st);
            else launch_fwd_g<DT, RM, false, false, false>(wp, op, scales, zeros, codes, cmin, cmax, nvec, G, qmax, sem, grid, st);
        } else if (!clip) {
            if (fast) launch_fwd_g<DT, RM, true, true, false>(wp, op, scales, zeros, codes, cmin, cmax, nvec, G, qmax, sem, grid, st);
            else launch_fwd_g<DT, RM, false, true, false>(wp, op, scales, zeros, codes, cmin, cmax, nvec, G, qmax, sem, grid, st);
        } else {
            if (fast) launch_fwd_g<DT, RM, true, true, true>(wp, op, scales, zeros, codes, cmin, cmax, nvec, G, qmax, sem, grid, st);
            else launch_fwd_g<DT, RM, false, true, true>(wp, op, scales, zeros, codes, cmin, cmax, nvec, G, qmax, sem, grid, st);
        }
    } else {
        int grid = grid_for(n / G, 1);
        int_generic_kernel<DT, RM, false><<<grid, kThreads, 0, st>>>(w, nullptr, dq, scales, zeros, codes, cmin, cmax, n / G, G, qmax, sem);
    }
    return check_launch("bdq_int_fakequant_fwd");
}

template <int DT, bool FAST>
static void launch_bwd_g(const uint8_t* wp, const uint8_t* gp, uint8_t* op, int64_t nvec, int G, float qmax, int sem,
                         int grid, cudaStream_t st) {
    if (G == 128) int_ste_bwd_kernel<DT, 128, FAST><<<grid, kThreads, 0, st>>>(wp, gp, op, nvec, qmax, sem);
    else if (G == 64) int_ste_bwd_kernel<DT, 64, FAST><<<grid, kThreads, 0, st>>>(wp, gp, op, nvec, qmax, sem);
    else int_ste_bwd_kernel<DT, 32, FAST><<<grid, kThreads, 0, st>>>(wp, gp, op, nvec, qmax, sem);
}

template <int DT>
static int launch_bwd(const void* w, const void* gy, void* gw, int64_t n, int G, float qmax, int sem, cudaStream_t st) {
    constexpr int VEC = VecN<DT>::v;
    const bool aligned = ((uintptr_t)w % 32 == 0) && ((uintptr_t)gy % 32 == 0) && ((uintptr_t)gw % 32 == 0);
    if (aligned && (G == 32 || G == 64 || G == 128) && G >= VEC) {
        int64_t nvec = n / VEC;
        int grid = grid_for((nvec + 31) / 32, kUnroll);
        const uint8_t *wp = (const uint8_t*)w, *gp = (const uint8_t*)gy;
        uint8_t* op = (uint8_t*)gw;
        if (qmax <= 63.f) launch_bwd_g<DT, true>(wp, gp, op, nvec, G, qmax, sem, grid, st);
        else launch_bwd_g<DT, false>(wp, gp, op, nvec, G, qmax, sem, grid, st);
    } else {
        int grid = grid_for(n / G, 1);
        int_generic_kernel<DT, BDQ_ROUND_HALF_AWAY, true><<<grid, kThreads, 0, st>>>(w, gy, gw, nullptr, nullptr, nullptr, nullptr, nullptr, n / G, G, qmax, sem);
    }
    return check_launch("bdq_int_ste_bwd");
}

}  // namespace bdq

using namespace bdq;

static int check_common(int64_t n, int group, int bits, int dtype) {
    if (n < 0 || group <= 0 || n % group != 0) { set_error("n_elem=%lld must be a non-negative multiple of group=%d", (long long)n, group); return BDQ_ERR_SHAPE; }
    if (bits < 1 || bits > 8) { set_error("bits=%d unsupported (1..8)", bits); return BDQ_ERR_UNSUPPORTED; }
    if (dtype < 0 || dtype > 2) { set_error("dtype=%d unknown", dtype); return BDQ_ERR_DTYPE; }
    return BDQ_OK;
}

extern "C" int bdq_int_fakequant_fwd(const void* w, void* dq, void* scales, void* zeros, uint8_t* codes,
                                     const void* clip_min, const void* clip_max, int64_t n_elem, int group, int bits,
                                     int dtype, int round_mode, int sem, bdq_stream_t stream) {
    int rc = check_common(n_elem, group, bits, dtype);
    if (rc) return rc;
    if ((clip_min == nullptr) != (clip_max == nullptr)) { set_error("clip_min and clip_max must be given together"); return BDQ_ERR_SHAPE; }
    if (n_elem == 0) return BDQ_OK;
    if (!w || !dq) { set_error("null tensor pointer"); return BDQ_ERR_SHAPE; }
    const float qmax = (float)((1 << bits) - 1);
    cudaStream_t st = (cudaStream_t)stream;
    if (!scales && !zeros && !codes &&
        launch_int_thr_fwd(w, dq, clip_min, clip_max, n_elem, group, bits, dtype, round_mode, sem, st))
        return check_launch("bdq_int_fakequant_fwd(thr)");
#define BDQ_FWD(DT)                                                                                                   \
    return round_mode == BDQ_ROUND_HALF_AWAY                                                                           \
               ? launch_fwd<DT, BDQ_ROUND_HALF_AWAY>(w, dq, scales, zeros, codes, clip_min, clip_max, n_elem, group, qmax, sem, st) \
               : launch_fwd<DT, BDQ_ROUND_HALF_EVEN>(w, dq, scales, zeros, codes, clip_min, clip_max, n_elem, group, qmax, sem, st)
    if (dtype == BDQ_F32) { BDQ_FWD(kF32); }
    if (dtype == BDQ_BF16) { BDQ_FWD(kBF16); }
    BDQ_FWD(kF16);
#undef BDQ_FWD
}

extern "C" int bdq_int_ste_bwd(const void* w, const void* gy, void* gw, int64_t n_elem, int group, int bits, int dtype,
                               int sem, bdq_stream_t stream) {
    int rc = check_common(n_elem, group, bits, dtype);
    if (rc) return rc;
    if (n_elem == 0) return BDQ_OK;
    if (!w || !gy || !gw) { set_error("null tensor pointer"); return BDQ_ERR_SHAPE; }
    const float qmax = (float)((1 << bits) - 1);
    cudaStream_t st = (cudaStream_t)stream;
    if (launch_int_thr_bwd(w, gy, gw, n_elem, group, bits, dtype, sem, st, 0, nullptr)) return check_launch("bdq_int_ste_bwd(thr)");
    if (dtype == BDQ_F32) return launch_bwd<kF32>(w, gy, gw, n_elem, group, qmax, sem, st);
    if (dtype == BDQ_BF16) return launch_bwd<kBF16>(w, gy, gw, n_elem, group, qmax, sem, st);
    return launch_bwd<kF16>(w, gy, gw, n_elem, group, qmax, sem, st);
}

// STE backward whose incoming gradient is the sum of `nsrc` tensors (1..8): gy[0] + gy[1] + ... accumulated in that order with
// one T rounding per addition.  The reduce-scatter of an FSDP2 gradient and the STE backward of the shard in one pass: gy[i]
// are this rank's rows of its own gradient and the inbox slots the other ranks pushed theirs into (bdq_rows_push).
extern "C" int bdq_int_ste_bwd_nsrc(const void* w, int nsrc, const void* const* gy, void* gw, int64_t n_elem, int group, int bits,
                                    int dtype, int sem, bdq_stream_t stream) {
    int rc = check_common(n_elem, group, bits, dtype);
    if (rc) return rc;
    if (nsrc < 1 || nsrc > 8 || !gy) { set_error("int_ste_bwd_nsrc: nsrc=%d must be 1..8", nsrc); return BDQ_ERR_SHAPE; }
    if (n_elem == 0) return BDQ_OK;
    if (!w || !gw) { set_error("null tensor pointer"); return BDQ_ERR_SHAPE; }
    for (int i = 0; i < nsrc; ++i)
        if (!gy[i]) { set_error("int_ste_bwd_nsrc: source %d is null", i); return BDQ_ERR_SHAPE; }
    if (nsrc == 1) return bdq_int_ste_bwd(w, gy[0], gw, n_elem, group, bits, dtype, sem, stream);
    if (launch_int_thr_bwd(w, gy[0], gw, n_elem, group, bits, dtype, sem, (cudaStream_t)stream, nsrc - 1, gy + 1))
        return check_launch("bdq_int_ste_bwd_nsrc");
    set_error("int_ste_bwd_nsrc: needs a 16-bit dtype, bits 2 or 3, group 32/64/128 and 32-byte aligned tensors");
    return BDQ_ERR_UNSUPPORTED;
}
