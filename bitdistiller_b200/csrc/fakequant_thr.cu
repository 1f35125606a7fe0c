// Threshold-formulated int2 / int3 group fake-quant forward and STE backward for bf16 / fp16 (sm_100a): the hot
// kernels of BASELINE config 2.  See fakequant_thr.cuh for the exactness argument.  Data movement: every lane owns
// 64 contiguous bytes (32 elements, two LDG.E.256 / STG.E.256), a group of G elements is G/32 adjacent lanes.
#include "fakequant_thr.cuh"
#include "fq_int_elem.cuh"

namespace bdq {

constexpr int kThrThreads = 256;

template <int LPG>
__device__ __forceinline__ uint32_t group_lane_mask(int lane) {
    return ((LPG == 32) ? 0xffffffffu : ((1u << LPG) - 1u)) << (lane & ~(LPG - 1));
}

// thresholds T_{k0} .. T_{k0+NTH-1} (n = lo + k) of this lane's group, computed one per lane and broadcast.
// Returns false if any of them could not be bracketed (the group then takes the op-by-op path).
template <int DT, int RM, int LPG, int NTH>
__device__ __forceinline__ bool group_thresholds(const IntQ& p, float lo, int k0, bool safe, int lane,
                                                 typename P2<DT>::h2* t2) {
    constexpr int R = (NTH + LPG - 1) / LPG;
    const int lig = lane & (LPG - 1);
    float tl[R];
    bool ok = safe;
#pragma unroll
    for (int r = 0; r < R; ++r) {
        const int k = lig + r * LPG;
        tl[r] = 0.f;
        if (k < NTH) {
            bool o;
            tl[r] = find_threshold<DT, RM>(lo + (float)(k0 + k), p, o);
            ok = ok && o;
        }
    }
    const uint32_t okm = __ballot_sync(0xffffffffu, ok);
    const uint32_t gm = group_lane_mask<LPG>(lane);
#pragma unroll
    for (int k = 0; k < NTH; ++k) {
        const float t = __shfl_sync(0xffffffffu, tl[k / LPG], (lane & ~(LPG - 1)) + (k % LPG));
        t2[k] = P2<DT>::dup(t);
    }
    return (okm & gm) == gm;
}

// Op-by-op fallback of one lane's 32 elements.  It re-reads its input from global memory and stores its result itself,
// so that no register array of the hot path ever has its address taken (that would park the whole tile in local memory).
template <int DT, int RM>
__device__ __noinline__ void int_fwd_slow_global(const void* w, void* dq, int64_t e0, float lo, float hi, IntQ p, float qmax) {
#pragma unroll 1
    for (int i = 0; i < 32; ++i) {
        const float x = fminf(fmaxf(load_scalar<DT>(w, e0 + i), lo), hi);
        float qu;
        const float c = quant_code<DT, RM>(x, p, qmax, qu);
        store_scalar<DT>(dq, e0 + i, (c - p.z) * p.s);
    }
}

template <int DT, int G, int RM, int NT, bool CLIP>
__global__ void __launch_bounds__(kThrThreads) int_thr_fwd_kernel(const uint8_t* __restrict__ w, uint8_t* __restrict__ dq,
                                                                  const void* __restrict__ cmin, const void* __restrict__ cmax,
                                                                  int64_t nchunk, int sem) {
    using P = P2<DT>;
    using h2 = typename P::h2;
    constexpr int LPG = G / 32;
    const float qmax = (float)NT;
    const int lane = threadIdx.x & 31;
    const int64_t warp = (int64_t)blockIdx.x * (kThrThreads / 32) + (threadIdx.x >> 5);
    const int64_t nwarps = (int64_t)gridDim.x * (kThrThreads / 32);
    const int64_t nwc = (nchunk + 31) / 32;
    for (int64_t wc = warp; wc < nwc; wc += nwarps) {
        const int64_t idx = wc * 32 + lane;
        const bool valid = idx < nchunk;
        U8 raw[2];
        if (valid) {
            raw[0] = ldg256(w + idx * 64);
            raw[1] = ldg256(w + idx * 64 + 32);
        } else {
#pragma unroll
            for (int i = 0; i < 8; ++i) raw[0].r[i] = raw[1].r[i] = 0;
        }
        float clo = -INFINITY, chi = INFINITY;
        if constexpr (CLIP) {
            if (valid) {
                const int64_t g = idx / LPG;
                clo = load_scalar<DT>(cmin, g);
                chi = load_scalar<DT>(cmax, g);
                const h2 l2 = P::dup(clo), u2 = P::dup(chi);
#pragma unroll
                for (int v = 0; v < 2; ++v)
#pragma unroll
                    for (int i = 0; i < 8; ++i) raw[v].r[i] = P::bits(__hmin2(__hmax2(P::from_bits(raw[v].r[i]), l2), u2));
            }
        }
        float mx, mn, mx1, mn1;
        vec_minmax16<DT>(raw[0], mx, mn);
        vec_minmax16<DT>(raw[1], mx1, mn1);
        mx = group_max<LPG>(fmaxf(mx, mx1));
        mn = group_min<LPG>(fminf(mn, mn1));
        const IntQ p = int_group_params<DT>(mx, mn, qmax, sem);
        const bool safe = fmaxf(fabsf(mx), fabsf(mn)) * p.y < 1e30f;
        h2 t2[NT];
        const bool fast = group_thresholds<DT, RM, LPG, NT>(p, -p.z, 1, safe, lane, t2);
        U8 out[2];
        if (fast) {
            const h2 lo2 = P::dup(-p.z), s2 = P::dup(p.s);
#pragma unroll
            for (int v = 0; v < 2; ++v)
#pragma unroll
                for (int i = 0; i < 8; ++i) {
                    const h2 x2 = P::from_bits(raw[v].r[i]);
                    h2 c2 = __hge2(x2, t2[0]);
#pragma unroll
                    for (int k = 1; k < NT; ++k) c2 = __hadd2(c2, __hge2(x2, t2[k]));
                    out[v].r[i] = P::bits(__hmul2(__hadd2(c2, lo2), s2));   // rT((c - z) * s)
                }
            if (valid) {
                stg256(dq + idx * 64, out[0]);
                stg256(dq + idx * 64 + 32, out[1]);
            }
        } else if (valid) {
            int_fwd_slow_global<DT, RM>(w, dq, idx * 32, clo, chi, p, qmax);
        }
    }
}

// Op-by-op backward fallback of one lane's 32 elements, through global memory (see int_fwd_slow_global).
// Phase 1: scale-path partial sum (x) and tie counts (y: low 16 bits #max, high 16 bits #min, as int bits).
template <int DT>
__device__ __noinline__ float2 int_bwd_slow_phase1(const void* w, const void* gy, int64_t e0, IntQ p, float qmax, float mx, float mn) {
    float Gs = 0.f;
    int cnt = 0;
#pragma unroll 1
    for (int i = 0; i < 32; ++i) {
        const float x = load_scalar<DT>(w, e0 + i), g = load_scalar<DT>(gy, e0 + i);
        float qu;
        const float c = quant_code<DT, BDQ_ROUND_HALF_AWAY>(x, p, qmax, qu);
        const bool m = (qu >= 0.f) && (qu <= qmax);   // clamp backward: closed interval
        Gs += g * ((c - p.z) - (m ? x * p.y : 0.f));
        cnt += (x == mx ? 1 : 0) + (x == mn ? 65536 : 0);
    }
    return make_float2(Gs, __int_as_float(cnt));
}
// Phase 2: gw = g * mask + [x == max] * gmax + [x == min] * gmin
template <int DT>
__device__ __noinline__ void int_bwd_slow_phase2(const void* w, const void* gy, void* gw, int64_t e0, IntQ p, float qmax, float mx,
                                                 float mn, float gmax, float gmin) {
#pragma unroll 1
    for (int i = 0; i < 32; ++i) {
        const float x = load_scalar<DT>(w, e0 + i), g = load_scalar<DT>(gy, e0 + i);
        float qu;
        quant_code<DT, BDQ_ROUND_HALF_AWAY>(x, p, qmax, qu);
        float r = ((qu >= 0.f) && (qu <= qmax)) ? g : 0.f;
        if (x == mx) r += gmax;
        if (x == mn) r += gmin;
        store_scalar<DT>(gw, e0 + i, r);
    }
}

// MS (multi-source): the incoming gradient is the sum of gy and ex.n further tensors of the same shape -- the other ranks'
// gradients of this rank's FSDP2 shard, pushed into a symmetric-memory inbox over NVLink (bdq_rows_push) -- accumulated in
// the given order with one T rounding per addition (HADD2: the correctly rounded sum of two T values, what a ring
// reduce-scatter in T computes; for two sources it IS the rounded exact sum).  The reduce-scatter and the STE backward are
// one pass: the reduced gradient is never materialised.
struct GySrcs {
    const uint8_t* p[7];
    int n;
};

template <int DT, int G, int NT, bool MS>
__global__ void __launch_bounds__(kThrThreads, 4) int_thr_bwd_kernel(const uint8_t* __restrict__ w, const uint8_t* __restrict__ gy,
                                                                  uint8_t* gw, int64_t nchunk, int sem, const GySrcs ex) {
    using P = P2<DT>;
    using h2 = typename P::h2;
    constexpr int LPG = G / 32;
    constexpr int RM = BDQ_ROUND_HALF_AWAY;
    const float qmax = (float)NT;
    const float inv_qmax = 1.0f / qmax;
    const int lane = threadIdx.x & 31;
    const int64_t warp = (int64_t)blockIdx.x * (kThrThreads / 32) + (threadIdx.x >> 5);
    const int64_t nwarps = (int64_t)gridDim.x * (kThrThreads / 32);
    const int64_t nwc = (nchunk + 31) / 32;
    for (int64_t wc = warp; wc < nwc; wc += nwarps) {
        const int64_t idx = wc * 32 + lane;
        const bool valid = idx < nchunk;
        U8 rw[2], rg[2];
        if (valid) {
            rw[0] = ldg256(w + idx * 64);
            rw[1] = ldg256(w + idx * 64 + 32);
            rg[0] = ldg256(gy + idx * 64);
            rg[1] = ldg256(gy + idx * 64 + 32);
        } else {
#pragma unroll
            for (int i = 0; i < 8; ++i) rw[0].r[i] = rw[1].r[i] = rg[0].r[i] = rg[1].r[i] = 0;
        }
        if constexpr (MS) {
            if (valid) {
#pragma unroll 1
                for (int sidx = 0; sidx < ex.n; ++sidx) {
                    const U8 a0 = ldg256(ex.p[sidx] + idx * 64), a1 = ldg256(ex.p[sidx] + idx * 64 + 32);
#pragma unroll
                    for (int i = 0; i < 8; ++i) {
                        rg[0].r[i] = P::bits(__hadd2(P::from_bits(rg[0].r[i]), P::from_bits(a0.r[i])));
                        rg[1].r[i] = P::bits(__hadd2(P::from_bits(rg[1].r[i]), P::from_bits(a1.r[i])));
                    }
                }
            }
        }
        float mx, mn, mx1, mn1;
        vec_minmax16<DT>(rw[0], mx, mn);
        vec_minmax16<DT>(rw[1], mx1, mn1);
        mx = group_max<LPG>(fmaxf(mx, mx1));
        mn = group_min<LPG>(fminf(mn, mn1));
        const IntQ p = int_group_params<DT>(mx, mn, qmax, sem);
        const bool safe = fmaxf(fabsf(mx), fabsf(mn)) * p.y < 1e30f;
        h2 t2[NT + 2];   // T_0 (mask low), T_1..T_NT (codes), T_{NT+1} (mask high)
        bool fast = group_thresholds<DT, RM, LPG, NT>(p, -p.z, 1, safe, lane, t2 + 1);
        {
            // The two clamp-mask thresholds are almost never needed: the group minimum normally quantises to code 0 and the
            // maximum to code NT, so the predicate (monotone in x) at the two extremes already says "no element is clamped"
            // and the mask is all ones (T_0 = -inf, T_{NT+1} = +inf).  Only otherwise are the two thresholds searched.
            float tlo = -INFINITY, thi = INFINITY;
            const bool need_lo = !thr_pred<DT, RM>(mn, p, -p.z);
            const bool need_hi = thr_pred<DT, RM>(mx, p, -p.z + (float)(NT + 1));
            if (need_lo) {
                bool o;
                tlo = find_threshold<DT, RM>(-p.z, p, o);
                fast = fast && o;
            }
            if (need_hi) {
                bool o;
                thi = find_threshold<DT, RM>(-p.z + (float)(NT + 1), p, o);
                fast = fast && o;
            }
            t2[0] = P::dup(tlo);
            t2[NT + 1] = P::dup(thi);
        }
        float Gs = 0.f;
        int cnt = 0;   // low 16 bits: #(x == max), high 16 bits: #(x == min)
        if (fast) {
            const h2 lo2 = P::dup(-p.z), mx2 = P::dup(mx), mn2 = P::dup(mn);
            const float ny = -p.y;
            h2 cx2 = P::dup(0.f), cn2 = P::dup(0.f);
#pragma unroll
            for (int v = 0; v < 2; ++v)
#pragma unroll
                for (int i = 0; i < 8; ++i) {
                    const h2 x2 = P::from_bits(rw[v].r[i]), g2 = P::from_bits(rg[v].r[i]);
                    h2 c2 = __hge2(x2, t2[1]);
#pragma unroll
                    for (int k = 2; k <= NT; ++k) c2 = __hadd2(c2, __hge2(x2, t2[k]));
                    const h2 cz2 = __hadd2(c2, lo2);                                    // c - z
                    const h2 m2 = __hmul2(__hge2(x2, t2[0]), __hlt2(x2, t2[NT + 1]));   // clamp pass-through mask
                    const h2 xm2 = __hmul2(x2, m2);
                    // scale-path cotangent: sum g * ((c - z) - m * x / s), fp32
                    Gs = fmaf(P::lo(g2), fmaf(P::lo(xm2), ny, P::lo(cz2)), Gs);
                    Gs = fmaf(P::hi(g2), fmaf(P::hi(xm2), ny, P::hi(cz2)), Gs);
                    cx2 = __hadd2(cx2, __heq2(x2, mx2));
                    cn2 = __hadd2(cn2, __heq2(x2, mn2));
                    rg[v].r[i] = P::bits(__hmul2(g2, m2));                              // direct STE term g * m
                }
            cnt = (int)(P::lo(cx2) + P::hi(cx2)) + ((int)(P::lo(cn2) + P::hi(cn2)) << 16);
        }
        // op-by-op fallback reads the gradient back from memory: with several sources the summed gradient of this lane's 32
        // elements is parked in its own slot of gw first (read and rewritten by this thread only)
        const uint8_t* gsl = gy;
        if constexpr (MS) {
            gsl = gw;
            if (!fast && valid) {
                stg256(gw + idx * 64, rg[0]);
                stg256(gw + idx * 64 + 32, rg[1]);
            }
        }
        if (!fast && valid) {
            const float2 r = int_bwd_slow_phase1<DT>(w, gsl, idx * 32, p, qmax, mx, mn);
            Gs = r.x;
            cnt = __float_as_int(r.y);
        }
        Gs = group_sum<LPG>(Gs);
        cnt = group_sum_i<LPG>(cnt);
        const float gs = (p.t1 >= 1e-5f) ? Gs * inv_qmax : 0.f;   // clamp(min=1e-5) cuts the scale path
        const float gmax = gs / (float)(cnt & 0xffff);
        const float gmin = -gs / (float)(cnt >> 16);
        if (fast) {
            // amax / amin backward (even tie split): add gmax / gmin to the elements equal to the extremes, as
            // hi + lo bf16 parts so that the two packed FMAs lose at most one final rounding each
            const float gxh = rT<DT>(gmax), gnh = rT<DT>(gmin);
            const h2 mx2 = P::dup(mx), mn2 = P::dup(mn);
            const h2 xh2 = P::dup(gxh), xl2 = P::dup(gmax - gxh), nh2 = P::dup(gnh), nl2 = P::dup(gmin - gnh);
#pragma unroll
            for (int v = 0; v < 2; ++v)
#pragma unroll
                for (int i = 0; i < 8; ++i) {
                    const h2 x2 = P::from_bits(rw[v].r[i]);
                    const h2 ex2 = __heq2(x2, mx2), en2 = __heq2(x2, mn2);
                    h2 r2 = P::from_bits(rg[v].r[i]);
                    r2 = __hfma2(ex2, xh2, r2);
                    r2 = __hfma2(ex2, xl2, r2);
                    r2 = __hfma2(en2, nh2, r2);
                    r2 = __hfma2(en2, nl2, r2);
                    rg[v].r[i] = P::bits(r2);
                }
            if (valid) {
                stg256(gw + idx * 64, rg[0]);
                stg256(gw + idx * 64 + 32, rg[1]);
            }
        } else if (valid) {
            int_bwd_slow_phase2<DT>(w, gsl, gw, idx * 32, p, qmax, mx, mn, gmax, gmin);
        }
    }
}

static int64_t thr_blocks(int64_t nchunk) { return ((nchunk + 31) / 32 + (kThrThreads / 32) - 1) / (kThrThreads / 32); }

template <int DT, int RM, int NT, bool CLIP>
static void thr_fwd_g(const uint8_t* w, uint8_t* dq, const void* cmin, const void* cmax, int64_t nchunk, int G, int sem,
                      cudaStream_t st) {
    const int64_t nb = thr_blocks(nchunk);
#define BDQ_L(GG) int_thr_fwd_kernel<DT, GG, RM, NT, CLIP><<<wave_grid(int_thr_fwd_kernel<DT, GG, RM, NT, CLIP>, kThrThreads, nb, 16), kThrThreads, 0, st>>>(w, dq, cmin, cmax, nchunk, sem)
    if (G == 128) BDQ_L(128);
    else if (G == 64) BDQ_L(64);
    else BDQ_L(32);
#undef BDQ_L
}

template <int DT, int NT>
static void thr_bwd_g(const uint8_t* w, const uint8_t* gy, uint8_t* gw, int64_t nchunk, int G, int sem, cudaStream_t st, const GySrcs& ex) {
    const int64_t nb = thr_blocks(nchunk);
#define BDQ_L(GG)                                                                                                                        \
    do {                                                                                                                                 \
        if (ex.n > 0)                                                                                                                    \
            int_thr_bwd_kernel<DT, GG, NT, true><<<wave_grid(int_thr_bwd_kernel<DT, GG, NT, true>, kThrThreads, nb, 16), kThrThreads, 0, st>>>(w, gy, gw, nchunk, sem, ex); \
        else                                                                                                                             \
            int_thr_bwd_kernel<DT, GG, NT, false><<<wave_grid(int_thr_bwd_kernel<DT, GG, NT, false>, kThrThreads, nb, 16), kThrThreads, 0, st>>>(w, gy, gw, nchunk, sem, ex); \
    } while (0)
    if (G == 128) BDQ_L(128);
    else if (G == 64) BDQ_L(64);
    else BDQ_L(32);
#undef BDQ_L
}

// Returns true if the threshold kernel was launched (16-bit dtype, bits 2 or 3, group 32/64/128, 32-byte aligned).
bool launch_int_thr_fwd(const void* w, void* dq, const void* cmin, const void* cmax, int64_t n, int G, int bits, int dtype,
                        int round_mode, int sem, cudaStream_t st) {
    if ((dtype != kBF16 && dtype != kF16) || (bits != 2 && bits != 3) || (G != 32 && G != 64 && G != 128)) return false;
    if (((uintptr_t)w % 32) || ((uintptr_t)dq % 32) || (n % 32)) return false;
    const uint8_t* wp = (const uint8_t*)w;
    uint8_t* op = (uint8_t*)dq;
    const int64_t nchunk = n / 32;
#define BDQ_THR_F(DT, RM, NT)                                                                  \
    do {                                                                                       \
        if (cmin) thr_fwd_g<DT, RM, NT, true>(wp, op, cmin, cmax, nchunk, G, sem, st);         \
        else thr_fwd_g<DT, RM, NT, false>(wp, op, cmin, cmax, nchunk, G, sem, st);             \
    } while (0)
#define BDQ_THR_F2(DT)                                                                         \
    do {                                                                                       \
        if (round_mode == BDQ_ROUND_HALF_AWAY) { if (bits == 2) BDQ_THR_F(DT, 1, 3); else BDQ_THR_F(DT, 1, 7); } \
        else { if (bits == 2) BDQ_THR_F(DT, 0, 3); else BDQ_THR_F(DT, 0, 7); }                 \
    } while (0)
    if (dtype == kBF16) BDQ_THR_F2(kBF16); else BDQ_THR_F2(kF16);
#undef BDQ_THR_F2
#undef BDQ_THR_F
    return true;
}

// n_extra further gradient sources (see GySrcs); 0 = the plain backward
bool launch_int_thr_bwd(const void* w, const void* gy, void* gw, int64_t n, int G, int bits, int dtype, int sem,
                        cudaStream_t st, int n_extra, const void* const* extra) {
    if ((dtype != kBF16 && dtype != kF16) || (bits != 2 && bits != 3) || (G != 32 && G != 64 && G != 128)) return false;
    if (((uintptr_t)w % 32) || ((uintptr_t)gy % 32) || ((uintptr_t)gw % 32) || (n % 32)) return false;
    if (n_extra < 0 || n_extra > 7) return false;
    GySrcs ex;
    ex.n = n_extra;
    for (int i = 0; i < 7; ++i) {
        ex.p[i] = i < n_extra ? (const uint8_t*)extra[i] : nullptr;
        if (i < n_extra && (!extra[i] || ((uintptr_t)extra[i] % 32))) return false;
    }
    const uint8_t *wp = (const uint8_t*)w, *gp = (const uint8_t*)gy;
    uint8_t* op = (uint8_t*)gw;
    const int64_t nchunk = n / 32;
    if (dtype == kBF16) { if (bits == 2) thr_bwd_g<kBF16, 3>(wp, gp, op, nchunk, G, sem, st, ex); else thr_bwd_g<kBF16, 7>(wp, gp, op, nchunk, G, sem, st, ex); }
    else { if (bits == 2) thr_bwd_g<kF16, 3>(wp, gp, op, nchunk, G, sem, st, ex); else thr_bwd_g<kF16, 7>(wp, gp, op, nchunk, G, sem, st, ex); }
    return true;
}

}  // namespace bdq
