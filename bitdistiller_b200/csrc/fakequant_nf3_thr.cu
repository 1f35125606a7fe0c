// Threshold-formulated NF3 group fake-quant forward for bf16 / fp16 (sm_100a): BASELINE config 3
// (SteN2F3Quantizer, quantization/quantizer.py:184-242, with the fused apply_clip of clip_utils.py:328-346).
//
// Exactness argument.  For one group (scales sp = |max|, sn = |min| fixed):
//   * q = rT(x / s) with s = sp for x >= 0 and sn for x < 0 is a monotone non-decreasing function of the T value x, and
//     the two reference threshold chains (quantizer.py:224-236) form ONE ascending staircase over the signed q;
//   * the straight-through value (y - q).detach() + q = rT(rT(lvl - q) + q) equals lvl for EVERY T value q in [-1, 1]
//     (checked exhaustively for bf16 and fp16: all levels satisfy Sterbenz' condition on their own interval), so the
//     forward output is  rT(lvl(x) * s): a step function of x with 7 thresholds and 8 per-group output values.
// Each x-space threshold T_k = min{ x : rT(x / s) >= thr_k } is computed per group from the exact product (midpoint below
// thr_k) * s and verified with the op-by-op predicate at the candidate and its predecessor; a group whose candidate
// fails the check takes the op-by-op path (nf3_elem).  The per-element work is 7 packed compares + 7 selects.
#include <cstring>

#include "fakequant_thr.cuh"
#include "fq_nf3_elem.cuh"

namespace bdq {

constexpr int kNf3ThrThreads = 256;

template <int LPG>
__device__ __forceinline__ uint32_t nf3_group_lane_mask(int lane) {
    return ((LPG == 32) ? 0xffffffffu : ((1u << LPG) - 1u)) << (lane & ~(LPG - 1));
}

// ascending staircase: k = 0..2 negative thresholds (neg_thr[2], [1], [0]) on x / sn, k = 3..6 positive ones on x / sp
struct Nf3Stair {
    float mid[7];   // midpoint between thr[k] and the next T value below it (exact in fp32)
    int incl[7];    // 1: thr[k] has an even significand, a quotient exactly at the midpoint rounds up to it
    float thr[7];   // in q space, rounded to T
    float lvl[8];   // level values in q space, rounded to T: -1, neg_lvl[2], neg_lvl[1], 0, pos_lvl[1..4]
};

static float host_prev_T(float v, int dt, int* even) {   // next T value below v (v != 0), and whether v's significand is even
    uint32_t b;
    memcpy(&b, &v, 4);
    const int sh = (dt == kBF16) ? 16 : 13;   // fp32 bits below the T significand (fp16 normal range; |thr| > 2^-14)
    *even = ((b >> sh) & 1u) == 0u;
    const uint32_t step = 1u << sh;
    b = (v > 0.f) ? b - step : b + step;      // towards -inf: magnitude down for positive, up for negative
    float r;
    memcpy(&r, &b, 4);
    return r;
}

static Nf3Stair make_stair(const Nf3Tables& t, int dt) {
    Nf3Stair s;
    s.thr[0] = t.neg_thr[2]; s.thr[1] = t.neg_thr[1]; s.thr[2] = t.neg_thr[0];
    for (int i = 0; i < 4; ++i) s.thr[3 + i] = t.pos_thr[i];
    s.lvl[0] = t.neg_lvl[3]; s.lvl[1] = t.neg_lvl[2]; s.lvl[2] = t.neg_lvl[1]; s.lvl[3] = 0.f;
    for (int i = 1; i < 5; ++i) s.lvl[3 + i] = t.pos_lvl[i];
    for (int k = 0; k < 7; ++k) {
        int even;
        const float pv = host_prev_T(s.thr[k], dt, &even);
        s.mid[k] = 0.5f * (s.thr[k] + pv);    // exact: adjacent T values
        s.incl[k] = even;
    }
    return s;
}

// smallest T value x with rT(x / s) >= thr.  rT(x / s) >= thr  <=>  x / s >= mid (or > mid), mid = the midpoint between
// thr and its predecessor in T, so the threshold is the smallest T value >= (>) mid * s -- a product that is exact in fp32
// (<= 12 x 11 significant bits).  `incl` = thr has an even significand (a tie at the midpoint rounds to thr).  The
// candidate is then VERIFIED with the exact op-by-op predicate at the candidate and at its predecessor; ok = false sends
// the group to the op-by-op path, so exactness does not rest on the argument above.
template <int DT>
__device__ __forceinline__ float nf3_find_threshold(float thr, float mid, bool incl, float s, float y, bool& ok) {
    const float v = mid * s;
    float c = rT<DT>(v);
    if (incl ? (c < v) : (c <= v)) c = t_step<DT>(c, 1);
    const float cp = t_step<DT>(c, -1);
    float q0;
    const bool pc = rT<DT>(div_fast(c, s, y, q0)) >= thr;
    const bool pp = rT<DT>(div_fast(cp, s, y, q0)) >= thr;
    ok = pc && !pp && (c != 0.f) && (cp != 0.f);
    return c;
}

template <int DT>
__device__ __noinline__ void nf3_fwd_slow_global(const void* w, void* dq, int64_t e0, float lo, float hi, Nf3Q p, Nf3Tables t) {
#pragma unroll 1
    for (int i = 0; i < 32; ++i) {
        const float x = fminf(fmaxf(load_scalar<DT>(w, e0 + i), lo), hi);
        float q, y;
        int code;
        store_scalar<DT>(dq, e0 + i, nf3_elem<DT>(x, p, t, q, y, code));
    }
}

template <int DT, int G, bool CLIP>
__global__ void __launch_bounds__(kNf3ThrThreads) nf3_thr_fwd_kernel(const uint8_t* __restrict__ w, uint8_t* __restrict__ dq,
                                                                     const void* __restrict__ cmin, const void* __restrict__ cmax,
                                                                     int64_t nchunk, const Nf3Tables tab, const Nf3Stair st) {
    using P = P2<DT>;
    using h2 = typename P::h2;
    constexpr int LPG = G / 32;
    constexpr int R = (7 + LPG - 1) / LPG;
    const int lane = threadIdx.x & 31;
    const int lig = lane & (LPG - 1);
    const int64_t warp = (int64_t)blockIdx.x * (kNf3ThrThreads / 32) + (threadIdx.x >> 5);
    const int64_t nwarps = (int64_t)gridDim.x * (kNf3ThrThreads / 32);
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
        const Nf3Q p = nf3_group_params(mx, mn);
        // thresholds of this lane's group: one (or R) per lane, then broadcast
        float tl[R];
        bool ok = !p.nan_group && (p.sp * p.yp > 0.5f) && (p.sn * p.yn > 0.5f) && (p.yp < 1e30f) && (p.yn < 1e30f);
#pragma unroll
        for (int r = 0; r < R; ++r) {
            const int k = lig + r * LPG;
            tl[r] = 0.f;
            if (k < 7) {
                bool o;
                tl[r] = nf3_find_threshold<DT>(st.thr[k], st.mid[k], st.incl[k] != 0, k < 3 ? p.sn : p.sp, k < 3 ? p.yn : p.yp, o);
                ok = ok && o;
            }
        }
        const uint32_t okm = __ballot_sync(0xffffffffu, ok);
        const uint32_t gm = nf3_group_lane_mask<LPG>(lane);
        const bool fast = (okm & gm) == gm;
        if (fast) {
            uint32_t t2[7], v2[8];
#pragma unroll
            for (int k = 0; k < 7; ++k)
                t2[k] = P::bits(P::dup(__shfl_sync(0xffffffffu, tl[k / LPG], (lane & ~(LPG - 1)) + (k % LPG))));
#pragma unroll
            for (int c = 0; c < 8; ++c) v2[c] = P::bits(P::dup(st.lvl[c] * (c < 3 ? p.sn : p.sp)));   // rT(lvl * s)
            U8 out[2];
#pragma unroll
            for (int v = 0; v < 2; ++v)
#pragma unroll
                for (int i = 0; i < 8; ++i) {
                    const h2 x2 = P::from_bits(raw[v].r[i]);
                    uint32_t o = v2[0];
#pragma unroll
                    for (int k = 0; k < 7; ++k) {
                        const uint32_t m = __hge2_mask(x2, P::from_bits(t2[k]));
                        asm("lop3.b32 %0, %1, %2, %3, 0xCA;" : "=r"(o) : "r"(m), "r"(v2[k + 1]), "r"(o));   // m ? v : o
                    }
                    out[v].r[i] = o;
                }
            if (valid) {
                stg256(dq + idx * 64, out[0]);
                stg256(dq + idx * 64 + 32, out[1]);
            }
        } else if (valid) {
            nf3_fwd_slow_global<DT>(w, dq, idx * 32, clo, chi, p, tab);
        }
    }
}

static int nf3_thr_grid(int64_t nchunk) {
    int64_t blocks = ((nchunk + 31) / 32 + (kNf3ThrThreads / 32) - 1) / (kNf3ThrThreads / 32);
    int64_t cap = (int64_t)sm_count() * 8;
    if (blocks > cap) blocks = cap;
    if (blocks < 1) blocks = 1;
    return (int)blocks;
}

template <int DT, bool CLIP>
static void nf3_thr_fwd_g(const uint8_t* w, uint8_t* dq, const void* cmin, const void* cmax, int64_t nchunk, int G,
                          const Nf3Tables& tab, cudaStream_t st) {
    const int grid = nf3_thr_grid(nchunk);
    const Nf3Stair stair = make_stair(tab, DT);
    if (G == 128) nf3_thr_fwd_kernel<DT, 128, CLIP><<<grid, kNf3ThrThreads, 0, st>>>(w, dq, cmin, cmax, nchunk, tab, stair);
    else if (G == 64) nf3_thr_fwd_kernel<DT, 64, CLIP><<<grid, kNf3ThrThreads, 0, st>>>(w, dq, cmin, cmax, nchunk, tab, stair);
    else nf3_thr_fwd_kernel<DT, 32, CLIP><<<grid, kNf3ThrThreads, 0, st>>>(w, dq, cmin, cmax, nchunk, tab, stair);
}

// Returns true if the threshold kernel was launched (16-bit dtype, group 32/64/128, 32-byte aligned, no aux outputs).
bool launch_nf3_thr_fwd(const void* w, void* dq, const void* cmin, const void* cmax, int64_t n, int G, int dtype, int sem,
                        cudaStream_t st) {
    if ((dtype != kBF16 && dtype != kF16) || (G != 32 && G != 64 && G != 128)) return false;
    if (((uintptr_t)w % 32) || ((uintptr_t)dq % 32) || (n % 32)) return false;
    const Nf3Tables tab = make_tables(dtype, sem);
    const uint8_t* wp = (const uint8_t*)w;
    uint8_t* op = (uint8_t*)dq;
    const int64_t nchunk = n / 32;
    if (dtype == kBF16) {
        if (cmin) nf3_thr_fwd_g<kBF16, true>(wp, op, cmin, cmax, nchunk, G, tab, st);
        else nf3_thr_fwd_g<kBF16, false>(wp, op, cmin, cmax, nchunk, G, tab, st);
    } else {
        if (cmin) nf3_thr_fwd_g<kF16, true>(wp, op, cmin, cmax, nchunk, G, tab, st);
        else nf3_thr_fwd_g<kF16, false>(wp, op, cmin, cmax, nchunk, G, tab, st);
    }
    return true;
}

}  // namespace bdq
