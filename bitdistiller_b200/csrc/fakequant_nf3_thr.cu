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
// thr_k) * s and then verified: with the fp32 quotient computed exactly as the op-by-op path computes it, the candidate
// must satisfy rT(q) >= thr_k and the T value just below it must not (rT(q) >= thr_k is tested as q >= midx, which is
// what round-to-nearest-even means).  A group whose candidate fails the check takes the op-by-op path (nf3_elem), so
// exactness never rests on the guess.  The per-element work is 7 packed compares + 7 selects, both on the ALU pipe,
// which bounds this kernel (ncu: ALU pipe 84 % active at 78 % of the HBM peak before the per-group work was trimmed).
#include <cstring>

#include "fakequant_thr.cuh"
#include "fq_nf3_elem.cuh"

namespace bdq {

constexpr int kNf3ThrThreads = 256;

template <int LPG>
__device__ __forceinline__ uint32_t nf3_group_lane_mask(int lane) {
    return ((LPG == 32) ? 0xffffffffu : ((1u << LPG) - 1u)) << (lane & ~(LPG - 1));
}

// ascending staircase: k = 0..2 negative thresholds (neg_thr[2], [1], [0]) on x / sn, k = 3..6 positive ones on x / sp.
// Per threshold thr (a T value; prev = the next T value below it, mid = their midpoint, exact in fp32):
//   rT(q) >= thr  <=>  q >= midx,   midx = mid if thr has an even significand (a tie rounds up to thr), else the next
//   fp32 value above mid -- plain round-to-nearest-even, no assumption about the data.
struct Nf3Step {
    float mid;       // candidate: smallest T value >= (or >) mid * s, computed on the fp32 bit pattern as
    float midx;      //   c = (bits(mid * s) + add) & mask   (ceil for positive values, truncation for negative ones)
    uint32_t add;
    int32_t pstep;   // bits(c) + pstep = the T value just below c
};
struct Nf3Stair {
    Nf3Step k[8];      // k[7]: dummy that always verifies (mid = midx = 1)
    uint32_t lvl2[8];  // level values as packed T pairs: -1, neg_lvl[2], neg_lvl[1], 0, pos_lvl[1..4]
    uint32_t mask;     // keeps the T significand of an fp32 pattern
};

static uint32_t host_f2u(float v) { uint32_t b; memcpy(&b, &v, 4); return b; }
static float host_u2f(uint32_t b) { float v; memcpy(&v, &b, 4); return v; }

static Nf3Stair make_stair(const Nf3Tables& t, int dt) {
    Nf3Stair s;
    const int sh = (dt == kBF16) ? 16 : 13;   // fp32 bits below the T significand (fp16: normal range only, |thr| > 2^-14)
    const uint32_t step = 1u << sh;
    s.mask = ~(step - 1u);
    float thr[7], lvl[8];
    thr[0] = t.neg_thr[2]; thr[1] = t.neg_thr[1]; thr[2] = t.neg_thr[0];
    for (int i = 0; i < 4; ++i) thr[3 + i] = t.pos_thr[i];
    lvl[0] = t.neg_lvl[3]; lvl[1] = t.neg_lvl[2]; lvl[2] = t.neg_lvl[1]; lvl[3] = 0.f;
    for (int i = 1; i < 5; ++i) lvl[3 + i] = t.pos_lvl[i];
    for (int k = 0; k < 7; ++k) {
        const uint32_t b = host_f2u(thr[k]);
        const bool neg = thr[k] < 0.f;
        const bool even = ((b >> sh) & 1u) == 0u;
        const float prev = host_u2f(neg ? b + step : b - step);   // towards -inf
        const float mid = 0.5f * (thr[k] + prev);                 // exact: adjacent T values
        s.k[k].mid = mid;
        s.k[k].midx = even ? mid : nextafterf(mid, INFINITY);
        s.k[k].add = neg ? (even ? 0u : 0xffffffffu) : (even ? step - 1u : step);
        s.k[k].pstep = neg ? (int32_t)step : -(int32_t)step;
    }
    s.k[7].mid = 1.f; s.k[7].midx = 1.f; s.k[7].add = 0u; s.k[7].pstep = -(int32_t)step;
    for (int c = 0; c < 8; ++c) {
        uint32_t h;
        if (dt == kBF16) h = host_f2u(lvl[c]) >> 16;
        else { const __half hv = __float2half_rn(lvl[c]); unsigned short u; memcpy(&u, &hv, 2); h = u; }
        s.lvl2[c] = h | (h << 16);
    }
    return s;
}

template <int DT>
__device__ __noinline__ void nf3_fwd_slow_global(const void* w, void* dq, int64_t e0, float lo, float hi, float mx, float mn,
                                                 const Nf3Tables& t) {
    const Nf3Q p = nf3_group_params(mx, mn);
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
                                                                     int64_t nchunk, const __grid_constant__ Nf3Tables tab,
                                                                     const __grid_constant__ Nf3Stair st) {
    using P = P2<DT>;
    using h2 = typename P::h2;
    constexpr int LPG = G / 32;
    constexpr int R = (7 + LPG - 1) / LPG;   // thresholds found per lane
    constexpr int NP = (R + 1) / 2;          // ... packed two per register
    const int lane = threadIdx.x & 31;
    const int lig = lane & (LPG - 1);
    const int gbase = lane & ~(LPG - 1);
    const int64_t warp = (int64_t)blockIdx.x * (kNf3ThrThreads / 32) + (threadIdx.x >> 5);
    const int64_t nwarps = (int64_t)gridDim.x * (kNf3ThrThreads / 32);
    const int64_t nwc = (nchunk + 31) / 32;
    Nf3Step my[R];   // this lane's thresholds: k = lig + r * LPG (k == 7, one lane per group: the always-true dummy)
#pragma unroll
    for (int r = 0; r < R; ++r) my[r] = st.k[lig + r * LPG];
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
        const float sp = fabsf(mx), sn = fabsf(mn);
        const float yp = rcp_approx(sp), yn = rcp_approx(sn);
        // scales in a range where the reciprocal is sane and (fp16) every candidate is a normal T value; a zero scale
        // (NaN group), NaN or inf fails these compares
        constexpr float kLo = (DT == kBF16) ? 1e-30f : 0.001953125f;   // 2^-9
        bool ok = (sp >= kLo) && (sp < 1e30f) && (sn >= kLo) && (sn < 1e30f);
        // thresholds of this lane's group: R per lane, verified, packed in pairs, then broadcast
        float tl[2 * NP];
#pragma unroll
        for (int r = 0; r < 2 * NP; ++r) tl[r] = 0.f;
#pragma unroll
        for (int r = 0; r < R; ++r) {
            const Nf3Step e = my[r];
            const bool neg = lig + r * LPG < 3;
            const float s = neg ? sn : sp, y = neg ? yn : yp;
            const uint32_t cb = (__float_as_uint(e.mid * s) + e.add) & st.mask;
            const float c = __uint_as_float(cb), cprev = __uint_as_float(cb + (uint32_t)e.pstep);
            float q0;
            const bool pc = div_fast(c, s, y, q0) >= e.midx;
            const bool pp = div_fast(cprev, s, y, q0) >= e.midx;
            ok = ok && pc && !pp;
            tl[r] = c;
        }
        uint32_t pk[NP], sh[LPG][NP];
#pragma unroll
        for (int j = 0; j < NP; ++j) pk[j] = P::bits(P::pack(tl[2 * j], tl[2 * j + 1]));   // exact: T values
#pragma unroll
        for (int l = 0; l < LPG; ++l)
#pragma unroll
            for (int j = 0; j < NP; ++j) sh[l][j] = (LPG == 1) ? pk[j] : __shfl_sync(0xffffffffu, pk[j], gbase + l);
        const uint32_t okm = __ballot_sync(0xffffffffu, ok);
        const uint32_t gm = nf3_group_lane_mask<LPG>(lane);
        const bool fast = (okm & gm) == gm;
        if (fast) {
            const h2 s2p = P::dup(sp), s2n = P::dup(sn);
            uint32_t v2[8];
#pragma unroll
            for (int c = 0; c < 8; ++c)   // rT(lvl * s): one rounding of an exact product either way
                v2[c] = (c == 3) ? 0u : P::bits(__hmul2(P::from_bits(st.lvl2[c]), c < 3 ? s2n : s2p));
            U8 out[2];
#pragma unroll
            for (int v = 0; v < 2; ++v)
#pragma unroll
                for (int i = 0; i < 8; ++i) {
                    const h2 x2 = P::from_bits(raw[v].r[i]);
                    uint32_t o = v2[0];
#pragma unroll
                    for (int k = 0; k < 7; ++k) {
                        const h2 tk = P::from_bits(sh[k % LPG][(k / LPG) / 2]);
                        const h2 t2 = ((k / LPG) & 1) ? P::dup_hi(tk) : P::dup_lo(tk);
                        const uint32_t m = __hge2_mask(x2, t2);
                        asm("lop3.b32 %0, %1, %2, %3, 0xCA;" : "=r"(o) : "r"(m), "r"(v2[k + 1]), "r"(o));   // m ? v : o
                    }
                    out[v].r[i] = o;
                }
            if (valid) {
                stg256(dq + idx * 64, out[0]);
                stg256(dq + idx * 64 + 32, out[1]);
            }
        } else if (valid) {
            nf3_fwd_slow_global<DT>(w, dq, idx * 32, clo, chi, mx, mn, tab);
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
