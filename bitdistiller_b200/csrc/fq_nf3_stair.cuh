// The NF3 staircase in x space (shared by fakequant_nf3_thr.cu and autoclip_tc.cu): per-threshold constants from which a group's
// seven exact x-space thresholds are computed and verified, and the eight level values as packed T pairs.
#pragma once
#include <cstring>

#include "bdq_common.cuh"
#include "fq_nf3_elem.cuh"
#include "fq_packed.cuh"

namespace bdq {

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

static inline uint32_t host_f2u(float v) { uint32_t b; memcpy(&b, &v, 4); return b; }
static inline float host_u2f(uint32_t b) { float v; memcpy(&v, &b, 4); return v; }

static inline Nf3Stair make_stair(const Nf3Tables& t, int dt) {
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


// One thread, one group: the seven verified x-space thresholds T[k] (k = 0..2 on x / sn, 3..6 on x / sp).  false = some
// candidate failed its check (or a scale is zero / NaN / out of range): the caller takes the op-by-op path.
template <int DT>
__device__ __forceinline__ bool nf3_thresholds_1t(const Nf3Stair& st, float sp, float sn, float (&T)[7]) {
    const float yp = rcp_approx(sp), yn = rcp_approx(sn);
    constexpr float kLo = (DT == kBF16) ? 1e-30f : 0.001953125f;   // as nf3_group_thresholds
    bool ok = (sp >= kLo) && (sp < 1e30f) && (sn >= kLo) && (sn < 1e30f);
#pragma unroll
    for (int k = 0; k < 7; ++k) {
        const Nf3Step e = st.k[k];
        const float s = k < 3 ? sn : sp, y = k < 3 ? yn : yp;
        const uint32_t cb = (__float_as_uint(e.mid * s) + e.add) & st.mask;
        const float c = __uint_as_float(cb), cprev = __uint_as_float(cb + (uint32_t)e.pstep);
        float q0;
        const bool pc = div_fast(c, s, y, q0) >= e.midx;
        const bool pp = div_fast(cprev, s, y, q0) >= e.midx;
        ok = ok && pc && !pp;
        T[k] = c;
    }
    return ok;
}

}  // namespace bdq
