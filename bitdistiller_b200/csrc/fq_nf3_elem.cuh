// NF3 element path (tables, group scales, one-element quantiser), shared by fakequant_nf3.cu and autoclip.cu.
#pragma once
#include "bdq_common.cuh"

namespace bdq {

// round_n2f3 decision thresholds and levels (quantizer.py:224-236), as doubles exactly as written there
struct Nf3Tables {
    float pos_thr[4], neg_thr[3], pos_lvl[5], neg_lvl[4];
};

static inline float host_round(float v, int dt) {
    if (dt == kBF16) return __bfloat162float(__float2bfloat16_rn(v));
    if (dt == kF16) return __half2float(__float2half_rn(v));
    return v;
}

static inline Nf3Tables make_tables(int dt, int sem) {
    const double pt[4] = {0.0916687622666359, 0.2826657369732857, 0.5024898052215576, 0.8114928305149078};
    const double nt[3] = {-0.1234657019376755, -0.39097706973552704, -0.7675113677978516};
    const double pl[5] = {0.0, 0.1833375245332718, 0.3819939494132996, 0.6229856610298157, 1.0};
    const double nl[4] = {0.0, -0.2469314038753510, -0.5350227355957031, -1.0};
    Nf3Tables t;
    // Both CPU and CUDA eager cast the python scalar to T before comparing (measured on the B200 box with
    // tools/dbg_nf3.py: SURVEY App. D-3's "CUDA compares in fp32" guess was wrong), so `sem` does not matter here.
    (void)sem;
    for (int i = 0; i < 4; ++i) t.pos_thr[i] = host_round((float)pt[i], dt);
    for (int i = 0; i < 3; ++i) t.neg_thr[i] = host_round((float)nt[i], dt);
    for (int i = 0; i < 5; ++i) t.pos_lvl[i] = host_round((float)pl[i], dt);
    for (int i = 0; i < 4; ++i) t.neg_lvl[i] = host_round((float)nl[i], dt);
    return t;
}

struct Nf3Q {
    float sp, sn, yp, yn;  // scales and their correctly rounded reciprocals
    bool nan_group;        // max == 0 or min == 0 exactly -> 0/0 poisons the whole group
};

__device__ __forceinline__ Nf3Q nf3_group_params(float mx, float mn) {
    Nf3Q p;
    p.sp = fabsf(mx);
    p.sn = fabsf(mn);
    p.yp = rcp_approx(p.sp);
    p.yn = rcp_approx(p.sn);
    p.nan_group = (p.sp == 0.f) || (p.sn == 0.f);
    return p;
}

// Quantise one element: returns the dequantised value; q = the normalised value, y = snapped STE value,
// code = level code.
template <int DT>
__device__ __forceinline__ float nf3_elem(float x, const Nf3Q& p, const Nf3Tables& t, float& q, float& y, int& code) {
    const bool pos = x >= 0.f;
    const float s = pos ? p.sp : p.sn;
    const float r = pos ? p.yp : p.yn;
    q = rT<DT>(div_exact<DT>(x, s, r));
    float lvl;
    if (pos) {
        if (q >= t.pos_thr[3]) { lvl = t.pos_lvl[4]; code = 4; }
        else if (q >= t.pos_thr[2]) { lvl = t.pos_lvl[3]; code = 3; }
        else if (q >= t.pos_thr[1]) { lvl = t.pos_lvl[2]; code = 2; }
        else if (q >= t.pos_thr[0]) { lvl = t.pos_lvl[1]; code = 1; }
        else { lvl = t.pos_lvl[0]; code = 0; }
    } else {
        if (q >= t.neg_thr[0]) { lvl = t.neg_lvl[0]; code = 0; }
        else if (q >= t.neg_thr[1]) { lvl = t.neg_lvl[1]; code = 5; }
        else if (q >= t.neg_thr[2]) { lvl = t.neg_lvl[2]; code = 6; }
        else { lvl = t.neg_lvl[3]; code = 7; }
    }
    y = rT<DT>(rT<DT>(lvl - q) + q);  // (y - q).detach() + q evaluated in T (quantizer.py:242)
    float out = y * s;                // the other sign's term is exactly 0 (rounded to T by the caller)
    if (p.nan_group) { out = __int_as_float(0x7fc00000); code = 255; }
    return out;
}

}  // namespace bdq
