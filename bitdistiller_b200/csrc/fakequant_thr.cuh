// Threshold formulation of the group asymmetric integer fake-quant for 16-bit dtypes and few levels (bits 2, 3).
//
// For one group (scale s, zero z fixed) the reference map  x -> c(x) = clamp(round(rT(x / s)) + z, 0, qmax) - z
// (quantization/quantizer.py:60-61,176-177) is a monotone non-decreasing step function of the T value x: division by
// s > 0, rounding to T and both integer roundings are monotone.  It is therefore fully described by the qmax
// thresholds  T_k = min{ x in T : round(rT(x/s)) >= lo + k },  lo = -z:
//        c(x) - lo = #{ k : x >= T_k },      mask(x) = [T_0 <= x < T_{qmax+1}]   (clamp pass-through, backward)
// Each T_k is found per group by evaluating the exact op-by-op predicate on 5 neighbouring T values around
// rT((lo + k - 0.5) * s) and is accepted only if the predicate flips inside that window (P(prev) false, P(T) true);
// otherwise the group takes the op-by-op path.  Exactness thus never depends on the guess.  The per-element work
// collapses from ~17 (fwd) / ~44 (bwd) instructions to packed compares and adds (fq_packed.cuh arguments apply:
// sums of small integers and products with {0,1} are exact in T).
#pragma once
#include "bdq_common.cuh"
#include "fq_packed.cuh"

namespace bdq {

// T value <-> monotone 16-bit pattern stepping (same sign only; |x| stays far from 0 here)
template <int DT>
__device__ __forceinline__ float t_step(float x, int j) {
    if constexpr (DT == kBF16) {
        const uint32_t b = __float_as_uint(x);
        const int d = (b & 0x80000000u) ? -j : j;
        return __uint_as_float(b + ((uint32_t)d << 16));
    } else {
        const unsigned short h = __half_as_ushort(__float2half_rn(x));
        const int d = (h & 0x8000u) ? -j : j;
        return __half2float(__ushort_as_half((unsigned short)(h + d)));
    }
}

// exact predicate: round(rT(x / s)) >= n   (x a T value, |x / s| small: the caller checked the group range)
template <int DT, int RM>
__device__ __forceinline__ bool thr_pred(float x, const IntQ& p, float n) {
    float q0;
    const float u = rT<DT>(div_fast(x, p.s, p.y, q0));
    return round_T<DT, RM>(u) >= n;
}

// smallest T value x with round(rT(x/s)) >= n; ok = false if it is not bracketed by the 5-point window
template <int DT, int RM>
__device__ __forceinline__ float find_threshold(float n, const IntQ& p, bool& ok) {
    const float xc = rT<DT>((n - 0.5f) * p.s);
    uint32_t pm = 0;
#pragma unroll
    for (int j = 0; j < 5; ++j) pm |= (thr_pred<DT, RM>(t_step<DT>(xc, j - 2), p, n) ? 1u : 0u) << j;
    const int i = __ffs(pm) - 1;                       // first true
    ok = (i >= 1) && (pm == ((0x1fu << i) & 0x1fu));   // false below, true from i on (monotone)
    return t_step<DT>(xc, i - 2);
}

}  // namespace bdq
