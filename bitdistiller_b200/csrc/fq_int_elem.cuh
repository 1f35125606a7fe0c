// Op-by-op element path of the integer fake-quant, shared by fakequant_int.cu and fakequant_thr.cu.
#pragma once
#include "bdq_common.cuh"
#include "fq_packed.cuh"

namespace bdq {

template <int DT, int RM>
__device__ __forceinline__ float quant_code(float x, const IntQ& p, float qmax, float& q_unclamped) {
    float u = rT<DT>(div_exact<DT>(x, p.s, p.y));
    float r;
    if constexpr (RM == BDQ_ROUND_HALF_AWAY) {
        // Round.forward (quantizer.py:14-17): sign(u) * floor(rT(|u| + 0.5)); the product and the later
        // r + z, c - z are exact in T for every bit width <= 8 (integers below 2^8), see DESIGN.md.
        float a = rT<DT>(fabsf(u) + 0.5f);
        r = copysignf(floorf(a), u);
    } else {
        r = rintf(u);  // torch.round: half to even
    }
    float q = r + p.z;
    q_unclamped = q;
    return fminf(fmaxf(q, 0.f), qmax);
}

// op-by-op path for one vector (groups whose quotients may overflow, or bits > 6); rarely executed
template <int DT, int RM>
__device__ __noinline__ void int_vec_slow(const U8& raw, const IntQ& p, float qmax, U8& out, float* czs) {
    constexpr int VEC = VecN<DT>::v;
    float f[VEC], o[VEC];
    unpack<DT>(raw, f);
#pragma unroll
    for (int i = 0; i < VEC; ++i) {
        float qu;
        float c = quant_code<DT, RM>(f[i], p, qmax, qu);
        czs[i] = c - p.z;
        o[i] = czs[i] * p.s;
    }
    pack<DT>(o, out);
}

}  // namespace bdq
