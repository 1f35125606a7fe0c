// Packed 2-wide (bf16x2 / half2) building blocks for the 16-bit fake-quant kernels.
//
// Why packed math is bit-exact here (each claim is also brute-forced in tests/test_fakequant_gpu.py through the
// oracle comparison, and the two integer-rounding identities were checked over every finite bf16 / fp16 value):
//   * eager PyTorch evaluates a 16-bit op as  rT(fp32_op(a, b)).  For add / sub / mul / min / max of two T values the
//     fp32 result is either exact or far from a T rounding boundary, so the single-rounding HADD2/HMUL2/HMNMX2
//     give the same bits (products of two <=11-bit significands are exact in fp32; sums only round in fp32 when
//     the exponent gap exceeds 13 bits, where both roundings return the larger operand).
//   * (clamp(r + z, 0, qmax) - z) == clamp(r, -z, qmax - z) for integer-valued r, z (exact integer arithmetic;
//     when |r + z| is too large to be exact in T it is outside [0, qmax] before and after rounding).
//   * Round.forward = sign(u) * floor(rT(|u| + 0.5)) equals sign(u) * floor(|u| + 0.5 + ulp_T(0.25)) for every
//     |u| < 128 (bf16) / 1024 (fp16): the only T value whose rT(|u| + 0.5) crosses an integer is 0.5 - ulp_T(0.25).
//     Larger |u| only occur with q far outside [0, qmax] for bits <= 6; bits 7..8 take the op-by-op path.
#pragma once
#include "bdq_common.cuh"

namespace bdq {

template <int DT>
struct P2;

template <>
struct P2<kBF16> {
    using h2 = __nv_bfloat162;
    static __device__ __forceinline__ h2 from_bits(uint32_t b) { return *reinterpret_cast<h2*>(&b); }
    static __device__ __forceinline__ uint32_t bits(h2 v) { return *reinterpret_cast<uint32_t*>(&v); }
    static __device__ __forceinline__ h2 pack(float a, float b) { return __floats2bfloat162_rn(a, b); }
    static __device__ __forceinline__ h2 dup(float a) { return __float2bfloat162_rn(a); }
    static __device__ __forceinline__ h2 dup_lo(h2 v) { return __low2bfloat162(v); }
    static __device__ __forceinline__ h2 dup_hi(h2 v) { return __high2bfloat162(v); }
    static __device__ __forceinline__ float lo(h2 v) { return __uint_as_float(bits(v) << 16); }
    static __device__ __forceinline__ float hi(h2 v) { return __uint_as_float(bits(v) & 0xffff0000u); }
    static constexpr float kRoundC = 0.5f + 0.001953125f;  // 0.5 + 2^-9
};

template <>
struct P2<kF16> {
    using h2 = __half2;
    static __device__ __forceinline__ h2 from_bits(uint32_t b) { return *reinterpret_cast<h2*>(&b); }
    static __device__ __forceinline__ uint32_t bits(h2 v) { return *reinterpret_cast<uint32_t*>(&v); }
    static __device__ __forceinline__ h2 pack(float a, float b) { return __floats2half2_rn(a, b); }
    static __device__ __forceinline__ h2 dup(float a) { return __float2half2_rn(a); }
    static __device__ __forceinline__ h2 dup_lo(h2 v) { return __low2half2(v); }
    static __device__ __forceinline__ h2 dup_hi(h2 v) { return __high2half2(v); }
    static __device__ __forceinline__ float lo(h2 v) { return __low2float(v); }
    static __device__ __forceinline__ float hi(h2 v) { return __high2float(v); }
    static constexpr float kRoundC = 0.5f + 0.000244140625f;  // 0.5 + 2^-12
};

// max / min over the 16 packed values of a 32-byte vector
template <int DT>
__device__ __forceinline__ void vec_minmax16(const U8& raw, float& mx, float& mn) {
    using P = P2<DT>;
    typename P::h2 a = P::from_bits(raw.r[0]), b = a;
#pragma unroll
    for (int i = 1; i < 8; ++i) {
        a = __hmax2(a, P::from_bits(raw.r[i]));
        b = __hmin2(b, P::from_bits(raw.r[i]));
    }
    mx = fmaxf(P::lo(a), P::hi(a));
    mn = fminf(P::lo(b), P::hi(b));
}

// quotient x / s rounded to fp32 (exact, see div_exact) without the range guard: the caller has checked the group
__device__ __forceinline__ float div_fast(float x, float s, float y, float& q0) {
    q0 = x * y;
    float r = __fmaf_rn(-s, q0, x);
    return __fmaf_rn(r, y, q0);
}

// integer rounding of a T-valued u (held in fp32): RM 0 = half-even (torch.round), 1 = Round.forward
template <int DT, int RM>
__device__ __forceinline__ float round_T(float u) {
    if constexpr (RM == BDQ_ROUND_HALF_AWAY) return copysignf(floorf(fabsf(u) + P2<DT>::kRoundC), u);
    else return rintf(u);
}

}  // namespace bdq
