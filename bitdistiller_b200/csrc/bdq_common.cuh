// Shared device helpers for the bdq kernels (sm_100a only).
#pragma once
#include <cstdlib>
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/bdq.h"

namespace bdq {

constexpr int kF32 = BDQ_F32, kBF16 = BDQ_BF16, kF16 = BDQ_F16;

void set_error(const char* fmt, ...);
int check_launch(const char* what);
int sm_count();
bool ensure_dyn_smem(const void* fn, int bytes);   // per (device, kernel): see runtime.cu

// Grid of a persistent, warp-strided streaming kernel: a whole number of waves (resident CTAs per SM x SMs x waves), so
// that no partial last wave leaves SMs idle.  Several short waves balance better than one long one (the SMs do not all
// see the same HBM latency); measured optimum on B200: 16 waves for the int kernels, 6 for the NF3 ones (heavier
// per-CTA prologue).  BDQ_FQ_WAVES overrides.
template <typename Kern>
static int wave_grid(Kern kern, int threads, int64_t blocks_needed, int default_waves) {
    // occupancy per kernel address, cached (all instantiations with one signature share this template instance); a benign
    // race at worst repeats the query
    static const void* seen[64];
    static int seen_occ[64];
    static int nseen = 0;
    int occ = 0;
    const void* key = reinterpret_cast<const void*>(kern);
    for (int i = 0; i < nseen && i < 64; ++i)
        if (seen[i] == key) { occ = seen_occ[i]; break; }
    if (occ < 1) {
        if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kern, threads, 0) != cudaSuccess || occ < 1) occ = 4;
        const int slot = nseen;
        if (slot < 64) { seen[slot] = key; seen_occ[slot] = occ; nseen = slot + 1; }
    }
    static int env_waves = -1;
    if (env_waves < 0) {
        const char* e = getenv("BDQ_FQ_WAVES");
        env_waves = (e && atoi(e) > 0) ? atoi(e) : 0;
    }
    int64_t grid = (int64_t)sm_count() * occ * (env_waves ? env_waves : default_waves);
    if (grid > blocks_needed) grid = blocks_needed;
    if (grid < 1) grid = 1;
    return (int)grid;
}

// ---- rounding of an fp32 "opmath" result to the logical tensor dtype (what eager PyTorch does between ops)
template <int DT>
__device__ __forceinline__ float rT(float x) {
    if constexpr (DT == kBF16) {
        // F2FP.BF16.F32.PACK_AB (ALU pipe) + mask instead of the scalar F2F (XU pipe)
        __nv_bfloat162 t = __floats2bfloat162_rn(0.f, x);
        return __uint_as_float(*reinterpret_cast<uint32_t*>(&t) & 0xffff0000u);
    } else if constexpr (DT == kF16) {
        return __half2float(__float2half_rn(x));
    } else {
        return x;
    }
}

template <int DT>
struct ElemBytes { static constexpr int v = (DT == kF32) ? 4 : 2; };

// ---- 256-bit global access (LDG.E.256 / STG.E.256 on sm_100a), streaming (no L1 allocate)
struct alignas(32) U8 { uint32_t r[8]; };

__device__ __forceinline__ U8 ldg256(const void* p) {
    U8 v;
    asm volatile("ld.global.nc.L1::no_allocate.v8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=r"(v.r[0]), "=r"(v.r[1]), "=r"(v.r[2]), "=r"(v.r[3]), "=r"(v.r[4]), "=r"(v.r[5]), "=r"(v.r[6]), "=r"(v.r[7])
                 : "l"(p));
    return v;
}
// coherent variant for kernels that rewrite the memory they read (in-place apply_clip): .nc promises read-only data
__device__ __forceinline__ U8 ld256(const void* p) {
    U8 v;
    asm volatile("ld.global.L1::no_allocate.v8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=r"(v.r[0]), "=r"(v.r[1]), "=r"(v.r[2]), "=r"(v.r[3]), "=r"(v.r[4]), "=r"(v.r[5]), "=r"(v.r[6]), "=r"(v.r[7])
                 : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void stg256(void* p, const U8& v) {
    asm volatile("st.global.L1::no_allocate.v8.b32 [%8], {%0,%1,%2,%3,%4,%5,%6,%7};"
                 :: "r"(v.r[0]), "r"(v.r[1]), "r"(v.r[2]), "r"(v.r[3]), "r"(v.r[4]), "r"(v.r[5]), "r"(v.r[6]), "r"(v.r[7]), "l"(p)
                 : "memory");
}

// elements per 32-byte vector
template <int DT>
struct VecN { static constexpr int v = 32 / ElemBytes<DT>::v; };

// unpack a 32-byte vector to floats (exact)
template <int DT>
__device__ __forceinline__ void unpack(const U8& raw, float* f) {
    if constexpr (DT == kF32) {
#pragma unroll
        for (int i = 0; i < 8; ++i) f[i] = __uint_as_float(raw.r[i]);
    } else if constexpr (DT == kBF16) {
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            f[2 * i] = __uint_as_float(raw.r[i] << 16);
            f[2 * i + 1] = __uint_as_float(raw.r[i] & 0xffff0000u);
        }
    } else {
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            float2 t = __half22float2(*reinterpret_cast<const __half2*>(&raw.r[i]));
            f[2 * i] = t.x;
            f[2 * i + 1] = t.y;
        }
    }
}

// pack floats to a 32-byte vector with round-to-nearest-even to DT
template <int DT>
__device__ __forceinline__ void pack(const float* f, U8& raw) {
    if constexpr (DT == kF32) {
#pragma unroll
        for (int i = 0; i < 8; ++i) raw.r[i] = __float_as_uint(f[i]);
    } else if constexpr (DT == kBF16) {
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            __nv_bfloat162 t = __floats2bfloat162_rn(f[2 * i], f[2 * i + 1]);
            raw.r[i] = *reinterpret_cast<uint32_t*>(&t);
        }
    } else {
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            __half2 t = __floats2half2_rn(f[2 * i], f[2 * i + 1]);
            raw.r[i] = *reinterpret_cast<uint32_t*>(&t);
        }
    }
}

template <int DT>
__device__ __forceinline__ float load_scalar(const void* p, int64_t i) {
    if constexpr (DT == kF32) return reinterpret_cast<const float*>(p)[i];
    else if constexpr (DT == kBF16) return __bfloat162float(reinterpret_cast<const __nv_bfloat16*>(p)[i]);
    else return __half2float(reinterpret_cast<const __half*>(p)[i]);
}
template <int DT>
__device__ __forceinline__ void store_scalar(void* p, int64_t i, float v) {
    if constexpr (DT == kF32) reinterpret_cast<float*>(p)[i] = v;
    else if constexpr (DT == kBF16) reinterpret_cast<__nv_bfloat16*>(p)[i] = __float2bfloat16_rn(v);
    else reinterpret_cast<__half*>(p)[i] = __float2half_rn(v);
}

__device__ __forceinline__ float rcp_approx(float x) {
    float y;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}

// ---- exact division a / b (IEEE RN in fp32).
// For 16-bit dtypes a and b carry <= 11 significant bits, so the true quotient is >= 2^-12 ulp away from any
// fp32 rounding boundary and one residual correction with y ~= 1/b (MUFU.RCP, <= 1 ulp) is exact:
//   q0 = a*y ; r = fma(-b, q0, a) (exact) ; q = RN(q0 + r*y)  -- the value q0 + r*y is within 2^-21 ulp of a/b.
// Falls back to __fdiv_rn when the quotient may overflow.
template <int DT>
__device__ __forceinline__ float div_exact(float a, float b, float y) {
    if constexpr (DT == kF32) {
        return __fdiv_rn(a, b);
    } else {
        float q0 = a * y;
        float r = __fmaf_rn(-b, q0, a);
        float q = __fmaf_rn(r, y, q0);
        if (!(fabsf(q0) < 1e30f)) q = __fdiv_rn(a, b);  // overflow range; tiny quotients round to 0 either way
        return q;
    }
}

// group reductions over LPG consecutive lanes (LPG power of two <= 32)
template <int LPG>
__device__ __forceinline__ float group_max(float v) {
#pragma unroll
    for (int o = LPG / 2; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}
template <int LPG>
__device__ __forceinline__ float group_min(float v) {
#pragma unroll
    for (int o = LPG / 2; o > 0; o >>= 1) v = fminf(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}
template <int LPG>
__device__ __forceinline__ float group_sum(float v) {
#pragma unroll
    for (int o = LPG / 2; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
template <int LPG>
__device__ __forceinline__ int group_sum_i(int v) {
#pragma unroll
    for (int o = LPG / 2; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

// Per-group scalars of the asymmetric integer quantiser (quantizer.py:38-43), op-by-op in DT.
struct IntQ {
    float s;     // scale (value in DT)
    float y;     // RN(1/s)
    float z;     // zero point, integer valued in [0, qmax]
    float t1;    // rT(max - min)
};

template <int DT>
__device__ __forceinline__ IntQ int_group_params(float mx, float mn, float qmax, int sem) {
    IntQ p;
    p.t1 = rT<DT>(mx - mn);
    float t2 = rT<DT>(fmaxf(p.t1, 1e-5f));
    // tensor / python scalar: CPU eager divides, CUDA eager multiplies by the fp32 reciprocal
    const float inv_q = 1.0f / qmax;  // correctly rounded (IEEE division, loop invariant)
    float sraw;
    if (sem) sraw = t2 * inv_q;
    else if constexpr (DT == kF32) sraw = __fdiv_rn(t2, qmax);
    else sraw = div_exact<DT>(t2, qmax, inv_q);
    p.s = rT<DT>(sraw);
    if constexpr (DT == kF32) p.y = __frcp_rn(p.s);
    else p.y = rcp_approx(p.s);
    float t3 = rT<DT>(div_exact<DT>(mn, p.s, p.y));
    float z = fminf(fmaxf(-rintf(t3), 0.f), qmax);
    p.z = z + 0.f;  // -0 -> +0
    return p;
}

}  // namespace bdq
