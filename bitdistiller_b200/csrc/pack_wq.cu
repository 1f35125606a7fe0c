// Real quantise + pack (sm_100a): WQLinear.from_linear, quantization/qmodule.py:70-146, as two fused kernels.
//
// pack_codes_kernel : one thread per output int32 word.  It reads the `pack` (16 for int2, 8 for int4) fake-quantised
//                     weights that land in the word with vector loads, recomputes the integer code exactly like the
//                     reference (round((w + z*s) / fp16(s)), every op rounded to its eager result dtype) and writes the
//                     word once.  HBM traffic = sizeof(T) per weight read + w_bit/8 B per weight written.
// pack_meta_kernel  : one thread per qzeros word / per 16 padded scales.
#include "bdq_common.cuh"

namespace bdq {

// promote(T, fp16): fp16/fp16 stays fp16, everything else computes and stays in fp32 (qmodule.py:96)
template <int DT>
__device__ __forceinline__ float rP(float x) {
    if constexpr (DT == kF16) return __half2float(__float2half_rn(x));
    else return x;
}

template <int DT>
__device__ __forceinline__ uint32_t code_of(float w, float s, float z, float s16) {
    const float sz = rT<DT>(z * s);            // scale_zeros = zeros * scales        (qmodule.py:79)
    const float a = rT<DT>(w + sz);            // weight + scale_zeros                (qmodule.py:96)
    const float q = rP<DT>(__fdiv_rn(a, s16)); // / awq_linear.scales (fp16)
    return (uint32_t)(int32_t)rintf(q);        // torch.round(...).to(torch.int)
}

template <int DT, int WBIT>
__global__ void __launch_bounds__(256) pack_codes_kernel(const uint8_t* __restrict__ w, const void* __restrict__ scales,
                                                         const void* __restrict__ zeros, uint32_t* __restrict__ qweight,
                                                         int64_t nwords, int64_t words_per_row, int64_t ng_per_row,
                                                         int group, bool vec_ok) {
    constexpr int PACK = 32 / WBIT;
    constexpr int EB = ElemBytes<DT>::v;
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= nwords) return;
    const int64_t r = i / words_per_row, c = i - r * words_per_row;
    const int64_t g = r * ng_per_row + (c * PACK) / group;
    const float s = load_scalar<DT>(scales, g), z = load_scalar<DT>(zeros, g);
    const float s16 = __half2float(__float2half_rn(s));
    float f[PACK];
    const uint8_t* src = w + i * (int64_t)(PACK * EB);
    if (vec_ok) {
        if constexpr (PACK * EB == 32) {
            U8 raw = ldg256(src);
            unpack<DT>(raw, f);
        } else if constexpr (PACK * EB == 64) {
            U8 r0 = ldg256(src), r1 = ldg256(src + 32);
            unpack<DT>(r0, f);
            unpack<DT>(r1, f + 8);
        } else {  // 16 bytes: 8 x 16-bit
            const uint4 v = __ldg(reinterpret_cast<const uint4*>(src));
            const uint32_t rr[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                if constexpr (DT == kBF16) {
                    f[2 * k] = __uint_as_float(rr[k] << 16);
                    f[2 * k + 1] = __uint_as_float(rr[k] & 0xffff0000u);
                } else {
                    float2 t = __half22float2(*reinterpret_cast<const __half2*>(&rr[k]));
                    f[2 * k] = t.x;
                    f[2 * k + 1] = t.y;
                }
            }
        }
    } else {
#pragma unroll
        for (int k = 0; k < PACK; ++k) f[k] = load_scalar<DT>(w, i * PACK + k);
    }
    uint32_t word = 0;
#pragma unroll
    for (int k = 0; k < PACK; ++k) word |= code_of<DT>(f[k], s, z, s16) << (k * WBIT);
    qweight[i] = word;
}

// one thread per (row, zeros word): packs up to PACK zeros and writes the PACK matching fp16 scales (zero padded)
template <int DT, int WBIT>
__global__ void __launch_bounds__(256) pack_meta_kernel(const void* __restrict__ scales, const void* __restrict__ zeros,
                                                        uint32_t* __restrict__ qzeros, __half* __restrict__ qscales,
                                                        int64_t rows, int64_t ng_per_row, int zw) {
    constexpr int PACK = 32 / WBIT;
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= rows * zw) return;
    const int64_t r = i / zw, c = i - r * zw;
    uint32_t word = 0;
#pragma unroll
    for (int k = 0; k < PACK; ++k) {
        const int64_t g = c * PACK + k;
        float s = 0.f;
        if (g < ng_per_row) {
            s = load_scalar<DT>(scales, r * ng_per_row + g);
            word |= ((uint32_t)(int32_t)load_scalar<DT>(zeros, r * ng_per_row + g)) << (k * WBIT);
        }
        qscales[(r * zw + c) * PACK + k] = __float2half_rn(s);
    }
    qzeros[i] = word;
}

template <int DT, int WBIT>
static int launch_pack(const void* w, const void* scales, const void* zeros, int32_t* qweight, int32_t* qzeros, void* qscales,
                       int64_t rows, int64_t cols, int group, int zw, cudaStream_t st) {
    constexpr int PACK = 32 / WBIT;
    const int64_t wpr = cols / PACK, nwords = rows * wpr, ngr = cols / group;
    const bool vec_ok = ((uintptr_t)w % 32 == 0);
    pack_codes_kernel<DT, WBIT><<<(unsigned)((nwords + 255) / 256), 256, 0, st>>>(
        (const uint8_t*)w, scales, zeros, (uint32_t*)qweight, nwords, wpr, ngr, group, vec_ok);
    int rc = check_launch("bdq_pack_wq(codes)");
    if (rc) return rc;
    pack_meta_kernel<DT, WBIT><<<(unsigned)((rows * zw + 255) / 256), 256, 0, st>>>(scales, zeros, (uint32_t*)qzeros,
                                                                                   (__half*)qscales, rows, ngr, zw);
    return check_launch("bdq_pack_wq(meta)");
}

}  // namespace bdq

using namespace bdq;

extern "C" int bdq_zeros_width(int64_t in_features, int group, int w_bit) {
    if (w_bit <= 0 || 32 % w_bit != 0 || group <= 0) return -1;
    const int pack = 32 / w_bit;
    const int mult = group >= 128 ? 1 : (group == 64 ? 2 : (group == 32 ? 4 : -1));
    if (mult < 0) return -1;
    int64_t bw = ((in_features / group) + pack - 1) / pack;
    bw = ((bw + mult - 1) / mult) * mult;
    return (int)bw;
}

extern "C" int bdq_pack_wq(const void* w_dq, const void* scales, const void* zeros, int32_t* qweight, int32_t* qzeros,
                           void* qscales, int64_t rows, int64_t cols, int group, int w_bit, int dtype,
                           bdq_stream_t stream) {
    if (w_bit != 2 && w_bit != 4) { set_error("w_bit=%d: only 4,2-bit are supported", w_bit); return BDQ_ERR_UNSUPPORTED; }
    if (dtype < 0 || dtype > 2) { set_error("dtype=%d unknown", dtype); return BDQ_ERR_DTYPE; }
    const int pack = 32 / w_bit;
    if (rows < 0 || cols < 0 || group <= 0 || cols % group != 0 || cols % pack != 0 || group % pack != 0) {
        set_error("pack: rows=%lld cols=%lld group=%d w_bit=%d: cols must be a multiple of group and of %d", (long long)rows,
                  (long long)cols, group, w_bit, pack);
        return BDQ_ERR_SHAPE;
    }
    const int zw = bdq_zeros_width(cols, group, w_bit);
    if (zw < 0) { set_error("group=%d unsupported by calculate_zeros_width", group); return BDQ_ERR_UNSUPPORTED; }
    if (rows == 0 || cols == 0) return BDQ_OK;
    if (!w_dq || !scales || !zeros || !qweight || !qzeros || !qscales) { set_error("null tensor pointer"); return BDQ_ERR_SHAPE; }
    cudaStream_t st = (cudaStream_t)stream;
#define BDQ_PACK(DT) \
    return w_bit == 2 ? launch_pack<DT, 2>(w_dq, scales, zeros, qweight, qzeros, qscales, rows, cols, group, zw, st) \
                      : launch_pack<DT, 4>(w_dq, scales, zeros, qweight, qzeros, qscales, rows, cols, group, zw, st)
    if (dtype == BDQ_F32) { BDQ_PACK(kF32); }
    if (dtype == BDQ_BF16) { BDQ_PACK(kBF16); }
    BDQ_PACK(kF16);
#undef BDQ_PACK
}
