// GPTQ-v2 2-bit export (sm_100a): train/convert_hf_to_gptq_manual.py:11-75 (get_hf_quant_params) and 85-157
// (convert_layer_to_gptq_format) as ONE pass over the weight instead of the reference's numpy round trip and
// 2 x 16 strided eager kernels per linear.
//
// One CTA owns a tile of 32 output channels x 512 input columns.  Each lane reads 16 consecutive weights of one row
// (one 32-byte vector for 16-bit weights), the G/16 lanes of a group reduce min/max with shuffles, every lane forms
// its 16 codes = one K-major int32 word, and the tile is transposed through shared memory so that qweight
// [IC/16, OC], scales [IC/G, OC] and qzeros [IC/G, OC/16] are all written as contiguous 32-channel runs.
// HBM traffic = sizeof(T) per weight read + 0.25 B per weight written (+ group metadata).
//
// Arithmetic is the reference's fp32 CPU arithmetic (the converter upcasts the weight and runs on the host):
// s = clamp(max - min, 1e-5) / 3 (IEEE division), z = clamp(-rne(min / s), 0, 3), code = clamp(rne(w / s) + z, 0, 3).
#include "fakequant_thr.cuh"
#include "fq_int_elem.cuh"

namespace bdq {

constexpr int kGqTileO = 32;    // output channels per CTA
constexpr int kGqTileK = 512;   // input columns per CTA = 32 lanes x 16
constexpr int kGqWarps = 8;

// 16 consecutive elements of one row as floats (exact widening)
template <int DT>
__device__ __forceinline__ void load16(const void* w, int64_t idx, bool vec_ok, float* f) {
    constexpr int EB = ElemBytes<DT>::v;
    const uint8_t* src = reinterpret_cast<const uint8_t*>(w) + idx * EB;
    if (vec_ok) {
        U8 r0 = ldg256(src);
        unpack<DT>(r0, f);
        if constexpr (DT == kF32) {
            U8 r1 = ldg256(src + 32);
            unpack<DT>(r1, f + 8);
        }
    } else {
#pragma unroll
        for (int i = 0; i < 16; ++i) f[i] = load_scalar<DT>(w, idx + i);
    }
}

// ---- threshold form for 16-bit weights (the converter upcasts them and computes in fp32): for one group the code
// c(x) = clamp(rne(x / s) + z, 0, 3) is a monotone step function of the T value x, so it is described by the three
// smallest T values T_k with rne(fdiv(x, s)) >= k - z.  Each is found by evaluating that exact fp32 predicate on the 5 T
// values around rT((k - z - 0.5) s) and accepted only if it flips inside the window (else the group goes op by op).
template <int DT>
__device__ __forceinline__ float gptq_threshold(float n, float s, bool& ok) {
    const float xc = rT<DT>((n - 0.5f) * s);
    uint32_t pm = 0;
#pragma unroll
    for (int j = 0; j < 5; ++j) pm |= (rintf(__fdiv_rn(t_step<DT>(xc, j - 2), s)) >= n ? 1u : 0u) << j;
    const int i = __ffs(pm) - 1;
    ok = (i >= 1) && (pm == ((0x1fu << i) & 0x1fu));
    return t_step<DT>(xc, i - 2);
}

template <int DT>
struct GqMagic;
template <>
struct GqMagic<kBF16> { static constexpr float v = 128.f; static constexpr uint32_t bits2 = 0x43004300u; };
template <>
struct GqMagic<kF16> { static constexpr float v = 1024.f; static constexpr uint32_t bits2 = 0x64006400u; };

// 8 two-bit fields at 2-bit pitch (16 bits) -> 4-bit pitch
__device__ __forceinline__ uint32_t spread2(uint32_t x) {
    x = (x | (x << 8)) & 0x00FF00FFu;
    x = (x | (x << 4)) & 0x0F0F0F0Fu;
    x = (x | (x << 2)) & 0x33333333u;
    return x;
}

// MODE 0: quantise `w` (T) here.  MODE 1: `w` holds uint8 codes [OC, IC] and `zin` uint8 zeros [OC, IC/G]
// (the two-step API, convert_layer_to_gptq_format on already unpacked parameters); scales are not touched.
template <int DT, int MODE>
__global__ void __launch_bounds__(kGqWarps * 32) gptq_pack_w2_kernel(const void* __restrict__ w,
                                                                     const uint8_t* __restrict__ zin,
                                                                     uint32_t* __restrict__ qweight,
                                                                     uint32_t* __restrict__ qzeros,
                                                                     __nv_bfloat16* __restrict__ scales, int64_t OC,
                                                                     int64_t IC, int group, bool vec_ok) {
    __shared__ uint32_t s_word[32][kGqTileO + 1];   // [k word][channel]
    __shared__ float s_scale[32][kGqTileO + 1];     // [group in tile][channel]
    __shared__ uint8_t s_zero[32][kGqTileO];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int64_t k0 = (int64_t)blockIdx.x * kGqTileK, o0 = (int64_t)blockIdx.y * kGqTileO;
    const int lpg = group >> 4;                      // lanes per group (power of two, 1..32)
    const int64_t k = k0 + 16 * lane;
    const bool k_ok = k < IC;
    const int lg = 31 - __clz(group);             // group is a power of two (checked by the host): shifts, not divisions
    const int64_t ng = IC >> lg;

#pragma unroll 1
    for (int j = 0; j < kGqTileO / kGqWarps; ++j) {
        const int ol = warp * (kGqTileO / kGqWarps) + j;
        const int64_t o = o0 + ol;
        const bool ok = k_ok && o < OC;              // warp-uniform in o; whole groups are in or out in k
        uint32_t word = 0;
        if constexpr (MODE == 0) {
            IntQ p;
            bool done = false;
            if constexpr (DT != kF32) {
                if (vec_ok) {                        // warp-uniform
                    using P = P2<DT>;
                    using h2 = typename P::h2;
                    U8 raw;
                    if (ok) raw = ldg256(reinterpret_cast<const uint8_t*>(w) + (o * IC + k) * 2);
                    else {
#pragma unroll
                        for (int i = 0; i < 8; ++i) raw.r[i] = 0u;
                    }
                    float mx, mn;
                    vec_minmax16<DT>(raw, mx, mn);
                    for (int d = lpg >> 1; d > 0; d >>= 1) {
                        mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, d));
                        mn = fminf(mn, __shfl_xor_sync(0xffffffffu, mn, d));
                    }
                    p = int_group_params<kF32>(mx, mn, 3.f, BDQ_SEM_CPU);
                    const int lig = lane & (lpg - 1), gbase = lane & ~(lpg - 1);
                    // every lane evaluates ONE threshold per round (k = lig + r * lpg): no divergence between the lanes of a
                    // group; one round when the group spans >= 4 lanes (G >= 64)
                    float tl0 = 0.f, tl1 = 0.f, tl2 = 0.f;
                    bool tok = fmaxf(fabsf(mx), fabsf(mn)) < 1e30f && p.s > 1e-30f;
                    const int rounds = lpg >= 4 ? 1 : (lpg == 2 ? 2 : 3);
#pragma unroll 1
                    for (int r = 0; r < rounds; ++r) {
                        const int kk = lig + r * lpg;
                        bool o1 = true;
                        float tv = 0.f;
                        if (kk < 3) tv = gptq_threshold<DT>((float)(kk + 1) - p.z, p.s, o1);
                        tok = tok && o1;
                        if (r == 0) tl0 = tv; else if (r == 1) tl1 = tv; else tl2 = tv;
                    }
                    const uint32_t okm = __ballot_sync(0xffffffffu, tok);
                    const uint32_t gm = (lpg == 32 ? 0xffffffffu : ((1u << lpg) - 1u)) << gbase;
                    h2 t2[3];
#pragma unroll
                    for (int kk = 0; kk < 3; ++kk) {
                        const int r = lpg >= 4 ? 0 : (kk >> (lg - 4));
                        const float src = r == 0 ? tl0 : (r == 1 ? tl1 : tl2);
                        t2[kk] = P::dup(__shfl_sync(0xffffffffu, src, gbase + (kk & (lpg - 1))));
                    }
                    if ((okm & gm) == gm) {
                        const h2 magic = P::dup(GqMagic<DT>::v);
                        uint32_t acc = 0u;
#pragma unroll
                        for (int i = 7; i >= 0; --i) {
                            const h2 x2 = P::from_bits(raw.r[i]);
                            h2 c2 = __hadd2(__hge2(x2, t2[0]), magic);
                            c2 = __hadd2(c2, __hge2(x2, t2[1]));
                            c2 = __hadd2(c2, __hge2(x2, t2[2]));
                            acc = acc * 4u + P::bits(c2);
                        }
                        acc -= GqMagic<DT>::bits2 * 21845u;              // even elements in the low half, odd in the high half
                        word = spread2(acc & 0xFFFFu) | (spread2(acc >> 16) << 2);
                    } else {
                        float f[16];
                        unpack<DT>(raw, f);
#pragma unroll 1
                        for (int i = 0; i < 16; ++i) {
                            float qu;
                            const float c = quant_code<kF32, BDQ_ROUND_HALF_EVEN>(f[i], p, 3.f, qu);
                            word |= (uint32_t)(int)c << (2 * i);
                        }
                    }
                    done = true;
                }
            }
            if (!done) {
                float f[16];
                if (ok) load16<DT>(w, o * IC + k, vec_ok, f);
                else {
#pragma unroll
                    for (int i = 0; i < 16; ++i) f[i] = 0.f;
                }
                float mx = f[0], mn = f[0];
#pragma unroll
                for (int i = 1; i < 16; ++i) { mx = fmaxf(mx, f[i]); mn = fminf(mn, f[i]); }
                for (int d = lpg >> 1; d > 0; d >>= 1) {
                    mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, d));
                    mn = fminf(mn, __shfl_xor_sync(0xffffffffu, mn, d));
                }
                p = int_group_params<kF32>(mx, mn, 3.f, BDQ_SEM_CPU);
#pragma unroll
                for (int i = 0; i < 16; ++i) {
                    float qu;
                    const float c = quant_code<kF32, BDQ_ROUND_HALF_EVEN>(f[i], p, 3.f, qu);
                    word |= (uint32_t)(int)c << (2 * i);
                }
            }
            if (ok && (lane & (lpg - 1)) == 0) {
                const int gl = lane >> (lg - 4);
                // scales leave get_hf_quant_params in the weight's dtype (line 64) and are then cast to bf16 (line 100)
                s_scale[gl][ol] = rT<DT>(p.s);
                s_zero[gl][ol] = (uint8_t)(int)p.z;
            }
        } else {
            if (ok) {
                const uint8_t* src = reinterpret_cast<const uint8_t*>(w) + o * IC + k;
                uint4 v;
                if (vec_ok) v = __ldg(reinterpret_cast<const uint4*>(src));
                else {
                    uint32_t t[4] = {0, 0, 0, 0};
                    for (int i = 0; i < 16; ++i) t[i >> 2] |= (uint32_t)src[i] << (8 * (i & 3));
                    v = make_uint4(t[0], t[1], t[2], t[3]);
                }
                const uint32_t t[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
                for (int i = 0; i < 16; ++i) word |= ((t[i >> 2] >> (8 * (i & 3))) & 3u) << (2 * i);
                if ((lane & (lpg - 1)) == 0) s_zero[lane >> (lg - 4)][ol] = zin[o * ng + (k >> lg)] & 3u;
            }
        }
        s_word[lane][ol] = word;
    }
    __syncthreads();

    // qweight [IC/16, OC]: row r of the tile = 32 consecutive channels
    const int64_t r0 = k0 >> 4, nrow = IC >> 4;
    for (int r = warp; r < 32; r += kGqWarps) {
        if (r0 + r < nrow && o0 + lane < OC) qweight[(r0 + r) * OC + o0 + lane] = s_word[r][lane];
    }
    // scales [IC/G, OC] bf16 and qzeros [IC/G, OC/16]
    const int gpt = kGqTileK >> lg;                 // groups per tile
    const int64_t g0 = k0 >> lg;
    for (int g = warp; g < gpt; g += kGqWarps) {
        if (g0 + g >= ng) break;
        const bool o_ok = o0 + lane < OC;
        if constexpr (MODE == 0) {
            if (o_ok) scales[(g0 + g) * OC + o0 + lane] = __float2bfloat16_rn(s_scale[g][lane]);
        }
        uint32_t z = o_ok ? (uint32_t)s_zero[g][lane] << (2 * (lane & 15)) : 0u;
#pragma unroll
        for (int d = 8; d > 0; d >>= 1) z |= __shfl_xor_sync(0xffffffffu, z, d);
        if ((lane & 15) == 0 && o_ok) qzeros[(g0 + g) * (OC >> 4) + ((o0 + lane) >> 4)] = z;
    }
}

}  // namespace bdq

using namespace bdq;

static int gptq_check(int64_t OC, int64_t IC, int group, int w_bit) {
    if (w_bit != 2) { set_error("gptq export: bits=%d, this conversion is designed for bits=2", w_bit); return BDQ_ERR_UNSUPPORTED; }
    if (OC < 0 || IC < 0) { set_error("gptq export: negative shape"); return BDQ_ERR_SHAPE; }
    if (group < 16 || group > kGqTileK || (group & (group - 1)) != 0) {
        set_error("gptq export: group=%d unsupported (power of two in [16, %d])", group, kGqTileK);
        return BDQ_ERR_UNSUPPORTED;
    }
    if (IC % group != 0) { set_error("gptq export: in_features=%lld not a multiple of group=%d", (long long)IC, group); return BDQ_ERR_SHAPE; }
    if (OC % 16 != 0) {
        set_error("gptq export: out_features (%lld) must be divisible by pack_factor (16) for zero points", (long long)OC);
        return BDQ_ERR_SHAPE;
    }
    return BDQ_OK;
}

extern "C" int bdq_gptq_export_w2(const void* w, int32_t* qweight, int32_t* qzeros, void* scales_bf16, int64_t OC,
                                  int64_t IC, int group, int w_bit, int dtype, bdq_stream_t stream) {
    if (dtype < 0 || dtype > 2) { set_error("dtype=%d unknown", dtype); return BDQ_ERR_DTYPE; }
    int rc = gptq_check(OC, IC, group, w_bit);
    if (rc) return rc;
    if (OC == 0 || IC == 0) return BDQ_OK;
    if (!w || !qweight || !qzeros || !scales_bf16) { set_error("null tensor pointer"); return BDQ_ERR_SHAPE; }
    const dim3 grid((unsigned)((IC + kGqTileK - 1) / kGqTileK), (unsigned)((OC + kGqTileO - 1) / kGqTileO));
    const bool vec_ok = ((uintptr_t)w % 32 == 0);   // IC % 16 == 0 keeps every 16-element run aligned with the base
    cudaStream_t st = (cudaStream_t)stream;
#define BDQ_GQ(DT)                                                                                                    \
    gptq_pack_w2_kernel<DT, 0><<<grid, kGqWarps * 32, 0, st>>>(w, nullptr, (uint32_t*)qweight, (uint32_t*)qzeros,     \
                                                               (__nv_bfloat16*)scales_bf16, OC, IC, group, vec_ok)
    if (dtype == BDQ_F32) BDQ_GQ(kF32);
    else if (dtype == BDQ_BF16) BDQ_GQ(kBF16);
    else BDQ_GQ(kF16);
#undef BDQ_GQ
    return check_launch("bdq_gptq_export_w2");
}

extern "C" int bdq_gptq_pack_w2(const uint8_t* codes, const uint8_t* zeros, int32_t* qweight, int32_t* qzeros, int64_t OC,
                                int64_t IC, int group, int w_bit, bdq_stream_t stream) {
    int rc = gptq_check(OC, IC, group, w_bit);
    if (rc) return rc;
    if (OC == 0 || IC == 0) return BDQ_OK;
    if (!codes || !zeros || !qweight || !qzeros) { set_error("null tensor pointer"); return BDQ_ERR_SHAPE; }
    const dim3 grid((unsigned)((IC + kGqTileK - 1) / kGqTileK), (unsigned)((OC + kGqTileO - 1) / kGqTileO));
    const bool vec_ok = ((uintptr_t)codes % 16 == 0);
    gptq_pack_w2_kernel<kF32, 1><<<grid, kGqWarps * 32, 0, (cudaStream_t)stream>>>(
        codes, zeros, (uint32_t*)qweight, (uint32_t*)qzeros, nullptr, OC, IC, group, vec_ok);
    return check_launch("bdq_gptq_pack_w2");
}
