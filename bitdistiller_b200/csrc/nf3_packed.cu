// Packed NF3 (3 bits per weight + two fp16 scales per group): pack, dequantise, small-shape decode GEMV.
// Format and its rationale: nf3_packed.cuh (0.375 B + 4/G B per weight = 0.40625 B/weight at G = 128).
// The bandwidth-bound decode GEMV (group 128, OC % 16 == 0) is gemv_nf3_stream.cu; the pair-lookup kernel below serves every
// other shape.
#include <cuda_fp16.h>

#include <algorithm>
#include <cstdlib>

#include "nf3_packed.cuh"

namespace bdq {

// codes as written by bdq_nf3_fakequant_fwd (fq_nf3_elem.cuh: 0 = zero, 1..4 = pos_lvl[1..4], 5..7 = neg_lvl[1..3]) -> e
__device__ __forceinline__ uint32_t nf3_code_to_e(uint32_t c) { return c == 0u ? 3u : (c <= 4u ? c + 3u : 7u - c); }

// one thread per 32 input channels
__global__ void __launch_bounds__(256) nf3_pack_kernel(const uint8_t* __restrict__ codes, uint32_t* __restrict__ qweight,
                                                       uint32_t* __restrict__ qhigh, int64_t nchunk) {
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (int64_t c = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; c < nchunk; c += stride) {
        const uint4 a = reinterpret_cast<const uint4*>(codes)[2 * c], b = reinterpret_cast<const uint4*>(codes)[2 * c + 1];
        const uint32_t in[8] = {a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w};
        uint32_t lo[2] = {0u, 0u}, hi = 0u;
#pragma unroll
        for (int i = 0; i < 32; ++i) {
            const uint32_t code = (in[i >> 2] >> (8 * (i & 3))) & 0xffu;
            const uint32_t e = code > 7u ? 3u : nf3_code_to_e(code);   // 255 (NaN group) packs as zero
            lo[i >> 4] |= (e & 3u) << (2 * (i & 15));
            hi |= (e >> 2) << nf3_hbit_pos(i);
        }
        qweight[2 * c] = lo[0];
        qweight[2 * c + 1] = lo[1];
        qhigh[c] = hi;
    }
}

// one thread per 32 input channels: 32 fp16 values = rT(lvl * s), the forward's own last two operations
__global__ void __launch_bounds__(256) nf3_dequant_kernel(const uint32_t* __restrict__ qweight, const uint32_t* __restrict__ qhigh,
                                                          const __half* __restrict__ scales, __half* __restrict__ out,
                                                          int64_t nchunk, int chunks_per_row, int chunks_per_group, int ngroups,
                                                          const Nf3Levels lv) {
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (int64_t c = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; c < nchunk; c += stride) {
        const int64_t row = c / chunks_per_row;
        const int g = (int)(c - row * chunks_per_row) / chunks_per_group;
        const __half2 s2 = *reinterpret_cast<const __half2*>(scales + (row * ngroups + g) * 2);   // (sp, sn)
        const float sp = __low2float(s2), sn = __high2float(s2);
        const uint32_t lo[2] = {qweight[2 * c], qweight[2 * c + 1]}, hi = qhigh[c];
        uint32_t o[16];
#pragma unroll
        for (int i = 0; i < 32; i += 2) {
            float v[2];
#pragma unroll
            for (int h = 0; h < 2; ++h) {
                const int k = i + h;
                const uint32_t e = ((lo[k >> 4] >> (2 * (k & 15))) & 3u) | (((hi >> nf3_hbit_pos(k)) & 1u) << 2);
                v[h] = lv.lvl[e] * (e < 3u ? sn : sp);   // exact product of two fp16 values; rounded once below
            }
            const __half2 p = __floats2half2_rn(v[0], v[1]);
            o[i >> 1] = *reinterpret_cast<const uint32_t*>(&p);
        }
        uint4* dst = reinterpret_cast<uint4*>(out + c * 32);
#pragma unroll
        for (int q = 0; q < 4; ++q) dst[q] = make_uint4(o[4 * q], o[4 * q + 1], o[4 * q + 2], o[4 * q + 3]);
    }
}

constexpr int kNgThreads = 256;

// Decode GEMV for the shapes the streaming kernel does not take (batch rows handled one at a time): PAIR lookup.  Two adjacent weights are
// 4 bits of the low plane and 2 bits of the high plane: those 6 bits, masked in place at bits 8..13, ARE the byte offset of
// a 64-entry table row (256 B = one 8-byte entry per lane, so a warp's lookup is conflict free whatever the codes are);
// the entry holds the two levels as half2, once for the sp-scaled (zero / positive) side and once for the sn-scaled
// (negative) side, and feeds two HFMA2 against the matching x pair.  Per pair: 2 shifts, 2 LOP3, 1 LDS.64, 2 HFMA2
// (~3.5 instructions per weight instead of ~12).  Half accumulation runs over 8 pairs per accumulator, then fp32.
// A warp walks kNg2Rows output channels per x chunk, so the 64 B of x per lane are loaded once per 4 x 12 B of weights.
constexpr int kNg2Rows = 4;

__device__ __forceinline__ uint2 lds64(uint32_t addr) {
    uint2 v;
    asm volatile("ld.shared.v2.u32 {%0, %1}, [%2];" : "=r"(v.x), "=r"(v.y) : "r"(addr));
    return v;
}

__global__ void __launch_bounds__(kNgThreads) nf3_gemv2_kernel(const __half* __restrict__ x, const uint32_t* __restrict__ qweight,
                                                               const uint32_t* __restrict__ qhigh, const __half* __restrict__ scales,
                                                               __half* __restrict__ out, int B, int IC, int OC, int chunks_per_group,
                                                               const Nf3Levels lv) {
    __shared__ __align__(16) uint8_t lut_raw[32768];          // a 16 KB table at a 16 KB boundary inside
    const int tid = threadIdx.x, lane = tid & 31;
    const uint32_t raw0 = (uint32_t)__cvta_generic_to_shared(lut_raw);
    const uint32_t base = (raw0 + 16383u) & ~16383u;
    uint8_t* lut = lut_raw + (base - raw0);
    for (int idx = tid >> 5; idx < 64; idx += kNgThreads / 32) {
        const int e0 = (idx & 3) | (((idx >> 4) & 1) << 2), e1 = ((idx >> 2) & 3) | (((idx >> 5) & 1) << 2);
        const __half2 pos = __floats2half2_rn(e0 >= 3 ? lv.lvl[e0] : 0.f, e1 >= 3 ? lv.lvl[e1] : 0.f);
        const __half2 neg = __floats2half2_rn(e0 < 3 ? lv.lvl[e0] : 0.f, e1 < 3 ? lv.lvl[e1] : 0.f);
        *reinterpret_cast<uint2*>(lut + idx * 256 + lane * 8) =
            make_uint2(*reinterpret_cast<const uint32_t*>(&pos), *reinterpret_cast<const uint32_t*>(&neg));
    }
    __syncthreads();
    const uint32_t lanebase = base + lane * 8;
    const int cpr = IC >> 5, ngroups = cpr / chunks_per_group;
    const int nrb = (OC + kNg2Rows - 1) / kNg2Rows;
    // split K over the warps of the CTA: KB blocks of 32 chunks, `slices` warps per row block, 8 / slices row blocks per CTA
    // (IC 4096: 4 warps x one 32-chunk block each, 2 row blocks = 8 channels per CTA pass)
    constexpr int kWarps = kNgThreads / 32;
    __shared__ float part[kWarps][kNg2Rows];
    const int KB = (cpr + 31) >> 5;
    const int slices = KB < kWarps ? KB : kWarps;
    const int rbpc = kWarps / slices;
    const int wid = tid >> 5, rbl = wid / slices, ks = wid - rbl * slices;
    const int nrbg = (nrb + rbpc - 1) / rbpc;
    for (int64_t item = blockIdx.x; item < (int64_t)nrbg * B; item += gridDim.x) {
        const int rbg = (int)(item % nrbg), b = (int)(item / nrbg);
        const int rb = rbg * rbpc + rbl;
        const bool active = rbl < rbpc && rb < nrb;
        const __half* xb = x + (size_t)b * IC;
        float acc[kNg2Rows];
#pragma unroll
        for (int r = 0; r < kNg2Rows; ++r) acc[r] = 0.f;
        for (int kb = ks; active && kb < KB; kb += slices) {
            const int c = kb * 32 + lane;
            if (c >= cpr) continue;
            uint32_t xr[16];
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                const uint4 xv = __ldg(reinterpret_cast<const uint4*>(xb + c * 32) + q);
                xr[4 * q] = xv.x; xr[4 * q + 1] = xv.y; xr[4 * q + 2] = xv.z; xr[4 * q + 3] = xv.w;
            }
            uint2 lo[kNg2Rows];
            uint32_t hi[kNg2Rows], sc[kNg2Rows];
#pragma unroll
            for (int r = 0; r < kNg2Rows; ++r) {
                const int row = min(rb * kNg2Rows + r, OC - 1);
                const size_t wi = (size_t)row * cpr + c;
                lo[r] = __ldg(reinterpret_cast<const uint2*>(qweight) + wi);
                hi[r] = __ldg(qhigh + wi);
                sc[r] = __ldg(reinterpret_cast<const uint32_t*>(scales) + (size_t)row * ngroups + c / chunks_per_group);
            }
#pragma unroll
            for (int r = 0; r < kNg2Rows; ++r) {
                __half2 ap[2], an[2];
                ap[0] = ap[1] = an[0] = an[1] = __float2half2_rn(0.f);
                const uint32_t low[2] = {lo[r].x, lo[r].y};
#pragma unroll
                for (int w = 0; w < 2; ++w)
#pragma unroll
                    for (int p = 0; p < 8; ++p) {
                        const int hb = 4 * p + 2 * w;   // nf3_hbit_pos(16w + 2p): the pair's two high bits are adjacent
                        const uint32_t t1 = (4 * p <= 8) ? (low[w] << (8 - 4 * p)) : (low[w] >> (4 * p - 8));
                        const uint32_t t2 = (hb <= 12) ? (hi[r] << (12 - hb)) : (hi[r] >> (hb - 12));
                        const uint32_t addr = (t2 & 0x3000u) | ((t1 & 0xF00u) | lanebase);
                        const uint2 l = lds64(addr);
                        const __half2 x2 = *reinterpret_cast<const __half2*>(&xr[8 * w + p]);
                        ap[p & 1] = __hfma2(*reinterpret_cast<const __half2*>(&l.x), x2, ap[p & 1]);
                        an[p & 1] = __hfma2(*reinterpret_cast<const __half2*>(&l.y), x2, an[p & 1]);
                    }
                const float2 p0 = __half22float2(ap[0]), p1 = __half22float2(ap[1]);
                const float2 n0 = __half22float2(an[0]), n1 = __half22float2(an[1]);
                const __half2 s2 = *reinterpret_cast<const __half2*>(&sc[r]);
                acc[r] = fmaf(__low2float(s2), (p0.x + p0.y) + (p1.x + p1.y), fmaf(__high2float(s2), (n0.x + n0.y) + (n1.x + n1.y), acc[r]));
            }
        }
#pragma unroll
        for (int r = 0; r < kNg2Rows; ++r) {
            float a = acc[r];
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) a += __shfl_xor_sync(0xffffffffu, a, o);
            if (lane == 0) part[wid][r] = a;
        }
        __syncthreads();
        if (tid < rbpc * kNg2Rows) {
            const int orb = tid / kNg2Rows, r = tid - orb * kNg2Rows;
            float a = 0.f;
            for (int j = 0; j < slices; ++j) a += part[orb * slices + j][r];
            const int row = (rbg * rbpc + orb) * kNg2Rows + r;
            if (row < OC) out[(size_t)b * OC + row] = __float2half_rn(a);
        }
        __syncthreads();
    }
}

}  // namespace bdq

using namespace bdq;

static bool nf3p_shape_ok(int64_t rows, int64_t cols, int group) {
    return rows > 0 && cols > 0 && group > 0 && group % 32 == 0 && cols % group == 0 && rows * (cols / 32) < (1ll << 40);
}

// codes: uint8 [rows, cols] as written by bdq_nf3_fakequant_fwd; qweight int32 [rows, cols/16]; qhigh int32 [rows, cols/32]
extern "C" int bdq_pack_nf3(const uint8_t* codes, int32_t* qweight, int32_t* qhigh, int64_t rows, int64_t cols, bdq_stream_t stream) {
    if (!codes || !qweight || !qhigh || rows <= 0 || cols <= 0 || cols % 32 != 0) {
        set_error("pack_nf3: rows=%lld cols=%lld (cols must be a multiple of 32), non-null pointers", (long long)rows, (long long)cols);
        return BDQ_ERR_SHAPE;
    }
    if ((uintptr_t)codes % 16) { set_error("pack_nf3: codes must be 16-byte aligned"); return BDQ_ERR_ALIGN; }
    const int64_t nchunk = rows * (cols / 32);
    const int grid = (int)std::min<int64_t>((nchunk + 255) / 256, (int64_t)sm_count() * 16);
    nf3_pack_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(codes, (uint32_t*)qweight, (uint32_t*)qhigh, nchunk);
    return check_launch("bdq_pack_nf3");
}

extern "C" int bdq_nf3_dequant(const int32_t* qweight, const int32_t* qhigh, const void* scales, void* out_f16, int64_t rows,
                               int64_t cols, int group, bdq_stream_t stream) {
    if (!qweight || !qhigh || !scales || !out_f16 || !nf3p_shape_ok(rows, cols, group)) {
        set_error("nf3_dequant: rows=%lld cols=%lld group=%d", (long long)rows, (long long)cols, group);
        return BDQ_ERR_SHAPE;
    }
    if (((uintptr_t)out_f16 % 16) || ((uintptr_t)scales % 4)) { set_error("nf3_dequant: out must be 16-byte aligned"); return BDQ_ERR_ALIGN; }
    const int64_t nchunk = rows * (cols / 32);
    const int grid = (int)std::min<int64_t>((nchunk + 255) / 256, (int64_t)sm_count() * 16);
    nf3_dequant_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>((const uint32_t*)qweight, (const uint32_t*)qhigh, (const __half*)scales,
                                                               (__half*)out_f16, nchunk, (int)(cols / 32), group / 32,
                                                               (int)(cols / group), nf3_levels_f16());
    return check_launch("bdq_nf3_dequant");
}

// out[b, oc] = fp16( sum_k x[b, k] * lvl[e(oc, k)] * (e < 3 ? sn : sp)[oc, k / G] ), fp32 accumulation
extern "C" int bdq_gemv_nf3(const void* x, const int32_t* qweight, const int32_t* qhigh, const void* scales, void* out, int64_t B,
                            int64_t IC, int64_t OC, int group, bdq_stream_t stream) {
    if (!x || !qweight || !qhigh || !scales || !out || B <= 0 || !nf3p_shape_ok(OC, IC, group) || IC > (1 << 24) || OC > (1 << 24)) {
        set_error("gemv_nf3: B=%lld IC=%lld OC=%lld group=%d", (long long)B, (long long)IC, (long long)OC, group);
        return BDQ_ERR_SHAPE;
    }
    if (((uintptr_t)x % 16) || ((uintptr_t)scales % 4)) { set_error("gemv_nf3: x must be 16-byte aligned"); return BDQ_ERR_ALIGN; }
    {
        const int rc = nf3_gemv_stream_launch(x, qweight, qhigh, scales, out, B, IC, OC, group, (cudaStream_t)stream);
        if (rc != BDQ_ERR_UNSUPPORTED) return rc;
    }
    const int64_t cpr2 = IC / 32, KB2 = (cpr2 + 31) / 32, slices2 = KB2 < 8 ? KB2 : 8, rbpc2 = 8 / slices2;
    const int64_t nrb2 = (OC + kNg2Rows - 1) / kNg2Rows, items2 = ((nrb2 + rbpc2 - 1) / rbpc2) * B;
    const int grid2 = (int)std::min<int64_t>(items2, (int64_t)sm_count() * 4);
    nf3_gemv2_kernel<<<grid2, kNgThreads, 0, (cudaStream_t)stream>>>((const __half*)x, (const uint32_t*)qweight, (const uint32_t*)qhigh,
                                                                     (const __half*)scales, (__half*)out, (int)B, (int)IC, (int)OC,
                                                                     group / 32, nf3_levels_f16());
    return check_launch("bdq_gemv_nf3");
}
