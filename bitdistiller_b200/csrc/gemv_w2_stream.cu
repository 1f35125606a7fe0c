// Batch-1 packed int2 decode GEMV, register-streaming variant (sm_100a).  Replaces gemv_kernel_g128_2bit /
// gemv_forward_cuda (inference/kernels/csrc/quantization/gemv_cuda.cu:202-272, 289-346) for rows == 1, and the
// fused-QKV / gate+up launches of inference/modules/fused_attn.py:304-340 and fused_mlp.py:44-114 (several weight
// matrices that share one activation vector go through ONE launch).
//
// A decode GEMV (1-27 MB of packed weights) is shorter than the bandwidth-delay product of the chip, so the kernel
// is organised around latency, not FLOPs:
//   * no shared-memory staging of weights and no CTA-wide barrier in the load path: every lane pulls the 2 x 32 B it
//     will consume itself (one quantisation group of channels g and g+8) straight into registers with LDG.256, kDepth
//     items deep.  (A per-warp cp.async ring was measured too: with one row per 4 lanes every LDGSTS.128 costs 32
//     shared-memory wavefronts instead of 4 and the MIO pipe becomes the bound -- profiles/r01_gemv_notes.md);
//   * two CTAs (16 warps) per SM: a warp alone issues one instruction per ~4 cycles (LOP3 -> HMMA dependencies), four
//     warps per scheduler are needed to fill the ALU pipe;
//   * the weights of the NEXT launch can be handed in as a prefetch hint: one lane per CTA issues
//     cp.async.bulk.prefetch.L2 for its slice at kernel start, so HBM keeps streaming across the launch boundary and
//     the next launch finds its weights in L2 (every byte still crosses HBM exactly once per token);
//   * all kDepth loads are issued BEFORE griddepcontrol.wait (programmatic dependent launch hides the launch latency;
//     x is read only after the dependency resolves);
//   * x is staged per warp (each warp only needs the K windows it owns), so the only CTA-wide barrier is the split-K
//     reduction at the end of a 16-channel block;
//   * 2-bit codes become fp16 with one LOP3 per two codes ((word & mask) | exponent = K + code, K in
//     {1024,256,64,16,4}) plus one shift per word; the dot products run on the tensor pipe
//     (mma.sync.m16n8k16, fp32 accumulate).  With batch 1 the N dimension of the MMA is used to keep the four
//     quantisation groups of a 512-wide K window apart: lane t of a quad holds group t of the window and the B
//     operand column n carries x only for group n (zero elsewhere), so D[row][n] is the group-n partial sum and the
//     scale / zero-point fix-up runs once per 4 groups instead of once per group:
//         out[oc] += s[oc,g] * (D[oc][g] - sum_k K(k) x[k] - z[oc,g] * sum_k x[k])
//   * the x-only sums are computed once per warp while x is staged in extraction order.
#include "bdq_common.cuh"

namespace bdq {

constexpr int kSThreads = 256;
constexpr int kSWarps = 8;
constexpr int kDepth = 3;              // items (16 channels x 512 input channels = 2 KB) in flight per warp
constexpr int kGrpStride = 272;        // 256 B of permuted x per group + 16 B so the 4 active lanes hit distinct banks
constexpr int kWinStride = 1152;      // 4 groups + (Kx, Sx) of the 4 groups at +1088, padded to a multiple of 128
constexpr int kMetaOff = 4 * kGrpStride;
constexpr int kZeroBytes = 384;        // per-warp zero block read by the lanes whose B column is empty
constexpr int kMaxMats = 4;

struct GemvMats {
    const uint8_t* qw[kMaxMats];
    const __half* sc[kMaxMats];
    const uint32_t* qz[kMaxMats];
    __half* out[kMaxMats];
    int rb_end[kMaxMats];   // cumulative count of 16-channel blocks
    const uint8_t* pf[kMaxMats];       // L2 prefetch hint: packed weights of the next launch (may be null)
    unsigned long long pf_bytes[kMaxMats];
};

struct ItemBuf {
    U8 a, b;              // packed codes: channel g / g+8, one group (128 codes) each
    uint32_t sa, sb;      // two fp16 scales each: channel g / g+8, groups 2(t&1), 2(t&1)+1 of the window
    uint32_t za, zb;      // the zero-point words holding those groups
};

__device__ __forceinline__ uint32_t lop_and_or_s(uint32_t a, uint32_t b, uint32_t c) {
    uint32_t r;
    asm("lop3.b32 %0, %1, %2, %3, 0xEA;" : "=r"(r) : "r"(a), "r"(b), "r"(c));   // (a & b) | c
    return r;
}

// additive offset of extraction e: codes e and e+8 of a word come out as (K + code)
__host__ __device__ constexpr float ext_ks(int e) {
    return (e == 0 || e == 5) ? 1024.f : (e == 1 || e == 6) ? 256.f : (e == 2 || e == 7) ? 64.f : (e == 3) ? 16.f : 4.f;
}

// h[e] = half2(K_e + code[e], K_e + code[e + 8]), e = 0..7: 8 LOP3 + 1 shift for 16 codes
__device__ __forceinline__ void extract16s(uint32_t v, uint32_t* h) {
    const uint32_t v1 = v >> 10;
    h[0] = lop_and_or_s(v, 0x00030003u, 0x64006400u);
    h[1] = lop_and_or_s(v, 0x000C000Cu, 0x5C005C00u);
    h[2] = lop_and_or_s(v, 0x00300030u, 0x54005400u);
    h[3] = lop_and_or_s(v, 0x00C000C0u, 0x4C004C00u);
    h[4] = lop_and_or_s(v, 0x03000300u, 0x44004400u);
    h[5] = lop_and_or_s(v1, 0x00030003u, 0x64006400u);
    h[6] = lop_and_or_s(v1, 0x000C000Cu, 0x5C005C00u);
    h[7] = lop_and_or_s(v1, 0x00300030u, 0x54005400u);
}

__device__ __forceinline__ void mma16816s(float* d, uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3, uint32_t b0, uint32_t b1) {
    asm("mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
        : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
        : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}

__device__ __forceinline__ uint4 ld_coherent_v4(const void* p) {
    uint4 v;
    asm volatile("ld.global.v4.b32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ uint32_t ldg_nc_u32(const void* p) {
    uint32_t v;
    asm volatile("ld.global.nc.b32 %0, [%1];" : "=r"(v) : "l"(p));
    return v;
}

// shared memory per warp: its windows of x (extraction order) + the zero block
__host__ __device__ inline int gemv_s_warp_bytes(int wpw) { return wpw * kWinStride + kZeroBytes; }
__host__ __device__ inline int gemv_s_smem_bytes(int wpw) { return kSWarps * gemv_s_warp_bytes(wpw) + 2 * kSWarps * 16 * 4; }

__device__ __forceinline__ unsigned long long gtime() {
    unsigned long long v;
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(v));
    return v;
}

// PROF: lane 0 of every warp records %globaltimer at the phase boundaries into prof[cta][warp][16] (dev tool only)
template <bool PDL, bool PROF>
__global__ void __launch_bounds__(kSThreads, 2) gemv_w2s_kernel(const __half* __restrict__ x, const GemvMats m, int IC,
                                                                int zw, int total_rb, int wpw,
                                                                unsigned long long* __restrict__ prof) {
    extern __shared__ __align__(128) uint8_t smem[];
    int pslot = 0;
    auto mark = [&]() {
        if constexpr (PROF) {
            if ((threadIdx.x & 31) == 0 && pslot < 16)
                prof[((size_t)blockIdx.x * kSWarps + (threadIdx.x >> 5)) * 16 + pslot] = gtime();
            ++pslot;
        }
    };
    mark();   // 0: start
    if constexpr (PDL) asm volatile("griddepcontrol.launch_dependents;");
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int g = lane >> 2, t = lane & 3;
    const int ngrp = IC >> 7, nwin = (ngrp + 3) >> 2;
    const int row_bytes = IC >> 2, zw16 = zw * 16;
    const int rb0 = (int)((int64_t)blockIdx.x * total_rb / gridDim.x);
    const int rb1 = (int)((int64_t)(blockIdx.x + 1) * total_rb / gridDim.x);
    const int mywin = warp < nwin ? (nwin - warp + kSWarps - 1) / kSWarps : 0;   // windows warp, warp+8, ...
    uint8_t* wbase = smem + (size_t)warp * gemv_s_warp_bytes(wpw);
    uint8_t* zblock = wbase + (size_t)wpw * kWinStride;
    float* red = reinterpret_cast<float*>(smem + (size_t)kSWarps * gemv_s_warp_bytes(wpw));   // [2][8][16]

    // ---- producer cursor: item = (16-channel block, window of this warp)
    int p_rb = rb0, p_j = 0, p_mi = 0;
    while (p_mi < kMaxMats - 1 && p_rb >= m.rb_end[p_mi]) ++p_mi;
    auto issue = [&](ItemBuf& b) {
        if (p_rb >= rb1) return;
        const int rstart = p_mi ? m.rb_end[p_mi - 1] : 0;
        const size_t r = (size_t)(p_rb - rstart) * 16 + g;
        const int win = warp + kSWarps * p_j;
        const int G = 4 * win + t;
        if (G < ngrp) {
            const uint8_t* pa = m.qw[p_mi] + r * row_bytes + (size_t)G * 32;
            b.a = ldg256(pa);
            b.b = ldg256(pa + (size_t)8 * row_bytes);
        } else {
#pragma unroll
            for (int i = 0; i < 8; ++i) b.a.r[i] = b.b.r[i] = 0u;
        }
        const int G0 = 4 * win + 2 * (t & 1);
        const __half* ps = m.sc[p_mi] + r * zw16 + G0;
        b.sa = ldg_nc_u32(ps);
        b.sb = ldg_nc_u32(ps + (size_t)8 * zw16);
        const uint32_t* pz = m.qz[p_mi] + r * zw + (G0 >> 4);
        b.za = ldg_nc_u32(pz);
        b.zb = ldg_nc_u32(pz + (size_t)8 * zw);
        if (++p_j == mywin) {
            p_j = 0;
            ++p_rb;
            while (p_mi < kMaxMats - 1 && p_rb >= m.rb_end[p_mi]) ++p_mi;
        }
    };
    ItemBuf buf[kDepth];
    if (mywin > 0) {
#pragma unroll
        for (int d = 0; d < kDepth; ++d) issue(buf[d]);   // weights first: nothing here depends on the previous kernel
    }
    // L2 prefetch of the next launch's weights: this CTA's slice, 8 KB per instruction, one lane
    if (tid == 0) {
#pragma unroll 1
        for (int i = 0; i < kMaxMats; ++i) {
            const unsigned long long nb = m.pf_bytes[i];
            if (nb == 0) continue;
            const unsigned long long per = (((nb + gridDim.x - 1) / gridDim.x) + 127) & ~127ull;
            unsigned long long o = per * blockIdx.x;
            const unsigned long long end = o + per < nb ? o + per : nb;
            for (; o < end; o += 8192) {
                const unsigned int n = (unsigned int)(end - o < 8192 ? (end - o) & ~15ull : 8192);
                if (n) asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(m.pf[i] + o), "r"(n) : "memory");
            }
        }
    }
    // zero block (read as the B operand of the columns this lane's quad does not own)
    for (int i = lane; i < kZeroBytes / 16; i += 32) reinterpret_cast<uint4*>(zblock)[i] = make_uint4(0, 0, 0, 0);

    // ---- stage this warp's windows of x in extraction order + the per-group x sums
    mark();   // 1: weight loads issued
    if constexpr (PDL) asm volatile("griddepcontrol.wait;" ::: "memory");
    mark();   // 2: dependency resolved
    for (int j = 0; j < mywin; ++j) {
        const int k0 = (warp + kSWarps * j) * 512 + lane * 16;
        uint32_t in[8] = {0, 0, 0, 0, 0, 0, 0, 0};
        if (k0 < IC) {
            const uint4 v0 = ld_coherent_v4(x + k0), v1 = ld_coherent_v4(x + k0 + 8);
            in[0] = v0.x; in[1] = v0.y; in[2] = v0.z; in[3] = v0.w;
            in[4] = v1.x; in[5] = v1.y; in[6] = v1.z; in[7] = v1.w;
        }
        uint32_t o[8];
        float bias = 0.f, sum = 0.f;
#pragma unroll
        for (int e = 0; e < 8; ++e) {
            const uint32_t lo = in[e >> 1], hi = in[(e + 8) >> 1];
            o[e] = (e & 1) ? __byte_perm(lo, hi, 0x7632) : __byte_perm(lo, hi, 0x5410);   // (x[e], x[e+8])
            const float2 f = __half22float2(*reinterpret_cast<const __half2*>(&o[e]));
            sum += f.x + f.y;
            bias = fmaf(ext_ks(e), f.x + f.y, bias);
        }
        uint8_t* wj = wbase + (size_t)j * kWinStride;
        uint4* dst = reinterpret_cast<uint4*>(wj + (lane >> 3) * kGrpStride + (lane & 7) * 32);
        dst[0] = make_uint4(o[0], o[1], o[2], o[3]);
        dst[1] = make_uint4(o[4], o[5], o[6], o[7]);
#pragma unroll
        for (int s = 1; s < 8; s <<= 1) {
            bias += __shfl_xor_sync(0xffffffffu, bias, s);
            sum += __shfl_xor_sync(0xffffffffu, sum, s);
        }
        if ((lane & 7) == 0) reinterpret_cast<float2*>(wj + kMetaOff)[lane >> 3] = make_float2(bias, sum);
    }
    __syncwarp();
    mark();   // 3: x staged

    // ---- consumer
    const bool active = (g == t);
    const uint8_t* bbase = active ? wbase + t * kGrpStride : zblock + 64;
    const int jstride = active ? kWinStride : 0;
    float fin_lo = 0.f, fin_hi = 0.f;
    int c_rb = rb0, c_j = 0, c_mi = 0, par = 0;
    while (c_mi < kMaxMats - 1 && c_rb >= m.rb_end[c_mi]) ++c_mi;

    auto flush = [&]() {
        fin_lo += __shfl_xor_sync(0xffffffffu, fin_lo, 1);
        fin_hi += __shfl_xor_sync(0xffffffffu, fin_hi, 1);
        float* my = red + (par * kSWarps + warp) * 16;
        if (t == 0) {
            my[g] = fin_lo;
            my[g + 8] = fin_hi;
        }
        fin_lo = fin_hi = 0.f;
        __syncthreads();
        if (tid < 16) {
            const float* rp = red + par * kSWarps * 16 + tid;
            float acc = 0.f;
#pragma unroll
            for (int w = 0; w < kSWarps; ++w) acc += rp[w * 16];
            const int rstart = c_mi ? m.rb_end[c_mi - 1] : 0;
            m.out[c_mi][(size_t)(c_rb - rstart) * 16 + tid] = __float2half_rn(acc);
        }
        par ^= 1;
        ++c_rb;
        while (c_mi < kMaxMats - 1 && c_rb >= m.rb_end[c_mi]) ++c_mi;
    };

    if (mywin == 0) {   // IC < 4096: this warp owns no window but still takes part in the block reductions
        for (int rb = rb0; rb < rb1; ++rb) flush();
        return;
    }

    auto consume = [&](const ItemBuf& b) {
        const uint8_t* bp = bbase + (size_t)c_j * jstride;
        float d[4][4];
#pragma unroll
        for (int c = 0; c < 4; ++c) d[c][0] = d[c][1] = d[c][2] = d[c][3] = 0.f;
#pragma unroll
        for (int wd = 0; wd < 8; ++wd) {
            uint32_t ha[8], hb[8];
            extract16s(b.a.r[wd], ha);
            extract16s(b.b.r[wd], hb);
            const uint4 q0 = *reinterpret_cast<const uint4*>(bp + wd * 32);
            const uint4 q1 = *reinterpret_cast<const uint4*>(bp + wd * 32 + 16);
            mma16816s(d[0], ha[0], hb[0], ha[1], hb[1], q0.x, q0.y);
            mma16816s(d[1], ha[2], hb[2], ha[3], hb[3], q0.z, q0.w);
            mma16816s(d[2], ha[4], hb[4], ha[5], hb[5], q1.x, q1.y);
            mma16816s(d[3], ha[6], hb[6], ha[7], hb[7], q1.z, q1.w);
        }
        if (t < 2) {   // columns 2t, 2t+1 = groups 2t, 2t+1 of the window
            const float4 mm = *reinterpret_cast<const float4*>(wbase + (size_t)c_j * kWinStride + kMetaOff + t * 16);
            const int sh = ((4 * (warp + kSWarps * c_j) + 2 * t) & 15) * 2;
            const float2 sl = __half22float2(*reinterpret_cast<const __half2*>(&b.sa));
            const float2 sh2 = __half22float2(*reinterpret_cast<const __half2*>(&b.sb));
            const float za0 = (float)((b.za >> sh) & 3u), za1 = (float)((b.za >> (sh + 2)) & 3u);
            const float zb0 = (float)((b.zb >> sh) & 3u), zb1 = (float)((b.zb >> (sh + 2)) & 3u);
            fin_lo = fmaf(sl.x, ((d[0][0] + d[1][0]) + (d[2][0] + d[3][0])) - fmaf(za0, mm.y, mm.x), fin_lo);
            fin_lo = fmaf(sl.y, ((d[0][1] + d[1][1]) + (d[2][1] + d[3][1])) - fmaf(za1, mm.w, mm.z), fin_lo);
            fin_hi = fmaf(sh2.x, ((d[0][2] + d[1][2]) + (d[2][2] + d[3][2])) - fmaf(zb0, mm.y, mm.x), fin_hi);
            fin_hi = fmaf(sh2.y, ((d[0][3] + d[1][3]) + (d[2][3] + d[3][3])) - fmaf(zb1, mm.w, mm.z), fin_hi);
        }
    };

    const int nitems = (rb1 - rb0) * mywin;
    for (int i = 0; i < nitems; i += kDepth) {
#pragma unroll
        for (int d = 0; d < kDepth; ++d) {
            if (i + d < nitems) {
                consume(buf[d]);
                issue(buf[d]);
                mark();   // 4..: item consumed
                if (++c_j == mywin) {
                    c_j = 0;
                    flush();
                    mark();   // block reduced + stored
                }
            }
        }
    }
}

}  // namespace bdq

using namespace bdq;

static unsigned long long* g_gemv_prof = nullptr;
static size_t g_gemv_prof_cap = 0, g_gemv_prof_used = 0;
// dev tool: device buffer of `cap` u64 that BDQ_GEMV_PROF launches fill with %globaltimer phase marks
extern "C" int bdq_gemv_set_prof(void* buf, int64_t cap) {
    g_gemv_prof = (unsigned long long*)buf;
    g_gemv_prof_cap = (size_t)cap;
    g_gemv_prof_used = 0;
    return BDQ_OK;
}

// Shared launcher (also used by bdq_gemv_w2 for batch 1).  Returns BDQ_ERR_UNSUPPORTED when the shape does not fit the
// streaming kernel so the caller can take the ring kernel instead.
int bdq_gemv_w2_stream_launch(const void* x, int nmat, const int32_t* const* qweight, const void* const* scales,
                              const int32_t* const* qzeros, void* const* out, const int64_t* OC, int64_t IC, int flags,
                              cudaStream_t st, int n_prefetch, const void* const* prefetch, const int64_t* prefetch_bytes) {
    if (nmat < 1 || nmat > kMaxMats || IC <= 0 || IC % 128 != 0 || IC > (1 << 20)) return BDQ_ERR_UNSUPPORTED;
    const int zw = bdq_zeros_width(IC, 128, 2);
    if (zw < 0) return BDQ_ERR_UNSUPPORTED;
    GemvMats m;
    int total_rb = 0;
    for (int i = 0; i < kMaxMats; ++i) {
        const int k = i < nmat ? i : nmat - 1;
        if (i < nmat) {
            if (OC[i] <= 0 || OC[i] % 16 != 0 || OC[i] > (1 << 24)) return BDQ_ERR_UNSUPPORTED;
            if (((uintptr_t)qweight[i] % 32) || ((uintptr_t)scales[i] % 4) || ((uintptr_t)qzeros[i] % 4) || !out[i])
                return BDQ_ERR_UNSUPPORTED;
            total_rb += (int)(OC[i] / 16);
        }
        m.qw[i] = (const uint8_t*)qweight[k];
        m.sc[i] = (const __half*)scales[k];
        m.qz[i] = (const uint32_t*)qzeros[k];
        m.out[i] = (__half*)out[k];
        m.rb_end[i] = total_rb;
        const bool pf = prefetch && prefetch_bytes && i < n_prefetch && prefetch[i] && prefetch_bytes[i] > 0 &&
                        ((uintptr_t)prefetch[i] % 16) == 0;
        m.pf[i] = pf ? (const uint8_t*)prefetch[i] : nullptr;
        m.pf_bytes[i] = pf ? (unsigned long long)prefetch_bytes[i] : 0ull;
    }
    if ((uintptr_t)x % 16) return BDQ_ERR_UNSUPPORTED;
    const int nwin = (int)((IC / 128 + 3) / 4);
    const int wpw = (nwin + kSWarps - 1) / kSWarps;
    const int smem = gemv_s_smem_bytes(wpw);
    if (smem > 112 * 1024) return BDQ_ERR_UNSUPPORTED;   // two CTAs (this launch + its PDL successor) per SM
    const bool pdl = (flags & BDQ_GEMV_PDL) != 0;
    const bool prof = (flags & BDQ_GEMV_PROF) != 0 && g_gemv_prof != nullptr;
    auto kern = prof ? (pdl ? gemv_w2s_kernel<true, true> : gemv_w2s_kernel<false, true>)
                     : (pdl ? gemv_w2s_kernel<true, false> : gemv_w2s_kernel<false, false>);
    static bool attr_done[4] = {false, false, false, false};
    const int ai = (prof ? 2 : 0) + (pdl ? 1 : 0);
    if (!attr_done[ai]) {
        if (cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 112 * 1024) != cudaSuccess)
            return check_launch("bdq_gemv_w2(stream attr)");
        attr_done[ai] = true;
    }
    // two CTAs (16 warps) per SM; BDQ_GEMV_CTAS1 = one (leaves room for a PDL successor's CTA to prefetch)
    int grid = sm_count() * ((flags & BDQ_GEMV_CTAS1) ? 1 : 2);
    if (grid > total_rb) grid = total_rb;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)grid);
    cfg.blockDim = dim3(kSThreads);
    cfg.dynamicSmemBytes = (size_t)smem;
    cfg.stream = st;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at;
    cfg.numAttrs = pdl ? 1 : 0;
    unsigned long long* pbuf = nullptr;
    if (prof) {   // consecutive profiled launches write consecutive [grid][8][16] slabs
        pbuf = g_gemv_prof + g_gemv_prof_used;
        g_gemv_prof_used += (size_t)grid * kSWarps * 16;
        if (g_gemv_prof_used > g_gemv_prof_cap) return BDQ_ERR_SHAPE;
    }
    cudaLaunchKernelEx(&cfg, kern, (const __half*)x, m, (int)IC, zw, total_rb, wpw, pbuf);
    return check_launch("bdq_gemv_w2(stream)");
}

extern "C" int bdq_gemv_w2_multi(const void* x, int nmat, const int32_t* const* qweight, const void* const* scales,
                                 const int32_t* const* qzeros, void* const* out, const int64_t* OC, int64_t IC, int group,
                                 int flags, int n_prefetch, const void* const* prefetch, const int64_t* prefetch_bytes,
                                 bdq_stream_t stream) {
    if (n_prefetch < 0 || n_prefetch > kMaxMats || (n_prefetch > 0 && (!prefetch || !prefetch_bytes))) {
        set_error("gemv_w2_multi: n_prefetch must be 0..%d with non-null hint arrays", kMaxMats);
        return BDQ_ERR_SHAPE;
    }
    if (nmat < 1 || nmat > kMaxMats || !x || !qweight || !scales || !qzeros || !out || !OC) {
        set_error("gemv_w2_multi: 1..%d matrices and non-null argument arrays required (nmat=%d)", kMaxMats, nmat);
        return BDQ_ERR_SHAPE;
    }
    for (int i = 0; i < nmat; ++i)
        if (!qweight[i] || !scales[i] || !qzeros[i] || !out[i] || OC[i] <= 0) {
            set_error("gemv_w2_multi: matrix %d has a null pointer or OC <= 0", i);
            return BDQ_ERR_SHAPE;
        }
    int rc = (group == 128 && !(flags & BDQ_GEMV_RING))
                 ? bdq_gemv_w2_stream_launch(x, nmat, qweight, scales, qzeros, out, OC, IC, flags, (cudaStream_t)stream,
                                             n_prefetch, prefetch, prefetch_bytes)
                 : BDQ_ERR_UNSUPPORTED;
    if (rc != BDQ_ERR_UNSUPPORTED) return rc;
    for (int i = 0; i < nmat; ++i) {   // shapes the fused kernel does not take: one launch per matrix
        rc = bdq_gemv_w2(x, qweight[i], scales[i], qzeros[i], out[i], 1, IC, OC[i], group, flags | BDQ_GEMV_RING, stream);
        if (rc) return rc;
    }
    return BDQ_OK;
}
