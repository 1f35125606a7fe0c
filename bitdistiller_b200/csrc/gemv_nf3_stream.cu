// Batch-1 packed NF3 decode GEMV on the INT8 tensor pipe (sm_100a), register-streaming like gemv_w2_stream.cu.
// North-star item (3) "packed 2-bit and nf3 weights"; the reference has no packed NF3 kernel (its NF3 path is the fake
// quantiser, quantization/quantizer.py:184-242), so parity is pinned on values: out = x @ dequant(pack(w)).T.
//
// NF3 levels are not an affine function of the code, so the digit-plane trick of the int2 kernel does not carry over.
// Instead the level LOOKUP runs as PRMT: an 8-entry int8 table is two registers, a PRMT selector is four 4-bit indices, and
// the packed format (nf3_packed.cuh) is laid out so that
//     sel = ((low_word >> 2p) & 0x33333333) | ((high_word << (2 - 2w - p)) & 0x44444444)
// IS eight selectors' worth of indices -- about half an ALU instruction per weight to form selectors, half a PRMT per
// weight to look levels up, nothing else per weight:
//   * per (channel, group) the 8 values lvl[e] * (e < 3 ? sn : sp) become a 16-bit fixed-point table relative to
//     smax = max(sp, sn): T = rint(lvl * s / smax * 32512) = 256 * Thi + Tlo with int8 limbs (4 registers per channel;
//     8 FFMA + 10 ALU instructions per 128 weights).  Both scales are folded in, so ONE lookup per limb serves both signs;
//   * x is staged exactly as in the int2 kernel: block fixed point per quantisation group, q = rint(x / unit),
//     |q| <= 32512, two int8 limbs, in the byte order the selectors produce;
//   * mma.sync.m16n8k32.s8.s8: A = the looked-up limb of 16 channels x 32 k-slots (lane (g, t) supplies channels g, g + 8
//     and 8 weights of ITS group: t-th group of the 512-wide K window), B column 2 * group + limb = that limb of x on the
//     k-slots of quad lane `group` only -- so lane (g, t) ends up with the four limb products of its own group and applies
//         out[oc] += smax[oc, G] / 32512 * unit[G] * (65536 HH + 256 (HL + LH) + LL)
//     once per group;  32 MMAs per 16 x 512 item;
//   * data movement as in gemv_w2_stream.cu: every lane pulls the 2 x (32 B low plane + 16 B high plane + 4 B scales) it
//     will consume itself straight into registers, kDepth items (3.3 KB each) deep per warp, 2 CTAs per SM; the only CTA
//     barrier is the split-K reduction of finished 16-channel blocks.
// Measured (B200, lm_head 151936 x 4096 = 253 MB, L2-cold): 77 us = 0.50 of the HBM peak (the CUDA-core pair lookup it
// replaces: 178 us, 0.22).  What bounds it (profiles/r02_nf3_gemv.md): ~480 instructions per item, of which 244 on the ALU
// pipe (2 cycles each per scheduler: 144 PRMT, 70 LOP3, 56 SHF) + 32 IMMA (8 cycles) + ~140 on the FMA pipe; the pure
// streaming skeleton of this kernel runs at 0.89 of peak, the compute alone at 0.58 with 16 warps per SM.  Tried and not
// kept: a TMA-fed shared-memory ring with 24 warps per SM (same time), IMAD.HI instead of SHF for the right shifts (4 cycles
// per scheduler, 5 % slower), L2 prefetch ahead of the register pipeline (7 % slower), deeper register pipeline (same).
// Value error: levels to 2^-16 of the group's larger scale, x to 2^-16 of the group's largest magnitude -- below the fp16
// rounding of the dequantised weights the reference-equivalent path multiplies with.  NaN scales / activations give NaN.
#include <cuda.h>

#include <algorithm>

#include "nf3_packed.cuh"

namespace bdq {
namespace nf3s {

constexpr int kThreads = 256;
constexpr int kDepth = 2;              // items (16 channels x 512 input channels = 3.3 KB) in flight per warp
constexpr int kWarps = 8;
constexpr int kSlotStride = 144;       // 128 B of int8 x limbs per (group, limb) + 16 B: the 8 active lanes hit distinct banks
constexpr int kMetaOff = 8 * kSlotStride;   // unit per group, 4 floats
constexpr int kWinStride = 1280;
constexpr int kZeroBytes = 128;
constexpr int kRedSlots = 4;

struct ItemBuf {
    U8 la, lb;          // low planes: channel g / g + 8, one group (128 codes x 2 bits) each
    uint4 ha, hb;       // high planes (128 bits)
    uint32_t sa, sb;    // (sp, sn) fp16 pair of this lane's group: channel g / g + 8
};

__device__ __forceinline__ void imma_s8(int* d, uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3, uint32_t b0, uint32_t b1) {
    asm("mma.sync.aligned.m16n8k32.row.col.s32.s8.s8.s32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
        : "+r"(d[0]), "+r"(d[1]), "+r"(d[2]), "+r"(d[3])
        : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}
__device__ __forceinline__ uint4 ldg128_stream(const void* p) {
    uint4 v;
    asm volatile("ld.global.nc.L1::no_allocate.v4.b32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p));
    return v;
}
__device__ __forceinline__ uint32_t ldg32_nc(const void* p) {
    uint32_t v;
    asm volatile("ld.global.nc.b32 %0, [%1];" : "=r"(v) : "l"(p));
    return v;
}
__device__ __forceinline__ uint32_t prmt(uint32_t a, uint32_t b, uint32_t s) {
    uint32_t d;
    asm("prmt.b32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(s));
    return d;
}

// left shifts as IMAD.SHL on the FMA pipe; right shifts stay SHF on the ALU pipe: IMAD.HI costs 4 cycles per scheduler against 2
// (tools/ubench/alu_rate.cu) and measured 5 % slower end to end
template <int N>
__device__ __forceinline__ uint32_t shl_fma(uint32_t x) {
    uint32_t d;
    asm("mul.lo.u32 %0, %1, %2;" : "=r"(d) : "r"(x), "r"(1u << N));
    return d;
}

__host__ __device__ inline int warp_bytes(int wpw) { return wpw * kWinStride + kZeroBytes; }
__host__ __device__ inline int smem_bytes(int wpw) { return kWarps * warp_bytes(wpw) + kRedSlots * kWarps * 64 * 4; }

struct LevelK {
    float c[8];   // lvl[e] * 32512
};

// 16-bit fixed-point level table of one (channel, group): T[e] = rint(c[e] * (e < 3 ? sn : sp) / max(sp, sn)), limbs packed as
// tab[0..1] = Thi[0..3], Thi[4..7]; tab[2..3] = Tlo[0..3], Tlo[4..7].  Returns max(sp, sn) / 32512 (NaN if a scale is NaN).
__device__ __forceinline__ float build_table(uint32_t sc, const LevelK& lk, uint32_t* tab) {
    const float2 s = __half22float2(*reinterpret_cast<const __half2*>(&sc));   // (sp, sn)
    const float smax = fmaxf(s.x, s.y);
    const float inv = smax > 0.f ? __fdividef(1.f, smax) : 0.f;
    const float rp = s.x * inv, rn = s.y * inv;
    uint32_t bits[8];
#pragma unroll
    for (int e = 0; e < 8; ++e) bits[e] = __float_as_uint(fmaf(lk.c[e], e < 3 ? rn : rp, 12583040.f));   // 0x4B400080 + T
#pragma unroll
    for (int h = 0; h < 2; ++h) {
        const uint32_t p01 = prmt(bits[4 * h], bits[4 * h + 1], 0x5140);       // (lo0, lo1, hi0, hi1)
        const uint32_t p23 = prmt(bits[4 * h + 2], bits[4 * h + 3], 0x5140);
        tab[h] = prmt(p01, p23, 0x7632);
        tab[2 + h] = prmt(p01, p23, 0x5410) ^ 0x80808080u;
    }
    const bool nan = (s.x != s.x) || (s.y != s.y);
    return nan ? __uint_as_float(0x7fc00000u) : smax * (1.f / 32512.f);
}

struct Args {
    const __half* x;
    const uint8_t* qw;
    const uint8_t* qh;
    const __half* sc;
    __half* out;
    int IC, rb_q, rb_r, wpw;
    LevelK lk;
};
// One item: the looked-up limbs of 16 channels x 512 input channels against this warp's staged x window.
// la / lb: low planes of channel g / g + 8 (this lane's group), ha / hb: high planes, sa / sb: (sp, sn).
__device__ __forceinline__ void consume_item(const U8& la, const U8& lb, const uint4& ha4, const uint4& hb4, uint32_t sa_, uint32_t sb_,
                                             const LevelK& lk, const uint8_t* bp, int sstride, float unit, float& fin_lo, float& fin_hi) {
    uint32_t ta[4], tb[4];
    const float fa = build_table(sa_, lk, ta), fb = build_table(sb_, lk, tb);
    int ah[4] = {0, 0, 0, 0}, al[4] = {0, 0, 0, 0};
    const uint32_t hA[4] = {ha4.x, ha4.y, ha4.z, ha4.w}, hB[4] = {hb4.x, hb4.y, hb4.z, hb4.w};
#pragma unroll
    for (int c = 0; c < 4; ++c) {          // high word c = 32 channels = low words 2c, 2c + 1
#pragma unroll
        for (int w = 0; w < 2; ++w) {
            const uint4 q = *reinterpret_cast<const uint4*>(bp + (2 * c + w) * sstride);   // x limbs of steps (w, p = 0), (w, p = 1)
            const uint32_t lwa = la.r[2 * c + w], lwb = lb.r[2 * c + w];
#pragma unroll
            for (int p = 0; p < 2; ++p) {
                constexpr int kSh[4] = {2, 1, 0, -1};   // 2 - 2w - p
                const int sh = kSh[2 * w + p];
                const uint32_t hsa = sh == 2 ? shl_fma<2>(hA[c]) : sh == 1 ? shl_fma<1>(hA[c]) : sh == 0 ? hA[c] : (hA[c] >> 1);
                const uint32_t hsb = sh == 2 ? shl_fma<2>(hB[c]) : sh == 1 ? shl_fma<1>(hB[c]) : sh == 0 ? hB[c] : (hB[c] >> 1);
                const uint32_t sa = ((p ? (lwa >> 2) : lwa) & 0x33333333u) | (hsa & 0x44444444u);
                const uint32_t sb = ((p ? (lwb >> 2) : lwb) & 0x33333333u) | (hsb & 0x44444444u);
                const uint32_t sa2 = (sa >> 16), sb2 = (sb >> 16);
                const uint32_t b0 = p ? q.z : q.x, b1 = p ? q.w : q.y;
                imma_s8(ah, prmt(ta[0], ta[1], sa), prmt(tb[0], tb[1], sb), prmt(ta[0], ta[1], sa2), prmt(tb[0], tb[1], sb2), b0, b1);
                imma_s8(al, prmt(ta[2], ta[3], sa), prmt(tb[2], tb[3], sb), prmt(ta[2], ta[3], sa2), prmt(tb[2], tb[3], sb2), b0, b1);
            }
        }
    }
    // this lane's columns 2t, 2t + 1 = (x hi limb, x lo limb) sums of group t: rows g ([0], [1]) and g + 8 ([2], [3])
    const float wa = fmaf((float)ah[0], 65536.f, fmaf((float)(ah[1] + al[0]), 256.f, (float)al[1]));
    const float wb = fmaf((float)ah[2], 65536.f, fmaf((float)(ah[3] + al[2]), 256.f, (float)al[3]));
    fin_lo = fmaf(fa * unit, wa, fin_lo);
    fin_hi = fmaf(fb * unit, wb, fin_hi);
}

__global__ void __launch_bounds__(kThreads, 2) gemv_nf3s_kernel(const Args a) {
    extern __shared__ __align__(128) uint8_t smem[];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int g = lane >> 2, t = lane & 3;
    const int IC = a.IC, wpw = a.wpw;
    const int ngrp = IC >> 7, nwin = (ngrp + 3) >> 2;
    const int lo_row = IC >> 2, hi_row = IC >> 3, sc_row = ngrp * 2;   // bytes, bytes, halves per channel
    const int rb0 = (int)blockIdx.x * a.rb_q + min((int)blockIdx.x, a.rb_r);
    const int rb1 = rb0 + a.rb_q + ((int)blockIdx.x < a.rb_r ? 1 : 0);
    const int mywin = warp < nwin ? (nwin - warp + kWarps - 1) / kWarps : 0;   // windows warp, warp + 8, ...
    uint8_t* wbase = smem + (size_t)warp * warp_bytes(wpw);
    uint8_t* zblock = wbase + (size_t)wpw * kWinStride;
    float* red = reinterpret_cast<float*>(smem + (size_t)kWarps * warp_bytes(wpw));   // [kRedSlots][8 warps][4 quad lanes][16]
    const int G0 = 4 * warp + t;
    const __half* sc = a.sc;

    // ---- producer cursors: item = (16-channel block, window of this warp)
    int p_rb = rb0, p_j = 0;
    auto advance = [&]() {
        if (++p_j == mywin) {
            p_j = 0;
            ++p_rb;
        }
    };
    auto issue = [&](ItemBuf& b) {
        if (p_rb >= rb1) return;
        const int G = G0 + 32 * p_j;
        const size_t r = (size_t)p_rb * 16 + g;
        if (G < ngrp) {
            const uint8_t* pl = a.qw + r * lo_row + (size_t)G * 32;
            const uint8_t* ph = a.qh + r * hi_row + (size_t)G * 16;
            const __half* ps = sc + r * sc_row + 2 * G;
            b.la = ldg256(pl);
            b.lb = ldg256(pl + (size_t)8 * lo_row);
            b.ha = ldg128_stream(ph);
            b.hb = ldg128_stream(ph + (size_t)8 * hi_row);
            b.sa = ldg32_nc(ps);
            b.sb = ldg32_nc(ps + (size_t)8 * sc_row);
        } else {
#pragma unroll
            for (int i = 0; i < 8; ++i) b.la.r[i] = b.lb.r[i] = 0u;
            b.ha = b.hb = make_uint4(0, 0, 0, 0);
            b.sa = b.sb = 0u;   // scales 0: table 0
        }
        advance();
    };
    ItemBuf buf[kDepth];
    if (mywin > 0) {
#pragma unroll
        for (int d = 0; d < kDepth; ++d) issue(buf[d]);   // weights first; x is staged while they are in flight
    }
    for (int i = lane; i < kZeroBytes / 16; i += 32) reinterpret_cast<uint4*>(zblock)[i] = make_uint4(0, 0, 0, 0);

    // ---- stage this warp's windows of x: per group block fixed point, two int8 limbs, selector byte order.
    // Lane l of a group's 8 lanes owns word l (activations 16l .. 16l+15 of the group) = MMA steps 2l (even activations of
    // the word) and 2l + 1 (odd ones); a step's 8 bytes are activations p, p + 2, .., p + 14.
    for (int j0 = 0; j0 < mywin; j0 += 4) {
        uint4 raw[4][2];
#pragma unroll
        for (int jj = 0; jj < 4; ++jj) {
            const int k0 = (warp + kWarps * (j0 + jj)) * 512 + lane * 16;
            raw[jj][0] = raw[jj][1] = make_uint4(0, 0, 0, 0);
            if (j0 + jj < mywin && k0 < IC) {
                raw[jj][0] = __ldg(reinterpret_cast<const uint4*>(a.x + k0));
                raw[jj][1] = __ldg(reinterpret_cast<const uint4*>(a.x + k0 + 8));
            }
        }
#pragma unroll
        for (int jj = 0; jj < 4; ++jj) {
            if (j0 + jj >= mywin) break;
            const uint32_t in[8] = {raw[jj][0].x, raw[jj][0].y, raw[jj][0].z, raw[jj][0].w,
                                    raw[jj][1].x, raw[jj][1].y, raw[jj][1].z, raw[jj][1].w};
            __half2 m2 = __habs2(*reinterpret_cast<const __half2*>(&in[0]));
#pragma unroll
            for (int e = 1; e < 8; ++e) m2 = __hmax2_nan(m2, __habs2(*reinterpret_cast<const __half2*>(&in[e])));
            uint32_t mx = *reinterpret_cast<const uint32_t*>(&m2);
            mx = max(mx & 0xffffu, mx >> 16);   // |.| bit patterns order like unsigned integers, NaN > Inf > finite
#pragma unroll
            for (int sft = 1; sft < 8; sft <<= 1) mx = max(mx, __shfl_xor_sync(0xffffffffu, mx, sft));
            const uint32_t ex = (mx >> 10) & 31u;
            const float gm = fmaxf(__half2float(__ushort_as_half((unsigned short)mx)), 5.9604645e-8f);
            const float inv_unit = __fdividef(32512.f, gm);
            const float unit = ex == 31u ? __uint_as_float(0x7fc00000u) : gm * (1.f / 32512.f);
            uint32_t bits[16];
#pragma unroll
            for (int e = 0; e < 8; ++e) {
                const float2 v = __half22float2(*reinterpret_cast<const __half2*>(&in[e]));
                bits[2 * e] = __float_as_uint(fmaf(v.x, inv_unit, 12583040.f));
                bits[2 * e + 1] = __float_as_uint(fmaf(v.y, inv_unit, 12583040.f));
            }
            uint32_t hi[4], lo[4];   // [2p + half]: activations 8 half + p + 2k, k = 0..3
#pragma unroll
            for (int p = 0; p < 2; ++p)
#pragma unroll
                for (int hf = 0; hf < 2; ++hf) {
                    const int a0 = 8 * hf + p;
                    const uint32_t p01 = prmt(bits[a0], bits[a0 + 2], 0x5140);
                    const uint32_t p23 = prmt(bits[a0 + 4], bits[a0 + 6], 0x5140);
                    lo[2 * p + hf] = prmt(p01, p23, 0x5410) ^ 0x80808080u;
                    hi[2 * p + hf] = prmt(p01, p23, 0x7632);
                }
            uint8_t* wj = wbase + (size_t)(j0 + jj) * kWinStride;
            const int grp = lane >> 3, w8 = lane & 7;   // slot = 2 * grp + limb; inside: 16 B per word = steps 2 w8, 2 w8 + 1
            uint8_t* dst = wj + (2 * grp) * kSlotStride + w8 * 16;
            *reinterpret_cast<uint4*>(dst) = make_uint4(hi[0], hi[1], hi[2], hi[3]);
            *reinterpret_cast<uint4*>(dst + kSlotStride) = make_uint4(lo[0], lo[1], lo[2], lo[3]);
            if (w8 == 0) reinterpret_cast<float*>(wj + kMetaOff)[grp] = unit;
        }
    }
    __syncwarp();

    // ---- consumer: B column n = 2 * group + limb carries x only on the k-slots of quad lane t = group
    const bool active = ((g >> 1) == t);
    const uint8_t* bbase = active ? wbase + g * kSlotStride : zblock;
    const int jstride = active ? kWinStride : 0;
    const int sstride = active ? 16 : 0;
    float fin_lo = 0.f, fin_hi = 0.f;
    int c_j = 0, rslot = 0, rb_base = rb0;

    auto block_reduce = [&]() {
        __syncthreads();
        if (tid < rslot * 16) {
            const int sl = tid >> 4, row = tid & 15;
            const float* rp = red + sl * kWarps * 64 + row;
            float acc = 0.f;
#pragma unroll
            for (int w = 0; w < kWarps * 4; ++w) acc += rp[w * 16];
            a.out[(size_t)(rb_base + sl) * 16 + row] = __float2half_rn(acc);
        }
        rb_base += rslot;
        rslot = 0;
        if (rb_base < rb1) __syncthreads();
    };
    auto flush = [&]() {
        float* my = red + ((rslot * kWarps + warp) * 4 + t) * 16;
        my[g] = fin_lo;
        my[g + 8] = fin_hi;
        fin_lo = fin_hi = 0.f;
        if (++rslot == kRedSlots) block_reduce();
    };
    if (mywin == 0) {
        for (int rb = rb0; rb < rb1; ++rb) flush();
        if (rslot) block_reduce();
        return;
    }
    auto unit_of = [&]() { return *reinterpret_cast<const float*>(wbase + (size_t)c_j * kWinStride + kMetaOff + t * 4); };
    auto next = [&]() {
        if (++c_j == mywin) {
            c_j = 0;
            flush();
        }
    };

    const int nitems = (rb1 - rb0) * mywin;
    for (int i = 0; i < nitems; i += kDepth) {
#pragma unroll
        for (int d = 0; d < kDepth; ++d) {
            if (i + d < nitems) {
                const ItemBuf& b = buf[d];
                consume_item(b.la, b.lb, b.ha, b.hb, b.sa, b.sb, a.lk, bbase + (size_t)c_j * jstride, sstride, unit_of(), fin_lo, fin_hi);
                issue(buf[d]);
                next();
            }
        }
    }
    if (rslot) block_reduce();
}

}  // namespace nf3s

int nf3_gemv_stream_launch(const void* x, const int32_t* qweight, const int32_t* qhigh, const void* scales, void* out, int64_t B,
                           int64_t IC, int64_t OC, int group, cudaStream_t stream) {
    static const bool off = getenv("BDQ_NF3_GEMV_STREAM") && atoi(getenv("BDQ_NF3_GEMV_STREAM")) == 0;   // A/B timing against the pair-lookup kernel
    if (off || group != 128 || IC % 128 != 0 || OC % 16 != 0 || ((uintptr_t)qweight % 32) || ((uintptr_t)qhigh % 16) || IC > (1 << 20))
        return BDQ_ERR_UNSUPPORTED;
    const int nwin = (int)((IC / 128 + 3) / 4);
    nf3s::Args a;
    a.qw = (const uint8_t*)qweight;
    a.qh = (const uint8_t*)qhigh;
    a.sc = (const __half*)scales;
    a.IC = (int)IC;
    a.wpw = (nwin + nf3s::kWarps - 1) / nf3s::kWarps;
    const Nf3Levels lv = nf3_levels_f16();
    for (int e = 0; e < 8; ++e) a.lk.c[e] = lv.lvl[e] * 32512.f;
    const int smem = nf3s::smem_bytes(a.wpw);
    if (smem > 112 * 1024) return BDQ_ERR_UNSUPPORTED;
    if (!ensure_dyn_smem(reinterpret_cast<const void*>(nf3s::gemv_nf3s_kernel), 112 * 1024)) return check_launch("bdq_gemv_nf3(attr)");
    const int total_rb = (int)(OC / 16);
    const int grid = std::min(total_rb, sm_count() * 2);
    a.rb_q = total_rb / grid;
    a.rb_r = total_rb % grid;
    for (int64_t b = 0; b < B; ++b) {   // batch rows one at a time (decode batches are 1..8 rows)
        a.x = (const __half*)x + b * IC;
        a.out = (__half*)out + b * OC;
        nf3s::gemv_nf3s_kernel<<<grid, nf3s::kThreads, smem, stream>>>(a);
    }
    return check_launch("bdq_gemv_nf3");
}

}  // namespace bdq
