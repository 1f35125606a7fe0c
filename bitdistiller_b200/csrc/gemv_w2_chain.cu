// Whole-token packed int2 decode chain in ONE persistent launch (sm_100a): the sequence of batch-1 GEMVs that
// WQLinear.forward / QuantLlamaMLP / QuantAttentionFused issue per decoded token (quantization/qmodule.py:166,
// inference/modules/fused_mlp.py:44-114, fused_attn.py:304-340 -> gemv_forward_cuda, gemv_cuda.cu:289-346), each phase =
// up to four matrices sharing one activation vector, phases executed in order with every phase seeing the outputs of all
// earlier ones -- exactly what one bdq_gemv_w2_multi launch per phase on one stream computes (same arithmetic, same bits).
//
// Why: a decode GEMV is 1-27 MB, shorter than the bandwidth-delay product of the chip; with one launch per phase ~3 us of
// every launch are serial (dependency resolution, x staging behind the launch's own weight requests, first item, block
// reduction + exit: profiles/r01_gemv_latency.md), 144 x 3 us = 430 of 690 us per Qwen3-8B token.  Here
//   * 2 x 148 CTAs stay resident for the whole token and walk a device-side list of phases;
//   * the dependency between phases is a grid barrier on an L2 counter (one red.release.gpu per CTA after its stores,
//     one ld.acquire.gpu poller per CTA) instead of a launch boundary;
//   * the weights do not depend on x: the register pipeline of the consumer (the bdq_gemv_w2_multi inner loop: LDG.256 per
//     lane, kDepth items = 6 KB per warp, INT8 digit-plane MMAs) has its own cursor that runs on into the following phases, so
//     the first items of phase p + 1 are in flight while phase p is reduced, stored and signalled and while the CTA waits;
//   * the whole plan lives in shared memory (256 B per phase): a phase change costs no dependent global load.
// An explicit HBM -> L2 prefetch (cp.async.bulk.prefetch.L2 of the next phases, first version of this file) made the token
// SLOWER (1014 vs 852 us: the prefetch requests queue in front of the demand loads, profiles/r02_chain_timeline.md) and was
// removed.  Every byte crosses HBM exactly once per token.
#include <cuda.h>
#include <cstring>
#include <mutex>

#include "bdq_common.cuh"

namespace bdq {
namespace chain {

constexpr int kThreads = 256;
constexpr int kWarps = 8;
constexpr int kDepth = 3;
constexpr int kSlotStride = 144;
constexpr int kMetaOff = 8 * kSlotStride;
constexpr int kWinStride = 1280;
constexpr int kZeroBytes = 256;
constexpr int kMaxMats = 4;
constexpr int kRedSlots = 8;
constexpr int kDescWords = 64;     // sizeof(Phase) / 4

struct __align__(16) Phase {       // one phase of the chain (device memory, written once by bdq_gemv_w2_chain_prepare)
    const uint8_t* qw[kMaxMats];
    const __half* sc[kMaxMats];
    const uint32_t* qz[kMaxMats];
    __half* out[kMaxMats];
    const __half* x;
    unsigned long long pad0;
    int rb_end[kMaxMats];
    unsigned int qw_bytes[kMaxMats];   // prefetch extents
    unsigned int sc_bytes[kMaxMats];
    unsigned int qz_bytes[kMaxMats];
    int IC, zw, wpw, silu;
    int rb_q, rb_r, nmat, wait;
    int pad1[4];
};
static_assert(sizeof(Phase) == kDescWords * 4, "Phase layout");

struct ItemBuf {
    U8 a, b;
    uint32_t sa, sb, za, zb;
};

__device__ __forceinline__ void imma16832(int* d, uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3, uint32_t b0, uint32_t b1) {
    asm("mma.sync.aligned.m16n8k32.row.col.s32.u8.s8.s32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
        : "+r"(d[0]), "+r"(d[1]), "+r"(d[2]), "+r"(d[3])
        : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}
// activations are written by other SMs during this launch: read them from L2 (.cg), never from this SM's L1
__device__ __forceinline__ uint4 ld_cg_v4(const void* p) {
    uint4 v;
    asm volatile("ld.global.cg.v4.b32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ uint32_t ldg_nc_u32(const void* p) {
    uint32_t v;
    asm volatile("ld.global.nc.b32 %0, [%1];" : "=r"(v) : "l"(p));
    return v;
}
__device__ __forceinline__ uint32_t ldg_nc_u16(const void* p) {
    uint16_t v;
    asm volatile("ld.global.nc.b16 %0, [%1];" : "=h"(v) : "l"(p));
    return v;
}
__device__ __forceinline__ unsigned int ld_acquire_u32(const unsigned int* p) {
    unsigned int v;
    asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void red_release_add(unsigned int* p, unsigned int v) {
    asm volatile("red.release.gpu.global.add.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ void prefetch_l2(const uint8_t* base, unsigned int nbytes, int part, int nparts) {
    // this CTA's 128-byte aligned slice of [base, base + nbytes), at most 16 KB per instruction
    if (!base || !nbytes) return;
    const unsigned int per = ((nbytes + nparts - 1) / nparts + 127u) & ~127u;
    unsigned int o = per * (unsigned int)part;
    const unsigned int end = o + per < nbytes ? o + per : nbytes;
    for (; o < end; o += 16384u) {
        const unsigned int n = end - o < 16384u ? (end - o) & ~15u : 16384u;
        if (n) asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(base + o), "r"(n) : "memory");
    }
}

__host__ __device__ inline int warp_bytes(int wpw) { return wpw * kWinStride + kZeroBytes; }
// [n_phases] plan copy | [8 warps] x windows + zero block | red
__host__ __device__ inline int smem_bytes(int wpw, int n_phases) {
    return n_phases * (int)sizeof(Phase) + kWarps * warp_bytes(wpw) + kRedSlots * kWarps * 64 * 4;
}

// counters: [n_phases] arrivals per phase, then [n_phases] = CTAs done with the whole chain (reset by the last one)
//
// Structure: ONE flat loop over this warp's items of ALL phases.  The producer side (register pipeline of kDepth items,
// LDG.256 per lane) has its own phase cursor and simply keeps going into the next phases -- the weights do not depend on
// x -- so the first items of phase p + 1 are in flight while phase p is still being consumed, reduced and signalled.
// The consumer side crosses a phase boundary with: flush + split-K reduction + stores, one CTA barrier, one
// red.release.gpu on the phase counter, [poll the counter of the phase it depends on], stage x.
__global__ void __launch_bounds__(kThreads, 2) gemv_w2_chain_kernel(const Phase* __restrict__ plan, int n_phases, int wpw_max,
                                                                     unsigned int* __restrict__ counters, int lookahead,
                                                                     unsigned long long* __restrict__ prof, int dbg) {
    extern __shared__ __align__(128) uint8_t smem[];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int g = lane >> 2, t = lane & 3;
    const int ncta = (int)gridDim.x, cta = (int)blockIdx.x;
    const Phase* splan = reinterpret_cast<const Phase*>(smem);
    uint8_t* wbase = smem + (size_t)n_phases * sizeof(Phase) + (size_t)warp * warp_bytes(wpw_max);
    uint8_t* zblock = wbase + (size_t)wpw_max * kWinStride;
    float* red = reinterpret_cast<float*>(smem + (size_t)n_phases * sizeof(Phase) + (size_t)kWarps * warp_bytes(wpw_max));

    for (int i = lane; i < kZeroBytes / 16; i += 32) reinterpret_cast<uint4*>(zblock)[i] = make_uint4(0, 0, 0, 0);
    for (int i = tid; i < n_phases * (int)(sizeof(Phase) / 16); i += kThreads)
        reinterpret_cast<uint4*>(smem)[i] = reinterpret_cast<const uint4*>(plan)[i];
    __syncthreads();

    // dev tool: %globaltimer marks of thread 0 of every CTA, prof[(cta * n_phases + p) * 8 + slot]
    auto mark = [&](int p, int slot) {
        if (prof && tid == 0) {
            unsigned long long v;
            asm volatile("mov.u64 %0, %globaltimer;" : "=l"(v));
            prof[((size_t)cta * n_phases + p) * 8 + slot] = v;
        }
    };
    auto rb_range = [&](const Phase& P, int& r0, int& r1) {
        const int rb_u = P.silu ? 2 : 1;
        r0 = (cta * P.rb_q + min(cta, P.rb_r)) * rb_u;
        r1 = r0 + (P.rb_q + (cta < P.rb_r ? 1 : 0)) * rb_u;
    };
    auto mat_of = [&](const Phase& P, int rb) {
        if (P.silu) return rb & 1;
        int mi = 0;
        while (mi < kMaxMats - 1 && rb >= P.rb_end[mi]) ++mi;
        return mi;
    };
    auto blk_of = [&](const Phase& P, int rb, int mi) { return P.silu ? (rb >> 1) : rb - (mi ? P.rb_end[mi - 1] : 0); };
    auto mywin_of = [&](const Phase& P) {
        const int nwin = ((P.IC >> 7) + 3) >> 2;
        return warp < nwin ? (nwin - warp + kWarps - 1) / kWarps : 0;
    };

    // ---------------------------------------------------------------- producer: cursor over (phase, block, window)
    const int G0 = 4 * warp + t;
    int q_p = -1, q_rb = 0, q_rb1 = 0, q_j = 0, q_mi = 0, q_mywin = 0, q_ngrp = 0, q_rowb = 0, q_zw = 0;
    const uint8_t* pa0 = nullptr;
    const __half* ps0 = nullptr;
    const uint32_t* pz0 = nullptr;
    auto q_set_base = [&]() {
        const Phase& P = splan[q_p];
        const size_t r = (size_t)blk_of(P, q_rb, q_mi) * 16 + g;
        pa0 = P.qw[q_mi] + r * q_rowb + (size_t)G0 * 32;
        ps0 = P.sc[q_mi] + r * (q_zw * 16) + G0;
        pz0 = P.qz[q_mi] + r * q_zw + (G0 >> 4);
    };
    auto q_next_phase = [&]() {   // move to the next phase in which this warp has items; false at the end of the chain
        while (++q_p < n_phases) {
            const Phase& P = splan[q_p];
            q_mywin = mywin_of(P);
            rb_range(P, q_rb, q_rb1);
            if (q_mywin > 0 && q_rb < q_rb1) {
                q_ngrp = P.IC >> 7;
                q_rowb = P.IC >> 2;
                q_zw = P.zw;
                q_j = 0;
                q_mi = mat_of(P, q_rb);
                q_set_base();
                return true;
            }
        }
        q_p = n_phases;
        return false;
    };
    auto issue = [&](ItemBuf& b) {
        if (q_p >= n_phases) return;
        if (G0 + 32 * q_j < q_ngrp) {
            const uint8_t* pa = pa0 + q_j * 1024;
            b.a = ldg256(pa);
            b.b = ldg256(pa + (size_t)8 * q_rowb);
        } else {
#pragma unroll
            for (int i = 0; i < 8; ++i) b.a.r[i] = b.b.r[i] = 0u;
        }
        const __half* ps = ps0 + 32 * q_j;
        b.sa = ldg_nc_u16(ps);
        b.sb = ldg_nc_u16(ps + (size_t)8 * (q_zw * 16));
        const uint32_t* pz = pz0 + 2 * q_j;
        b.za = ldg_nc_u32(pz);
        b.zb = ldg_nc_u32(pz + (size_t)8 * q_zw);
        if (++q_j == q_mywin) {
            q_j = 0;
            if (++q_rb >= q_rb1) {
                q_next_phase();
            } else {
                const Phase& P = splan[q_p];
                if (P.silu || q_rb >= P.rb_end[q_mi]) {
                    q_mi = mat_of(P, q_rb);
                    q_set_base();
                } else {
                    pa0 += (size_t)16 * q_rowb;
                    ps0 += (size_t)16 * (q_zw * 16);
                    pz0 += (size_t)16 * q_zw;
                }
            }
        }
    };

    // ---------------------------------------------------------------- consumer state
    int p = 0;                         // phase being consumed (CTA-uniform)
    int c_left = 0, c_j = 0, c_mywin = 0, rslot = 0, rb_base = 0, c_rb1 = 0;
    float fin_lo = 0.f, fin_hi = 0.f;
    const bool active = ((g >> 1) == t);
    const uint8_t* bbase = active ? wbase + g * kSlotStride : zblock;
    const int jstride = active ? kWinStride : 0;

    auto block_reduce = [&]() {
        const Phase& D = splan[p];
        __syncthreads();
        if (tid < rslot * 16) {
            const int sl = tid >> 4, row = tid & 15;
            const float* rp = red + sl * kWarps * 64 + row;
            float acc = 0.f;   // same serial order as bdq_gemv_w2_multi: the chain is bit-equal to the launch sequence
#pragma unroll
            for (int w = 0; w < kWarps * 4; ++w) acc += rp[w * 16];
            const int rb = rb_base + sl;
            const int mi = mat_of(D, rb);
            __half hv = __float2half_rn(acc);
            bool store = true;
            if (D.silu) {
                const float up = __shfl_down_sync(0xffffffffu, __half2float(hv), 16);
                const float gt = __half2float(hv);
                const float sg = __half2float(__float2half_rn(gt / (1.f + __expf(-gt))));
                hv = __float2half_rn(sg * up);
                store = (sl & 1) == 0;
            }
            const size_t oi = (size_t)blk_of(D, rb, mi) * 16 + row;
            if (store) D.out[mi][oi] = hv;
        }
        rb_base += rslot;
        rslot = 0;
        if (rb_base < c_rb1) __syncthreads();
    };
    auto flush = [&]() {
        float* my = red + ((rslot * kWarps + warp) * 4 + t) * 16;
        my[g] = fin_lo;
        my[g + 8] = fin_hi;
        fin_lo = fin_hi = 0.f;
        if (++rslot == kRedSlots) block_reduce();
    };
    auto consume = [&](const ItemBuf& b) {
        const uint8_t* bp = bbase + (size_t)c_j * jstride;
        int acc[4][4];
#pragma unroll
        for (int i = 0; i < 4; ++i) acc[i][0] = acc[i][1] = acc[i][2] = acc[i][3] = 0;
#pragma unroll
        for (int p2 = 0; p2 < 4; ++p2) {
            const uint4 q0 = *reinterpret_cast<const uint4*>(bp + p2 * 32);
            const uint4 q1 = *reinterpret_cast<const uint4*>(bp + p2 * 32 + 16);
            const uint32_t a0 = b.a.r[2 * p2], a1 = b.b.r[2 * p2], a2 = b.a.r[2 * p2 + 1], a3 = b.b.r[2 * p2 + 1];
            imma16832(acc[0], a0 & 0x03030303u, a1 & 0x03030303u, a2 & 0x03030303u, a3 & 0x03030303u, q0.x, q0.y);
            imma16832(acc[1], a0 & 0x0C0C0C0Cu, a1 & 0x0C0C0C0Cu, a2 & 0x0C0C0C0Cu, a3 & 0x0C0C0C0Cu, q0.z, q0.w);
            imma16832(acc[2], a0 & 0x30303030u, a1 & 0x30303030u, a2 & 0x30303030u, a3 & 0x30303030u, q1.x, q1.y);
            imma16832(acc[3], a0 & 0xC0C0C0C0u, a1 & 0xC0C0C0C0u, a2 & 0xC0C0C0C0u, a3 & 0xC0C0C0C0u, q1.z, q1.w);
        }
        const float2 mt = *reinterpret_cast<const float2*>(wbase + (size_t)c_j * kWinStride + kMetaOff + t * 8);
        const int sq64 = __float_as_int(mt.y);
        const int sh = ((4 * (warp + kWarps * c_j) + t) & 15) * 2;
        const int za = (int)((b.za >> sh) & 3u), zb = (int)((b.zb >> sh) & 3u);
        int u[4];
#pragma unroll
        for (int c = 0; c < 4; ++c) u[c] = ((acc[0][c] * 4 + acc[1][c]) * 4 + acc[2][c]) * 4 + acc[3][c];
        const int wa = u[0] * 256 + u[1] - za * sq64, wb = u[2] * 256 + u[3] - zb * sq64;
        const float sa = __half2float(__ushort_as_half((unsigned short)b.sa)) * mt.x;
        const float sb = __half2float(__ushort_as_half((unsigned short)b.sb)) * mt.x;
        fin_lo = fmaf(sa, (float)wa, fin_lo);
        fin_hi = fmaf(sb, (float)wb, fin_hi);
    };

    // stage this warp's windows of x as block fixed point (two int8 limbs per activation), see gemv_w2_stream.cu
    auto stage_x = [&](const Phase& D) {
        const __half* x = D.x;
        const int IC = D.IC;
        for (int j0 = 0; j0 < c_mywin; j0 += 4) {
            uint4 raw[4][2];
#pragma unroll
            for (int jj = 0; jj < 4; ++jj) {
                const int k0 = (warp + kWarps * (j0 + jj)) * 512 + lane * 16;
                raw[jj][0] = raw[jj][1] = make_uint4(0, 0, 0, 0);
                if (j0 + jj < c_mywin && k0 < IC) {
                    raw[jj][0] = ld_cg_v4(x + k0);
                    raw[jj][1] = ld_cg_v4(x + k0 + 8);
                }
            }
#pragma unroll
            for (int jj = 0; jj < 4; ++jj) {
                if (j0 + jj >= c_mywin) break;
                const uint32_t in[8] = {raw[jj][0].x, raw[jj][0].y, raw[jj][0].z, raw[jj][0].w,
                                        raw[jj][1].x, raw[jj][1].y, raw[jj][1].z, raw[jj][1].w};
                __half2 m2 = __habs2(*reinterpret_cast<const __half2*>(&in[0]));
#pragma unroll
                for (int e = 1; e < 8; ++e) m2 = __hmax2_nan(m2, __habs2(*reinterpret_cast<const __half2*>(&in[e])));
                uint32_t mx = *reinterpret_cast<const uint32_t*>(&m2);
                mx = max(mx & 0xffffu, mx >> 16);
#pragma unroll
                for (int sft = 1; sft < 8; sft <<= 1) mx = max(mx, __shfl_xor_sync(0xffffffffu, mx, sft));
                const uint32_t ex = (mx >> 10) & 31u;
                const float gm = fmaxf(__half2float(__ushort_as_half((unsigned short)mx)), 5.9604645e-8f);   // group max |x| (exact)
                const float inv_unit = __fdividef(32512.f, gm);
                const float unit = ex == 31u ? __uint_as_float(0x7fc00000u) : gm * (1.f / 32512.f);
                uint32_t bits[16], qsum = 0;
#pragma unroll
                for (int e = 0; e < 8; ++e) {
                    const float2 v = __half22float2(*reinterpret_cast<const __half2*>(&in[e]));
                    bits[2 * e] = __float_as_uint(fmaf(v.x, inv_unit, 12583040.f));
                    bits[2 * e + 1] = __float_as_uint(fmaf(v.y, inv_unit, 12583040.f));
                    qsum += bits[2 * e] + bits[2 * e + 1];
                }
                uint32_t hi[4], lo[4];
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    const uint32_t p01 = __byte_perm(bits[i], bits[i + 4], 0x5140);
                    const uint32_t p23 = __byte_perm(bits[i + 8], bits[i + 12], 0x5140);
                    lo[i] = __byte_perm(p01, p23, 0x5410) ^ 0x80808080u;
                    hi[i] = __byte_perm(p01, p23, 0x7632);
                }
                uint8_t* wj = wbase + (size_t)(j0 + jj) * kWinStride;
                const int grp = lane >> 3, w = lane & 7;
                uint8_t* dst = wj + (2 * grp) * kSlotStride + (w >> 1) * 32 + (w & 1) * 4;
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    *reinterpret_cast<uint32_t*>(dst + i * 8) = hi[i];
                    *reinterpret_cast<uint32_t*>(dst + kSlotStride + i * 8) = lo[i];
                }
#pragma unroll
                for (int sft = 1; sft < 8; sft <<= 1) qsum += __shfl_xor_sync(0xffffffffu, qsum, sft);
                if (w == 0)
                    reinterpret_cast<float2*>(wj + kMetaOff)[grp] =
                        make_float2(unit * 0.015625f, __int_as_float((int)(qsum - 0xA0004000u) * 64));
            }
        }
        __syncwarp();
    };

    // wait for the dependency of phase p, stage x, set the consumer counters
    auto begin_phase = [&]() {
        const Phase& D = splan[p];
        mark(p, 0);
        int r0;
        rb_range(D, r0, c_rb1);
        c_mywin = mywin_of(D);
        rb_base = r0;
        rslot = 0;
        c_j = 0;
        c_left = (c_rb1 - r0) * c_mywin;
        if (D.wait && p > 0 && !(dbg & 1)) {
            if (tid == 0) {
                const unsigned int* c = counters + (p - 1);
                while (ld_acquire_u32(c) < (unsigned int)ncta) {}
            }
            __syncthreads();
        }
        mark(p, 2);
        stage_x(D);
        mark(p, 3);
        if (c_mywin == 0)   // IC < 4096: this warp owns no window but still takes part in the block reductions
            for (int rb = r0; rb < c_rb1; ++rb) flush();
    };
    // all items of phase p consumed: reduce what is parked, publish
    auto end_phase = [&]() {
        if (rslot) block_reduce();
        mark(p, 4);
        __syncthreads();                           // every thread's stores of this phase precede ...
        if (tid == 0) red_release_add(counters + p, 1u);   // ... this gpu-scope release (cumulative over the CTA barrier)
        mark(p, 5);
    };

    ItemBuf buf[kDepth];
    q_next_phase();
#pragma unroll
    for (int d = 0; d < kDepth; ++d) issue(buf[d]);
    begin_phase();
    bool done = false;
    while (!done) {
#pragma unroll
        for (int d = 0; d < kDepth; ++d) {
            while (c_left == 0) {
                end_phase();
                if (++p >= n_phases) { done = true; break; }
                begin_phase();
            }
            if (done) break;
            consume(buf[d]);
            issue(buf[d]);
            --c_left;
            if (++c_j == c_mywin) {
                c_j = 0;
                flush();
            }
        }
    }

    // the last CTA out resets the counters for the next token (every CTA has passed all of its waits by then)
    if (tid == 0) {
        __threadfence();
        const unsigned int prev = atomicAdd(counters + n_phases, 1u);
        if (prev == (unsigned int)ncta - 1) {
            for (int i = 0; i <= n_phases; ++i) counters[i] = 0u;
            __threadfence();
        }
    }
    (void)lookahead;
}

}  // namespace chain
}  // namespace bdq

using namespace bdq;

static unsigned long long* g_chain_prof = nullptr;
static int64_t g_chain_prof_cap = 0;
// dev tool: device buffer of `cap` u64; every later run writes [grid][n_phases][8] %globaltimer marks into it
extern "C" int bdq_gemv_w2_chain_set_prof(void* buf, int64_t cap) {
    g_chain_prof = (unsigned long long*)buf;
    g_chain_prof_cap = cap;
    return BDQ_OK;
}

// phases one run can hold (the plan is copied into shared memory next to the x windows; two CTAs per SM)
extern "C" int bdq_gemv_w2_chain_max_phases(int64_t max_IC) {
    if (max_IC <= 0 || max_IC % 128 != 0) return 0;
    const int nwin = (int)((max_IC / 128 + 3) / 4);
    const int wpw = (nwin + chain::kWarps - 1) / chain::kWarps;
    const int fixed = chain::smem_bytes(wpw, 0);
    return fixed >= 112 * 1024 ? 0 : (112 * 1024 - fixed) / (int)sizeof(chain::Phase);
}

extern "C" int64_t bdq_gemv_w2_chain_workspace_bytes(int n_phases) {
    if (n_phases < 1) return 0;
    return (int64_t)n_phases * (int64_t)sizeof(chain::Phase) + (int64_t)(n_phases + 1) * 4 + 256;
}

// Builds the device-side plan (synchronous, once per model): phases[i] as documented in include/bdq.h.
extern "C" int bdq_gemv_w2_chain_prepare(const bdq_gemv_phase_t* phases, int n_phases, void* workspace, int64_t workspace_bytes,
                                         bdq_stream_t stream) {
    if (!phases || n_phases < 1 || !workspace) { set_error("gemv_w2_chain: null arguments / no phases"); return BDQ_ERR_SHAPE; }
    if (workspace_bytes < bdq_gemv_w2_chain_workspace_bytes(n_phases) || ((uintptr_t)workspace % 256)) {
        set_error("gemv_w2_chain: workspace must be 256-byte aligned and >= bdq_gemv_w2_chain_workspace_bytes(n_phases)");
        return BDQ_ERR_SHAPE;
    }
    const int grid = sm_count() * 2;
    chain::Phase* host = (chain::Phase*)calloc((size_t)n_phases, sizeof(chain::Phase));
    if (!host) { set_error("gemv_w2_chain: out of host memory"); return BDQ_ERR_CUDA; }
    int rc = BDQ_OK;
    for (int p = 0; p < n_phases && rc == BDQ_OK; ++p) {
        const bdq_gemv_phase_t& s = phases[p];
        chain::Phase& d = host[p];
        const int64_t IC = s.IC;
        const bool silu = (s.flags & BDQ_GEMV_SILU_MUL) != 0;
        if (s.nmat < 1 || s.nmat > chain::kMaxMats || !s.x || IC <= 0 || IC % 128 != 0 || IC > (1 << 20) || ((uintptr_t)s.x % 16) ||
            s.group != 128 || (silu && (s.nmat != 2 || s.OC[0] != s.OC[1]))) {
            set_error("gemv_w2_chain: phase %d: need 1..4 matrices, group 128, IC %% 128 == 0, 16-byte aligned x", p);
            rc = BDQ_ERR_UNSUPPORTED;
            break;
        }
        const int zw = bdq_zeros_width(IC, 128, 2);
        int total_rb = 0;
        for (int i = 0; i < chain::kMaxMats; ++i) {
            const int k = i < s.nmat ? i : s.nmat - 1;
            if (i < s.nmat) {
                if (s.OC[i] <= 0 || s.OC[i] % 16 != 0 || !s.qweight[i] || !s.scales[i] || !s.qzeros[i] || (!s.out[i] && !(silu && i == 1)) ||
                    ((uintptr_t)s.qweight[i] % 32) || ((uintptr_t)s.scales[i] % 4) || ((uintptr_t)s.qzeros[i] % 4)) {
                    set_error("gemv_w2_chain: phase %d matrix %d: OC %% 16, non-null, 32-byte aligned qweight required", p, i);
                    rc = BDQ_ERR_UNSUPPORTED;
                    break;
                }
                total_rb += (int)(s.OC[i] / 16);
                d.qw_bytes[i] = (unsigned int)(s.OC[i] * (IC / 4));
                d.sc_bytes[i] = (unsigned int)(s.OC[i] * zw * 32);
                d.qz_bytes[i] = (unsigned int)(s.OC[i] * zw * 4);
            }
            d.qw[i] = (const uint8_t*)s.qweight[k];
            d.sc[i] = (const __half*)s.scales[k];
            d.qz[i] = (const uint32_t*)s.qzeros[k];
            d.out[i] = (__half*)s.out[k];
            d.rb_end[i] = total_rb;
        }
        const int nwin = (int)((IC / 128 + 3) / 4);
        d.x = (const __half*)s.x;
        d.IC = (int)IC;
        d.zw = zw;
        d.wpw = (nwin + chain::kWarps - 1) / chain::kWarps;
        d.silu = silu ? 1 : 0;
        const int units = silu ? total_rb / 2 : total_rb;
        d.rb_q = units / grid;
        d.rb_r = units % grid;
        d.nmat = s.nmat;
        d.wait = (s.flags & BDQ_GEMV_CHAIN_NOWAIT) ? 0 : 1;
    }
    if (rc == BDQ_OK) {
        cudaStream_t st = (cudaStream_t)stream;
        uint8_t* ws = (uint8_t*)workspace;
        const size_t plan_bytes = (size_t)n_phases * sizeof(chain::Phase);
        if (cudaMemcpyAsync(ws, host, plan_bytes, cudaMemcpyHostToDevice, st) != cudaSuccess ||
            cudaMemsetAsync(ws + ((plan_bytes + 255) & ~(size_t)255), 0, (size_t)(n_phases + 1) * 4, st) != cudaSuccess ||
            cudaStreamSynchronize(st) != cudaSuccess)
            rc = check_launch("bdq_gemv_w2_chain_prepare");
    }
    free(host);
    return rc;
}

// One persistent cooperative launch for the whole chain.  `wpw_max` / `n_phases` must describe the prepared plan.
extern "C" int bdq_gemv_w2_chain_run(const void* workspace, int n_phases, int64_t max_IC, int lookahead, bdq_stream_t stream) {
    if (!workspace || n_phases < 1 || max_IC <= 0 || max_IC % 128 != 0) { set_error("gemv_w2_chain_run: bad arguments"); return BDQ_ERR_SHAPE; }
    const int nwin = (int)((max_IC / 128 + 3) / 4);
    const int wpw = (nwin + chain::kWarps - 1) / chain::kWarps;
    const int smem = chain::smem_bytes(wpw, n_phases);
    if (smem > 112 * 1024) {
        set_error("gemv_w2_chain_run: %d phases at max IC %lld need %d B of shared memory (> 112 KB, two CTAs per SM): split the chain",
                  n_phases, (long long)max_IC, smem);
        return BDQ_ERR_UNSUPPORTED;
    }
    const void* kern = reinterpret_cast<const void*>(chain::gemv_w2_chain_kernel);
    if (!ensure_dyn_smem(kern, 112 * 1024)) return check_launch("bdq_gemv_w2_chain(attr)");
    int occ = 0;
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, chain::gemv_w2_chain_kernel, chain::kThreads, (size_t)smem) != cudaSuccess || occ < 2) {
        set_error("gemv_w2_chain_run: the kernel needs two resident CTAs per SM (occupancy %d)", occ);
        return BDQ_ERR_UNSUPPORTED;
    }
    const size_t plan_bytes = (size_t)n_phases * sizeof(chain::Phase);
    const chain::Phase* plan = (const chain::Phase*)workspace;
    unsigned int* counters = (unsigned int*)((const uint8_t*)workspace + ((plan_bytes + 255) & ~(size_t)255));
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)(sm_count() * 2));
    cfg.blockDim = dim3(chain::kThreads);
    cfg.dynamicSmemBytes = (size_t)smem;
    cfg.stream = (cudaStream_t)stream;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeCooperative;   // all CTAs co-resident or the launch fails (the grid barrier spins)
    at[0].val.cooperative = 1;
    cfg.attrs = at;
    cfg.numAttrs = 1;
    static int dbg = -1;   // BDQ_CHAIN_DEBUG (dev experiments only): 1 = skip the barrier waits (wrong results), 2 = no L2 prefetch
    if (dbg < 0) { const char* e = getenv("BDQ_CHAIN_DEBUG"); dbg = e ? atoi(e) : 0; }
    unsigned long long* prof = (g_chain_prof && g_chain_prof_cap >= (int64_t)sm_count() * 2 * n_phases * 8) ? g_chain_prof : nullptr;
    cudaLaunchKernelEx(&cfg, chain::gemv_w2_chain_kernel, plan, n_phases, wpw, counters, lookahead, prof, dbg);
    return check_launch("bdq_gemv_w2_chain_run");
}

// ------------------------------------------------------------------------------------------------------------------------
// bdq_gemv_w2_chain_static: the same persistent kernel for callers that cannot allocate device memory (no counterpart in
// the reference -- include/bdq.h).  The kernel's scratch -- the barrier
// counters and the device-side plan, which bdq_gemv_w2_chain_prepare / _run keep in a caller-provided workspace -- lives in
// a file-scope __device__ array of BDQ_GEMV_CHAIN_STATIC_MAX_PHASES phases (16.5 KB, held by every GPU that loads the
// library until exit).  No cudaMalloc / cudaFree / host allocation anywhere on this path: the plan is built in a stack
// array, copied with cudaMemcpyAsync and the counters are cleared with cudaMemsetAsync on the caller's stream.
// All calls on a device share the one array: two calls that overlap on the GPU (different streams) race on it.
// The existing entry points above are untouched; the phase validation below repeats theirs on purpose.
namespace bdq {
namespace chain {

constexpr int kStaticMaxPhases = BDQ_GEMV_CHAIN_STATIC_MAX_PHASES;
constexpr int kStaticCounterBytes = 512;   // (kStaticMaxPhases + 1) u32 counters, padded so that the plan stays 256-aligned
constexpr int kStaticBytes = kStaticCounterBytes + kStaticMaxPhases * (int)sizeof(Phase);
static_assert((kStaticMaxPhases + 1) * 4 <= kStaticCounterBytes, "counter block");

static __device__ __align__(256) unsigned char g_static_scratch[kStaticBytes];

constexpr int kMaxDevices = 64;
static std::mutex g_static_mu;
static unsigned char* g_static_ptr[kMaxDevices];   // device address of g_static_scratch on device i (the host-side name is NOT one)

// once per device, with that device current
static unsigned char* static_scratch_ptr() {
    int dev = -1;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= kMaxDevices) {
        set_error("gemv_w2_chain_static: cudaGetDevice failed / device index beyond %d", kMaxDevices - 1);
        return nullptr;
    }
    std::lock_guard<std::mutex> lk(g_static_mu);
    if (!g_static_ptr[dev]) {
        unsigned char* p = nullptr;
        const cudaError_t e = cudaGetSymbolAddress((void**)&p, g_static_scratch);
        if (e != cudaSuccess || !p) {
            set_error("gemv_w2_chain_static: cudaGetSymbolAddress: %s", cudaGetErrorString(e));
            (void)cudaGetLastError();
            return nullptr;
        }
        g_static_ptr[dev] = p;
    }
    return g_static_ptr[dev];
}

// one bdq_gemv_phase_t -> one device-side Phase (same rules as bdq_gemv_w2_chain_prepare)
static int fill_phase_static(const bdq_gemv_phase_t& s, int p, int grid, Phase& d) {
    memset(&d, 0, sizeof(d));
    const int64_t IC = s.IC;
    const bool silu = (s.flags & BDQ_GEMV_SILU_MUL) != 0;
    if (s.nmat < 1 || s.nmat > kMaxMats || !s.x || IC <= 0 || IC % 128 != 0 || IC > (1 << 20) || ((uintptr_t)s.x % 16) || s.group != 128 ||
        (silu && (s.nmat != 2 || s.OC[0] != s.OC[1]))) {
        set_error("gemv_w2_chain_static: phase %d: need 1..4 matrices, group 128, IC %% 128 == 0, 16-byte aligned x", p);
        return BDQ_ERR_UNSUPPORTED;
    }
    const int zw = bdq_zeros_width(IC, 128, 2);
    int total_rb = 0;
    for (int i = 0; i < kMaxMats; ++i) {
        const int k = i < s.nmat ? i : s.nmat - 1;
        if (i < s.nmat) {
            if (s.OC[i] <= 0 || s.OC[i] % 16 != 0 || !s.qweight[i] || !s.scales[i] || !s.qzeros[i] || (!s.out[i] && !(silu && i == 1)) ||
                ((uintptr_t)s.qweight[i] % 32) || ((uintptr_t)s.scales[i] % 4) || ((uintptr_t)s.qzeros[i] % 4)) {
                set_error("gemv_w2_chain_static: phase %d matrix %d: OC %% 16, non-null, 32-byte aligned qweight required", p, i);
                return BDQ_ERR_UNSUPPORTED;
            }
            total_rb += (int)(s.OC[i] / 16);
            d.qw_bytes[i] = (unsigned int)(s.OC[i] * (IC / 4));
            d.sc_bytes[i] = (unsigned int)(s.OC[i] * zw * 32);
            d.qz_bytes[i] = (unsigned int)(s.OC[i] * zw * 4);
        }
        d.qw[i] = (const uint8_t*)s.qweight[k];
        d.sc[i] = (const __half*)s.scales[k];
        d.qz[i] = (const uint32_t*)s.qzeros[k];
        d.out[i] = (__half*)s.out[k];
        d.rb_end[i] = total_rb;
    }
    const int nwin = (int)((IC / 128 + 3) / 4);
    d.x = (const __half*)s.x;
    d.IC = (int)IC;
    d.zw = zw;
    d.wpw = (nwin + kWarps - 1) / kWarps;
    d.silu = silu ? 1 : 0;
    const int units = silu ? total_rb / 2 : total_rb;
    d.rb_q = units / grid;
    d.rb_r = units % grid;
    d.nmat = s.nmat;
    d.wait = (s.flags & BDQ_GEMV_CHAIN_NOWAIT) ? 0 : 1;
    return BDQ_OK;
}

}  // namespace chain
}  // namespace bdq

extern "C" int bdq_gemv_w2_chain_static(const bdq_gemv_phase_t* phases, int n_phases, int lookahead, bdq_stream_t stream) {
    if (!phases || n_phases < 1) { set_error("gemv_w2_chain_static: null arguments / no phases"); return BDQ_ERR_SHAPE; }
    if (n_phases > chain::kStaticMaxPhases) {
        set_error("gemv_w2_chain_static: %d phases exceed the static scratch (BDQ_GEMV_CHAIN_STATIC_MAX_PHASES = %d): split the chain "
                  "or use bdq_gemv_w2_chain_prepare / _run with a caller-provided workspace", n_phases, chain::kStaticMaxPhases);
        return BDQ_ERR_SHAPE;
    }
    const int grid = sm_count() * 2;
    chain::Phase host[chain::kStaticMaxPhases];   // 16 KB of stack: no host allocation either
    int wpw = 1;
    for (int p = 0; p < n_phases; ++p) {
        const int rc = chain::fill_phase_static(phases[p], p, grid, host[p]);
        if (rc != BDQ_OK) return rc;
        wpw = host[p].wpw > wpw ? host[p].wpw : wpw;
    }
    const int smem = chain::smem_bytes(wpw, n_phases);
    if (smem > 112 * 1024) {
        set_error("gemv_w2_chain_static: %d phases with these in_features need %d B of shared memory (> 112 KB, two CTAs per SM): "
                  "split the chain", n_phases, smem);
        return BDQ_ERR_UNSUPPORTED;
    }
    unsigned char* scratch = chain::static_scratch_ptr();
    if (!scratch) return BDQ_ERR_CUDA;
    const void* kern = reinterpret_cast<const void*>(chain::gemv_w2_chain_kernel);
    if (!ensure_dyn_smem(kern, 112 * 1024)) return check_launch("bdq_gemv_w2_chain_static(attr)");
    int occ = 0;
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, chain::gemv_w2_chain_kernel, chain::kThreads, (size_t)smem) != cudaSuccess || occ < 2) {
        set_error("gemv_w2_chain_static: the kernel needs two resident CTAs per SM (occupancy %d)", occ);
        return BDQ_ERR_UNSUPPORTED;
    }
    cudaStream_t st = (cudaStream_t)stream;
    unsigned int* counters = reinterpret_cast<unsigned int*>(scratch);
    const chain::Phase* plan = reinterpret_cast<const chain::Phase*>(scratch + chain::kStaticCounterBytes);
    // the array keeps the previous call's contents: the counters must start at zero, the plan is overwritten.  `host` is
    // pageable stack memory, so cudaMemcpyAsync has staged it when it returns (stream-ordered behind the previous call).
    cudaError_t e = cudaMemsetAsync(counters, 0, chain::kStaticCounterBytes, st);
    if (e == cudaSuccess)
        e = cudaMemcpyAsync(scratch + chain::kStaticCounterBytes, host, (size_t)n_phases * sizeof(chain::Phase), cudaMemcpyHostToDevice, st);
    if (e != cudaSuccess) {
        set_error("gemv_w2_chain_static: staging the plan: %s", cudaGetErrorString(e));
        (void)cudaGetLastError();
        return BDQ_ERR_CUDA;
    }
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)grid);
    cfg.blockDim = dim3(chain::kThreads);
    cfg.dynamicSmemBytes = (size_t)smem;
    cfg.stream = st;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeCooperative;
    at[0].val.cooperative = 1;
    cfg.attrs = at;
    cfg.numAttrs = 1;
    cudaLaunchKernelEx(&cfg, chain::gemv_w2_chain_kernel, plan, n_phases, wpw, counters, lookahead, (unsigned long long*)nullptr, 0);
    return check_launch("bdq_gemv_w2_chain_static");
}
