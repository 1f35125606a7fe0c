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
//     one ld.acquire.gpu poller per CTA): ~0.4 us instead of a launch boundary;
//   * the weights do not depend on x: one lane per CTA keeps `lookahead` phases of packed weights flowing HBM -> L2 with
//     cp.async.bulk.prefetch.L2 while the CTAs wait / stage x, so HBM streams across the barriers and the register
//     pipeline of the consumer (the bdq_gemv_w2_multi inner loop: LDG.256 per lane, INT8 digit-plane MMAs) reads L2 hits;
//   * the first items of a phase are requested BEFORE the barrier wait.
// Every byte still crosses HBM exactly once per token.
#include <cuda.h>
#include <cstring>

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
__host__ __device__ inline int smem_bytes(int wpw) { return kWarps * warp_bytes(wpw) + kRedSlots * kWarps * 64 * 4 + 2 * kDescWords * 4; }

// counters: [n_phases] arrivals per phase, then [n_phases] = CTAs done with the whole chain (reset by the last one)
__global__ void __launch_bounds__(kThreads, 2) gemv_w2_chain_kernel(const Phase* __restrict__ plan, int n_phases, int wpw_max,
                                                                     unsigned int* __restrict__ counters, int lookahead) {
    extern __shared__ __align__(128) uint8_t smem[];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int g = lane >> 2, t = lane & 3;
    const int ncta = (int)gridDim.x, cta = (int)blockIdx.x;
    uint8_t* wbase = smem + (size_t)warp * warp_bytes(wpw_max);
    uint8_t* zblock = wbase + (size_t)wpw_max * kWinStride;
    float* red = reinterpret_cast<float*>(smem + (size_t)kWarps * warp_bytes(wpw_max));
    uint32_t* sdesc = reinterpret_cast<uint32_t*>(red + kRedSlots * kWarps * 64);   // [2][kDescWords]

    for (int i = lane; i < kZeroBytes / 16; i += 32) reinterpret_cast<uint4*>(zblock)[i] = make_uint4(0, 0, 0, 0);
    if (tid < kDescWords) sdesc[tid] = reinterpret_cast<const uint32_t*>(plan)[tid];
    // HBM -> L2 for the first `lookahead` phases (phase p + lookahead is requested when phase p starts)
    if (tid == kThreads - 1) {
        for (int q = 0; q < lookahead && q < n_phases; ++q) {
            const Phase* P = plan + q;
            for (int i = 0; i < P->nmat; ++i) {
                prefetch_l2(P->qw[i], P->qw_bytes[i], cta, ncta);
                prefetch_l2(reinterpret_cast<const uint8_t*>(P->sc[i]), P->sc_bytes[i], cta, ncta);
                prefetch_l2(reinterpret_cast<const uint8_t*>(P->qz[i]), P->qz_bytes[i], cta, ncta);
            }
        }
    }
    __syncthreads();

    for (int p = 0; p < n_phases; ++p) {
        const Phase& D = *reinterpret_cast<const Phase*>(sdesc + (p & 1) * kDescWords);
        uint32_t next_word = 0;
        if (tid < kDescWords && p + 1 < n_phases) next_word = reinterpret_cast<const uint32_t*>(plan + p + 1)[tid];
        const int IC = D.IC, zw = D.zw, silu = D.silu;
        const int ngrp = IC >> 7, nwin = (ngrp + 3) >> 2;
        const int row_bytes = IC >> 2, zw16 = zw * 16;
        const int rb_u = silu ? 2 : 1;
        const int rb0 = (cta * D.rb_q + min(cta, D.rb_r)) * rb_u;
        const int rb1 = rb0 + (D.rb_q + (cta < D.rb_r ? 1 : 0)) * rb_u;
        auto mat_of = [&](int rb) {
            if (silu) return rb & 1;
            int mi = 0;
            while (mi < kMaxMats - 1 && rb >= D.rb_end[mi]) ++mi;
            return mi;
        };
        auto rstart_of = [&](int mi) { return mi ? D.rb_end[mi - 1] : 0; };
        auto blk_of = [&](int rb, int mi) { return silu ? (rb >> 1) : rb - rstart_of(mi); };
        const int mywin = warp < nwin ? (nwin - warp + kWarps - 1) / kWarps : 0;

        // ---- producer cursor (register pipeline of kDepth items, see gemv_w2_stream.cu)
        int p_rb = rb0, p_j = 0, p_mi = mat_of(rb0);
        const int G0 = 4 * warp + t;
        const uint8_t* pa0 = nullptr;
        const __half* ps0 = nullptr;
        const uint32_t* pz0 = nullptr;
        auto set_base = [&]() {
            const size_t r = (size_t)blk_of(p_rb, p_mi) * 16 + g;
            pa0 = D.qw[p_mi] + r * row_bytes + (size_t)G0 * 32;
            ps0 = D.sc[p_mi] + r * zw16 + G0;
            pz0 = D.qz[p_mi] + r * zw + (G0 >> 4);
        };
        if (rb0 < rb1) set_base();
        auto issue = [&](ItemBuf& b) {
            if (p_rb >= rb1) return;
            if (G0 + 32 * p_j < ngrp) {
                const uint8_t* pa = pa0 + p_j * 1024;
                b.a = ldg256(pa);
                b.b = ldg256(pa + (size_t)8 * row_bytes);
            } else {
#pragma unroll
                for (int i = 0; i < 8; ++i) b.a.r[i] = b.b.r[i] = 0u;
            }
            const __half* ps = ps0 + 32 * p_j;
            b.sa = ldg_nc_u16(ps);
            b.sb = ldg_nc_u16(ps + (size_t)8 * zw16);
            const uint32_t* pz = pz0 + 2 * p_j;
            b.za = ldg_nc_u32(pz);
            b.zb = ldg_nc_u32(pz + (size_t)8 * zw);
            if (++p_j == mywin) {
                p_j = 0;
                ++p_rb;
                if (silu || p_rb >= D.rb_end[p_mi]) {
                    p_mi = mat_of(p_rb);
                    if (p_rb < rb1) set_base();
                } else {
                    pa0 += (size_t)16 * row_bytes;
                    ps0 += (size_t)16 * zw16;
                    pz0 += (size_t)16 * zw;
                }
            }
        };
        ItemBuf buf[kDepth];
        if (mywin > 0) {
#pragma unroll
            for (int d = 0; d < kDepth; ++d) issue(buf[d]);   // nothing here depends on the previous phase
        }
        // keep HBM streaming: request phase p + lookahead into L2 (one lane per CTA, its slice of every tensor)
        if (tid == kThreads - 1 && p + lookahead < n_phases) {
            const Phase* P = plan + p + lookahead;
            const int nm = P->nmat;
            for (int i = 0; i < nm; ++i) {
                prefetch_l2(P->qw[i], P->qw_bytes[i], cta, ncta);
                prefetch_l2(reinterpret_cast<const uint8_t*>(P->sc[i]), P->sc_bytes[i], cta, ncta);
                prefetch_l2(reinterpret_cast<const uint8_t*>(P->qz[i]), P->qz_bytes[i], cta, ncta);
            }
        }

        // ---- dependency: every CTA has stored its outputs of phase p - 1
        if (D.wait && p > 0) {
            if (tid == 0) {
                const unsigned int* c = counters + (p - 1);
                while (ld_acquire_u32(c) < (unsigned int)ncta) {}
            }
            __syncthreads();
        }

        // ---- stage this warp's windows of x as block fixed point (two int8 limbs per activation), see gemv_w2_stream.cu
        const __half* x = D.x;
        for (int j0 = 0; j0 < mywin; j0 += 4) {
            uint4 raw[4][2];
#pragma unroll
            for (int jj = 0; jj < 4; ++jj) {
                const int k0 = (warp + kWarps * (j0 + jj)) * 512 + lane * 16;
                raw[jj][0] = raw[jj][1] = make_uint4(0, 0, 0, 0);
                if (j0 + jj < mywin && k0 < IC) {
                    raw[jj][0] = ld_cg_v4(x + k0);
                    raw[jj][1] = ld_cg_v4(x + k0 + 8);
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
                mx = max(mx & 0xffffu, mx >> 16);
#pragma unroll
                for (int sft = 1; sft < 8; sft <<= 1) mx = max(mx, __shfl_xor_sync(0xffffffffu, mx, sft));
                const uint32_t ex = (mx >> 10) & 31u;
                const uint32_t exc = ex < 1u ? 1u : ex;
                const float inv_unit = __uint_as_float((154u - exc) << 23);
                const float unit = ex == 31u ? __uint_as_float(0x7fc00000u) : __uint_as_float((exc + 100u) << 23);
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

        // ---- consumer
        const bool active = ((g >> 1) == t);
        const uint8_t* bbase = active ? wbase + g * kSlotStride : zblock;
        const int jstride = active ? kWinStride : 0;
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
                const int rb = rb_base + sl;
                const int mi = mat_of(rb);
                __half hv = __float2half_rn(acc);
                bool store = true;
                if (silu) {
                    const float up = __shfl_down_sync(0xffffffffu, __half2float(hv), 16);
                    const float gt = __half2float(hv);
                    const float sg = __half2float(__float2half_rn(gt / (1.f + __expf(-gt))));
                    hv = __float2half_rn(sg * up);
                    store = (sl & 1) == 0;
                }
                const size_t oi = (size_t)blk_of(rb, mi) * 16 + row;
                if (store) D.out[mi][oi] = hv;
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

        if (mywin == 0) {
            for (int rb = rb0; rb < rb1; ++rb) flush();
        } else {
            const int nitems = (rb1 - rb0) * mywin;
            for (int i = 0; i < nitems; i += kDepth) {
#pragma unroll
                for (int d = 0; d < kDepth; ++d) {
                    if (i + d < nitems) {
                        consume(buf[d]);
                        issue(buf[d]);
                        if (++c_j == mywin) {
                            c_j = 0;
                            flush();
                        }
                    }
                }
            }
        }
        if (rslot) block_reduce();

        // ---- phase done for this CTA: publish (stores of all threads -> CTA barrier -> one gpu-scope release)
        __syncthreads();
        if (tid == 0) {
            __threadfence();
            red_release_add(counters + p, 1u);
        }
        if (tid < kDescWords) sdesc[((p + 1) & 1) * kDescWords + tid] = next_word;
        __syncthreads();
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
}

}  // namespace chain
}  // namespace bdq

using namespace bdq;

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
    const int smem = chain::smem_bytes(wpw);
    if (smem > 112 * 1024) { set_error("gemv_w2_chain_run: IC too large for two CTAs per SM"); return BDQ_ERR_UNSUPPORTED; }
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
    if (lookahead < 1) lookahead = 2;
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
    cudaLaunchKernelEx(&cfg, chain::gemv_w2_chain_kernel, plan, n_phases, wpw, counters, lookahead);
    return check_launch("bdq_gemv_w2_chain_run");
}
