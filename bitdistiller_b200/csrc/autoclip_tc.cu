// Clip-ratio grid search on the 5th-generation tensor cores (sm_100a): auto_2clip_layer, quantization/autoclip.py:13-83,
// for 16-bit weights / activations, q_group_size 128 or 64 (template KH), up to 512 sampled tokens (the reference's n_sample_token) and up to
// 10 x 10 candidates.  bdq_autoclip dispatches here; every other shape keeps the mma.sync kernel of autoclip.cu.
//
// A CTA owns one quantisation group g and 64 output channels, exactly like autoclip.cu, but the contraction
//     out[c, t] = sum_k Q_c(clamp(w, lo_c, hi_c))[k] * X[t, g*128 + k]          (autoclip.py:47-65)
// is ONE tcgen05 accumulator per channel: M = 128 rows (100 candidates + the unquantised weight) x N = 512 tokens x K = 128.
//   * X[:, g] (512 tokens x 128 = 128 KB) is staged ONCE per CTA by TMA (cp.async.bulk.tensor, SWIZZLE_128B, four boxes of
//     256 tokens x 64 k) and is the B operand of every MMA of the CTA;
//   * per channel the candidate rows are generated on the CUDA cores with the reference's exact T roundings (threshold form
//     of fakequant_thr.cuh where it verifies, op by op otherwise) and written as a K-major SWIZZLE_128B A tile (32 KB,
//     double-buffered); one lane issues 16 x tcgen05.mma (M128 N256 K16, kind::f16, fp32 accumulate) into two 256-column
//     TMEM accumulators (tokens 0-255 / 256-511 = all 512 TMEM columns), so the epilogue of one half overlaps the MMAs of
//     the other half and of the next channel;
//   * epilogue: tcgen05.ld 32x32b, (out[c, t] - out[w, t])^2 summed over tokens per candidate row.  A warp can only read
//     its own 32 TMEM lanes, so the weight row is replicated into every lane quarter (row 32q + 25; candidates 25q + r live
//     in row 32q + r) and each warp broadcasts its copy through 128 B of shared memory;
//   * warp roles: 1 x (TMA + MMA issue), 4 x (weight load, extremes, the 100 clamp / quantiser parameter sets incl. the
//     verified thresholds -- a ~1000-cycle dependent chain, run one channel ahead), 8 x (candidate rows), 8 x (epilogue +
//     first-minimum selection), all connected by mbarriers; nothing but the per-channel argmin is CTA-serial.
// Same results contract as autoclip.cu: fp32 error of exact products, first smallest candidate in (i_p, i_n) order.
#include <cuda.h>
#include <cstring>

#include "bdq_common.cuh"
#include "fq_int_elem.cuh"
#include "fq_nf3_elem.cuh"
#include "fakequant_thr.cuh"
#include "fq_nf3_stair.cuh"

namespace bdq {
namespace actc {

constexpr int kOT = 64;                 // output channels per CTA
// quantisation group = K = 64 * KH (KH = k halves of 64: 2 for group 128, 1 for group 64 -- the train scripts' group size)
constexpr int kTok = 512;               // tokens per CTA = N (two halves of 256)
constexpr int kRowsUsed = 104;          // 4 quarters x (25 candidates + the weight row)
constexpr int kWarpMma = 0, kWarpParam0 = 1, kNParamWarps = 4, kWarpRow0 = 5, kNRowWarps = 8, kWarpEpi0 = 13, kNEpiWarps = 8;
constexpr int kThreads = 32 * (kWarpEpi0 + kNEpiWarps);   // 672
template <int KH> struct Geo {
    static constexpr int kG = 64 * KH;
    static constexpr int kXBytes = kTok * kG * 2;         // 128 / 64 KB: [k half][512 tokens][128 B]
    static constexpr int kABytes = 128 * kG * 2;          // 32 / 16 KB:  [k half][128 rows][128 B]
};
constexpr int kCandFloats = 12;
constexpr int kCandP = 20;              // packed parameter words per slot (see the parameter role)

template <int KH> struct SmemT {   // offsets from the 1024-aligned base
    static constexpr int x = 0;
    static constexpr int a = x + Geo<KH>::kXBytes;              // 2 buffers
    static constexpr int cand = a + 2 * Geo<KH>::kABytes;       // [2][kRowsUsed][12] floats
    static constexpr int candp = cand + 2 * kRowsUsed * kCandFloats * 4;  // [2][kRowsUsed][kCandP] u32: duplicated T pairs + mode word
    static constexpr int wfp = candp + 2 * kRowsUsed * kCandP * 4;      // [2][64] packed T pairs (the raw weight group)
    static constexpr int ext = wfp + 2 * 256;                   // [kOT][2] floats: group max / min per channel
    static constexpr int errh = ext + kOT * 8;                  // [2 parity][2 halves][128] floats
    static constexpr int owb = errh + 2 * 2 * 128 * 4;          // [kNEpiWarps][2][32] floats
    static constexpr int wred = owb + kNEpiWarps * 256;         // [kNParamWarps][2] floats
    static constexpr int fac = wred + 64;                       // [16] floats: (float)(1 - i / n_grid), the python-double shrink factors
    static constexpr int bars = fac + 64;                       // mbarriers
    static constexpr int tmem_slot = bars + 16 * 8;
    static constexpr int total = tmem_slot + 16 + 1024;         // + alignment slack
};
// barrier indices
constexpr int kBarX = 0, kBarAFull = 1, kBarAEmpty = 3, kBarAccFull = 5, kBarAccEmpty = 7, kBarPFull = 9, kBarPEmpty = 11;

static_assert(SmemT<2>::total <= 227 * 1024, "autoclip_tc shared memory");

struct Args {
    const void* w;
    void* best_max;
    void* best_min;
    int32_t* best_idx;
    float* best_err;
    int64_t co, ci, n_tok;
    int bits, quant_type, n_grid, nc, sem;
    Nf3Stair stair;    // NF3 candidates in threshold form
    long long* prof;   // dev tool (BDQ_AUTOCLIP_TC_PROF): cycles every role of CTA (0, 0) spent waiting, see tools/prof_autoclip_roles.py
    Nf3Tables tb;
};

__device__ __forceinline__ uint32_t s_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* b, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(s_u32(b)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* b, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(s_u32(b)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* b) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(s_u32(b)) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* b, uint32_t parity) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "AW_%=:\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1, %2;\n\t"
        "@p bra AD_%=;\n\t"
        "bra AW_%=;\n\t"
        "AD_%=:\n\t}" ::"r"(s_u32(b)), "r"(parity), "r"(0x989680u) : "memory");   // suspend-time hint: the warp sleeps, it does not spin
}
// K-major SWIZZLE_128B shared-memory matrix descriptor (see gemm_w2.cu:gm_desc)
__device__ __forceinline__ uint64_t sdesc(uint32_t smem_addr) {
    return (uint64_t)((smem_addr >> 4) & 0x3FFFu) | (1ull << 16) | ((uint64_t)(1024 >> 4) << 32) | (1ull << 46) | (2ull << 61);
}
__device__ __forceinline__ void mma_f16(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
        : "memory");
}
__device__ __forceinline__ void mma_commit(uint64_t* bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(s_u32(bar)) : "memory");
}
__device__ __forceinline__ void tma_load_2d(void* dst, const CUtensorMap* map, int c0, int c1, uint64_t* bar) {
    asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];" ::"r"(s_u32(dst)),
                 "l"(map), "r"(c0), "r"(c1), "r"(s_u32(bar))
                 : "memory");
}
__device__ __forceinline__ void bar_sync(int id, int nthreads) { asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(nthreads) : "memory"); }

// row of the 128-row accumulator <-> candidate: quarter q holds candidates 25q .. 25q+24 in rows 32q + r, the weight in 32q + 25
__device__ __forceinline__ int row_of_slot(int slot) { return 32 * (slot / 26) + (slot % 26); }   // slot in [0, 104)

template <int DT, int KH>
__global__ void __launch_bounds__(kThreads, 1) autoclip_tc_kernel(const __grid_constant__ CUtensorMap xmap, const Args a) {
    constexpr bool F16 = (DT == kF16);
    constexpr int kG = Geo<KH>::kG, kXBytes = Geo<KH>::kXBytes, kABytes = Geo<KH>::kABytes;
    using Smem = SmemT<KH>;
    // kind::f16 instruction descriptor: D fp32 (bit 4), A / B format bf16 (1) or f16 (0) in [7,10) / [10,13), K-major both,
    // N >> 3 in [17,23), M >> 4 in [24,29)
    constexpr uint32_t kIdesc = (1u << 4) | (F16 ? 0u : ((1u << 7) | (1u << 10))) | ((uint32_t)(256 >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);
    extern __shared__ uint8_t smem_raw[];
    // align on the 32-bit shared address and ADD the offset: the pointer keeps its shared address space (LDS / STS, not generic
    // LD / ST -- the integer round trip through uintptr_t made every shared access of the first version a generic one)
    uint8_t* smem = smem_raw + ((1024u - (s_u32(smem_raw) & 1023u)) & 1023u);
    uint8_t* Xs = smem + Smem::x;
    uint8_t* As = smem + Smem::a;
    float* cand = reinterpret_cast<float*>(smem + Smem::cand);
    uint32_t* candp = reinterpret_cast<uint32_t*>(smem + Smem::candp);
    uint32_t* wfp = reinterpret_cast<uint32_t*>(smem + Smem::wfp);
    float* ext = reinterpret_cast<float*>(smem + Smem::ext);
    float* errh = reinterpret_cast<float*>(smem + Smem::errh);
    float* owb = reinterpret_cast<float*>(smem + Smem::owb);
    float* wred = reinterpret_cast<float*>(smem + Smem::wred);
    float* facs = reinterpret_cast<float*>(smem + Smem::fac);
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + Smem::bars);
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(smem + Smem::tmem_slot);

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int g = blockIdx.x;
    const int64_t o0 = (int64_t)blockIdx.y * kOT;
    const int no = (int)min((int64_t)kOT, a.co - o0);
    const int64_t ng = a.ci / kG;
    const int nc = a.nc, nc2 = nc * nc;
    const float qmax = (float)((1 << a.bits) - 1);

    // ---- setup: barriers, TMEM (all 512 columns), zeroed A buffers (rows 26..31 of every quarter stay zero for good)
    if (tid == 0) {
        mbar_init(&bars[kBarX], 1);
        for (int i = 0; i < 2; ++i) {
            mbar_init(&bars[kBarAFull + i], kNRowWarps * 32);
            mbar_init(&bars[kBarAEmpty + i], 1);
            mbar_init(&bars[kBarAccFull + i], 1);
            mbar_init(&bars[kBarAccEmpty + i], 4 * 32);
            mbar_init(&bars[kBarPFull + i], kNParamWarps * 32);
            mbar_init(&bars[kBarPEmpty + i], kNRowWarps * 32);
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("prefetch.tensormap [%0];" ::"l"(&xmap) : "memory");
    }
    if (tid < 16) facs[tid] = (float)(1.0 - (double)tid / (double)a.n_grid);   // org * (1 - i / n_grid): python double, cast to fp32 opmath
    for (int i = tid; i < 2 * kABytes / 16; i += kThreads) reinterpret_cast<uint4*>(As)[i] = make_uint4(0, 0, 0, 0);
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    if (warp == kWarpMma) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(s_u32(tmem_slot)), "r"(512) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem_d = *tmem_slot;
    // dev tool: per-role wait accounting of CTA (0, 0), lane 0 of each warp: prof[warp * 8 + slot] += cycles waited
    const bool profiling = a.prof != nullptr && blockIdx.x == 0 && blockIdx.y == 0 && lane == 0;
    long long t_role0 = profiling ? clock64() : 0;
    auto wait_t = [&](uint64_t* b, uint32_t parity, int slot) {
        if (profiling) {
            const long long t0 = clock64();
            mbar_wait(b, parity);
            a.prof[warp * 8 + slot] += clock64() - t0;
        } else {
            mbar_wait(b, parity);
        }
    };

    if (warp == kWarpMma) {
        // =========================== TMA (X once) + MMA issue ===========================
        if (lane == 0) {
            mbar_expect_tx(&bars[kBarX], kXBytes);
            for (int kh = 0; kh < KH; ++kh)
                for (int th = 0; th < 2; ++th)   // box = 64 k x 256 tokens; rows past n_tok are zero-filled by the TMA unit
                    tma_load_2d(Xs + kh * (kXBytes / KH) + th * 256 * 128, &xmap, g * kG + kh * 64, th * 256, &bars[kBarX]);
            mbar_wait(&bars[kBarX], 0);
            const uint32_t xa = s_u32(Xs), aa = s_u32(As);
            for (int ol = 0; ol < no; ++ol) {
                const int ab = ol & 1;
                wait_t(&bars[kBarAFull + ab], (uint32_t)((ol >> 1) & 1), 0);
                asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
#pragma unroll
                for (int h = 0; h < 2; ++h) {
                    if (ol > 0) {
                        wait_t(&bars[kBarAccEmpty + h], (uint32_t)((ol - 1) & 1), 1 + h);
                        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                    }
#pragma unroll
                    for (int kh = 0; kh < KH; ++kh) {
                        const uint64_t adesc = sdesc(aa + ab * kABytes + kh * (kABytes / KH));
                        const uint64_t bdesc = sdesc(xa + kh * (kXBytes / KH) + h * 256 * 128);
#pragma unroll
                        for (int k = 0; k < 4; ++k)   // 32 bytes (>> 4 = 2) inside the swizzle row per K = 16
                            mma_f16(tmem_d + h * 256, adesc + 2 * k, bdesc + 2 * k, kIdesc, (kh | k) != 0 ? 1u : 0u);
                    }
                    mma_commit(&bars[kBarAccFull + h]);
                }
                mma_commit(&bars[kBarAEmpty + ab]);
            }
        }
    } else if (warp < kWarpRow0) {
        // =========================== parameters: one channel ahead of the row generators ===========================
        const int pt = tid - kWarpParam0 * 32;   // 0..127: element of the group (group 64: each element twice) / accumulator slot
        const int pe = pt & (kG - 1);
        float vnext = load_scalar<DT>(a.w, o0 * a.ci + (int64_t)g * kG + pe);
        for (int ol = 0; ol < no; ++ol) {
            const int pb = ol & 1;
            const float v = vnext;
            if (ol + 1 < no) vnext = load_scalar<DT>(a.w, (o0 + ol + 1) * a.ci + (int64_t)g * kG + pe);   // hide the global latency
            const float mxw = group_max<32>(v), mnw = group_min<32>(v);
            if (ol >= 2) wait_t(&bars[kBarPEmpty + pb], (uint32_t)(((ol >> 1) - 1) & 1), 0);   // the row warps are done with this buffer
            if (lane == 0) { wred[(warp - kWarpParam0) * 2] = mxw; wred[(warp - kWarpParam0) * 2 + 1] = mnw; }
            // raw T pairs of the group for the row generators
            const float vn = __shfl_down_sync(0xffffffffu, v, 1);
            if ((pt & 1) == 0 && pt < kG) {
                if constexpr (F16) wfp[pb * 64 + (pt >> 1)] = (uint32_t)__half_as_ushort(__float2half_rn(v)) | ((uint32_t)__half_as_ushort(__float2half_rn(vn)) << 16);
                else wfp[pb * 64 + (pt >> 1)] = (__float_as_uint(v) >> 16) | (__float_as_uint(vn) & 0xffff0000u);
            }
            bar_sync(1, kNParamWarps * 32);
            float mx = wred[0], mn = wred[1];
#pragma unroll
            for (int i = 1; i < kNParamWarps; ++i) { mx = fmaxf(mx, wred[2 * i]); mn = fminf(mn, wred[2 * i + 1]); }
            if (pt == 0) { ext[ol * 2] = mx; ext[ol * 2 + 1] = mn; }
            if (pt < kRowsUsed) {
                float* cp = cand + (pb * kRowsUsed + pt) * kCandFloats;
                const int q = pt / 26, r = pt - q * 26;
                const int c = 25 * q + r;
                cp[11] = 0.f;
                if (r == 25) {
                    cp[4] = -INFINITY; cp[5] = INFINITY; cp[7] = 1.f;   // weight row: pass-through
                } else if (c >= nc2) {
                    cp[7] = 2.f;                                        // unused slot: the row is left alone
                } else {
                    const int ip = c / nc, in = c - ip * nc;
                    // org * (1 - i / n_grid): python double factor, cast to fp32 opmath, product rounded to T (autoclip.py:50,52)
                    const float hi = rT<DT>(mx * facs[ip]);
                    const float lo = rT<DT>(mn * facs[in]);
                    const float cmx = fminf(fmaxf(mx, lo), hi), cmn = fminf(fmaxf(mn, lo), hi);   // extremes of clamp(w, lo, hi)
                    cp[4] = lo; cp[5] = hi; cp[7] = 0.f;
                    if (a.quant_type == BDQ_QUANT_INT) {
                        const IntQ p = int_group_params<DT>(cmx, cmn, qmax, a.sem);
                        cp[0] = p.s; cp[1] = p.y; cp[2] = p.z; cp[3] = p.t1;
                        if (a.bits == 2) {
                            bool ok = fmaxf(fabsf(cmx), fabsf(cmn)) * p.y < 1e30f;
#pragma unroll
                            for (int k = 0; k < 3; ++k) {
                                bool o2;
                                cp[8 + k] = find_threshold<DT, BDQ_ROUND_HALF_EVEN>(-p.z + (float)(k + 1), p, o2);
                                ok = ok && o2;
                            }
                            cp[11] = ok ? 1.f : 0.f;
                        }
                    } else {
                        const Nf3Q p = nf3_group_params(cmx, cmn);
                        cp[0] = p.sp; cp[1] = p.sn; cp[2] = p.yp; cp[3] = p.yn; cp[6] = p.nan_group ? 1.f : 0.f;
                        // NF3 in threshold form (fakequant_nf3_thr.cu): the candidate's seven verified x-space thresholds and
                        // eight output values rT(level * scale); a candidate that fails the check is generated op by op
                        using P = P2<DT>;
                        float T[7];
                        const bool ok = nf3_thresholds_1t<DT>(a.stair, p.sp, p.sn, T) && !p.nan_group;
                        cp[11] = ok ? 2.f : 0.f;
                        uint32_t* dp = candp + (pb * kRowsUsed + pt) * kCandP;
#pragma unroll
                        for (int k = 0; k < 5; ++k) dp[2 + k] = P::bits(P::dup(T[k]));
                        dp[8] = P::bits(P::dup(T[5]));
                        dp[9] = P::bits(P::dup(T[6]));
                        const typename P::h2 s2p = P::dup(p.sp), s2n = P::dup(p.sn);
#pragma unroll
                        for (int c = 0; c < 8; ++c)
                            dp[10 + c] = (c == 3) ? 0u : P::bits(__hmul2(P::from_bits(a.stair.lvl2[c]), c < 3 ? s2n : s2p));
                    }
                }
            }
            if (pt < kRowsUsed) {   // the same parameters as duplicated T pairs + a mode word: the row generators read LDS.128s
                using P = P2<DT>;
                const float* cp = cand + (pb * kRowsUsed + pt) * kCandFloats;
                // mode: 0 int threshold form, 1 pass-through (weight row), 2 unused slot, 3 op by op, 4 NF3 threshold form
                const uint32_t mode = cp[7] == 1.f ? 1u : (cp[7] == 2.f ? 2u : (cp[11] == 1.f ? 0u : (cp[11] == 2.f ? 4u : 3u)));
                uint32_t* dp = candp + (pb * kRowsUsed + pt) * kCandP;
                dp[0] = P::bits(P::dup(cp[4]));
                dp[1] = P::bits(P::dup(cp[5]));
                dp[7] = mode;
                if (mode != 4u) {
                    dp[2] = P::bits(P::dup(cp[8]));
                    dp[3] = P::bits(P::dup(cp[9]));
                    dp[4] = P::bits(P::dup(cp[10]));
                    dp[5] = P::bits(P::dup(-cp[2]));
                    dp[6] = P::bits(P::dup(cp[0]));
                }
            }
            mbar_arrive(&bars[kBarPFull + pb]);
            bar_sync(1, kNParamWarps * 32);   // wred is rewritten by the next channel
        }
    } else if (warp < kWarpEpi0) {
        // =========================== candidate rows -> K-major SWIZZLE_128B A tile ===========================
        const int rt = tid - kWarpRow0 * 32;      // 0..255
        constexpr int kChunks = kG / 8;           // 16-byte chunks per row: 16 / 8
        const int jt = rt & (kChunks - 1);        // 16-byte chunk of the row: k = 8 jt .. 8 jt + 7
        const int kh = jt >> 3, ch = jt & 7;
        using P = P2<DT>;
        using h2 = typename P::h2;
        for (int ol = 0; ol < no; ++ol) {
            const int pb = ol & 1, ab = ol & 1;
            wait_t(&bars[kBarPFull + pb], (uint32_t)((ol >> 1) & 1), 0);
            if (ol >= 2) wait_t(&bars[kBarAEmpty + ab], (uint32_t)(((ol >> 1) - 1) & 1), 1);   // the MMAs that read this buffer retired
            const uint4 wraw = *reinterpret_cast<const uint4*>(wfp + pb * 64 + jt * 4);         // this thread's 8 weights (4 T pairs)
            const uint32_t wp[4] = {wraw.x, wraw.y, wraw.z, wraw.w};
            uint8_t* abase = As + ab * kABytes + kh * (kABytes / KH);
            for (int slot = rt / kChunks; slot < kRowsUsed; slot += (kNRowWarps * 32) / kChunks) {
                const uint4 pa = *reinterpret_cast<const uint4*>(candp + (pb * kRowsUsed + slot) * kCandP);
                const uint4 pq = *reinterpret_cast<const uint4*>(candp + (pb * kRowsUsed + slot) * kCandP + 4);
                const uint32_t mode = pq.w;
                if (mode == 2u) continue;
                const int row = row_of_slot(slot);
                uint32_t o4[4];
                if (mode == 0u) {   // threshold form: clamp, count the thresholds passed, (count - z) * s
                    const h2 lo2 = P::from_bits(pa.x), hi2 = P::from_bits(pa.y), t1 = P::from_bits(pa.z), t2 = P::from_bits(pa.w);
                    const h2 t3 = P::from_bits(pq.x), nz2 = P::from_bits(pq.y), s2 = P::from_bits(pq.z);
#pragma unroll
                    for (int i = 0; i < 4; ++i) {
                        const h2 x2 = __hmin2(__hmax2(P::from_bits(wp[i]), lo2), hi2);
                        const h2 c2 = __hadd2(__hadd2(__hge2(x2, t1), __hge2(x2, t2)), __hge2(x2, t3));
                        o4[i] = P::bits(__hmul2(__hadd2(c2, nz2), s2));   // rT((c - z) * s)
                    }
                } else if (mode == 4u) {   // NF3 threshold form: clamp, 7 packed compares + 7 bit selects per pair
                    const uint4 pr = *reinterpret_cast<const uint4*>(candp + (pb * kRowsUsed + slot) * kCandP + 8);    // T5, T6, v0, v1
                    const uint4 pv = *reinterpret_cast<const uint4*>(candp + (pb * kRowsUsed + slot) * kCandP + 12);   // v2 .. v5
                    const uint2 pw = *reinterpret_cast<const uint2*>(candp + (pb * kRowsUsed + slot) * kCandP + 16);   // v6, v7
                    const h2 lo2 = P::from_bits(pa.x), hi2 = P::from_bits(pa.y);
                    const uint32_t tk[7] = {pa.z, pa.w, pq.x, pq.y, pq.z, pr.x, pr.y};
                    const uint32_t vk[8] = {pr.z, pr.w, pv.x, pv.y, pv.z, pv.w, pw.x, pw.y};
#pragma unroll
                    for (int i = 0; i < 4; ++i) {
                        const h2 x2 = __hmin2(__hmax2(P::from_bits(wp[i]), lo2), hi2);
                        uint32_t o = vk[0];
#pragma unroll
                        for (int k = 0; k < 7; ++k) {
                            const uint32_t m = __hge2_mask(x2, P::from_bits(tk[k]));
                            asm("lop3.b32 %0, %1, %2, %3, 0xCA;" : "=r"(o) : "r"(m), "r"(vk[k + 1]), "r"(o));   // m ? v : o
                        }
                        o4[i] = o;
                    }
                } else if (mode == 1u) {
#pragma unroll
                    for (int i = 0; i < 4; ++i) o4[i] = wp[i];
                } else {
                    const float* cp = cand + (pb * kRowsUsed + slot) * kCandFloats;
#pragma unroll
                    for (int i = 0; i < 4; ++i) {
                        float v[2] = {P::lo(P::from_bits(wp[i])), P::hi(P::from_bits(wp[i]))};
#pragma unroll
                        for (int e = 0; e < 2; ++e) {
                            const float xw = fminf(fmaxf(v[e], cp[4]), cp[5]);
                            if (a.quant_type == BDQ_QUANT_INT) {
                                IntQ p; p.s = cp[0]; p.y = cp[1]; p.z = cp[2]; p.t1 = cp[3];
                                float qu;
                                const float cc = quant_code<DT, BDQ_ROUND_HALF_EVEN>(xw, p, qmax, qu);
                                v[e] = rT<DT>((cc - p.z) * p.s);
                            } else {
                                Nf3Q p; p.sp = cp[0]; p.sn = cp[1]; p.yp = cp[2]; p.yn = cp[3]; p.nan_group = cp[6] != 0.f;
                                float q, y; int code;
                                v[e] = rT<DT>(nf3_elem<DT>(xw, p, a.tb, q, y, code));
                            }
                        }
                        o4[i] = P::bits(P::pack(v[0], v[1]));
                    }
                }
                *reinterpret_cast<uint4*>(abase + row * 128 + ((ch ^ (row & 7)) << 4)) = make_uint4(o4[0], o4[1], o4[2], o4[3]);
            }
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");   // generic-proxy writes -> visible to the MMA (async proxy)
            mbar_arrive(&bars[kBarAFull + ab]);
            mbar_arrive(&bars[kBarPEmpty + pb]);
        }
    } else {
        // =========================== epilogue: sum_t (out[c,t] - out[w,t])^2 per row, argmin per channel ===========================
        const int ew = warp - kWarpEpi0;          // 0..7
        const int quarter = warp & 3;             // TMEM lanes [32 * (warp % 4), +32)
        const int half = ew >> 2;                 // tokens [256 * half, +256)
        const int et = tid - kWarpEpi0 * 32;      // 0..255
        float* myow = owb + ew * 64;
        const float inv_tok = 1.0f / (float)a.n_tok;
        for (int ol = 0; ol < no; ++ol) {
            wait_t(&bars[kBarAccFull + half], (uint32_t)(ol & 1), 0);
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
            float e0 = 0.f, e1 = 0.f, e2 = 0.f, e3 = 0.f;
            // 16 chunks of 16 tokens, software-pipelined: the TMEM load of chunk cb + 1 is in flight while chunk cb is reduced
            const uint32_t tbase = tmem_d + ((uint32_t)(quarter * 32) << 16) + (uint32_t)(half * 256);
            uint32_t ra[16], rb[16];
            auto tld = [&](uint32_t* r, int cb) {
                asm volatile(
                    "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
                    : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
                      "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
                    : "r"(tbase + (uint32_t)(cb * 16)));
            };
            auto reduce = [&](const uint32_t* r, float* ow) {
                if (lane == 25) {
#pragma unroll
                    for (int i = 0; i < 4; ++i)
                        reinterpret_cast<uint4*>(ow)[i] = make_uint4(r[4 * i], r[4 * i + 1], r[4 * i + 2], r[4 * i + 3]);
                }
                __syncwarp();
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    const float4 w4 = reinterpret_cast<const float4*>(ow)[i];
                    const float d0 = __uint_as_float(r[4 * i]) - w4.x, d1 = __uint_as_float(r[4 * i + 1]) - w4.y;
                    const float d2 = __uint_as_float(r[4 * i + 2]) - w4.z, d3 = __uint_as_float(r[4 * i + 3]) - w4.w;
                    e0 = fmaf(d0, d0, e0);
                    e1 = fmaf(d1, d1, e1);
                    e2 = fmaf(d2, d2, e2);
                    e3 = fmaf(d3, d3, e3);
                }
            };
            tld(ra, 0);
#pragma unroll 2
            for (int cb = 0; cb < 16; cb += 2) {
                asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
                tld(rb, cb + 1);
                reduce(ra, myow);                 // two broadcast buffers: no WAR hazard between consecutive chunks
                asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
                if (cb + 2 < 16) tld(ra, cb + 2);
                reduce(rb, myow + 32);
            }
            asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
            mbar_arrive(&bars[kBarAccEmpty + half]);            // this half of TMEM may be overwritten
            errh[((ol & 1) * 2 + half) * 128 + quarter * 32 + lane] = (e0 + e1) + (e2 + e3);
            bar_sync(2, kNEpiWarps * 32);
            // ---- first smallest candidate (strict '<' from 1e9 in (i_p, i_n) order): one warp, alternating by channel
            if (ew == (ol & 7)) {
                const float* eh = errh + (ol & 1) * 256;
                float be = 1e9f;
                int bi = -1;
                for (int c = lane; c < nc2; c += 32) {
                    const int row = 32 * (c / 25) + (c % 25);
                    const float e = (eh[row] + eh[128 + row]) * inv_tok;
                    if (e < be) { be = e; bi = c; }   // NaN never wins; ascending c keeps the first minimum
                }
#pragma unroll
                for (int s = 16; s > 0; s >>= 1) {
                    const float oe = __shfl_xor_sync(0xffffffffu, be, s);
                    const int oi = __shfl_xor_sync(0xffffffffu, bi, s);
                    const bool take = (oi >= 0) && (bi < 0 || oe < be || (oe == be && oi < bi));
                    if (take) { be = oe; bi = oi; }
                }
                if (lane == 0) {
                    const int64_t o = o0 + ol;
                    const float mx = ext[ol * 2], mn = ext[ol * 2 + 1];
                    float bmax = mx, bmin = mn;
                    if (bi >= 0) {
                        const int ip = bi / nc, in = bi - ip * nc;
                        bmax = rT<DT>(mx * facs[ip]);
                        bmin = rT<DT>(mn * facs[in]);
                    }
                    store_scalar<DT>(a.best_max, o * ng + g, bmax);
                    store_scalar<DT>(a.best_min, o * ng + g, bmin);
                    if (a.best_idx) a.best_idx[o * ng + g] = bi;
                    if (a.best_err) a.best_err[o * ng + g] = be;
                }
            }
            (void)et;
        }
    }
    if (profiling) a.prof[warp * 8 + 7] = clock64() - t_role0;   // lifetime of the role
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == kWarpMma) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_d), "r"(512) : "memory");
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static EncodeTiledFn get_encode() {
    static EncodeTiledFn fn = nullptr;
    if (!fn) {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult qres;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres) == cudaSuccess &&
            qres == cudaDriverEntryPointSuccess)
            fn = (EncodeTiledFn)p;
    }
    return fn;
}

}  // namespace actc

static long long* g_actc_prof = nullptr;
// Returns BDQ_ERR_UNSUPPORTED when the shape is not this kernel's (the caller then runs autoclip.cu's kernel).
int launch_autoclip_tc(const void* w, const void* x, void* best_max, void* best_min, int32_t* best_idx, float* best_err, int64_t co,
                       int64_t ci, int64_t n_tok, int64_t xs, int group, int bits, int quant_type, int n_grid, int nc, int dtype, int sem,
                       cudaStream_t st) {
    using namespace actc;
    if (dtype == BDQ_F32 || (group != 128 && group != 64) || n_tok > kTok || nc * nc > 100 || (xs % 8) != 0 || ((uintptr_t)x % 16) != 0 ||
        ci / group > 65535 * 32 || (co + kOT - 1) / kOT > 65535)
        return BDQ_ERR_UNSUPPORTED;
    {   // BDQ_AUTOCLIP_TC=0 forces the mma.sync kernel (A/B comparisons in the tests; read per call)
        const char* e = getenv("BDQ_AUTOCLIP_TC");
        if (e && atoi(e) == 0) return BDQ_ERR_UNSUPPORTED;
    }
    EncodeTiledFn enc = get_encode();
    if (!enc) return BDQ_ERR_UNSUPPORTED;
    CUtensorMap map;
    const cuuint64_t dims[2] = {(cuuint64_t)ci, (cuuint64_t)n_tok};
    const cuuint64_t strides[1] = {(cuuint64_t)xs * 2};
    const cuuint32_t box[2] = {64, 256};
    const cuuint32_t estr[2] = {1, 1};
    CUresult r = enc(&map, dtype == BDQ_F16 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT16 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(x), dims,
                     strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                     CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) return BDQ_ERR_UNSUPPORTED;
    Args a;
    a.w = w; a.best_max = best_max; a.best_min = best_min; a.best_idx = best_idx; a.best_err = best_err;
    a.co = co; a.ci = ci; a.n_tok = n_tok;
    a.bits = bits; a.quant_type = quant_type; a.n_grid = n_grid; a.nc = nc; a.sem = sem;
    a.tb = make_tables(dtype, sem);
    a.stair = make_stair(a.tb, dtype);
    a.prof = g_actc_prof;
    dim3 grid((unsigned)(ci / group), (unsigned)((co + kOT - 1) / kOT));
#define BDQ_ACTC(DTT, KHH)                                                                                                                  \
    do {                                                                                                                                     \
        if (!ensure_dyn_smem(reinterpret_cast<const void*>(autoclip_tc_kernel<DTT, KHH>), SmemT<KHH>::total)) return check_launch("bdq_autoclip(tc attr)"); \
        autoclip_tc_kernel<DTT, KHH><<<grid, kThreads, SmemT<KHH>::total, st>>>(map, a);                                                     \
    } while (0)
    if (dtype == BDQ_BF16) { if (group == 128) BDQ_ACTC(kBF16, 2); else BDQ_ACTC(kBF16, 1); }
    else { if (group == 128) BDQ_ACTC(kF16, 2); else BDQ_ACTC(kF16, 1); }
#undef BDQ_ACTC
    return check_launch("bdq_autoclip(tcgen05)");
}

}  // namespace bdq

// dev tool: device buffer of 32 x 8 int64 (zeroed by the caller) that later tcgen05 autoclip launches accumulate wait cycles into
extern "C" int bdq_autoclip_tc_set_prof(void* buf) {
    bdq::g_actc_prof = (long long*)buf;
    return BDQ_OK;
}
