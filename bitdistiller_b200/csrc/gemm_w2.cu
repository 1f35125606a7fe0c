// Packed int2 prefill GEMM on the 5th-generation tensor cores (sm_100a).  Replaces quant_matmul_v2 / quant_matmul_kernel
// (quantization/triton_kernels.py:14-134) as called from WQLinear.forward (quantization/qmodule.py:148-166), without the
// three .T.contiguous() copies of the packed layer the reference makes on every call.
//
//   out[m, n] = fp16( sum_k x[m, k] * fp16( scales[n, k/G] * (code[n, k] - zero[n, k/G]) ) )      fp32 accumulate
// (the dequantised weight is rounded to fp16 before the product exactly like the Triton kernel, triton_kernels.py:89-93).
//
// One CTA computes a (128 * MT tokens) x (NT output channels) tile, MT in {1, 2}, NT in {256, 128} (host-side choice):
//   warp 0      : TMA producer -- x tiles [128 x 64] fp16, SWIZZLE_128B, cp.async.bulk.tensor + mbarrier complete_tx
//   warp 1      : allocates 256 TMEM columns, then one elected lane issues tcgen05.mma (M128 N256 K16, kind::f16,
//                 A and B from shared-memory descriptors, fp32 accumulator in TMEM) and tcgen05.commit
//   warps 2..9  : dequant producers -- thread j owns output channel j of the tile: one 32-byte load = 128 codes = one
//                 quantisation group, expanded in registers (mask|exponent trick, natural k order) to fp16, written as
//                 two 128-byte rows of the K-major SWIZZLE_128B B tile, fence.proxy.async, mbarrier arrive.  The packed
//                 weights are never materialised in fp16 anywhere but in this ring.
//                 After the k loop the same warps drain TMEM (tcgen05.ld 32x32b) and store fp16 rows of the output.
#include <cuda.h>
#include <cstring>

#include "bdq_common.cuh"

namespace bdq {

constexpr int kGmM = 128, kGmK = 64;     // CTA tile: 128 tokens x NT (256 or 128) output channels; kGmK fp16 = one 128-byte swizzle row
// MT = token tiles of 128 rows per CTA (1 or 2).  With MT = 2 the dequantised B stage feeds two accumulators (2 x 256 TMEM
// columns = all of TMEM): the dequant work, which bounds the MT = 1 kernel, is amortised over twice the MMAs.
constexpr int kGmABytes = kGmM * kGmK * 2;           // 16 KB per token tile
constexpr int kGmThreads = 32 * 10;
// NT = 128 serves grids that would leave most SMs idle with 256-wide tiles (k/v_proj: 64 -> 128 CTAs); the 256 producer
// threads then split the k blocks (even / odd pairs) instead of the channels.
template <int MT, int NT> struct GmCfg {
    static constexpr int kBBytes = NT * kGmK * 2;        // 32 KB / 16 KB
    static constexpr int kStageBytes = MT * kGmABytes + kBBytes;
    static constexpr int kStages = (196 * 1024) / kStageBytes > 6 ? 6 : (196 * 1024) / kStageBytes;   // 4 / 3 (NT 256), 6 / 4 (NT 128)
    static constexpr int kSmem = kStages * kStageBytes + 1024 /*alignment slack*/ + 256 /*barriers*/;
};

__device__ __forceinline__ uint32_t gm_smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void gm_mbar_init(uint64_t* b, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(gm_smem_u32(b)), "r"(count));
}
__device__ __forceinline__ void gm_mbar_expect_tx(uint64_t* b, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(gm_smem_u32(b)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void gm_mbar_arrive(uint64_t* b) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(gm_smem_u32(b)) : "memory");
}
__device__ __forceinline__ void gm_mbar_wait(uint64_t* b, uint32_t parity) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "GW_%=:\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1, %2;\n\t"
        "@p bra GD_%=;\n\t"
        "bra GW_%=;\n\t"
        "GD_%=:\n\t}" ::"r"(gm_smem_u32(b)), "r"(parity), "r"(0x989680u) : "memory");   // suspend-time hint: the warp sleeps, it does not spin
}

// K-major, SWIZZLE_128B shared-memory matrix descriptor (cute::UMMA::SmemDescriptor layout): start address >> 4 in
// bits [0,14), LBO (unused for swizzled K-major, 1) in [16,30), SBO = 1024 B (one 8-row swizzle atom) in [32,46),
// version 1 in [46,48), layout type 2 (SWIZZLE_128B) in [61,64).
__device__ __forceinline__ uint64_t gm_desc(uint32_t smem_addr) {
    return (uint64_t)((smem_addr >> 4) & 0x3FFFu) | (1ull << 16) | ((uint64_t)(1024 >> 4) << 32) | (1ull << 46) | (2ull << 61);
}
// kind::f16 instruction descriptor: D fp32 (bit 4), A/B fp16 (0), both K-major, N >> 3 in [17,23), M >> 4 in [24,29)
template <int NT> struct GmIdesc { static constexpr uint32_t v = (1u << 4) | ((uint32_t)(NT >> 3) << 17) | ((uint32_t)(kGmM >> 4) << 24); };

__device__ __forceinline__ void gm_mma(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
        : "memory");
}
__device__ __forceinline__ void gm_commit(uint64_t* bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(gm_smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void gm_tma_load_2d(void* dst, const CUtensorMap* map, int c0, int c1, uint64_t* bar) {
    asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];" ::"r"(
                     gm_smem_u32(dst)),
                 "l"(map), "r"(c0), "r"(c1), "r"(gm_smem_u32(bar))
                 : "memory");
}

__device__ __forceinline__ uint32_t gm_lop(uint32_t a, uint32_t b, uint32_t c) {
    uint32_t r;
    asm("lop3.b32 %0, %1, %2, %3, 0xEA;" : "=r"(r) : "r"(a), "r"(b), "r"(c));   // (a & b) | c
    return r;
}

// 16 consecutive 2-bit codes of `v` -> 8 half2 (c0,c1) (c2,c3) ... (c14,c15), each value = s * (code - z) rounded to fp16.
// u = (a & 0xffff) | (a << 14) puts code 2i in bits [4i,4i+2) of the low half and code 2i+1 at the same bits of the
// high half, so one (u & mask) | exponent yields the pair (K + c_2i, K + c_2i+1) with K = 1024 / 64 / 4 for i = 0, 1, 2
// (i = 3 comes from u >> 8 at the K = 64 position); (K + c) - (K + z) is exact in fp16 and the product rounds once.
__device__ __forceinline__ void gm_dequant16(uint32_t v, __half2 s2, __half2 kz1024, __half2 kz64, __half2 kz4, uint32_t* out) {
    const uint32_t ua = gm_lop(v, 0x0000FFFFu, v << 14);
    const uint32_t hb = v >> 16;
    const uint32_t ub = hb | (hb << 14);
    const uint32_t ua8 = ua >> 8, ub8 = ub >> 8;
    uint32_t h[8];
    h[0] = gm_lop(ua, 0x00030003u, 0x64006400u);
    h[1] = gm_lop(ua, 0x00300030u, 0x54005400u);
    h[2] = gm_lop(ua, 0x03000300u, 0x44004400u);
    h[3] = gm_lop(ua8, 0x00300030u, 0x54005400u);
    h[4] = gm_lop(ub, 0x00030003u, 0x64006400u);
    h[5] = gm_lop(ub, 0x00300030u, 0x54005400u);
    h[6] = gm_lop(ub, 0x03000300u, 0x44004400u);
    h[7] = gm_lop(ub8, 0x00300030u, 0x54005400u);
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        const __half2 kz = ((i & 3) == 0) ? kz1024 : (((i & 3) == 2) ? kz4 : kz64);
        const __half2 r = __hmul2(__hsub2(*reinterpret_cast<__half2*>(&h[i]), kz), s2);
        out[i] = *reinterpret_cast<const uint32_t*>(&r);
    }
}

struct GemmPeers {   // tensor parallel: this rank's column slice inside the peers' output matrices
    __half* p[7];
    int n;
};

// BTMA = true: plain fp16 GEMM out = x @ W^T with W [OC, IC] fp16 row-major (bdq_gemm_f16): the B tile is a second TMA load
// (box 64 k x 256 channels, the same K-major SWIZZLE_128B layout the dequant producers write), the producer warps only
// run the epilogue.  Used for weight formats whose in-register decode is too expensive to hide behind the MMAs (packed NF3).
template <int MT, bool BTMA, int NT>
__global__ void __launch_bounds__(kGmThreads, 1) gemm_w2_kernel(const __grid_constant__ CUtensorMap xmap,
                                                                const __grid_constant__ CUtensorMap wmap,
                                                                const uint32_t* __restrict__ qweight,
                                                                const __half* __restrict__ scales,
                                                                const uint32_t* __restrict__ qzeros, __half* __restrict__ out,
                                                                int M, int IC, int OC, int G, int zw, int ldo,
                                                                const GemmPeers peers) {
    constexpr int kGmStages = GmCfg<MT, NT>::kStages, kGmStageBytes = GmCfg<MT, NT>::kStageBytes, kGmBBytes = GmCfg<MT, NT>::kBBytes;
    constexpr int kGmN = NT;
    extern __shared__ uint8_t gm_smem_raw[];
    // align on the 32-bit shared address and ADD the offset, so that the pointer keeps its shared address space (STS / LDS
    // instead of generic ST / LD: the round trip through uintptr_t made all 280 stores of this kernel generic ones)
    uint8_t* smem = gm_smem_raw + ((1024u - (gm_smem_u32(gm_smem_raw) & 1023u)) & 1023u);
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + kGmStages * kGmStageBytes);
    uint64_t* full_a = bars;                    // [stages] TMA landed
    uint64_t* full_b = bars + kGmStages;        // [stages] dequant written (one arrival per producer warp)
    uint64_t* empty = bars + 2 * kGmStages;     // [stages] MMAs that read the stage retired
    uint64_t* tmem_full = bars + 3 * kGmStages;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 3 * kGmStages + 1);

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int n0 = blockIdx.x * kGmN, m0 = blockIdx.y * (kGmM * MT);
    const int nkb = IC / kGmK;

    if (tid == 0) {
        for (int s = 0; s < kGmStages; ++s) {
            gm_mbar_init(&full_a[s], 1);
            gm_mbar_init(&full_b[s], BTMA ? 1 : kGmN / 32);   // one arrival per producer warp
            gm_mbar_init(&empty[s], 1);
        }
        gm_mbar_init(tmem_full, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("prefetch.tensormap [%0];" ::"l"(&xmap) : "memory");
        if constexpr (BTMA) asm volatile("prefetch.tensormap [%0];" ::"l"(&wmap) : "memory");
    }
    if (warp == 1) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(gm_smem_u32(tmem_slot)), "r"(NT * MT) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem_d = *tmem_slot;

    if (warp == 0) {
        // ================= TMA producer: x tile [128 * MT tokens x 64 k] per stage (one box) =================
        if (lane == 0) {
            for (int kb = 0; kb < nkb; ++kb) {
                const int s = kb % kGmStages;
                if (kb >= kGmStages) gm_mbar_wait(&empty[s], ((kb / kGmStages) - 1) & 1);
                gm_mbar_expect_tx(&full_a[s], MT * kGmABytes);
                gm_tma_load_2d(smem + s * kGmStageBytes, &xmap, kb * kGmK, m0, &full_a[s]);
                if constexpr (BTMA) {
                    gm_mbar_expect_tx(&full_b[s], kGmBBytes);
                    gm_tma_load_2d(smem + s * kGmStageBytes + MT * kGmABytes, &wmap, kb * kGmK, n0, &full_b[s]);
                }
            }
        }
    } else if (warp == 1) {
        // ================= MMA issuer =================
        if (lane == 0) {
            for (int kb = 0; kb < nkb; ++kb) {
                const int s = kb % kGmStages;
                const uint32_t ph = (kb / kGmStages) & 1;
                gm_mbar_wait(&full_a[s], ph);
                gm_mbar_wait(&full_b[s], ph);
                asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                const uint32_t a_addr = gm_smem_u32(smem + s * kGmStageBytes);
                const uint64_t bdesc = gm_desc(a_addr + MT * kGmABytes);
#pragma unroll
                for (int mt = 0; mt < MT; ++mt) {
                    const uint64_t adesc = gm_desc(a_addr + mt * kGmABytes);
#pragma unroll
                    for (int k = 0; k < kGmK / 16; ++k)   // advance 32 bytes (>> 4 = 2) inside the swizzle row per K = 16
                        gm_mma(tmem_d + mt * NT, adesc + 2 * k, bdesc + 2 * k, GmIdesc<NT>::v, (kb | k) != 0 ? 1u : 0u);
                }
                gm_commit(&empty[s]);                 // frees the stage when these MMAs have read it
            }
            gm_commit(tmem_full);                     // accumulator complete
        }
    } else {
        // ================= dequant producers: thread j <-> output channel n0 + j =================
        const int jt = tid - 64;
        // NT = 256: thread jt owns channel jt and every k block.  NT = 128: channel jt % 128 and the pairs of k blocks gp with
        // gp % 2 == jt / 128 (a 32-byte load is one pair), so the two halves of the producer warps fill different stages.
        // (16 producer warps with one k block per 16-byte load were measured too: 10 % slower on every shape.)
        const int j = (NT == 256) ? jt : (jt & 127);
        constexpr int kPairStep = (NT == 256) ? 1 : 2;
        const int gp0 = (NT == 256) ? 0 : (jt >> 7);
        if constexpr (!BTMA) {
        const int oc = min(n0 + j, OC - 1);   // a partial last tile re-reads the last channel; its columns are never stored
        const int wpr = IC >> 4;
        const uint32_t* wrow = qweight + (size_t)oc * wpr;
        const __half* srow = scales + (size_t)oc * zw * 16;
        const uint32_t* zrow = qzeros + (size_t)oc * zw;
        const int npair = nkb >> 1;   // 128 codes (two k blocks) per 32-byte load
        const int gshift = (G == 128) ? 1 : 0;   // k blocks per quantisation group: 2 (G = 128) or 1 (G = 64)
        // software prefetch one pair ahead: packed words, the pair's (up to) two scales and its zero-point word(s)
        U8 cur = ldg256(wrow + (gp0 < npair ? gp0 : 0) * 8), nxt = cur;
        __half csc[2], nsc[2];
        uint32_t czw[2], nzw[2];
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            const int grp = ((gp0 < npair ? gp0 : 0) * 2 + h) >> gshift;
            csc[h] = srow[grp];
            czw[h] = __ldg(zrow + (grp >> 4));
            nsc[h] = csc[h];
            nzw[h] = czw[h];
        }
        for (int gp = gp0; gp < npair; gp += kPairStep) {
            if (gp + kPairStep < npair) {
                nxt = ldg256(wrow + (gp + kPairStep) * 8);
#pragma unroll
                for (int h = 0; h < 2; ++h) {
                    const int grp = ((gp + kPairStep) * 2 + h) >> gshift;
                    nsc[h] = srow[grp];
                    nzw[h] = __ldg(zrow + (grp >> 4));
                }
            }
#pragma unroll
            for (int h = 0; h < 2; ++h) {
                const int kb = gp * 2 + h;
                const int grp = kb >> gshift;
                const float z = (float)((czw[h] >> ((grp & 15) * 2)) & 3u);
                const __half2 s2 = __half2half2(csc[h]);
                const __half2 kz1024 = __float2half2_rn(1024.f + z), kz64 = __float2half2_rn(64.f + z), kz4 = __float2half2_rn(4.f + z);
                uint32_t o[32];
#pragma unroll
                for (int w4 = 0; w4 < 4; ++w4) gm_dequant16(cur.r[h * 4 + w4], s2, kz1024, kz64, kz4, o + w4 * 8);
                const int s = kb % kGmStages;
                if (kb >= kGmStages) gm_mbar_wait(&empty[s], ((kb / kGmStages) - 1) & 1);
                uint8_t* brow = smem + s * kGmStageBytes + MT * kGmABytes + j * 128;
#pragma unroll
                for (int c = 0; c < 8; ++c)   // 16-byte chunk c of row j lives at chunk position c ^ (j % 8)
                    *reinterpret_cast<uint4*>(brow + ((c ^ (j & 7)) << 4)) = make_uint4(o[4 * c], o[4 * c + 1], o[4 * c + 2], o[4 * c + 3]);
            }
            // ONE proxy fence for the pair of rows: the fence (MEMBAR + FENCE.VIEW.ASYNC) is where the producer warps spend a fifth of
            // their time (ncu), so it is paid once per two k blocks; the first stage's arrival waits for the second row's stores
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");   // generic-proxy writes -> visible to the MMA (async proxy)
            __syncwarp();
            if (lane == 0) {   // one arrival per warp and stage
                gm_mbar_arrive(&full_b[(gp * 2) % kGmStages]);
                gm_mbar_arrive(&full_b[(gp * 2 + 1) % kGmStages]);
            }
            cur = nxt;
#pragma unroll
            for (int h = 0; h < 2; ++h) { csc[h] = nsc[h]; czw[h] = nzw[h]; }
        }
        }   // !BTMA
        (void)j; (void)gp0;
        // ================= epilogue: TMEM -> registers -> fp16 -> global =================
        gm_mbar_wait(tmem_full, 0);
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        const int lane_grp = warp & 3;               // this warp may only touch TMEM lanes [32*(warp%4), +32)
        const int col_half = (warp - 2) >> 2;        // 0: columns 0 .. NT/2-1, 1: the other half
        constexpr int kCb = NT / 64;                 // 32-column chunks per warp and token tile
#pragma unroll 1
        for (int cbi = 0; cbi < kCb * MT; ++cbi) {
            const int mt = cbi / kCb, cb = cbi % kCb;
            const int row = m0 + mt * kGmM + lane_grp * 32 + lane;   // token
            const int col = col_half * (NT / 2) + cb * 32;
            uint32_t r[32];
            const uint32_t taddr = tmem_d + ((uint32_t)(lane_grp * 32) << 16) + (uint32_t)(mt * NT + col);
            asm volatile(
                "tcgen05.ld.sync.aligned.32x32b.x32.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];"
                : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
                  "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]),
                  "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]),
                  "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
                : "r"(taddr));
            asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
            if (row < M && n0 + col + 32 <= OC) {
                uint32_t pk[16];
#pragma unroll
                for (int i = 0; i < 16; ++i) {
                    const __half2 h2 = __floats2half2_rn(__uint_as_float(r[2 * i]), __uint_as_float(r[2 * i + 1]));
                    pk[i] = *reinterpret_cast<const uint32_t*>(&h2);
                }
                const size_t oi = (size_t)row * ldo + n0 + col;   // ldo = row pitch of the output (OC, or the full width under TP)
                uint4* dst = reinterpret_cast<uint4*>(out + oi);
#pragma unroll
                for (int i = 0; i < 4; ++i) dst[i] = make_uint4(pk[4 * i], pk[4 * i + 1], pk[4 * i + 2], pk[4 * i + 3]);
                for (int pr = 0; pr < peers.n; ++pr) {   // tensor parallel: the same tile into every peer's output (NVLink stores)
                    uint4* pd = reinterpret_cast<uint4*>(peers.p[pr] + oi);
#pragma unroll
                    for (int i = 0; i < 4; ++i) pd[i] = make_uint4(pk[4 * i], pk[4 * i + 1], pk[4 * i + 2], pk[4 * i + 3]);
                }
            }
        }
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 1) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_d), "r"(NT * MT) : "memory");
}

// cuTensorMapEncodeTiled through the runtime's driver entry point lookup (libbdq.so must load on machines without a driver)
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static EncodeTiledFn get_encode_tiled() {
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

int launch_gemm_w2(const void* x, const int32_t* qweight, const void* scales, const int32_t* qzeros, void* out, int64_t M,
                   int64_t IC, int64_t OC, int group, int zw, cudaStream_t st, int64_t ldo, int n_peer, void* const* peer_out,
                   const void* w_f16 = nullptr) {
    GemmPeers peers;
    memset(&peers, 0, sizeof(peers));
    peers.n = n_peer;
    for (int i = 0; i < n_peer; ++i) peers.p[i] = (__half*)peer_out[i];
    EncodeTiledFn enc = get_encode_tiled();
    if (!enc) { set_error("gemm_w2: cuTensorMapEncodeTiled is not available from this driver"); return BDQ_ERR_CUDA; }
    CUtensorMap map;
    const cuuint64_t dims[2] = {(cuuint64_t)IC, (cuuint64_t)M};          // innermost first
    const cuuint64_t strides[1] = {(cuuint64_t)IC * 2};                   // bytes, dims 1..rank-1
    // Tile choice: 128 * MT tokens x NT channels per CTA, by estimated efficiency = CTAs / (waves * SMs) x the tile's own
    // efficiency, measured on full grids at M = 2048 (profiles/r02_gemm_tiles.txt): (2, 256) 1.0, (2, 128) 0.77, (1, 256) 0.68,
    // (1, 128) 0.57 -- the smaller the tile, the more often x is re-read from L2 per MAC, which is what bounds them.
    // Qwen3-8B at M = 2048: everything keeps (2, 256) except k/v_proj alone, 64 CTAs of (1, 256) -> 128 of (1, 128)
    // (0.30 -> 0.51 of the sustained bf16 peak).  BDQ_GEMM_TILE=<MT><NT/128> (e.g. 21) forces one (dev).
    int MT = 1, NT = 256;
    {
        static const int forced = getenv("BDQ_GEMM_TILE") ? atoi(getenv("BDQ_GEMM_TILE")) : 0;
        const double tile_eff[2][2] = {{0.57, 0.68}, {0.77, 1.00}};   // [MT - 1][NT / 128 - 1]
        double best = -1.0;
        for (int mt = 2; mt >= 1; --mt)
            for (int nt = 256; nt >= 128; nt -= 128) {
                if (mt == 2 && M <= kGmM) continue;
                const int64_t tn = (OC + nt - 1) / nt, tm = (M + mt * kGmM - 1) / (mt * kGmM);
                const int64_t ctas = tn * tm, waves = (ctas + sm_count() - 1) / sm_count();
                const double useful = (double)OC * (double)M / ((double)(tn * nt) * (double)(tm * mt * kGmM));   // padding of partial tiles
                const double e = (double)ctas / (double)(waves * sm_count()) * tile_eff[mt - 1][nt / 128 - 1] * useful;
                if (e > best * 1.02) { best = e; MT = mt; NT = nt; }   // ties go to the larger tile
            }
        if (forced == 11 || forced == 12 || forced == 21 || forced == 22) { MT = forced / 10; NT = 128 * (forced % 10); if (M <= kGmM) MT = 1; }
    }
    const cuuint32_t box[2] = {(cuuint32_t)kGmK, (cuuint32_t)(kGmM * MT)};
    const cuuint32_t estr[2] = {1, 1};
    CUresult r = enc(&map, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, const_cast<void*>(x), dims, strides, box, estr,
                     CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                     CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) { set_error("gemm_w2: cuTensorMapEncodeTiled failed (%d)", (int)r); return BDQ_ERR_CUDA; }
    CUtensorMap wmap;
    memset(&wmap, 0, sizeof(wmap));
    if (w_f16) {   // plain fp16 weights [OC, IC]: B tiles by TMA
        const cuuint64_t wd[2] = {(cuuint64_t)IC, (cuuint64_t)OC};
        const cuuint64_t ws[1] = {(cuuint64_t)IC * 2};
        const cuuint32_t wb[2] = {(cuuint32_t)kGmK, (cuuint32_t)NT};
        r = enc(&wmap, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, const_cast<void*>(w_f16), wd, ws, wb, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        if (r != CUDA_SUCCESS) { set_error("gemm_f16: cuTensorMapEncodeTiled failed (%d)", (int)r); return BDQ_ERR_CUDA; }
    }
    dim3 grid((unsigned)((OC + NT - 1) / NT), (unsigned)((M + kGmM * MT - 1) / (kGmM * MT)));
#define BDQ_GM(MTT, BT, NTT)                                                                                                         \
    do {                                                                                                                              \
        if (!ensure_dyn_smem(reinterpret_cast<const void*>(gemm_w2_kernel<MTT, BT, NTT>), GmCfg<MTT, NTT>::kSmem))                    \
            return check_launch("bdq_gemm_w2(attr)");                                                                                 \
        gemm_w2_kernel<MTT, BT, NTT><<<grid, kGmThreads, GmCfg<MTT, NTT>::kSmem, st>>>(                                               \
            map, wmap, (const uint32_t*)qweight, (const __half*)scales, (const uint32_t*)qzeros, (__half*)out, (int)M, (int)IC,       \
            (int)OC, group, zw, (int)ldo, peers);                                                                                     \
    } while (0)
#define BDQ_GM_B(BT)                                                      \
    do {                                                                  \
        if (MT == 2 && NT == 256) BDQ_GM(2, BT, 256);                     \
        else if (MT == 1 && NT == 256) BDQ_GM(1, BT, 256);                \
        else if (MT == 2) BDQ_GM(2, BT, 128);                             \
        else BDQ_GM(1, BT, 128);                                          \
    } while (0)
    if (w_f16) BDQ_GM_B(true);
    else BDQ_GM_B(false);
#undef BDQ_GM_B
#undef BDQ_GM
    return check_launch(w_f16 ? "bdq_gemm_f16" : "bdq_gemm_w2");
}

}  // namespace bdq

using namespace bdq;

extern "C" int bdq_gemv_w2(const void* x, const int32_t* qweight, const void* scales, const int32_t* qzeros, void* out,
                           int64_t B, int64_t IC, int64_t OC, int group, int flags, bdq_stream_t stream);

extern "C" int bdq_gemm_w2(const void* x, const int32_t* qweight, const void* scales, const int32_t* qzeros, void* out,
                           int64_t M, int64_t IC, int64_t OC, int group, bdq_stream_t stream) {
    if (M < 0 || IC <= 0 || OC <= 0 || group <= 0 || IC % group != 0 || IC % 16 != 0) {
        set_error("gemm_w2: M=%lld IC=%lld OC=%lld group=%d", (long long)M, (long long)IC, (long long)OC, group);
        return BDQ_ERR_SHAPE;
    }
    const int zw = bdq_zeros_width(IC, group, 2);
    if (zw < 0) { set_error("gemm_w2: group=%d unsupported", group); return BDQ_ERR_UNSUPPORTED; }
    if (M == 0) return BDQ_OK;
    if (!x || !qweight || !scales || !qzeros || !out) { set_error("null tensor pointer"); return BDQ_ERR_SHAPE; }
    const bool fast = (group == 128 || group == 64) && IC % 128 == 0 && OC % 32 == 0 && ((uintptr_t)x % 16 == 0) &&
                      ((uintptr_t)qweight % 32 == 0) && ((uintptr_t)out % 16 == 0) && M < (1ll << 31) && IC < (1 << 24);
    if (fast) return launch_gemm_w2(x, qweight, scales, qzeros, out, M, IC, OC, group, zw, (cudaStream_t)stream, OC, 0, nullptr);
    // shapes the tcgen05 tile does not cover: stream the weights once per 8 rows through the GEMV kernel (correct, slow)
    return bdq_gemv_w2(x, qweight, scales, qzeros, out, M, IC, OC, group, 0, stream);
}

// Tensor-parallel prefill GEMM with the all-gather fused into the epilogue: this rank's [M, OC_local] tile block is
// written at row pitch `ldo` (the full output width) into its own output and, over NVLink peer memory, into every
// peer's output at the same position.  out / peer_out[p] already point at this rank's first column.  tcgen05 shapes only.
extern "C" int bdq_gemm_w2_tp(const void* x, const int32_t* qweight, const void* scales, const int32_t* qzeros, void* out,
                              int64_t M, int64_t IC, int64_t OC_local, int group, int64_t ldo, int n_peer, void* const* peer_out,
                              bdq_stream_t stream) {
    if (M <= 0 || IC <= 0 || OC_local <= 0 || ldo < OC_local || n_peer < 0 || n_peer > 7 || (n_peer > 0 && !peer_out) || !x ||
        !qweight || !scales || !qzeros || !out) {
        set_error("gemm_w2_tp: bad arguments (M=%lld IC=%lld OC_local=%lld ldo=%lld n_peer=%d)", (long long)M, (long long)IC,
                  (long long)OC_local, (long long)ldo, n_peer);
        return BDQ_ERR_SHAPE;
    }
    const int zw = bdq_zeros_width(IC, group, 2);
    const bool fast = zw >= 0 && (group == 128 || group == 64) && IC % 128 == 0 && OC_local % 32 == 0 && ((uintptr_t)x % 16 == 0) &&
                      ((uintptr_t)qweight % 32 == 0) && ((uintptr_t)out % 16 == 0) && ldo % 8 == 0 && M < (1ll << 31) && IC < (1 << 24);
    if (!fast) { set_error("gemm_w2_tp: needs group 64/128, IC %% 128 == 0, OC_local %% 32 == 0"); return BDQ_ERR_UNSUPPORTED; }
    for (int i = 0; i < n_peer; ++i)
        if (!peer_out[i] || ((uintptr_t)peer_out[i] % 16)) { set_error("gemm_w2_tp: peer pointer %d null or misaligned", i); return BDQ_ERR_ALIGN; }
    return launch_gemm_w2(x, qweight, scales, qzeros, out, M, IC, OC_local, group, zw, (cudaStream_t)stream, ldo, n_peer, peer_out);
}

// Plain fp16 GEMM on the same tcgen05 kernel: out [M, OC] = x [M, IC] @ w [OC, IC]^T, fp32 accumulate, both operands by TMA.
// This library's own dense GEMM for weight formats that are dequantised to fp16 first (WQLinearNF3's prefill: bdq_nf3_dequant
// + this, no cuBLAS); the reference has no counterpart (its NF3 never leaves the fake-quantiser).
extern "C" int bdq_gemm_f16(const void* x, const void* w, void* out, int64_t M, int64_t IC, int64_t OC, bdq_stream_t stream) {
    if (M < 0 || IC <= 0 || OC <= 0) { set_error("gemm_f16: M=%lld IC=%lld OC=%lld", (long long)M, (long long)IC, (long long)OC); return BDQ_ERR_SHAPE; }
    if (M == 0) return BDQ_OK;
    if (!x || !w || !out) { set_error("null tensor pointer"); return BDQ_ERR_SHAPE; }
    if (IC % 64 != 0 || OC % 32 != 0 || ((uintptr_t)x % 16) || ((uintptr_t)w % 16) || ((uintptr_t)out % 16) || M >= (1ll << 31) || IC >= (1 << 24)) {
        set_error("gemm_f16: needs IC %% 64 == 0, OC %% 32 == 0 and 16-byte aligned tensors");
        return BDQ_ERR_UNSUPPORTED;
    }
    return launch_gemm_w2(x, nullptr, nullptr, nullptr, out, M, IC, OC, 128, 0, (cudaStream_t)stream, OC, 0, nullptr, w);
}
