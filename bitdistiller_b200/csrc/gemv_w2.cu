// Packed int2 decode GEMV (sm_100a).  Replaces gemv_kernel_g128_2bit / gemv_forward_cuda
// (inference/kernels/csrc/quantization/gemv_cuda.cu:202-272, 289-346).
//
// HBM bound at 0.26758 B per weight.  A decode GEMV (1-13 MB) is smaller than the bandwidth-delay product of the chip,
// so the design is about keeping bytes in flight, not about FLOPs:
//   * persistent CTAs, one per SM, each owning a contiguous slab of ~OC/148 output channels (perfect HBM balance);
//   * the slab streams through a ring of shared-memory stages with 16-byte cp.async (LDGSTS, L1 bypass), every
//     thread issuing its share: one stage = 16 channels x 4096 input channels of packed codes, rows padded by 32 B so
//     the 64-bit fragment reads are bank-conflict free; scales and zero points ride in a small side ring.  (A first
//     version used one producer lane issuing cp.async.bulk per row: measured 190 cycles per 1 KB row copy, i.e.
//     1.2 TB/s chip-wide -- see profiles/r01_gemv_notes.md);
//   * the first stages are requested before anything else, so under programmatic dependent launch (BDQ_GEMV_PDL)
//     the weight stream of launch n+1 starts while launch n is still computing (two CTAs fit per SM);
//     griddepcontrol.wait comes only before x is read;
//   * 8 consumer warps split the K range of a stage by quantisation group.  2-bit codes become fp16 without a shift
//     per code: (word & mask) | exponent yields (K + code), K in {4,16,64}, for three mantissa positions; two shifted
//     copies of the word cover all 16 codes (10 ALU ops per 16 codes).  The multiply-accumulate runs on the tensor
//     pipe (mma.sync.m16n8k16, fp32 accumulate) with A = 16 channels x 16 codes and B = up to 8 activation rows, so
//     batch <= 8 costs the same weight traffic and instruction count as batch 1.  Offsets and zero points leave in
//     fp32 once per group:  out[b,oc] += s[oc,g] * (acc - sum_k K(k) x[b,k] - z[oc,g] * sum_k x[b,k]);
//     the two x-only sums are computed once per CTA while x is staged (permuted to extraction order) in smem;
//   * split-K partials of the 8 warps meet in shared memory once per 16-channel block (no atomics, no second pass).
#include "bdq_common.cuh"

namespace bdq {

constexpr int kCW = 8;                        // consumer warps
constexpr int kGemvThreads = kCW * 32;
constexpr int kSliceBytes = 1024;             // packed bytes of one row per stage (4096 codes = 32 groups)
constexpr int kRowPad = kSliceBytes + 32;     // smem row pitch
constexpr int kStageBytes = 16 * kRowPad;
constexpr int kMaxStages = 8;

// extraction e (0..7) yields codes p_e and p_e + 8 with p = {2,3,4,5,6,7,0,1}; ext_k(e) is its additive offset K
__host__ __device__ constexpr float ext_k(int e) { return (e == 0 || e == 3 || e == 6) ? 64.f : ((e == 2 || e == 5) ? 4.f : 16.f); }

__device__ __forceinline__ uint32_t lop_and_or(uint32_t a, uint32_t b, uint32_t c) {
    uint32_t r;
    asm("lop3.b32 %0, %1, %2, %3, 0xEA;" : "=r"(r) : "r"(a), "r"(b), "r"(c));  // (a & b) | c
    return r;
}

// 8 half2 = (K_e + code[p_e], K_e + code[p_e + 8]) for e = 0..7
__device__ __forceinline__ void extract16(uint32_t v, uint32_t* h) {
    const uint32_t v1 = v >> 6, v2 = v << 4;
    h[0] = lop_and_or(v, 0x00300030u, 0x54005400u);   // code 2 -> 64 + c
    h[1] = lop_and_or(v, 0x00C000C0u, 0x4C004C00u);   // code 3 -> 16 + c
    h[2] = lop_and_or(v, 0x03000300u, 0x44004400u);   // code 4 ->  4 + c
    h[3] = lop_and_or(v1, 0x00300030u, 0x54005400u);  // code 5
    h[4] = lop_and_or(v1, 0x00C000C0u, 0x4C004C00u);  // code 6
    h[5] = lop_and_or(v1, 0x03000300u, 0x44004400u);  // code 7
    h[6] = lop_and_or(v2, 0x00300030u, 0x54005400u);  // code 0
    h[7] = lop_and_or(v2, 0x00C000C0u, 0x4C004C00u);  // code 1
}

__device__ __forceinline__ void mma16816(float* d, uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3, uint32_t b0, uint32_t b1) {
    asm("mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
                 : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}

// coherent 128-bit load (x may have been written by the kernel this launch programmatically depends on)
__device__ __forceinline__ uint4 ld_global_v4(const void* p) {
    uint4 v;
    asm volatile("ld.global.v4.b32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p) : "memory");
    return v;
}

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void cp_async16(void* dst, const void* src) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(smem_u32(dst)), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
__device__ __forceinline__ void cp_async_wait(int n) {   // wait until at most n groups are pending
    switch (n) {
        case 0: asm volatile("cp.async.wait_group 0;" ::: "memory"); break;
        case 1: asm volatile("cp.async.wait_group 1;" ::: "memory"); break;
        case 2: asm volatile("cp.async.wait_group 2;" ::: "memory"); break;
        case 3: asm volatile("cp.async.wait_group 3;" ::: "memory"); break;
        case 4: asm volatile("cp.async.wait_group 4;" ::: "memory"); break;
        case 5: asm volatile("cp.async.wait_group 5;" ::: "memory"); break;
        default: asm volatile("cp.async.wait_group 6;" ::: "memory"); break;
    }
}

struct GemvSmem {   // byte offsets inside dynamic shared memory (all multiples of 32)
    int xs, xmeta, red, ring, stages, total;
    int x_stride, meta_slot, meta_zoff, nmeta;
};

__host__ __device__ inline GemvSmem gemv_layout(int B, int IC, int zw, int nstages) {
    GemvSmem L;
    L.x_stride = IC * 2 + 32;
    L.meta_zoff = 16 * zw * 32;                               // scales of 16 rows: 16 * (zw*16) halves
    L.meta_slot = ((L.meta_zoff + 16 * zw * 4 + 16 + 31) / 32) * 32;   // + zero words (16-byte-aligned superset)
    L.nmeta = nstages + 1;
    L.xs = 0;
    L.xmeta = L.xs + B * L.x_stride;
    L.red = L.xmeta + (IC / 128) * 8 * 8;
    L.ring = L.red + 2 * kCW * 128 * 4;
    L.stages = L.ring + L.nmeta * L.meta_slot;
    L.total = L.stages + nstages * kStageBytes;
    return L;
}

template <bool PDL>
__global__ void __launch_bounds__(kGemvThreads, 2) gemv_w2_g128_kernel(
    const __half* __restrict__ x, const uint8_t* __restrict__ qweight, const uint8_t* __restrict__ scales,
    const uint8_t* __restrict__ qzeros, __half* __restrict__ out, int B, int IC, int OC, int zw, int rows_per_cta,
    int nstages) {
    extern __shared__ __align__(128) uint8_t smem[];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const GemvSmem L = gemv_layout(B, IC, zw, nstages);
    if constexpr (PDL) asm volatile("griddepcontrol.launch_dependents;");

    const int row_begin = blockIdx.x * rows_per_cta;
    const int row_end = min(OC, row_begin + rows_per_cta);
    if (row_begin >= row_end) return;
    const int row_bytes = IC >> 2;
    const int nsl = (row_bytes + kSliceBytes - 1) / kSliceBytes;     // K slices per block
    const int nblk = (row_end - row_begin + 15) >> 4;
    const int nunits = nblk * nsl;

    // request the next unit (16 channels x one K slice; with the first slice of a block also its scales / zero
    // points).  Unit coordinates advance incrementally: no divisions in the steady state.
    int i_u = 0, i_slot = 0, i_b = 0, i_sl = 0, i_meta = 0;
    auto issue = [&]() {
        if (i_u < nunits) {
            const int r = row_begin + i_b * 16;
            const int nrows = min(16, row_end - r);
            const int kbytes = min(kSliceBytes, row_bytes - i_sl * kSliceBytes);
            uint8_t* dst = smem + L.stages + i_slot * kStageBytes;
            const uint8_t* src = qweight + (size_t)r * row_bytes + (size_t)i_sl * kSliceBytes;
            if (kbytes == kSliceBytes) {
                for (int c = tid; c < nrows * 64; c += kGemvThreads)
                    cp_async16(dst + (c >> 6) * kRowPad + (c & 63) * 16, src + (size_t)(c >> 6) * row_bytes + (c & 63) * 16);
            } else {
                const int cpr = kbytes >> 4;
                for (int c = tid; c < nrows * cpr; c += kGemvThreads) {
                    const int row = c / cpr, col = c - row * cpr;
                    cp_async16(dst + row * kRowPad + col * 16, src + (size_t)row * row_bytes + col * 16);
                }
            }
            if (i_sl == 0) {
                uint8_t* mslot = smem + L.ring + i_meta * L.meta_slot;
                const size_t z0a = ((size_t)r * zw * 4) & ~(size_t)15;
                const int zchunks = (int)(((((size_t)(r + nrows) * zw * 4 + 15) & ~(size_t)15) - z0a) >> 4);
                const int schunks = nrows * zw * 2;
                const uint8_t* ssrc = scales + (size_t)r * zw * 32;
                for (int c = tid; c < schunks + zchunks; c += kGemvThreads) {
                    if (c < schunks) cp_async16(mslot + c * 16, ssrc + c * 16);
                    else cp_async16(mslot + L.meta_zoff + (c - schunks) * 16, qzeros + z0a + (size_t)(c - schunks) * 16);
                }
            }
            ++i_u;
            if (++i_slot == nstages) i_slot = 0;
            if (++i_sl == nsl) {
                i_sl = 0;
                ++i_b;
                if (++i_meta == L.nmeta) i_meta = 0;
            }
        }
        cp_async_commit();
    };
    for (int u = 0; u < nstages - 1; ++u) issue();   // weights first: nothing here depends on the previous kernel

    const int g = lane >> 2, t = lane & 3;
    const int wpr = IC >> 4, ngrp = IC >> 7;
    uint8_t* xs = smem + L.xs;
    float2* xmeta = reinterpret_cast<float2*>(smem + L.xmeta);   // [ngrp][8] (sum K x, sum x) per activation row
    float* red = reinterpret_cast<float*>(smem + L.red);         // [2][kCW][16][8]

    // ---- stage x (permuted to extraction order) and the per-group x sums
    if constexpr (PDL) asm volatile("griddepcontrol.wait;" ::: "memory");
    for (int n = 0; n < B; ++n) {
        for (int kb = tid; kb < ((wpr + 31) & ~31); kb += kGemvThreads) {   // whole warps stay in the loop (shuffles)
            const bool valid = kb < wpr;   // wpr % 8 == 0: every 8-lane group is all-valid or all-invalid
            const uint4* src = reinterpret_cast<const uint4*>(x + (size_t)n * IC + (size_t)(valid ? kb : 0) * 16);
            const uint4 v0 = ld_global_v4(src), v1 = ld_global_v4(src + 1);
            const uint32_t in[8] = {v0.x, v0.y, v0.z, v0.w, v1.x, v1.y, v1.z, v1.w};   // (x0,x1) (x2,x3) ... (x14,x15)
            uint32_t o[8];
            float bias = 0.f, sum = 0.f;
#pragma unroll
            for (int e = 0; e < 8; ++e) {
                const int p = (e < 6) ? e + 2 : e - 6;
                const uint32_t lo = in[p >> 1], hi = in[(p + 8) >> 1];
                o[e] = (p & 1) ? __byte_perm(lo, hi, 0x7632) : __byte_perm(lo, hi, 0x5410);   // (x[p], x[p+8])
                const float2 f = __half22float2(*reinterpret_cast<const __half2*>(&o[e]));
                sum += f.x + f.y;
                bias += ext_k(e) * (f.x + f.y);
            }
            if (valid) {
                uint4* dst = reinterpret_cast<uint4*>(xs + (size_t)n * L.x_stride + (size_t)kb * 32);
                dst[0] = make_uint4(o[0], o[1], o[2], o[3]);
                dst[1] = make_uint4(o[4], o[5], o[6], o[7]);
            }
#pragma unroll
            for (int s = 1; s < 8; s <<= 1) {   // 8 consecutive k-blocks = one group of 128
                bias += __shfl_xor_sync(0xffffffffu, bias, s);
                sum += __shfl_xor_sync(0xffffffffu, sum, s);
            }
            if (valid && (kb & 7) == 0) xmeta[(size_t)(kb >> 3) * 8 + n] = make_float2(bias, sum);
        }
    }
    for (int i = tid; i < ngrp * 8; i += kGemvThreads)
        if ((i & 7) >= B) xmeta[i] = make_float2(0.f, 0.f);
    const bool has_b = g < B;
    const uint8_t* xrow = xs + (size_t)(has_b ? g : 0) * L.x_stride;
    float fin[4] = {0.f, 0.f, 0.f, 0.f};
    int slot = 0, b = 0, sl = 0, c_meta = 0;
    for (int u = 0; u < nunits; ++u) {
        const int ngs = min(kSliceBytes, row_bytes - sl * kSliceBytes) >> 5;   // groups in this slice
        cp_async_wait(nstages - 2);   // this thread's share of unit u has landed
        __syncthreads();              // everyone's share has; everyone is done with unit u-1 (and with x staging)
        issue();                      // refill the slot unit u-1 used
        const uint8_t* st = smem + L.stages + slot * kStageBytes;
        const uint8_t* mslot = smem + L.ring + c_meta * L.meta_slot;
        const int zskew = (int)(((size_t)(row_begin + b * 16) * zw * 4) & 15);
        // this warp's 4 consecutive groups of the slice: gi = 4*warp + j
        uint2 wa[4], wb[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const int gi = warp * 4 + j;
            if (gi < ngs) {
                wa[j] = *reinterpret_cast<const uint2*>(st + g * kRowPad + gi * 32 + t * 8);
                wb[j] = *reinterpret_cast<const uint2*>(st + (g + 8) * kRowPad + gi * 32 + t * 8);
            } else {
                wa[j] = wb[j] = make_uint2(0, 0);
            }
        }
        if (warp * 4 < ngs) {
            const int grp0 = sl * 32 + warp * 4;
            // 4 scales per row in one 64-bit read, the zero points of the 4 groups sit in one word
            const uint2 s4a = *reinterpret_cast<const uint2*>(mslot + (size_t)(g * (zw * 16) + grp0) * 2);
            const uint2 s4b = *reinterpret_cast<const uint2*>(mslot + (size_t)((g + 8) * (zw * 16) + grp0) * 2);
            const uint32_t* zq = reinterpret_cast<const uint32_t*>(mslot + L.meta_zoff + zskew);
            const uint32_t zwa = zq[g * zw + (grp0 >> 4)] >> ((grp0 & 15) * 2);
            const uint32_t zwb = zq[(g + 8) * zw + (grp0 >> 4)] >> ((grp0 & 15) * 2);
            float d[4][4];
#pragma unroll
            for (int j = 0; j < 4; ++j) d[j][0] = d[j][1] = d[j][2] = d[j][3] = 0.f;
#pragma unroll
            for (int uu = 0; uu < 2; ++uu) {
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    uint32_t ha[8], hb[8];
                    extract16(uu ? wa[j].y : wa[j].x, ha);
                    extract16(uu ? wb[j].y : wb[j].x, hb);
                    const int kb = (grp0 + j) * 8 + t * 2 + uu;
                    // lanes whose column g >= B read row 0: their output columns are never stored
                    const uint4* bp = reinterpret_cast<const uint4*>(xrow + (size_t)kb * 32);
                    const uint4 b0 = bp[0], b1 = bp[1];
                    mma16816(d[j], ha[0], hb[0], ha[1], hb[1], b0.x, b0.y);
                    mma16816(d[j], ha[2], hb[2], ha[3], hb[3], b0.z, b0.w);
                    mma16816(d[j], ha[4], hb[4], ha[5], hb[5], b1.x, b1.y);
                    mma16816(d[j], ha[6], hb[6], ha[7], hb[7], b1.z, b1.w);
                }
            }
            // per-group fix-up: remove the K offsets and the zero point, apply the scale
            const __half2 sa01 = *reinterpret_cast<const __half2*>(&s4a.x), sa23 = *reinterpret_cast<const __half2*>(&s4a.y);
            const __half2 sb01 = *reinterpret_cast<const __half2*>(&s4b.x), sb23 = *reinterpret_cast<const __half2*>(&s4b.y);
            const float2 fa01 = __half22float2(sa01), fa23 = __half22float2(sa23);
            const float2 fb01 = __half22float2(sb01), fb23 = __half22float2(sb23);
            const float s_lo[4] = {fa01.x, fa01.y, fa23.x, fa23.y}, s_hi[4] = {fb01.x, fb01.y, fb23.x, fb23.y};
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                if (warp * 4 + j >= ngs) break;
                const float z_lo = (float)((zwa >> (2 * j)) & 3u), z_hi = (float)((zwb >> (2 * j)) & 3u);
                const float4 m = *reinterpret_cast<const float4*>(&xmeta[(size_t)(grp0 + j) * 8 + t * 2]);   // cols 2t, 2t+1
                fin[0] = fmaf(s_lo[j], d[j][0] - fmaf(z_lo, m.y, m.x), fin[0]);
                fin[1] = fmaf(s_lo[j], d[j][1] - fmaf(z_lo, m.w, m.z), fin[1]);
                fin[2] = fmaf(s_hi[j], d[j][2] - fmaf(z_hi, m.y, m.x), fin[2]);
                fin[3] = fmaf(s_hi[j], d[j][3] - fmaf(z_hi, m.w, m.z), fin[3]);
            }
        }

        if (sl == nsl - 1) {
            // ---- cross-warp (split-K) reduction and store of one 16-channel block
            float* my = red + (size_t)((b & 1) * kCW + warp) * 128;
            *reinterpret_cast<float2*>(&my[g * 8 + t * 2]) = make_float2(fin[0], fin[1]);
            *reinterpret_cast<float2*>(&my[(g + 8) * 8 + t * 2]) = make_float2(fin[2], fin[3]);
            fin[0] = fin[1] = fin[2] = fin[3] = 0.f;
            __syncthreads();
            if (tid < 128) {
                const int row = tid & 15, col = tid >> 4;   // consecutive threads -> consecutive output channels
                const int r = row_begin + b * 16 + row;
                if (col < B && r < row_end) {
                    const float* rp = red + (size_t)(b & 1) * kCW * 128 + row * 8 + col;
                    float acc = 0.f;
#pragma unroll
                    for (int w = 0; w < kCW; ++w) acc += rp[w * 128];
                    out[(size_t)col * OC + r] = __float2half_rn(acc);
                }
            }
        }
        if (++slot == nstages) slot = 0;
        if (++sl == nsl) {
            sl = 0;
            ++b;
            if (++c_meta == L.nmeta) c_meta = 0;
        }
    }
}

// Generic fallback (any group that is a multiple of 16): one warp per output channel, CUDA-core math.
__global__ void __launch_bounds__(256) gemv_w2_generic_kernel(const __half* __restrict__ x, const uint32_t* __restrict__ qweight,
                                                              const __half* __restrict__ scales, const uint32_t* __restrict__ qzeros,
                                                              __half* __restrict__ out, int IC, int OC, int group, int zw) {
    const int lane = threadIdx.x & 31;
    const int oc = blockIdx.x * 8 + (threadIdx.x >> 5);
    const int b = blockIdx.y;
    if (oc >= OC) return;
    const int wpr = IC >> 4;
    float acc = 0.f;
    for (int wi = lane; wi < wpr; wi += 32) {
        const uint32_t v = __ldg(qweight + (size_t)oc * wpr + wi);
        const int grp = (wi * 16) / group;
        const float s = __half2float(scales[(size_t)oc * zw * 16 + grp]);
        const float z = (float)((__ldg(qzeros + (size_t)oc * zw + (grp >> 4)) >> ((grp & 15) * 2)) & 3u);
        float part = 0.f;
#pragma unroll
        for (int k = 0; k < 16; ++k)
            part += ((float)((v >> (2 * k)) & 3u) - z) * __half2float(x[(size_t)b * IC + wi * 16 + k]);
        acc = fmaf(s, part, acc);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
    if (lane == 0) out[(size_t)b * OC + oc] = __float2half_rn(acc);
}

}  // namespace bdq

using namespace bdq;

int bdq_gemv_w2_stream_launch(const void* x, int nmat, const int32_t* const* qweight, const void* const* scales,
                              const int32_t* const* qzeros, void* const* out, const int64_t* OC, int64_t IC, int flags,
                              cudaStream_t st, int n_prefetch, const void* const* prefetch, const int64_t* prefetch_bytes,
                              int n_peer, void* const* peer_out);

extern "C" int bdq_gemv_w2(const void* x, const int32_t* qweight, const void* scales, const int32_t* qzeros, void* out,
                           int64_t B, int64_t IC, int64_t OC, int group, int flags, bdq_stream_t stream) {
    if (B < 0 || IC <= 0 || OC <= 0 || group <= 0 || IC % group != 0 || IC % 16 != 0 || group % 16 != 0) {
        set_error("gemv_w2: B=%lld IC=%lld OC=%lld group=%d: IC must be a multiple of group, group of 16", (long long)B,
                  (long long)IC, (long long)OC, group);
        return BDQ_ERR_SHAPE;
    }
    const int zw = bdq_zeros_width(IC, group, 2);
    if (zw < 0) { set_error("gemv_w2: group=%d unsupported", group); return BDQ_ERR_UNSUPPORTED; }
    if (B == 0) return BDQ_OK;
    if (!x || !qweight || !scales || !qzeros || !out) { set_error("null tensor pointer"); return BDQ_ERR_SHAPE; }
    if (IC > (1 << 20) || OC > (1 << 24)) { set_error("gemv_w2: shape too large"); return BDQ_ERR_SHAPE; }
    cudaStream_t st = (cudaStream_t)stream;
    if (B == 1 && group == 128 && !(flags & BDQ_GEMV_RING)) {   // batch 1: register-streaming kernel (gemv_w2_stream.cu)
        const int32_t* qw1[1] = {qweight};
        const void* sc1[1] = {scales};
        const int32_t* qz1[1] = {qzeros};
        void* out1[1] = {out};
        const int64_t oc1[1] = {OC};
        const int rc = bdq_gemv_w2_stream_launch(x, 1, qw1, sc1, qz1, out1, oc1, IC, flags, st, 0, nullptr, nullptr, 0, nullptr);
        if (rc != BDQ_ERR_UNSUPPORTED) return rc;
    }
    const bool fast = group == 128 && OC % 16 == 0 && ((uintptr_t)x % 16 == 0) && ((uintptr_t)qweight % 16 == 0) &&
                      ((uintptr_t)scales % 16 == 0) && ((uintptr_t)qzeros % 16 == 0);
    if (!fast) {
        dim3 grid((unsigned)((OC + 7) / 8), (unsigned)B);
        gemv_w2_generic_kernel<<<grid, 256, 0, st>>>((const __half*)x, (const uint32_t*)qweight, (const __half*)scales,
                                                     (const uint32_t*)qzeros, (__half*)out, (int)IC, (int)OC, group, zw);
        return check_launch("bdq_gemv_w2(generic)");
    }
    const bool pdl = (flags & BDQ_GEMV_PDL) != 0;
    auto kern = pdl ? gemv_w2_g128_kernel<true> : gemv_w2_g128_kernel<false>;
    if (!ensure_dyn_smem(reinterpret_cast<const void*>(kern), 227 * 1024)) return check_launch("bdq_gemv_w2(attr)");
    const int nsm = sm_count();
    for (int64_t b0 = 0; b0 < B;) {
        // activation rows per launch: as many (<= 8) as leave room for at least 2 ring stages
        int nb = (int)((B - b0) < 8 ? (B - b0) : 8);
        int nstages = 0;
        for (; nb >= 1; --nb) {
            // prefer a footprint that lets two CTAs share an SM (so a PDL successor can prefetch), else take the whole SM
            for (int s = 4; s >= 2 && !nstages; --s)
                if (gemv_layout(nb, (int)IC, zw, s).total <= 113 * 1024) nstages = s;
            for (int s = kMaxStages; s >= 2 && !nstages; --s)
                if (gemv_layout(nb, (int)IC, zw, s).total <= 227 * 1024) nstages = s;
            if (nstages) break;
        }
        if (!nstages) { set_error("gemv_w2: IC=%lld does not fit in shared memory", (long long)IC); return BDQ_ERR_SHAPE; }
        const GemvSmem L = gemv_layout(nb, (int)IC, zw, nstages);
        // two CTAs per SM when the footprint allows it (more warps to hide the dependent-issue latency), >= 8 channels each
        int grid = L.total <= 113 * 1024 ? 2 * nsm : nsm;
        if ((int64_t)grid * 8 > OC) grid = (int)((OC + 7) / 8);
        int rows_per_cta = (int)((OC + grid - 1) / grid);
        grid = (int)((OC + rows_per_cta - 1) / rows_per_cta);
        cudaLaunchConfig_t cfg = {};
        cfg.gridDim = dim3((unsigned)grid);
        cfg.blockDim = dim3(kGemvThreads);
        cfg.dynamicSmemBytes = (size_t)L.total;
        cfg.stream = st;
        cudaLaunchAttribute at[1];
        at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
        at[0].val.programmaticStreamSerializationAllowed = 1;
        cfg.attrs = at;
        cfg.numAttrs = pdl ? 1 : 0;
        const __half* xp = (const __half*)x + b0 * IC;
        __half* op = (__half*)out + b0 * OC;
        cudaLaunchKernelEx(&cfg, kern, xp, (const uint8_t*)qweight, (const uint8_t*)scales, (const uint8_t*)qzeros, op, nb,
                           (int)IC, (int)OC, zw, rows_per_cta, nstages);
        int rc = check_launch("bdq_gemv_w2");
        if (rc) return rc;
        b0 += nb;
    }
    return BDQ_OK;
}
