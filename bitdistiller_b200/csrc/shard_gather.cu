// Quantise-before-all-gather for dim-0 (FSDP2-style) weight shards.
//
// The reference's QAT step all-gathers the bf16 shards of a weight and then fake-quantises the full weight on every rank
// (QLinear.forward, quantization/qlinear.py:63-78 under FSDP2).  Groups lie along the last dimension, so a dim-0 shard
// holds whole groups: the owner can quantise its shard ONCE and ship the 2-bit codes plus one (scale, zero) pair per group
// -- 0.28 B/weight instead of 2 -- and every rank rebuilds the dequantised weight  dq = rT((c - z) * s)  with exactly the
// last two operations of the reference (quantizer.py:177), so the result is bit-equal to fake-quantising the gathered
// weight.
//
//   bdq_int2_shard_encode: thresholds per group as in fakequant_thr.cu, codes extracted with a magic-number add, and the
//                          payload stored to this rank's slot of up to 8 gather buffers (its own + peer-mapped NVLink
//                          addresses): the all-gather is the kernel's store phase, no collective runs.
//   bdq_int2_shard_decode: payload -> dq, 8 B + 4 B/group read and 64 B written per lane.
//
// Wire format (this build's own; the reference has none).  Per 32 consecutive elements one uint2 = two 32-bit words, one
// per 16 elements: element j of the 16 (pair i = j / 2, half h = j % 2) sits at bits [16h + 2i, 16h + 2i + 2).  Per group
// one (scale, zero) pair of T values (4 bytes).
#include "fakequant_thr.cuh"
#include "fq_int_elem.cuh"

namespace bdq {

constexpr int kGatherThreads = 256;
constexpr int kMaxDst = 8;

struct GatherDst {
    uint8_t* codes[kMaxDst];
    uint8_t* sz[kMaxDst];
    int n;
};

// several shards in one launch (the tensors of one decoder layer): chunk (= 32 elements) index space is the concatenation
struct GatherSegs {
    const uint8_t* w[kMaxDst];
    int64_t begin[kMaxDst];      // first chunk of the segment in the concatenated index space (multiple of G / 32)
    int64_t code_off[kMaxDst];   // byte offset of the segment's codes / (scale, zero) pairs inside every destination
    int64_t sz_off[kMaxDst];
    int n;
};

template <int DT>
struct Magic;
template <>
struct Magic<kBF16> { static constexpr float v = 128.f; static constexpr uint32_t bits2 = 0x43004300u; };   // ulp(128) = 1 in bf16
template <>
struct Magic<kF16> { static constexpr float v = 1024.f; static constexpr uint32_t bits2 = 0x64006400u; };   // ulp(1024) = 1 in fp16

// op-by-op codes of one lane's 32 elements (groups whose thresholds could not be bracketed); re-reads global memory
template <int DT, int RM>
__device__ __noinline__ uint2 encode_slow_global(const void* w, int64_t e0, IntQ p) {
    uint32_t word[2] = {0u, 0u};
#pragma unroll 1
    for (int e = 0; e < 32; ++e) {
        float qu;
        const float c = quant_code<DT, RM>(load_scalar<DT>(w, e0 + e), p, 3.f, qu);
        const int j = e & 15;
        word[e >> 4] |= ((uint32_t)c & 3u) << (16 * (j & 1) + 2 * (j >> 1));
    }
    return make_uint2(word[0], word[1]);
}

template <int LPG>
__device__ __forceinline__ uint32_t lane_group_mask(int lane) {
    return ((LPG == 32) ? 0xffffffffu : ((1u << LPG) - 1u)) << (lane & ~(LPG - 1));
}

template <int DT, int G, int RM, bool MULTI>
__global__ void __launch_bounds__(kGatherThreads) int2_encode_kernel(const uint8_t* __restrict__ w, GatherDst dst, int64_t nchunk,
                                                                     int sem, GatherSegs segs) {
    using P = P2<DT>;
    using h2 = typename P::h2;
    constexpr int LPG = G / 32;
    constexpr int NT = 3;
    const int lane = threadIdx.x & 31;
    const int lig = lane & (LPG - 1);
    const int64_t warp = (int64_t)blockIdx.x * (kGatherThreads / 32) + (threadIdx.x >> 5);
    const int64_t nwarps = (int64_t)gridDim.x * (kGatherThreads / 32);
    const int64_t nwc = (nchunk + 31) / 32;
    for (int64_t wc = warp; wc < nwc; wc += nwarps) {
        const int64_t gidx = wc * 32 + lane;
        const bool valid = gidx < nchunk;
        int64_t idx = gidx, coff = 0, soff = 0;
        if constexpr (MULTI) {
            int sgi = 0;
            for (int q = 1; q < segs.n; ++q) sgi = gidx >= segs.begin[q] ? q : sgi;
            idx = gidx - segs.begin[sgi];
            w = segs.w[sgi];
            coff = segs.code_off[sgi];
            soff = segs.sz_off[sgi];
        }
        U8 raw[2];
        if (valid) {
            raw[0] = ldg256(w + idx * 64);
            raw[1] = ldg256(w + idx * 64 + 32);
        } else {
#pragma unroll
            for (int i = 0; i < 8; ++i) raw[0].r[i] = raw[1].r[i] = 0;
        }
        float mx, mn, mx1, mn1;
        vec_minmax16<DT>(raw[0], mx, mn);
        vec_minmax16<DT>(raw[1], mx1, mn1);
        mx = group_max<LPG>(fmaxf(mx, mx1));
        mn = group_min<LPG>(fminf(mn, mn1));
        const IntQ p = int_group_params<DT>(mx, mn, 3.f, sem);
        const bool safe = fmaxf(fabsf(mx), fabsf(mn)) * p.y < 1e30f;
        // thresholds T_1..T_3 (n = -z + k) of the group: computed by its first lanes (R per lane when G < 128), broadcast
        constexpr int R = (NT + LPG - 1) / LPG;
        float tl[R];
        bool ok = safe;
#pragma unroll
        for (int r = 0; r < R; ++r) {
            const int k = lig + r * LPG;
            tl[r] = 0.f;
            if (k < NT) {
                bool o;
                tl[r] = find_threshold<DT, RM>(-p.z + (float)(1 + k), p, o);
                ok = ok && o;
            }
        }
        const uint32_t okm = __ballot_sync(0xffffffffu, ok);
        const uint32_t gm = lane_group_mask<LPG>(lane);
        h2 t2[NT];
#pragma unroll
        for (int k = 0; k < NT; ++k) t2[k] = P::dup(__shfl_sync(0xffffffffu, tl[k / LPG], (lane & ~(LPG - 1)) + (k % LPG)));
        uint2 cw = make_uint2(0u, 0u);
        if ((okm & gm) == gm) {
            const h2 magic = P::dup(Magic<DT>::v);
            uint32_t word[2];
#pragma unroll
            for (int v = 0; v < 2; ++v) {
                uint32_t acc = 0u;
#pragma unroll
                for (int i = 7; i >= 0; --i) {
                    const h2 x2 = P::from_bits(raw[v].r[i]);
                    h2 c2 = __hadd2(__hge2(x2, t2[0]), magic);      // magic + c: the code is the low 2 bits of each half
                    c2 = __hadd2(c2, __hge2(x2, t2[1]));
                    c2 = __hadd2(c2, __hge2(x2, t2[2]));
                    acc = acc * 4u + P::bits(c2);                   // Horner, unmasked: the magic bits are linear in acc too
                }
                // acc = sum_i (code_i + magic2) * 4^i (mod 2^32); both 16-bit code sums are < 2^16, so removing the
                // magic term leaves exactly the two code fields
                word[v] = acc - Magic<DT>::bits2 * 21845u;
            }
            cw = make_uint2(word[0], word[1]);
        } else if (valid) {
            cw = encode_slow_global<DT, RM>(w, idx * 32, p);
        }
        if (valid) {
            const h2 szp = P::pack(p.s, p.z);
            const int64_t g = idx / LPG;
#pragma unroll 1
            for (int d = 0; d < dst.n; ++d) {
                *reinterpret_cast<uint2*>(dst.codes[d] + coff + idx * 8) = cw;
                if (lig == 0) *reinterpret_cast<uint32_t*>(dst.sz[d] + soff + g * 4) = P::bits(szp);
            }
        }
    }
}

template <int DT, int G>
__global__ void __launch_bounds__(kGatherThreads) int2_decode_kernel(const uint8_t* __restrict__ codes, const uint8_t* __restrict__ sz,
                                                                     uint8_t* __restrict__ dq, int64_t nchunk) {
    using P = P2<DT>;
    using h2 = typename P::h2;
    constexpr int LPG = G / 32;
    const int lane = threadIdx.x & 31;
    const int64_t warp = (int64_t)blockIdx.x * (kGatherThreads / 32) + (threadIdx.x >> 5);
    const int64_t nwarps = (int64_t)gridDim.x * (kGatherThreads / 32);
    const int64_t nwc = (nchunk + 31) / 32;
    for (int64_t wc = warp; wc < nwc; wc += nwarps) {
        const int64_t idx = wc * 32 + lane;
        if (idx >= nchunk) continue;
        const uint2 cw = __ldg(reinterpret_cast<const uint2*>(codes + idx * 8));
        const h2 szp = P::from_bits(__ldg(reinterpret_cast<const uint32_t*>(sz + (idx / LPG) * 4)));
        const h2 s2 = P::dup_lo(szp);
        // (magic + c) - (magic + z) = c - z exactly (integers below 2^8); then rT((c - z) * s) as quantizer.py:177
        const h2 off2 = P::dup(-(Magic<DT>::v + P::hi(szp)));
        const uint32_t word[2] = {cw.x, cw.y};
        U8 out[2];
#pragma unroll
        for (int v = 0; v < 2; ++v)
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                const uint32_t k = ((word[v] >> (2 * i)) & 0x00030003u) | Magic<DT>::bits2;
                out[v].r[i] = P::bits(__hmul2(__hadd2(P::from_bits(k), off2), s2));
            }
        stg256(dq + idx * 64, out[0]);
        stg256(dq + idx * 64 + 32, out[1]);
    }
}

static int check_shard(int64_t n_elem, int group, int dtype) {
    if (dtype != BDQ_BF16 && dtype != BDQ_F16) { set_error("shard gather: dtype must be bf16 or fp16"); return BDQ_ERR_DTYPE; }
    if (group != 32 && group != 64 && group != 128) { set_error("shard gather: group must be 32, 64 or 128"); return BDQ_ERR_UNSUPPORTED; }
    if (n_elem < 0 || n_elem % group != 0) { set_error("shard gather: n_elem=%lld is not a multiple of group=%d", (long long)n_elem, group); return BDQ_ERR_SHAPE; }
    return BDQ_OK;
}

static int64_t gather_blocks(int64_t nchunk) { return ((nchunk + 31) / 32 + (kGatherThreads / 32) - 1) / (kGatherThreads / 32); }

template <int DT, int RM, bool MULTI>
static void encode_g2(const uint8_t* w, const GatherDst& dst, int64_t nchunk, int G, int sem, const GatherSegs& segs, cudaStream_t st) {
    const int64_t nb = gather_blocks(nchunk);
#define BDQ_L(GG) int2_encode_kernel<DT, GG, RM, MULTI><<<wave_grid(int2_encode_kernel<DT, GG, RM, MULTI>, kGatherThreads, nb, 16), kGatherThreads, 0, st>>>(w, dst, nchunk, sem, segs)
    if (G == 128) BDQ_L(128);
    else if (G == 64) BDQ_L(64);
    else BDQ_L(32);
#undef BDQ_L
}

template <int DT, int RM>
static void encode_g(const uint8_t* w, const GatherDst& dst, int64_t nchunk, int G, int sem, const GatherSegs* segs, cudaStream_t st) {
    if (segs) encode_g2<DT, RM, true>(w, dst, nchunk, G, sem, *segs, st);
    else {
        GatherSegs none;
        none.n = 0;
        encode_g2<DT, RM, false>(w, dst, nchunk, G, sem, none, st);
    }
}

template <int DT>
static void decode_g(const uint8_t* codes, const uint8_t* sz, uint8_t* dq, int64_t nchunk, int G, cudaStream_t st) {
    const int64_t nb = gather_blocks(nchunk);
#define BDQ_L(GG) int2_decode_kernel<DT, GG><<<wave_grid(int2_decode_kernel<DT, GG>, kGatherThreads, nb, 16), kGatherThreads, 0, st>>>(codes, sz, dq, nchunk)
    if (G == 128) BDQ_L(128);
    else if (G == 64) BDQ_L(64);
    else BDQ_L(32);
#undef BDQ_L
}

}  // namespace bdq

using namespace bdq;

static int shard_encode_impl(const void* w, const GatherSegs* segs, void* codes, void* sz, int64_t n_elem, int group, int dtype,
                             int round_mode, int sem, int n_peer, void* const* peer_codes, void* const* peer_sz, bdq_stream_t stream,
                             const char* what) {
    if (round_mode != BDQ_ROUND_HALF_EVEN && round_mode != BDQ_ROUND_HALF_AWAY) { set_error("%s: bad round_mode %d", what, round_mode); return BDQ_ERR_UNSUPPORTED; }
    if (n_peer < 0 || n_peer > kMaxDst - 1) { set_error("%s: n_peer=%d, at most %d", what, n_peer, kMaxDst - 1); return BDQ_ERR_SHAPE; }
    if (n_elem == 0) return BDQ_OK;
    if (!codes || !sz || (n_peer > 0 && (!peer_codes || !peer_sz))) { set_error("null tensor pointer"); return BDQ_ERR_SHAPE; }
    GatherDst dst;
    dst.n = 1 + n_peer;
    dst.codes[0] = (uint8_t*)codes;
    dst.sz[0] = (uint8_t*)sz;
    for (int i = 0; i < n_peer; ++i) {
        dst.codes[1 + i] = (uint8_t*)peer_codes[i];
        dst.sz[1 + i] = (uint8_t*)peer_sz[i];
    }
    for (int d = 0; d < dst.n; ++d)
        if (!dst.codes[d] || !dst.sz[d] || ((uintptr_t)dst.codes[d] % 8) || ((uintptr_t)dst.sz[d] % 4)) {
            set_error("%s: destination %d null or misaligned (codes 8 B, scale/zero 4 B)", what, d);
            return BDQ_ERR_ALIGN;
        }
    cudaStream_t st = (cudaStream_t)stream;
    const int64_t nchunk = n_elem / 32;
    const uint8_t* wp = (const uint8_t*)w;
    if (dtype == BDQ_BF16) {
        if (round_mode == BDQ_ROUND_HALF_AWAY) encode_g<kBF16, BDQ_ROUND_HALF_AWAY>(wp, dst, nchunk, group, sem, segs, st);
        else encode_g<kBF16, BDQ_ROUND_HALF_EVEN>(wp, dst, nchunk, group, sem, segs, st);
    } else {
        if (round_mode == BDQ_ROUND_HALF_AWAY) encode_g<kF16, BDQ_ROUND_HALF_AWAY>(wp, dst, nchunk, group, sem, segs, st);
        else encode_g<kF16, BDQ_ROUND_HALF_EVEN>(wp, dst, nchunk, group, sem, segs, st);
    }
    return check_launch(what);
}

extern "C" int bdq_int2_shard_encode(const void* w, void* codes, void* sz, int64_t n_elem, int group, int dtype, int round_mode,
                                     int sem, int n_peer, void* const* peer_codes, void* const* peer_sz, bdq_stream_t stream) {
    int rc = check_shard(n_elem, group, dtype);
    if (rc) return rc;
    if (n_elem > 0 && !w) { set_error("null tensor pointer"); return BDQ_ERR_SHAPE; }
    if ((uintptr_t)w % 32) { set_error("shard encode: w must be 32-byte aligned"); return BDQ_ERR_ALIGN; }
    return shard_encode_impl(w, nullptr, codes, sz, n_elem, group, dtype, round_mode, sem, n_peer, peer_codes, peer_sz, stream,
                             "bdq_int2_shard_encode");
}

extern "C" int bdq_int2_shard_encode_multi(int nseg, const void* const* w, const int64_t* n_elem, const int64_t* code_off,
                                           const int64_t* sz_off, void* codes, void* sz, int group, int dtype, int round_mode, int sem,
                                           int n_peer, void* const* peer_codes, void* const* peer_sz, bdq_stream_t stream) {
    if (nseg < 1 || nseg > kMaxDst || !w || !n_elem || !code_off || !sz_off) { set_error("shard encode multi: nseg=%d (1..%d), non-null arrays", nseg, kMaxDst); return BDQ_ERR_SHAPE; }
    GatherSegs segs;
    segs.n = 0;
    int64_t total = 0;
    for (int i = 0; i < nseg; ++i) {
        int rc = check_shard(n_elem[i], group, dtype);
        if (rc) return rc;
        if (n_elem[i] == 0) continue;
        if (!w[i]) { set_error("null tensor pointer"); return BDQ_ERR_SHAPE; }
        if (((uintptr_t)w[i] % 32) || (code_off[i] % 8) || (sz_off[i] % 4) || code_off[i] < 0 || sz_off[i] < 0) {
            set_error("shard encode multi: segment %d misaligned (w 32 B, code offset 8 B, scale/zero offset 4 B)", i);
            return BDQ_ERR_ALIGN;
        }
        segs.w[segs.n] = (const uint8_t*)w[i];
        segs.begin[segs.n] = total / 32;
        segs.code_off[segs.n] = code_off[i];
        segs.sz_off[segs.n] = sz_off[i];
        ++segs.n;
        total += n_elem[i];
    }
    if (total == 0) return BDQ_OK;
    return shard_encode_impl(nullptr, &segs, codes, sz, total, group, dtype, round_mode, sem, n_peer, peer_codes, peer_sz, stream,
                             "bdq_int2_shard_encode_multi");
}

extern "C" int bdq_int2_shard_decode(const void* codes, const void* sz, void* dq, int64_t n_elem, int group, int dtype,
                                     bdq_stream_t stream) {
    int rc = check_shard(n_elem, group, dtype);
    if (rc) return rc;
    if (n_elem == 0) return BDQ_OK;
    if (!codes || !sz || !dq) { set_error("null tensor pointer"); return BDQ_ERR_SHAPE; }
    if (((uintptr_t)dq % 32) || ((uintptr_t)codes % 8) || ((uintptr_t)sz % 4)) { set_error("shard decode: misaligned pointer (dq 32 B, codes 8 B, scale/zero 4 B)"); return BDQ_ERR_ALIGN; }
    cudaStream_t st = (cudaStream_t)stream;
    const int64_t nchunk = n_elem / 32;
    if (dtype == BDQ_BF16) decode_g<kBF16>((const uint8_t*)codes, (const uint8_t*)sz, (uint8_t*)dq, nchunk, group, st);
    else decode_g<kF16>((const uint8_t*)codes, (const uint8_t*)sz, (uint8_t*)dq, nchunk, group, st);
    return check_launch("bdq_int2_shard_decode");
}


// ---- gradient push (the scatter half of the fused reduce-scatter + STE backward, bdq_int_ste_bwd_nsrc) ----------------------
// Up to 64 contiguous segments copied in one launch; the destinations are normally peer-mapped (symmetric memory) inbox slots
// on the other GPUs, so the kernel's stores ARE the NVLink transfer (no collective, no staging copy).  32-byte vectors.
namespace bdq {
constexpr int kPushMaxSeg = 64;
struct PushSegs {
    const uint8_t* src[kPushMaxSeg];
    uint8_t* dst[kPushMaxSeg];
    unsigned int nvec[kPushMaxSeg];   // 32-byte vectors
};
__global__ void __launch_bounds__(256) rows_push_kernel(const __grid_constant__ PushSegs sg) {
    const int seg = blockIdx.y;
    const uint8_t* __restrict__ src = sg.src[seg];
    uint8_t* __restrict__ dst = sg.dst[seg];
    const unsigned int n = sg.nvec[seg];
    for (unsigned int i = blockIdx.x * 256u + threadIdx.x; i < n; i += gridDim.x * 256u) stg256(dst + (size_t)i * 32, ldg256(src + (size_t)i * 32));
}
}  // namespace bdq

extern "C" int bdq_rows_push(int nseg, const void* const* src, void* const* dst, const int64_t* nbytes, bdq_stream_t stream) {
    using namespace bdq;
    if (nseg < 0 || (nseg > 0 && (!src || !dst || !nbytes))) { set_error("rows_push: bad arguments"); return BDQ_ERR_SHAPE; }
    cudaStream_t st = (cudaStream_t)stream;
    for (int s0 = 0; s0 < nseg; s0 += kPushMaxSeg) {
        const int n = nseg - s0 < kPushMaxSeg ? nseg - s0 : kPushMaxSeg;
        PushSegs sg;
        memset(&sg, 0, sizeof(sg));
        int64_t maxvec = 0;
        for (int i = 0; i < n; ++i) {
            const int64_t nb = nbytes[s0 + i];
            if (nb < 0 || nb % 32 != 0 || nb / 32 > 0xffffffffll || !src[s0 + i] || !dst[s0 + i] || ((uintptr_t)src[s0 + i] % 32) ||
                ((uintptr_t)dst[s0 + i] % 32)) {
                set_error("rows_push: segment %d must be non-null, 32-byte aligned and a multiple of 32 bytes", s0 + i);
                return BDQ_ERR_ALIGN;
            }
            sg.src[i] = (const uint8_t*)src[s0 + i];
            sg.dst[i] = (uint8_t*)dst[s0 + i];
            sg.nvec[i] = (unsigned int)(nb / 32);
            if (nb / 32 > maxvec) maxvec = nb / 32;
        }
        if (maxvec == 0) continue;
        int64_t bx = (maxvec + 255) / 256;
        // CTAs per SM in total: 2 x 256 threads already reach the NVLink rate (671 vs 698 GB/s at 8, tools/bench_push.py on 2 x B200)
        // and leave the SMs to the HBM-bound kernel this transfer is meant to overlap with (BDQ_PUSH_CTAS_PER_SM overrides)
        int per_sm = 2;
        { const char* e = getenv("BDQ_PUSH_CTAS_PER_SM"); if (e && atoi(e) > 0) per_sm = atoi(e); }
        const int64_t cap = (int64_t)sm_count() * per_sm / n + 1;
        if (bx > cap) bx = cap;
        rows_push_kernel<<<dim3((unsigned)bx, (unsigned)n), 256, 0, st>>>(sg);
        const int rc = check_launch("bdq_rows_push");
        if (rc) return rc;
    }
    return BDQ_OK;
}

namespace bdq {
struct PeerFlags {
    unsigned int* flags[8];   // flags[r] = rank r's flag array (8 x u32), peer-mapped into this process
};
}  // namespace bdq

// ---- cross-GPU barrier on peer-mapped flags --------------------------------------------------------------------------
// One tiny kernel per barrier: lane k stores the epoch into slot [my rank] of rank k's flag array (peer-mapped, NVLink) with
// system-scope release semantics, then spins on slot [k] of its own array until rank k's store of the same epoch arrives.
// Epochs only grow, so the flags are never reset.  Stream order makes the kernel a barrier for the work around it: everything
// this rank launched before it (e.g. the payload / gradient peer stores) has completed when its flags go out, and nothing
// launched after it starts before every peer's flag has arrived.  Replaces torch symmetric memory's barrier on the hot path
// (measured on 2 x B200: ~40 us per barrier there, a handful of microseconds here).
namespace bdq {
__global__ void __launch_bounds__(32) peer_barrier_kernel(const PeerFlags pf, int me, int n, unsigned int epoch) {
    const int k = threadIdx.x;
    if (k < n && k != me) {
        __threadfence_system();
        unsigned int* dst = pf.flags[k] + me;
        asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(dst), "r"(epoch) : "memory");
        const unsigned int* src = pf.flags[me] + k;
        unsigned int v;
        do {
            asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(src) : "memory");
        } while ((int)(v - epoch) < 0);   // wrap-safe: epochs are compared as a signed distance
    }
    __syncwarp();
    __threadfence_system();
}
}  // namespace bdq

extern "C" int bdq_peer_barrier(void* const* flags, int my_rank, int n_ranks, unsigned int epoch, bdq_stream_t stream) {
    using namespace bdq;
    if (!flags || n_ranks < 1 || n_ranks > 8 || my_rank < 0 || my_rank >= n_ranks) {
        set_error("peer_barrier: 1..8 ranks, flags[n_ranks] required");
        return BDQ_ERR_SHAPE;
    }
    PeerFlags pf;
    memset(&pf, 0, sizeof(pf));
    for (int i = 0; i < n_ranks; ++i) {
        if (!flags[i] || ((uintptr_t)flags[i] % 4)) { set_error("peer_barrier: flags[%d] null or misaligned", i); return BDQ_ERR_ALIGN; }
        pf.flags[i] = (unsigned int*)flags[i];
    }
    if (n_ranks == 1) return BDQ_OK;
    peer_barrier_kernel<<<1, 32, 0, (cudaStream_t)stream>>>(pf, my_rank, n_ranks, epoch);
    return check_launch("bdq_peer_barrier");
}
