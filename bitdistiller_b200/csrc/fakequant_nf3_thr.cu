// Threshold-formulated NF3 group fake-quant forward for bf16 / fp16 (sm_100a): BASELINE config 3
// (SteN2F3Quantizer, quantization/quantizer.py:184-242, with the fused apply_clip of clip_utils.py:328-346).
//
// Exactness argument.  For one group (scales sp = |max|, sn = |min| fixed):
//   * q = rT(x / s) with s = sp for x >= 0 and sn for x < 0 is a monotone non-decreasing function of the T value x, and
//     the two reference threshold chains (quantizer.py:224-236) form ONE ascending staircase over the signed q;
//   * the straight-through value (y - q).detach() + q = rT(rT(lvl - q) + q) equals lvl for EVERY T value q in [-1, 1]
//     (checked exhaustively for bf16 and fp16: all levels satisfy Sterbenz' condition on their own interval), so the
//     forward output is  rT(lvl(x) * s): a step function of x with 7 thresholds and 8 per-group output values.
// Each x-space threshold T_k = min{ x : rT(x / s) >= thr_k } is computed per group from the exact product (midpoint below
// thr_k) * s and then verified: with the fp32 quotient computed exactly as the op-by-op path computes it, the candidate
// must satisfy rT(q) >= thr_k and the T value just below it must not (rT(q) >= thr_k is tested as q >= midx, which is
// what round-to-nearest-even means).  A group whose candidate fails the check takes the op-by-op path (nf3_elem), so
// exactness never rests on the guess.  The per-element work is 7 packed compares + 7 selects, both on the ALU pipe,
// which bounds this kernel (ncu: ALU pipe 84 % active at 78 % of the HBM peak before the per-group work was trimmed).
#include <cstring>

#include "fakequant_thr.cuh"
#include "fq_nf3_elem.cuh"
#include "fq_nf3_stair.cuh"

namespace bdq {

constexpr int kNf3ThrThreads = 256;

template <int LPG>
__device__ __forceinline__ uint32_t nf3_group_lane_mask(int lane) {
    return ((LPG == 32) ? 0xffffffffu : ((1u << LPG) - 1u)) << (lane & ~(LPG - 1));
}

// Thresholds of the lane's group (scales sp, sn): every lane finds and verifies R of the 7 (my[r]: k = lig + r * LPG),
// packs them in pairs and the group's lanes exchange them: threshold k ends up in sh[k % LPG][(k / LPG) / 2], half
// (k / LPG) & 1, on every lane of the group.  Returns this lane's verdict (all of the group's lanes must agree).
template <int DT, int LPG>
__device__ __forceinline__ bool nf3_group_thresholds(const Nf3Step (&my)[(7 + LPG - 1) / LPG], uint32_t mask, float sp, float sn,
                                                     int lig, int gbase, uint32_t (&sh)[LPG][((7 + LPG - 1) / LPG + 1) / 2]) {
    using P = P2<DT>;
    constexpr int R = (7 + LPG - 1) / LPG;
    constexpr int NP = (R + 1) / 2;
    const float yp = rcp_approx(sp), yn = rcp_approx(sn);
    // scales in a range where the reciprocal is sane and (fp16) every candidate is a normal T value; a zero scale
    // (NaN group), NaN or inf fails these compares
    constexpr float kLo = (DT == kBF16) ? 1e-30f : 0.001953125f;   // 2^-9
    bool ok = (sp >= kLo) && (sp < 1e30f) && (sn >= kLo) && (sn < 1e30f);
    float tl[2 * NP];
#pragma unroll
    for (int r = 0; r < 2 * NP; ++r) tl[r] = 0.f;
#pragma unroll
    for (int r = 0; r < R; ++r) {
        const Nf3Step e = my[r];
        const bool neg = lig + r * LPG < 3;
        const float s = neg ? sn : sp, y = neg ? yn : yp;
        const uint32_t cb = (__float_as_uint(e.mid * s) + e.add) & mask;
        const float c = __uint_as_float(cb), cprev = __uint_as_float(cb + (uint32_t)e.pstep);
        float q0;
        const bool pc = div_fast(c, s, y, q0) >= e.midx;
        const bool pp = div_fast(cprev, s, y, q0) >= e.midx;
        ok = ok && pc && !pp;
        tl[r] = c;
    }
    uint32_t pk[NP];
#pragma unroll
    for (int j = 0; j < NP; ++j) pk[j] = P::bits(P::pack(tl[2 * j], tl[2 * j + 1]));   // exact: T values
#pragma unroll
    for (int l = 0; l < LPG; ++l)
#pragma unroll
        for (int j = 0; j < NP; ++j) sh[l][j] = (LPG == 1) ? pk[j] : __shfl_sync(0xffffffffu, pk[j], gbase + l);
    return ok;
}

// packed level select: 7 compares + 7 bit-selects per pair; v2[c] = the output value of level c (packed pair)
template <int DT, int LPG>
__device__ __forceinline__ uint32_t nf3_select(uint32_t x, const uint32_t (&sh)[LPG][((7 + LPG - 1) / LPG + 1) / 2],
                                               const uint32_t (&v2)[8]) {
    using P = P2<DT>;
    using h2 = typename P::h2;
    const h2 x2 = P::from_bits(x);
    uint32_t o = v2[0];
#pragma unroll
    for (int k = 0; k < 7; ++k) {
        const h2 tk = P::from_bits(sh[k % LPG][(k / LPG) / 2]);
        const h2 t2 = ((k / LPG) & 1) ? P::dup_hi(tk) : P::dup_lo(tk);
        const uint32_t m = __hge2_mask(x2, t2);
        asm("lop3.b32 %0, %1, %2, %3, 0xCA;" : "=r"(o) : "r"(m), "r"(v2[k + 1]), "r"(o));   // m ? v : o
    }
    return o;
}

template <int DT>
__device__ __noinline__ void nf3_fwd_slow_global(const void* w, void* dq, int64_t e0, float lo, float hi, float mx, float mn,
                                                 const Nf3Tables& t) {
    const Nf3Q p = nf3_group_params(mx, mn);
#pragma unroll 1
    for (int i = 0; i < 32; ++i) {
        const float x = fminf(fmaxf(load_scalar<DT>(w, e0 + i), lo), hi);
        float q, y;
        int code;
        store_scalar<DT>(dq, e0 + i, nf3_elem<DT>(x, p, t, q, y, code));
    }
}

template <int DT, int G, bool CLIP>
__global__ void __launch_bounds__(kNf3ThrThreads) nf3_thr_fwd_kernel(const uint8_t* __restrict__ w, uint8_t* __restrict__ dq,
                                                                     const void* __restrict__ cmin, const void* __restrict__ cmax,
                                                                     int64_t nchunk, const __grid_constant__ Nf3Tables tab,
                                                                     const __grid_constant__ Nf3Stair st) {
    using P = P2<DT>;
    using h2 = typename P::h2;
    constexpr int LPG = G / 32;
    constexpr int R = (7 + LPG - 1) / LPG;   // thresholds found per lane
    constexpr int NP = (R + 1) / 2;          // ... packed two per register
    const int lane = threadIdx.x & 31;
    const int lig = lane & (LPG - 1);
    const int gbase = lane & ~(LPG - 1);
    const int64_t warp = (int64_t)blockIdx.x * (kNf3ThrThreads / 32) + (threadIdx.x >> 5);
    const int64_t nwarps = (int64_t)gridDim.x * (kNf3ThrThreads / 32);
    const int64_t nwc = (nchunk + 31) / 32;
    Nf3Step my[R];   // this lane's thresholds: k = lig + r * LPG (k == 7, one lane per group: the always-true dummy)
#pragma unroll
    for (int r = 0; r < R; ++r) my[r] = st.k[lig + r * LPG];
    for (int64_t wc = warp; wc < nwc; wc += nwarps) {
        const int64_t idx = wc * 32 + lane;
        const bool valid = idx < nchunk;
        U8 raw[2];
        if (valid) {
            raw[0] = ldg256(w + idx * 64);
            raw[1] = ldg256(w + idx * 64 + 32);
        } else {
#pragma unroll
            for (int i = 0; i < 8; ++i) raw[0].r[i] = raw[1].r[i] = 0;
        }
        float clo = -INFINITY, chi = INFINITY;
        if constexpr (CLIP) {
            if (valid) {
                const int64_t g = idx / LPG;
                clo = load_scalar<DT>(cmin, g);
                chi = load_scalar<DT>(cmax, g);
                // The clamp itself is only needed when a bound lies on the wrong side of zero.  With clo <= 0 <= chi the extremes
                // of the clamped group are the clamped extremes (clamp is monotone), an element above chi sits on the top step of
                // the staircase whether it is clamped to chi = max -> q = 1 or not, one below clo on the bottom step: the level
                // selection on the raw elements is the selection on the clamped ones, and 32 packed min / max per lane go away.
                if (!(clo <= 0.f && chi >= 0.f)) {   // uniform per group
                    const h2 l2 = P::dup(clo), u2 = P::dup(chi);
#pragma unroll
                    for (int v = 0; v < 2; ++v)
#pragma unroll
                        for (int i = 0; i < 8; ++i) raw[v].r[i] = P::bits(__hmin2(__hmax2(P::from_bits(raw[v].r[i]), l2), u2));
                }
            }
        }
        float mx, mn, mx1, mn1;
        vec_minmax16<DT>(raw[0], mx, mn);
        vec_minmax16<DT>(raw[1], mx1, mn1);
        mx = group_max<LPG>(fmaxf(mx, mx1));
        mn = group_min<LPG>(fminf(mn, mn1));
        if constexpr (CLIP) {
            mx = fminf(fmaxf(mx, clo), chi);
            mn = fminf(fmaxf(mn, clo), chi);
        }
        const float sp = fabsf(mx), sn = fabsf(mn);
        uint32_t sh[LPG][NP];
        const bool ok = nf3_group_thresholds<DT, LPG>(my, st.mask, sp, sn, lig, gbase, sh);
        const uint32_t okm = __ballot_sync(0xffffffffu, ok);
        const uint32_t gm = nf3_group_lane_mask<LPG>(lane);
        const bool fast = (okm & gm) == gm;
        if (fast) {
            const h2 s2p = P::dup(sp), s2n = P::dup(sn);
            uint32_t v2[8];
#pragma unroll
            for (int c = 0; c < 8; ++c)   // rT(lvl * s): one rounding of an exact product either way
                v2[c] = (c == 3) ? 0u : P::bits(__hmul2(P::from_bits(st.lvl2[c]), c < 3 ? s2n : s2p));
            U8 out[2];
#pragma unroll
            for (int v = 0; v < 2; ++v)
#pragma unroll
                for (int i = 0; i < 8; ++i) out[v].r[i] = nf3_select<DT, LPG>(raw[v].r[i], sh, v2);
            if (valid) {
                stg256(dq + idx * 64, out[0]);
                stg256(dq + idx * 64 + 32, out[1]);
            }
        } else if (valid) {
            nf3_fwd_slow_global<DT>(w, dq, idx * 32, clo, chi, mx, mn, tab);
        }
    }
}

static int64_t nf3_thr_blocks(int64_t nchunk) { return ((nchunk + 31) / 32 + (kNf3ThrThreads / 32) - 1) / (kNf3ThrThreads / 32); }

template <int DT, bool CLIP>
static void nf3_thr_fwd_g(const uint8_t* w, uint8_t* dq, const void* cmin, const void* cmax, int64_t nchunk, int G,
                          const Nf3Tables& tab, cudaStream_t st) {
    const int64_t nb = nf3_thr_blocks(nchunk);
    const Nf3Stair stair = make_stair(tab, DT);
#define BDQ_L(GG) nf3_thr_fwd_kernel<DT, GG, CLIP><<<wave_grid(nf3_thr_fwd_kernel<DT, GG, CLIP>, kNf3ThrThreads, nb, 6), kNf3ThrThreads, 0, st>>>(w, dq, cmin, cmax, nchunk, tab, stair)
    if (G == 128) BDQ_L(128);
    else if (G == 64) BDQ_L(64);
    else BDQ_L(32);
#undef BDQ_L
}

// Returns true if the threshold kernel was launched (16-bit dtype, group 32/64/128, 32-byte aligned, no aux outputs).
bool launch_nf3_thr_fwd(const void* w, void* dq, const void* cmin, const void* cmax, int64_t n, int G, int dtype, int sem,
                        cudaStream_t st) {
    if ((dtype != kBF16 && dtype != kF16) || (G != 32 && G != 64 && G != 128)) return false;
    if (((uintptr_t)w % 32) || ((uintptr_t)dq % 32) || (n % 32)) return false;
    const Nf3Tables tab = make_tables(dtype, sem);
    const uint8_t* wp = (const uint8_t*)w;
    uint8_t* op = (uint8_t*)dq;
    const int64_t nchunk = n / 32;
    if (dtype == kBF16) {
        if (cmin) nf3_thr_fwd_g<kBF16, true>(wp, op, cmin, cmax, nchunk, G, tab, st);
        else nf3_thr_fwd_g<kBF16, false>(wp, op, cmin, cmax, nchunk, G, tab, st);
    } else {
        if (cmin) nf3_thr_fwd_g<kF16, true>(wp, op, cmin, cmax, nchunk, G, tab, st);
        else nf3_thr_fwd_g<kF16, false>(wp, op, cmin, cmax, nchunk, G, tab, st);
    }
    return true;
}

// ------------------------------------------------------------------------------------------------ STE backward
// SteN2F3Quantizer backward: gw = gy, plus at the group's max (min) element(s)
//   sign(max) * sum_{x >= 0} gy * (y - x / sp) / #max      (sign(min) * sum_{x < 0} gy * (y - x / sn) / #min),
// y = the snapped level.  With the forward's x-space thresholds y is one packed select per pair, and the four group
// sums  sum gy*y, sum_{x>=0} gy*y, sum gy*x, sum_{x>=0} gy*x  are dot products along the group: they run on the tensor
// pipe (mma.sync m16n8k16, fp32 accumulate; products of two T values are exact in fp32) instead of 5 fp32 ops per
// element on the FMA/ALU pipes.  Fragment trick: a warp chunk is 8 groups x 128, group i = lanes 4i..4i+3, each lane
// owning 32 consecutive elements.  A rows 0-7 = gy of group i (rows 8-15 = gy masked to x >= 0), B column n = y (or x)
// of group n, both taken from the lane's OWN registers: lane 4i+t supplies A[i][k-slice t] and B[k-slice t][i], so the
// diagonal D[i][i] (D[8+i][i]) is exactly group i's dot product whatever k order the fragment layout implies.  The 56
// off-diagonal products are wasted; the tensor pipe is otherwise idle.  Tie counts (#max, #min) the same way with
// A = [x == max], [x == min] as 1.0 / 0.0 and B = 1.
template <int DT>
__device__ __forceinline__ void mma_16816(float (&d)[4], uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3, uint32_t b0,
                                          uint32_t b1) {
    if constexpr (DT == kBF16)
        asm("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
            : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3]) : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
    else
        asm("mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
            : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3]) : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}

// gy pair (+ tmax where x == mx) (+ tmin where x == mn), evaluated in fp32 and rounded once
template <int DT>
__device__ __forceinline__ uint32_t nf3_bwd_fix(uint32_t g, uint32_t x, uint32_t mx2, uint32_t mn2, float tmax, float tmin) {
    using P = P2<DT>;
    float lo = P::lo(P::from_bits(g)), hi = P::hi(P::from_bits(g));
    if constexpr (DT == kBF16)
        asm("{ .reg .pred p, q, r, s;\n setp.eq.bf16x2 p|q, %4, %5;\n setp.eq.bf16x2 r|s, %4, %6;\n"
            " @p add.f32 %0, %0, %2;\n @q add.f32 %1, %1, %2;\n @r add.f32 %0, %0, %3;\n @s add.f32 %1, %1, %3;\n }"
            : "+f"(lo), "+f"(hi) : "f"(tmax), "f"(tmin), "r"(x), "r"(mx2), "r"(mn2));
    else
        asm("{ .reg .pred p, q, r, s;\n setp.eq.f16x2 p|q, %4, %5;\n setp.eq.f16x2 r|s, %4, %6;\n"
            " @p add.f32 %0, %0, %2;\n @q add.f32 %1, %1, %2;\n @r add.f32 %0, %0, %3;\n @s add.f32 %1, %1, %3;\n }"
            : "+f"(lo), "+f"(hi) : "f"(tmax), "f"(tmin), "r"(x), "r"(mx2), "r"(mn2));
    return P::bits(P::pack(lo, hi));
}

// op-by-op path for one lane's 32 elements (its group = 4 lanes); called by the whole warp
template <int DT>
__device__ __noinline__ void nf3_bwd_slow_global(const void* w, const void* gy, void* gw, int64_t e0, bool valid, float mx,
                                                 float mn, const Nf3Tables& t, int n_el = 32) {
    const Nf3Q p = nf3_group_params(mx, mn);
    float Gp = 0.f, Gn = 0.f;
    int cnt = 0;
    if (valid) {
#pragma unroll 1
        for (int i = 0; i < n_el; ++i) {
            const float x = load_scalar<DT>(w, e0 + i), g = load_scalar<DT>(gy, e0 + i);
            float q, y;
            int code;
            nf3_elem<DT>(x, p, t, q, y, code);
            if (x >= 0.f) Gp += g * (y - x * p.yp);
            else Gn += g * (y - x * p.yn);
            cnt += (x == mx ? 1 : 0) + (x == mn ? 65536 : 0);
        }
    }
    Gp = group_sum<4>(Gp);
    Gn = group_sum<4>(Gn);
    cnt = group_sum_i<4>(cnt);
    const float sgx = mx > 0.f ? 1.f : (mx < 0.f ? -1.f : 0.f);
    const float sgn = mn > 0.f ? 1.f : (mn < 0.f ? -1.f : 0.f);
    const float tmax = sgx * Gp / (float)(cnt & 0xffff), tmin = sgn * Gn / (float)(cnt >> 16);
    if (valid) {
#pragma unroll 1
        for (int i = 0; i < n_el; ++i) {
            const float x = load_scalar<DT>(w, e0 + i);
            float r = load_scalar<DT>(gy, e0 + i);
            if (x == mx) r += tmax;
            if (x == mn) r += tmin;
            if (p.nan_group) r = __int_as_float(0x7fc00000);
            store_scalar<DT>(gw, e0 + i, r);
        }
    }
}

// MS: the incoming gradient is gy + ex.p[0] + ... + ex.p[ex.n - 1] (fused reduce-scatter for FSDP2 shards, see fakequant_thr.cu:
// accumulated in that order with one rounding to T per addition; the op-by-op fallback reads the sum back from its own slot of gw)
struct Nf3GySrcs {
    const uint8_t* p[7];
    int n;
};

// EPL = elements per lane (32 / 16 / 8): a group is always one quad of lanes, so G = 4 * EPL = 128 / 64 / 32 share the fragment
// trick above unchanged (G = 64 is the group size of every train script of the reference, scripts/train/train.sh:52).
template <int EPL>
__device__ __forceinline__ void nf3_lane_load(const uint8_t* p, int64_t idx, uint32_t* r) {
    if constexpr (EPL == 32) {
        const U8 a = ldg256(p + idx * 64), b = ldg256(p + idx * 64 + 32);
#pragma unroll
        for (int i = 0; i < 8; ++i) { r[i] = a.r[i]; r[8 + i] = b.r[i]; }
    } else if constexpr (EPL == 16) {
        const U8 a = ldg256(p + idx * 32);
#pragma unroll
        for (int i = 0; i < 8; ++i) r[i] = a.r[i];
    } else {
        asm volatile("ld.global.nc.L1::no_allocate.v4.b32 {%0,%1,%2,%3}, [%4];" : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "l"(p + idx * 16));
    }
}
template <int EPL>
__device__ __forceinline__ void nf3_lane_store(uint8_t* p, int64_t idx, const uint32_t* r) {
    if constexpr (EPL == 32) {
        U8 a, b;
#pragma unroll
        for (int i = 0; i < 8; ++i) { a.r[i] = r[i]; b.r[i] = r[8 + i]; }
        stg256(p + idx * 64, a);
        stg256(p + idx * 64 + 32, b);
    } else if constexpr (EPL == 16) {
        U8 a;
#pragma unroll
        for (int i = 0; i < 8; ++i) a.r[i] = r[i];
        stg256(p + idx * 32, a);
    } else {
        asm volatile("st.global.v4.b32 [%0], {%1,%2,%3,%4};" ::"l"(p + idx * 16), "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]) : "memory");
    }
}

template <int DT, bool MS, int EPL>
__global__ void __launch_bounds__(kNf3ThrThreads) nf3_thr_bwd_kernel(const uint8_t* __restrict__ w, const uint8_t* __restrict__ gy,
                                                                     uint8_t* gw, int64_t nchunk,
                                                                     const __grid_constant__ Nf3Tables tab,
                                                                     const __grid_constant__ Nf3Stair st, const Nf3GySrcs ex) {
    using P = P2<DT>;
    using h2 = typename P::h2;
    constexpr int LPG = 4, R = 2, NP = 1, NR = EPL / 2;
    const int lane = threadIdx.x & 31;
    const int lig = lane & 3, gbase = lane & ~3, grp = lane >> 2;
    const int64_t warp = (int64_t)blockIdx.x * (kNf3ThrThreads / 32) + (threadIdx.x >> 5);
    const int64_t nwarps = (int64_t)gridDim.x * (kNf3ThrThreads / 32);
    const int64_t nwc = (nchunk + 31) / 32;
    const uint32_t ones = (DT == kBF16) ? 0x3f803f80u : 0x3c003c00u;
    Nf3Step my[R];
#pragma unroll
    for (int r = 0; r < R; ++r) my[r] = st.k[lig + r * LPG];
    for (int64_t wc = warp; wc < nwc; wc += nwarps) {
        const int64_t idx = wc * 32 + lane;
        const bool valid = idx < nchunk;
        uint32_t rx[NR], rg[NR];
        if (valid) {
            nf3_lane_load<EPL>(w, idx, rx);
            nf3_lane_load<EPL>(gy, idx, rg);
        } else {
#pragma unroll
            for (int i = 0; i < NR; ++i) rx[i] = rg[i] = 0;
        }
        if constexpr (MS) {
            if (valid) {
#pragma unroll 1
                for (int sidx = 0; sidx < ex.n; ++sidx) {
                    uint32_t a[NR];
                    nf3_lane_load<EPL>(ex.p[sidx], idx, a);
#pragma unroll
                    for (int i = 0; i < NR; ++i) rg[i] = P::bits(__hadd2(P::from_bits(rg[i]), P::from_bits(a[i])));
                }
            }
        }
        h2 a2 = P::from_bits(rx[0]), b2 = a2;
#pragma unroll
        for (int i = 1; i < NR; ++i) {
            a2 = __hmax2(a2, P::from_bits(rx[i]));
            b2 = __hmin2(b2, P::from_bits(rx[i]));
        }
        const float mx = group_max<LPG>(fmaxf(P::lo(a2), P::hi(a2)));
        const float mn = group_min<LPG>(fminf(P::lo(b2), P::hi(b2)));
        const float sp = fabsf(mx), sn = fabsf(mn);
        uint32_t sh[LPG][NP];
        bool ok = nf3_group_thresholds<DT, LPG>(my, st.mask, sp, sn, lig, gbase, sh);
        ok = ok && (mx > 0.f) && (mn < 0.f);
        if (__all_sync(0xffffffffu, ok || !valid)) {
            uint32_t lv[8];
#pragma unroll
            for (int c = 0; c < 8; ++c) lv[c] = st.lvl2[c];
            const uint32_t mx2 = P::bits(P::dup(mx)), mn2 = P::bits(P::dup(mn));
            const h2 zero2 = P::from_bits(0u);
            float dy[4] = {0.f, 0.f, 0.f, 0.f}, dx[4] = {0.f, 0.f, 0.f, 0.f}, dc[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
            for (int i = 0; i < NR; i += 2) {
                const uint32_t x0 = rx[i], x1 = rx[i + 1], g0 = rg[i], g1 = rg[i + 1];
                const uint32_t y0 = nf3_select<DT, LPG>(x0, sh, lv), y1 = nf3_select<DT, LPG>(x1, sh, lv);
                const uint32_t p0 = g0 & __hge2_mask(P::from_bits(x0), zero2), p1 = g1 & __hge2_mask(P::from_bits(x1), zero2);
                mma_16816<DT>(dy, g0, p0, g1, p1, y0, y1);
                mma_16816<DT>(dx, g0, p0, g1, p1, x0, x1);
                const uint32_t e0 = P::bits(__heq2(P::from_bits(x0), P::from_bits(mx2)));
                const uint32_t e1 = P::bits(__heq2(P::from_bits(x1), P::from_bits(mx2)));
                const uint32_t f0 = P::bits(__heq2(P::from_bits(x0), P::from_bits(mn2)));
                const uint32_t f1 = P::bits(__heq2(P::from_bits(x1), P::from_bits(mn2)));
                mma_16816<DT>(dc, e0, f0, e1, f1, ones, ones);
            }
            // D[i][i] sits in lane 4i + (i >> 1), register i & 1 (rows 0-7) / 2 + (i & 1) (rows 8-15)
            const int src = gbase + (grp >> 1);
            const bool odd = grp & 1;
            const float sy_all = __shfl_sync(0xffffffffu, odd ? dy[1] : dy[0], src);
            const float sy_pos = __shfl_sync(0xffffffffu, odd ? dy[3] : dy[2], src);
            const float sx_all = __shfl_sync(0xffffffffu, odd ? dx[1] : dx[0], src);
            const float sx_pos = __shfl_sync(0xffffffffu, odd ? dx[3] : dx[2], src);
            // every column of the count product equals the group's count: no exchange
            const float tmax = __fdividef(sy_pos - __fdividef(sx_pos, sp), dc[0]);
            const float tmin = -__fdividef((sy_all - sy_pos) - __fdividef(sx_all - sx_pos, sn), dc[2]);
            uint32_t out[NR];
#pragma unroll
            for (int i = 0; i < NR; ++i) out[i] = nf3_bwd_fix<DT>(rg[i], rx[i], mx2, mn2, tmax, tmin);
            if (valid) nf3_lane_store<EPL>(gw, idx, out);
        } else {
            const uint8_t* gsl = gy;
            if constexpr (MS) {   // park the summed gradient of this lane's elements in its own slot of gw (read and rewritten by this thread only)
                gsl = gw;
                if (valid) nf3_lane_store<EPL>(gw, idx, rg);
            }
            nf3_bwd_slow_global<DT>(w, gsl, gw, idx * EPL, valid, mx, mn, tab, EPL);
        }
    }
}

bool launch_nf3_thr_bwd_ms(const void* w, const void* gy, void* gw, int64_t n, int G, int dtype, int sem, cudaStream_t st, int n_extra,
                           const void* const* extra) {
    if ((dtype != kBF16 && dtype != kF16) || (G != 128 && G != 64 && G != 32)) return false;
    if (((uintptr_t)w % 32) || ((uintptr_t)gy % 32) || ((uintptr_t)gw % 32) || (n % G)) return false;
    if (n_extra < 0 || n_extra > 7) return false;
    Nf3GySrcs ex;
    ex.n = n_extra;
    for (int i = 0; i < 7; ++i) {
        ex.p[i] = i < n_extra ? (const uint8_t*)extra[i] : nullptr;
        if (i < n_extra && (!extra[i] || ((uintptr_t)extra[i] % 32))) return false;
    }
    const Nf3Tables tab = make_tables(dtype, sem);
    const int64_t nchunk = n / (G / 4);
    const int64_t nb = nf3_thr_blocks(nchunk);
#define BDQ_NB(DTT, MSS, EPLL)                                                                                                               \
    nf3_thr_bwd_kernel<DTT, MSS, EPLL><<<wave_grid(nf3_thr_bwd_kernel<DTT, MSS, EPLL>, kNf3ThrThreads, nb, 6), kNf3ThrThreads, 0, st>>>(       \
        (const uint8_t*)w, (const uint8_t*)gy, (uint8_t*)gw, nchunk, tab, make_stair(tab, DTT), ex)
#define BDQ_NB_G(DTT, MSS)                       \
    do {                                         \
        if (G == 128) BDQ_NB(DTT, MSS, 32);      \
        else if (G == 64) BDQ_NB(DTT, MSS, 16);  \
        else BDQ_NB(DTT, MSS, 8);                \
    } while (0)
    if (dtype == kBF16) { if (n_extra) BDQ_NB_G(kBF16, true); else BDQ_NB_G(kBF16, false); }
    else { if (n_extra) BDQ_NB_G(kF16, true); else BDQ_NB_G(kF16, false); }
#undef BDQ_NB_G
#undef BDQ_NB
    return true;
}

bool launch_nf3_thr_bwd(const void* w, const void* gy, void* gw, int64_t n, int G, int dtype, int sem, cudaStream_t st) {
    return launch_nf3_thr_bwd_ms(w, gy, gw, n, G, dtype, sem, st, 0, nullptr);
}

}  // namespace bdq

// ================================================================================================ NF3 shard payload
// Quantise-before-all-gather for NF3 (QLinear's class default quant_type is "ste-n2f3", quantization/qlinear.py:51): the
// owner of a dim-0 shard quantises it once and ships 3-bit level indices + the two scales of every group -- 0.40625 B/weight
// at group 128 instead of 2 -- and every rank rebuilds  dq = rT(lvl[e] * s)  (s = sp for e >= 3, sn for e < 3), the forward's
// own last operation, so the result is bit-equal to SteN2F3Quantizer.forward on the gathered weight.
// Wire format (this build's own): per 32 consecutive elements three uint32 words (lo0, lo1, hi) = the packed-NF3 split of
// nf3_packed.cu (bits [2j, 2j+2) of lo(j / 16) = low two bits of the staircase index e of element j, bit j of hi = its high
// bit); per group one (sp, sn) pair of T values.  A group whose max or min is exactly 0 (the reference's 0/0: the whole group
// is NaN) ships NaN scales.
namespace bdq {

constexpr int kNf3MaxDst = 8;
struct Nf3Dst {
    uint8_t* codes[kNf3MaxDst];
    uint8_t* sc[kNf3MaxDst];
    int n;
};

__device__ __forceinline__ uint32_t nf3_code_to_stair(int c) { return c == 0 ? 3u : (c <= 4 ? (uint32_t)c + 3u : 7u - (uint32_t)c); }

// op-by-op level indices of one lane's 32 elements (groups whose thresholds could not be verified)
template <int DT>
__device__ __noinline__ uint3 nf3_encode_slow_global(const void* w, int64_t e0, float mx, float mn, const Nf3Tables& t) {
    const Nf3Q p = nf3_group_params(mx, mn);
    uint32_t lo[2] = {0u, 0u}, hi = 0u;
#pragma unroll 1
    for (int j = 0; j < 32; ++j) {
        float q, y;
        int code;
        nf3_elem<DT>(load_scalar<DT>(w, e0 + j), p, t, q, y, code);
        const uint32_t e = code > 7 ? 3u : nf3_code_to_stair(code);
        lo[j >> 4] |= (e & 3u) << (2 * (j & 15));
        hi |= (e >> 2) << j;
    }
    return make_uint3(lo[0], lo[1], hi);
}

template <int DT, int G>
__global__ void __launch_bounds__(kNf3ThrThreads) nf3_shard_encode_kernel(const uint8_t* __restrict__ w, Nf3Dst dst, int64_t nchunk,
                                                                          const __grid_constant__ Nf3Tables tab,
                                                                          const __grid_constant__ Nf3Stair st) {
    using P = P2<DT>;
    constexpr int LPG = G / 32;
    constexpr int R = (7 + LPG - 1) / LPG;
    constexpr int NP = (R + 1) / 2;
    const int lane = threadIdx.x & 31;
    const int lig = lane & (LPG - 1);
    const int gbase = lane & ~(LPG - 1);
    const int64_t warp = (int64_t)blockIdx.x * (kNf3ThrThreads / 32) + (threadIdx.x >> 5);
    const int64_t nwarps = (int64_t)gridDim.x * (kNf3ThrThreads / 32);
    const int64_t nwc = (nchunk + 31) / 32;
    Nf3Step my[R];
#pragma unroll
    for (int r = 0; r < R; ++r) my[r] = st.k[lig + r * LPG];
    // "value" of level c for the select = its index, laid out so that one shift + or puts both elements of a pair side by
    // side: low half (c & 3) | (c >> 2) << 8, high half (c & 3) | (c >> 2) << 7
    uint32_t v2e[8];
#pragma unroll
    for (int c = 0; c < 8; ++c) v2e[c] = ((uint32_t)(c & 3) | ((uint32_t)(c >> 2) << 8)) | (((uint32_t)(c & 3) | ((uint32_t)(c >> 2) << 7)) << 16);
    for (int64_t wc = warp; wc < nwc; wc += nwarps) {
        const int64_t idx = wc * 32 + lane;
        const bool valid = idx < nchunk;
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
        const float sp = fabsf(mx), sn = fabsf(mn);
        uint32_t sh[LPG][NP];
        const bool ok = nf3_group_thresholds<DT, LPG>(my, st.mask, sp, sn, lig, gbase, sh);
        const uint32_t okm = __ballot_sync(0xffffffffu, ok);
        const uint32_t gm = nf3_group_lane_mask<LPG>(lane);
        uint3 cw = make_uint3(0u, 0u, 0u);
        if ((okm & gm) == gm) {
            uint32_t lo[2], hi = 0u;
#pragma unroll
            for (int v = 0; v < 2; ++v) {
                uint32_t la = 0u, ha = 0u;
#pragma unroll
                for (int i = 7; i >= 0; --i) {
                    const uint32_t r = nf3_select<DT, LPG>(raw[v].r[i], sh, v2e);
                    const uint32_t u = r | (r >> 14);          // bits 0-3: low codes of the pair, bits 8-9: their high bits
                    la = la * 16u + (u & 0xFu);
                    ha = ha * 4u + ((u >> 8) & 3u);
                }
                lo[v] = la;
                hi |= ha << (16 * v);
            }
            cw = make_uint3(lo[0], lo[1], hi);
        } else if (valid) {
            cw = nf3_encode_slow_global<DT>(w, idx * 32, mx, mn, tab);
        }
        if (valid) {
            const bool nan_group = (sp == 0.f) || (sn == 0.f) || (sp != sp) || (sn != sn);
            const float qnan = __int_as_float(0x7fc00000);
            const uint32_t scp = P::bits(P::pack(nan_group ? qnan : sp, nan_group ? qnan : sn));
            const int64_t g = idx / LPG;
#pragma unroll 1
            for (int d = 0; d < dst.n; ++d) {
                uint32_t* cp = reinterpret_cast<uint32_t*>(dst.codes[d] + idx * 12);
                cp[0] = cw.x;
                cp[1] = cw.y;
                cp[2] = cw.z;
                if (lig == 0) *reinterpret_cast<uint32_t*>(dst.sc[d] + g * 4) = scp;
            }
        }
    }
}

template <int DT, int G>
__global__ void __launch_bounds__(kNf3ThrThreads) nf3_shard_decode_kernel(const uint8_t* __restrict__ codes, const uint8_t* __restrict__ sc,
                                                                          uint8_t* __restrict__ dq, int64_t nchunk,
                                                                          const __grid_constant__ Nf3Stair st) {
    using P = P2<DT>;
    using h2 = typename P::h2;
    constexpr int LPG = G / 32;
    const int lane = threadIdx.x & 31;
    const int64_t warp = (int64_t)blockIdx.x * (kNf3ThrThreads / 32) + (threadIdx.x >> 5);
    const int64_t nwarps = (int64_t)gridDim.x * (kNf3ThrThreads / 32);
    const int64_t nwc = (nchunk + 31) / 32;
    for (int64_t wc = warp; wc < nwc; wc += nwarps) {
        const int64_t idx = wc * 32 + lane;
        if (idx >= nchunk) continue;
        const uint32_t* cp = reinterpret_cast<const uint32_t*>(codes + idx * 12);
        const uint32_t lo[2] = {__ldg(cp), __ldg(cp + 1)}, hi = __ldg(cp + 2);
        const h2 scp = P::from_bits(__ldg(reinterpret_cast<const uint32_t*>(sc + (idx / LPG) * 4)));
        const h2 s2p = P::dup_lo(scp), s2n = P::dup_hi(scp);
        // the group's 8 output values rT(lvl * s) as a 16-byte table (v0,v1) (v2,v3) (v4,v5) (v6,v7); level 3 is exactly 0
        // unless the group is NaN (0 * NaN)
        uint32_t tb[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const h2 a = __hmul2(P::from_bits(st.lvl2[2 * j]), (2 * j) < 3 ? s2n : s2p);
            const h2 b = __hmul2(P::from_bits(st.lvl2[2 * j + 1]), (2 * j + 1) < 3 ? s2n : s2p);
            tb[j] = (P::bits(a) & 0xffffu) | (P::bits(b) & 0xffff0000u);
        }
        U8 out[2];
#pragma unroll
        for (int v = 0; v < 2; ++v)
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                const uint32_t t4 = (lo[v] >> (4 * i)) & 0xFu;                   // low codes of elements 2i, 2i+1
                const uint32_t sel = 0x1010u + (t4 & 3u) * 0x22u + (t4 >> 2) * 0x2200u;   // byte selectors of the two halfwords
                const uint32_t x = __byte_perm(tb[0], tb[1], sel), y = __byte_perm(tb[2], tb[3], sel);
                const uint32_t hb = (hi >> (16 * v + 2 * i)) & 3u;
                const uint32_t m = (hb & 1u) * 0xffffu + (hb >> 1) * 0xffff0000u;
                out[v].r[i] = (x & ~m) | (y & m);
            }
        stg256(dq + idx * 64, out[0]);
        stg256(dq + idx * 64 + 32, out[1]);
    }
}

static int64_t nf3_shard_blocks(int64_t nchunk) { return ((nchunk + 31) / 32 + (kNf3ThrThreads / 32) - 1) / (kNf3ThrThreads / 32); }

}  // namespace bdq

using namespace bdq;

static int nf3_shard_check(int64_t n_elem, int group, int dtype) {
    if (dtype != BDQ_BF16 && dtype != BDQ_F16) { set_error("nf3 shard: dtype must be bf16 or fp16"); return BDQ_ERR_DTYPE; }
    if (group != 32 && group != 64 && group != 128) { set_error("nf3 shard: group must be 32, 64 or 128"); return BDQ_ERR_UNSUPPORTED; }
    if (n_elem < 0 || n_elem % group != 0) { set_error("nf3 shard: n_elem=%lld is not a multiple of group=%d", (long long)n_elem, group); return BDQ_ERR_SHAPE; }
    return BDQ_OK;
}

extern "C" int bdq_nf3_shard_encode(const void* w, void* codes, void* scales, int64_t n_elem, int group, int dtype, int sem, int n_peer,
                                    void* const* peer_codes, void* const* peer_scales, bdq_stream_t stream) {
    int rc = nf3_shard_check(n_elem, group, dtype);
    if (rc) return rc;
    if (n_peer < 0 || n_peer > kNf3MaxDst - 1) { set_error("nf3 shard encode: n_peer=%d, at most %d", n_peer, kNf3MaxDst - 1); return BDQ_ERR_SHAPE; }
    if (n_elem == 0) return BDQ_OK;
    if (!w || !codes || !scales || (n_peer > 0 && (!peer_codes || !peer_scales))) { set_error("null tensor pointer"); return BDQ_ERR_SHAPE; }
    if ((uintptr_t)w % 32) { set_error("nf3 shard encode: w must be 32-byte aligned"); return BDQ_ERR_ALIGN; }
    Nf3Dst dst;
    dst.n = 1 + n_peer;
    dst.codes[0] = (uint8_t*)codes;
    dst.sc[0] = (uint8_t*)scales;
    for (int i = 0; i < n_peer; ++i) { dst.codes[1 + i] = (uint8_t*)peer_codes[i]; dst.sc[1 + i] = (uint8_t*)peer_scales[i]; }
    for (int d = 0; d < dst.n; ++d)
        if (!dst.codes[d] || !dst.sc[d] || ((uintptr_t)dst.codes[d] % 4) || ((uintptr_t)dst.sc[d] % 4)) {
            set_error("nf3 shard encode: destination %d null or misaligned (4 B)", d);
            return BDQ_ERR_ALIGN;
        }
    const Nf3Tables tab = make_tables(dtype, sem);
    const Nf3Stair stair = make_stair(tab, dtype);
    const int64_t nchunk = n_elem / 32, nb = nf3_shard_blocks(nchunk);
    cudaStream_t st = (cudaStream_t)stream;
    const uint8_t* wp = (const uint8_t*)w;
#define BDQ_L(DT, GG) nf3_shard_encode_kernel<DT, GG><<<wave_grid(nf3_shard_encode_kernel<DT, GG>, kNf3ThrThreads, nb, 6), kNf3ThrThreads, 0, st>>>(wp, dst, nchunk, tab, stair)
    if (dtype == BDQ_BF16) { if (group == 128) BDQ_L(kBF16, 128); else if (group == 64) BDQ_L(kBF16, 64); else BDQ_L(kBF16, 32); }
    else { if (group == 128) BDQ_L(kF16, 128); else if (group == 64) BDQ_L(kF16, 64); else BDQ_L(kF16, 32); }
#undef BDQ_L
    return check_launch("bdq_nf3_shard_encode");
}

extern "C" int bdq_nf3_shard_decode(const void* codes, const void* scales, void* dq, int64_t n_elem, int group, int dtype,
                                    bdq_stream_t stream) {
    int rc = nf3_shard_check(n_elem, group, dtype);
    if (rc) return rc;
    if (n_elem == 0) return BDQ_OK;
    if (!codes || !scales || !dq) { set_error("null tensor pointer"); return BDQ_ERR_SHAPE; }
    if (((uintptr_t)dq % 32) || ((uintptr_t)codes % 4) || ((uintptr_t)scales % 4)) { set_error("nf3 shard decode: misaligned pointer (dq 32 B, payload 4 B)"); return BDQ_ERR_ALIGN; }
    const Nf3Stair stair = make_stair(make_tables(dtype, 0), dtype);
    const int64_t nchunk = n_elem / 32, nb = nf3_shard_blocks(nchunk);
    cudaStream_t st = (cudaStream_t)stream;
#define BDQ_L(DT, GG) nf3_shard_decode_kernel<DT, GG><<<wave_grid(nf3_shard_decode_kernel<DT, GG>, kNf3ThrThreads, nb, 16), kNf3ThrThreads, 0, st>>>((const uint8_t*)codes, (const uint8_t*)scales, (uint8_t*)dq, nchunk, stair)
    if (dtype == BDQ_BF16) { if (group == 128) BDQ_L(kBF16, 128); else if (group == 64) BDQ_L(kBF16, 64); else BDQ_L(kBF16, 32); }
    else { if (group == 128) BDQ_L(kF16, 128); else if (group == 64) BDQ_L(kF16, 64); else BDQ_L(kF16, 32); }
#undef BDQ_L
    return check_launch("bdq_nf3_shard_decode");
}
