// Clip-ratio grid search (sm_100a).  Replaces auto_2clip_layer, quantization/autoclip.py:13-83.
//
// A CTA owns one quantisation group g (K = G input channels) and a tile of 64 output channels.  The calibration
// activations X[:, g] (up to 512 tokens x G, 16-bit) are staged once in shared memory and reused by every channel of
// the tile.  For each (channel, group) pair all nc*nc candidate (max, min) clamps are generated in registers with the
// exact T roundings of the reference (clamp -> pseudo_quantize_tensor / n2f3) and written as rows of an A tile in
// shared memory; row nc*nc holds the unquantised weight itself.  The contraction  out[c, t] = sum_k A[c, k] X[t, k]
// runs on the tensor cores (mma.sync m16n8k16, fp32 accumulate, ldmatrix-fed, padded pitches are bank-conflict free),
// the epilogue forms (out[c, t] - out[w, t])^2, reduces over tokens, and the first smallest candidate wins.
// fp32 weights/activations are contracted as bf16 hi/lo splits (q_hi x_hi + q_hi x_lo + q_lo x_hi, K tripled).
// The reference materialises a [oc_batch, 512, ci] broadcast product per candidate (1 GiB) -- nothing of that size
// exists here: HBM traffic is one read of W and X.
#include "bdq_common.cuh"
#include "fq_int_elem.cuh"
#include "fq_nf3_elem.cuh"
#include "fakequant_thr.cuh"

namespace bdq {

constexpr int kAcMaxWarps = 16;
constexpr int kAcRows = 112;   // candidate rows (7 MMA row tiles): nc*nc candidates + the weight row
constexpr int kAcOT = 64;      // output channels per CTA

template <int DT>
struct AcCfg {
    static constexpr int KMUL = (DT == kF32) ? 3 : 1;      // K expansion of the hi/lo split
    static constexpr int TCH = (DT == kF32) ? 128 : 512;   // tokens staged per pass
    static constexpr bool F16 = (DT == kF16);              // MMA operand type: f16, else bf16
    // 16 warps for the 16-bit dtypes: the kernel holds one CTA per SM (the staged X tile fills shared memory) and its
    // phases are separated by barriers, so the warp count is what hides latency; 32 tokens per warp also halves the
    // accumulator registers (128 registers per thread at 512 threads = the whole register file)
    static constexpr int WARPS = (DT == kF32) ? 8 : 16;
    static constexpr int THREADS = WARPS * 32;
};

__device__ __forceinline__ void ldsm_x4(uint32_t* r, const void* p) {
    asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"((uint32_t)__cvta_generic_to_shared(p)));
}
template <bool F16>
__device__ __forceinline__ void mma_ac(float* d, const uint32_t* a, uint32_t b0, uint32_t b1) {
    if constexpr (F16)
        asm("mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
            : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3]) : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
    else
        asm("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
            : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3]) : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}

// 16-bit operand encoding of a value (exact for T = bf16 / f16 values)
template <bool F16>
__device__ __forceinline__ unsigned short enc16(float v) {
    if constexpr (F16) return __half_as_ushort(__float2half_rn(v));
    else return __bfloat16_as_ushort(__float2bfloat16_rn(v));
}
__device__ __forceinline__ float bf16_hi(float v) { return __bfloat162float(__float2bfloat16_rn(v)); }

struct AcSmem {
    int xs, as, errw, errsum, cand, wf, wred, total, pitch;
};
__host__ __device__ inline AcSmem ac_layout(int G, int kmul, int tch) {
    AcSmem L;
    L.pitch = G * kmul * 2 + 16;              // bytes; (pitch / 4) % 32 == 4 -> ldmatrix rows hit distinct banks
    L.xs = 0;
    L.as = L.xs + tch * L.pitch;
    L.errw = L.as + kAcRows * L.pitch;
    L.errsum = L.errw + kAcMaxWarps * kAcRows * 4;
    L.cand = L.errsum + kAcOT * kAcRows * 4;
    L.wf = L.cand + kAcRows * 12 * 4;
    L.wred = L.wf + 128 * 4;
    L.total = L.wred + 64;
    return L;
}

struct AcArgs {
    const void* w;
    const void* x;
    void* best_max;
    void* best_min;
    int32_t* best_idx;
    float* best_err;
    int64_t co, ci, n_tok, xs;
    int G, bits, quant_type, n_grid, nc, sem;
    Nf3Tables tb;
};

template <int DT>
__global__ void __launch_bounds__(AcCfg<DT>::THREADS, 1) autoclip_kernel(const AcArgs a) {
    using C = AcCfg<DT>;
    constexpr bool F16 = C::F16;
    constexpr int kAcThreads = C::THREADS, kAcWarps = C::WARPS;
    constexpr int NTW = C::TCH / (8 * kAcWarps);   // 8-token MMA column tiles per warp
    extern __shared__ __align__(128) uint8_t smem[];
    const int G = a.G, KX = G * C::KMUL;
    const AcSmem L = ac_layout(G, C::KMUL, C::TCH);
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int gq = lane >> 2, tq = lane & 3;
    const int g = blockIdx.x;                         // quantisation group
    const int64_t o0 = (int64_t)blockIdx.y * kAcOT;   // first output channel of the tile
    const int no = (int)min((int64_t)kAcOT, a.co - o0);
    const int64_t ng = a.ci / G;
    const int nc = a.nc, nc2 = nc * nc;               // row nc2 = the weight itself
    const float qmax = (float)((1 << a.bits) - 1);
    uint8_t* Xs = smem + L.xs;
    uint8_t* As = smem + L.as;
    float* errw = reinterpret_cast<float*>(smem + L.errw);
    float* errsum = reinterpret_cast<float*>(smem + L.errsum);
    float* cand = reinterpret_cast<float*>(smem + L.cand);   // [row][12]: int: s, y, z, t1, lo, hi, -, pass, T1, T2, T3, thr_ok | nf3: sp, sn, yp, yn, lo, hi, nan, pass
    float* wf = reinterpret_cast<float*>(smem + L.wf);
    float* wred = reinterpret_cast<float*>(smem + L.wred);

    for (int i = tid; i < kAcOT * kAcRows; i += kAcThreads) errsum[i] = 0.f;
    for (int i = tid; i < (kAcRows * L.pitch) / 4; i += kAcThreads) reinterpret_cast<uint32_t*>(As)[i] = 0u;

    for (int64_t t0 = 0; t0 < a.n_tok; t0 += C::TCH) {
        __syncthreads();
        // ---- stage X[t0 : t0 + TCH, g*G : (g+1)*G] (zero rows past n_tok)
        for (int i = tid; i < C::TCH * (G / 2); i += kAcThreads) {
            const int t = i / (G / 2), j = i - t * (G / 2);
            float v0 = 0.f, v1 = 0.f;
            if (t0 + t < a.n_tok) {
                const int64_t e = (t0 + t) * a.xs + (int64_t)g * G + 2 * j;
                v0 = load_scalar<DT>(a.x, e);
                v1 = load_scalar<DT>(a.x, e + 1);
            }
            uint8_t* row = Xs + (size_t)t * L.pitch;
            if constexpr (DT == kF32) {
                const float h0 = bf16_hi(v0), h1 = bf16_hi(v1);
                const uint32_t hi = enc16<false>(h0) | ((uint32_t)enc16<false>(h1) << 16);
                const uint32_t lo = enc16<false>(v0 - h0) | ((uint32_t)enc16<false>(v1 - h1) << 16);
                reinterpret_cast<uint32_t*>(row)[j] = hi;                  // pairs with q_hi
                reinterpret_cast<uint32_t*>(row + G * 2)[j] = lo;          // pairs with q_hi
                reinterpret_cast<uint32_t*>(row + G * 4)[j] = hi;          // pairs with q_lo
            } else {
                reinterpret_cast<uint32_t*>(row)[j] = enc16<F16>(v0) | ((uint32_t)enc16<F16>(v1) << 16);
            }
        }

        // 16-bit dtypes: this warp's X fragments (its 32 tokens x K <= 128) do not depend on the channel: load them once per
        // staged chunk and keep them in registers (64), so the contraction issues one ldmatrix per 4 MMAs instead of 3
        constexpr int KS = (DT == kF32) ? 1 : 8, NP = (DT == kF32) ? 1 : NTW / 2;
        uint32_t bfr[KS][NP][4];
        if constexpr (DT != kF32) {
            __syncthreads();
            const uint8_t* brow = Xs + (size_t)(warp * 8 * NTW + (lane & 7) + ((lane >> 4) << 3)) * L.pitch + ((lane >> 3) & 1) * 16;
#pragma unroll
            for (int ks = 0; ks < KS; ++ks)
#pragma unroll
                for (int np = 0; np < NP; ++np) {
                    if (ks * 16 < KX) ldsm_x4(bfr[ks][np], brow + (size_t)np * 16 * L.pitch + ks * 32);
                    else bfr[ks][np][0] = bfr[ks][np][1] = bfr[ks][np][2] = bfr[ks][np][3] = 0u;
                }
        }
        for (int ol = 0; ol < no; ++ol) {
            const int64_t o = o0 + ol;
            // ---- (a) the weight group and its extremes
            if (tid < G) {
                const float v = load_scalar<DT>(a.w, o * a.ci + (int64_t)g * G + tid);
                wf[tid] = v;
                const float mxw = group_max<32>(v), mnw = group_min<32>(v);
                if (lane == 0) { wred[warp * 2] = mxw; wred[warp * 2 + 1] = mnw; }
            }
            __syncthreads();
            // ---- (b) per-candidate clamp bounds and quantiser parameters
            if (tid <= nc2) {
                float mx = wred[0], mn = wred[1];
                for (int i = 1; i < G / 32; ++i) { mx = fmaxf(mx, wred[2 * i]); mn = fminf(mn, wred[2 * i + 1]); }
                float* cp = cand + tid * 12;
                cp[11] = 0.f;
                if (tid == nc2) {
                    cp[4] = -INFINITY; cp[5] = INFINITY; cp[7] = 1.f;   // weight row: pass-through
                } else {
                    const int ip = tid / nc, in = tid - ip * nc;
                    // org * (1 - i / n_grid): python double factor, cast to fp32 opmath, product rounded to T (autoclip.py:50,52)
                    const float hi = rT<DT>(mx * (float)(1.0 - (double)ip / (double)a.n_grid));
                    const float lo = rT<DT>(mn * (float)(1.0 - (double)in / (double)a.n_grid));
                    const float cmx = fminf(fmaxf(mx, lo), hi), cmn = fminf(fmaxf(mn, lo), hi);   // extremes of clamp(w, lo, hi)
                    cp[4] = lo; cp[5] = hi; cp[7] = 0.f;
                    if (a.quant_type == BDQ_QUANT_INT) {
                        const IntQ p = int_group_params<DT>(cmx, cmn, qmax, a.sem);
                        cp[0] = p.s; cp[1] = p.y; cp[2] = p.z; cp[3] = p.t1;
                        if constexpr (DT != kF32) {
                            // 2-bit candidates in threshold form (fakequant_thr.cuh: exact, verified thresholds): the row
                            // is then written with packed compares instead of the op-by-op quantiser
                            if (a.bits == 2) {
                                bool ok = fmaxf(fabsf(cmx), fabsf(cmn)) * p.y < 1e30f;
#pragma unroll
                                for (int k = 0; k < 3; ++k) {
                                    bool o;
                                    cp[8 + k] = find_threshold<DT, BDQ_ROUND_HALF_EVEN>(-p.z + (float)(k + 1), p, o);
                                    ok = ok && o;
                                }
                                cp[11] = ok ? 1.f : 0.f;
                            }
                        }
                    } else {
                        const Nf3Q p = nf3_group_params(cmx, cmn);
                        cp[0] = p.sp; cp[1] = p.sn; cp[2] = p.yp; cp[3] = p.yn; cp[6] = p.nan_group ? 1.f : 0.f;
                    }
                }
            }
            __syncthreads();
            // ---- (c) candidate rows of A: Q(clamp(w, lo, hi)) with the reference's T roundings
            // (16 threads per row: the row's parameters are read once, not once per element pair)
            for (int c = tid >> 4; c <= nc2; c += kAcThreads / 16) {
              const float* cp = cand + c * 12;
              if constexpr (DT != kF32) {
                  if (cp[11] != 0.f) {   // threshold form: clamp, count the thresholds passed, (count - z) * s
                      using P = P2<DT>;
                      using h2 = typename P::h2;
                      const h2 lo2 = P::dup(cp[4]), hi2 = P::dup(cp[5]), t1 = P::dup(cp[8]), t2 = P::dup(cp[9]), t3 = P::dup(cp[10]);
                      const h2 nz2 = P::dup(-cp[2]), s2 = P::dup(cp[0]);
                      uint32_t* arow = reinterpret_cast<uint32_t*>(As + (size_t)c * L.pitch);
                      for (int j = tid & 15; j < G / 2; j += 16) {
                          const h2 x2 = __hmin2(__hmax2(P::pack(wf[2 * j], wf[2 * j + 1]), lo2), hi2);
                          const h2 c2 = __hadd2(__hadd2(__hge2(x2, t1), __hge2(x2, t2)), __hge2(x2, t3));
                          arow[j] = P::bits(__hmul2(__hadd2(c2, nz2), s2));   // rT((c - z) * s)
                      }
                      continue;
                  }
              }
              for (int j = tid & 15; j < G / 2; j += 16) {
                float v[2];
#pragma unroll
                for (int e = 0; e < 2; ++e) {
                    const float xw = fminf(fmaxf(wf[2 * j + e], cp[4]), cp[5]);
                    if (cp[7] != 0.f) {
                        v[e] = xw;
                    } else if (a.quant_type == BDQ_QUANT_INT) {
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
                uint8_t* row = As + (size_t)c * L.pitch;
                if constexpr (DT == kF32) {
                    const float h0 = bf16_hi(v[0]), h1 = bf16_hi(v[1]);
                    const uint32_t hi = enc16<false>(h0) | ((uint32_t)enc16<false>(h1) << 16);
                    const uint32_t lo = enc16<false>(v[0] - h0) | ((uint32_t)enc16<false>(v[1] - h1) << 16);
                    reinterpret_cast<uint32_t*>(row)[j] = hi;
                    reinterpret_cast<uint32_t*>(row + G * 2)[j] = hi;
                    reinterpret_cast<uint32_t*>(row + G * 4)[j] = lo;
                } else {
                    reinterpret_cast<uint32_t*>(row)[j] = enc16<F16>(v[0]) | ((uint32_t)enc16<F16>(v[1]) << 16);
                }
              }
            }
            __syncthreads();
            // ---- (d) tensor-core contraction: this warp owns tokens [warp*8*NTW, +8*NTW) of the staged chunk
            const int wt = nc2 >> 4, wr = nc2 & 15;          // row tile / row of the weight row
            float ow[NTW][2];
#pragma unroll 1
            for (int mi = 0; mi < kAcRows / 16; ++mi) {
                const int mt = (mi == 0) ? wt : ((mi <= wt) ? mi - 1 : mi);   // weight-row tile first
                float acc[NTW][4];
#pragma unroll
                for (int n = 0; n < NTW; ++n) acc[n][0] = acc[n][1] = acc[n][2] = acc[n][3] = 0.f;
                const uint8_t* arow = As + (size_t)(mt * 16 + (lane & 15)) * L.pitch + (lane >> 4) * 16;
                const uint8_t* brow = Xs + (size_t)(warp * 8 * NTW + (lane & 7) + ((lane >> 4) << 3)) * L.pitch + ((lane >> 3) & 1) * 16;
                if constexpr (DT != kF32) {
#pragma unroll
                    for (int ks = 0; ks < KS; ++ks) {
                        if (ks * 16 < KX) {
                            uint32_t af[4];
                            ldsm_x4(af, arow + ks * 32);
#pragma unroll
                            for (int np = 0; np < NP; ++np) {
                                mma_ac<F16>(acc[2 * np], af, bfr[ks][np][0], bfr[ks][np][1]);
                                mma_ac<F16>(acc[2 * np + 1], af, bfr[ks][np][2], bfr[ks][np][3]);
                            }
                        }
                    }
                } else {
#pragma unroll 4
                    for (int k0 = 0; k0 < KX; k0 += 16) {
                        uint32_t af[4];
                        ldsm_x4(af, arow + k0 * 2);
#pragma unroll
                        for (int n = 0; n < NTW; n += 2) {
                            uint32_t bf[4];
                            ldsm_x4(bf, brow + (size_t)n * 8 * L.pitch + k0 * 2);
                            mma_ac<F16>(acc[n], af, bf[0], bf[1]);
                            mma_ac<F16>(acc[n + 1], af, bf[2], bf[3]);
                        }
                    }
                }
                if (mi == 0) {
                    // out[w, t] of this warp's tokens lives in the lanes with gq == wr % 8: broadcast along the quad column
#pragma unroll
                    for (int n = 0; n < NTW; ++n) {
                        const float v0 = (wr < 8) ? acc[n][0] : acc[n][2], v1 = (wr < 8) ? acc[n][1] : acc[n][3];
                        ow[n][0] = __shfl_sync(0xffffffffu, v0, ((wr & 7) << 2) + tq);
                        ow[n][1] = __shfl_sync(0xffffffffu, v1, ((wr & 7) << 2) + tq);
                    }
                }
                float e_lo = 0.f, e_hi = 0.f;
#pragma unroll
                for (int n = 0; n < NTW; ++n) {
                    const float d0 = acc[n][0] - ow[n][0], d1 = acc[n][1] - ow[n][1];
                    const float d2 = acc[n][2] - ow[n][0], d3 = acc[n][3] - ow[n][1];
                    e_lo = fmaf(d0, d0, fmaf(d1, d1, e_lo));
                    e_hi = fmaf(d2, d2, fmaf(d3, d3, e_hi));
                }
                e_lo += __shfl_xor_sync(0xffffffffu, e_lo, 1);
                e_lo += __shfl_xor_sync(0xffffffffu, e_lo, 2);
                e_hi += __shfl_xor_sync(0xffffffffu, e_hi, 1);
                e_hi += __shfl_xor_sync(0xffffffffu, e_hi, 2);
                if (tq == 0) {
                    errw[warp * kAcRows + mt * 16 + gq] = e_lo;
                    errw[warp * kAcRows + mt * 16 + gq + 8] = e_hi;
                }
            }
            __syncthreads();
            // ---- (e) fold the 8 token blocks into the running per-candidate error of this channel
            if (tid < kAcRows) {
                float s = 0.f;
#pragma unroll
                for (int w8 = 0; w8 < kAcWarps; ++w8) s += errw[w8 * kAcRows + tid];
                errsum[ol * kAcRows + tid] += s;
            }
        }
    }
    __syncthreads();
    // ---- first smallest candidate per channel (strict '<' from 1e9 in (i_p, i_n) order), one warp per channel
    const float inv_tok = 1.0f / (float)a.n_tok;
    for (int ol = warp; ol < no; ol += kAcWarps) {
        const int64_t o = o0 + ol;
        float be = 1e9f;
        int bi = -1;
        for (int c = lane; c < nc2; c += 32) {
            const float e = errsum[ol * kAcRows + c] * inv_tok;
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
            // extremes of the group (recomputed: wred belongs to the last channel processed)
            float mx = -INFINITY, mn = INFINITY;
            for (int k = 0; k < G; ++k) {
                const float v = load_scalar<DT>(a.w, o * a.ci + (int64_t)g * G + k);
                mx = fmaxf(mx, v);
                mn = fminf(mn, v);
            }
            float bmax = mx, bmin = mn;
            if (bi >= 0) {
                const int ip = bi / nc, in = bi - ip * nc;
                bmax = rT<DT>(mx * (float)(1.0 - (double)ip / (double)a.n_grid));
                bmin = rT<DT>(mn * (float)(1.0 - (double)in / (double)a.n_grid));
            }
            store_scalar<DT>(a.best_max, o * ng + g, bmax);
            store_scalar<DT>(a.best_min, o * ng + g, bmin);
            if (a.best_idx) a.best_idx[o * ng + g] = bi;
            if (a.best_err) a.best_err[o * ng + g] = be;
        }
    }
}

template <int DT>
static int launch_autoclip(const AcArgs& a, cudaStream_t st) {
    using C = AcCfg<DT>;
    const AcSmem L = ac_layout(a.G, C::KMUL, C::TCH);
    if (!ensure_dyn_smem(reinterpret_cast<const void*>(autoclip_kernel<DT>), 227 * 1024)) return check_launch("bdq_autoclip(attr)");
    dim3 grid((unsigned)(a.ci / a.G), (unsigned)((a.co + kAcOT - 1) / kAcOT));
    autoclip_kernel<DT><<<grid, C::THREADS, (size_t)L.total, st>>>(a);
    return check_launch("bdq_autoclip");
}

// autoclip_tc.cu: tcgen05 + TMA kernel for 16-bit dtypes, group 128, <= 512 tokens
int launch_autoclip_tc(const void* w, const void* x, void* best_max, void* best_min, int32_t* best_idx, float* best_err, int64_t co,
                       int64_t ci, int64_t n_tok, int64_t xs, int group, int bits, int quant_type, int n_grid, int nc, int dtype, int sem,
                       cudaStream_t st);

}  // namespace bdq

using namespace bdq;

extern "C" int bdq_autoclip(const void* w, const void* x, void* best_max, void* best_min, int32_t* best_idx, float* best_err,
                            int64_t co, int64_t ci, int64_t n_tok, int64_t x_row_stride, int group, int bits, int quant_type,
                            int n_grid, double max_shrink, int dtype, int sem, bdq_stream_t stream) {
    if (dtype < 0 || dtype > 2) { set_error("dtype=%d unknown", dtype); return BDQ_ERR_DTYPE; }
    if (group != 32 && group != 64 && group != 128) { set_error("autoclip: group=%d unsupported (32, 64, 128)", group); return BDQ_ERR_UNSUPPORTED; }
    if (co < 0 || ci <= 0 || ci % group != 0 || n_tok <= 0 || x_row_stride < ci) {
        set_error("autoclip: co=%lld ci=%lld n_tok=%lld stride=%lld group=%d", (long long)co, (long long)ci, (long long)n_tok,
                  (long long)x_row_stride, group);
        return BDQ_ERR_SHAPE;
    }
    if (quant_type != BDQ_QUANT_INT && quant_type != BDQ_QUANT_NF3) { set_error("autoclip: quant_type=%d", quant_type); return BDQ_ERR_UNSUPPORTED; }
    if (quant_type == BDQ_QUANT_INT && (bits < 1 || bits > 8)) { set_error("autoclip: bits=%d", bits); return BDQ_ERR_UNSUPPORTED; }
    const int nc = (int)(max_shrink * n_grid);
    if (n_grid <= 0 || nc < 1 || nc * nc > kAcRows - 1) { set_error("autoclip: n_grid=%d max_shrink=%g -> %d candidates per side (1..10)", n_grid, max_shrink, nc); return BDQ_ERR_UNSUPPORTED; }
    if (co == 0) return BDQ_OK;
    if (!w || !x || !best_max || !best_min) { set_error("null tensor pointer"); return BDQ_ERR_SHAPE; }
    if (ci / group > 65535 * 32 || (co + kAcOT - 1) / kAcOT > 65535) { set_error("autoclip: shape too large"); return BDQ_ERR_SHAPE; }
    {   // Blackwell-native path first (tcgen05 MMA, TMA-staged X); BDQ_ERR_UNSUPPORTED = not its shape
        const int rc = launch_autoclip_tc(w, x, best_max, best_min, best_idx, best_err, co, ci, n_tok, x_row_stride, group, bits,
                                          quant_type, n_grid, nc, dtype, sem, (cudaStream_t)stream);
        if (rc != BDQ_ERR_UNSUPPORTED) return rc;
    }
    AcArgs a;
    a.w = w; a.x = x; a.best_max = best_max; a.best_min = best_min; a.best_idx = best_idx; a.best_err = best_err;
    a.co = co; a.ci = ci; a.n_tok = n_tok; a.xs = x_row_stride;
    a.G = group; a.bits = bits; a.quant_type = quant_type; a.n_grid = n_grid; a.nc = nc; a.sem = sem;
    a.tb = make_tables(dtype, sem);
    cudaStream_t st = (cudaStream_t)stream;
    if (dtype == BDQ_F32) return launch_autoclip<kF32>(a, st);
    if (dtype == BDQ_BF16) return launch_autoclip<kBF16>(a, st);
    return launch_autoclip<kF16>(a, st);
}
