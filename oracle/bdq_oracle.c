/*
 * bdq_oracle.c -- CPU restatement of BitDistiller's low-bit weight-quantisation hot path.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing in the product path (bitdistiller_b200/) may link or call
 * this file; only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
 * legs use it, and there only as the checker / the timed CPU baseline.
 *
 * Parity pin: every function here is checked against outputs of the UNMODIFIED reference Python
 * (imported from /root/reference/quantization in the build container) stored under tests/golden/
 * by tests/golden/make_golden.py.  The reference itself ships no golden vectors.
 *
 * Conventions: tensors travel as float32 arrays whose values are representable in the logical
 * dtype T (0 = fp32, 1 = bf16, 2 = fp16); rT() rounds an fp32 opmath result to T exactly like a
 * PyTorch eager op does between two kernels.  `sem` selects eager-device semantics:
 *   0 = CPU eager  (tensor / python_scalar is a true division)
 *   1 = CUDA eager (tensor / python_scalar multiplies by the fp32 reciprocal)
 * (scalar compares cast the python scalar to T on BOTH devices -- measured, tools/dbg_nf3.py)
 * All citations are file:line under /root/reference.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#define ORC_F32 0
#define ORC_BF16 1
#define ORC_F16 2

static inline float bf16_rn(float f) {
    uint32_t u;
    memcpy(&u, &f, 4);
    if ((u & 0x7fffffffu) > 0x7f800000u) return f; /* NaN */
    u += 0x7fffu + ((u >> 16) & 1u);
    u &= 0xffff0000u;
    memcpy(&f, &u, 4);
    return f;
}

static inline float f16_rn(float f) { return (float)(_Float16)f; }

static inline float rT(float f, int dt) {
    if (dt == ORC_BF16) return bf16_rn(f);
    if (dt == ORC_F16) return f16_rn(f);
    return f;
}

/* tensor / python scalar, quantization/quantizer.py:42 ("/ max_int") */
static inline float div_scalar(float a, float q, int sem) {
    if (sem == 1) return a * (1.0f / q);
    return a / q;
}

/* Round.forward, quantization/quantizer.py:14-17: sign(x) * floor(abs(x) + 0.5), each op rounded to T */
static inline float round_half_away(float u, int dt) {
    float sg = (u > 0.f) ? 1.f : ((u < 0.f) ? -1.f : 0.f);
    float a = rT(fabsf(u) + 0.5f, dt);
    return rT(sg * floorf(a), dt);
}

/*
 * Group asymmetric integer fake-quant.
 * pseudo_quantize_tensor, quantization/quantizer.py:25-69 (round_mode 0: torch.round, half-even)
 * SteInt2AsymQuantizer / SteInt3AsymQuantizer .forward, quantizer.py:119-182 (round_mode 1: Round.apply)
 * Optional clamp to [clip_min[g], clip_max[g]] first = apply_clip, quantization/clip_utils.py:343-345.
 * Outputs (any may be NULL): dq[n], scales[n/G], zeros[n/G], codes[n] (uint8), mask[n] (uint8, clamp pass-through).
 */
void orc_int_fakequant(const float *w, int64_t n, int G, int bits, int dt, int round_mode, int sem,
                       const float *clip_min, const float *clip_max,
                       float *dq, float *scales, float *zeros, uint8_t *codes, uint8_t *mask) {
    const float qmax = (float)((1 << bits) - 1);
    float *buf = (float *)malloc(sizeof(float) * (size_t)G);
    for (int64_t g = 0; g < n / G; ++g) {
        const float *x0 = w + g * G;
        for (int k = 0; k < G; ++k) {
            float v = x0[k];
            if (clip_min) v = fminf(fmaxf(v, clip_min[g]), clip_max[g]); /* torch.clamp(x, lo, hi) */
            buf[k] = v;
        }
        float mx = buf[0], mn = buf[0];
        for (int k = 1; k < G; ++k) { mx = fmaxf(mx, buf[k]); mn = fminf(mn, buf[k]); }
        float t1 = rT(mx - mn, dt);
        float t2 = rT(fmaxf(t1, 1e-5f), dt);
        float s = rT(div_scalar(t2, qmax, sem), dt);              /* quantizer.py:42 */
        float t3 = rT(mn / s, dt);
        float z = fminf(fmaxf(-rintf(t3), 0.f), qmax);             /* quantizer.py:43 */
        if (z == 0.f) z = 0.f;                                     /* -0 -> +0, numerically identical */
        if (scales) scales[g] = s;
        if (zeros) zeros[g] = z;
        for (int k = 0; k < G; ++k) {
            float u = rT(buf[k] / s, dt);
            float r = round_mode ? round_half_away(u, dt) : rintf(u);
            float q = rT(r + z, dt);
            float c = fminf(fmaxf(q, 0.f), qmax);
            if (mask) mask[g * G + k] = (q >= 0.f && q <= qmax) ? 1 : 0;
            if (codes) codes[g * G + k] = (uint8_t)c;
            if (dq) dq[g * G + k] = rT(rT(c - z, dt) * s, dt);     /* quantizer.py:60-61 */
        }
    }
    free(buf);
}

/*
 * Backward of Ste{Int2,Int3}AsymQuantizer (autograd over quantizer.py:166-177), closed form of
 * SURVEY.md App. B evaluated in double:  gW = g*m + [x==mx]*(Gs/qmax)/#max - [x==mn]*(Gs/qmax)/#min,
 * Gs = sum_j g_j*((c_j - z) - m_j*x_j/s); the scale path is cut when (mx-mn) < 1e-5 (clamp(min=1e-5)).
 */
void orc_int_ste_bwd(const float *w, const float *gy, int64_t n, int G, int bits, int dt, int sem,
                     double *gw) {
    const float qmax = (float)((1 << bits) - 1);
    float *buf = (float *)malloc(sizeof(float) * (size_t)G);
    for (int64_t g = 0; g < n / G; ++g) {
        for (int k = 0; k < G; ++k) {
            buf[k] = w[g * G + k];
        }
        float mx = buf[0], mn = buf[0];
        for (int k = 1; k < G; ++k) { mx = fmaxf(mx, buf[k]); mn = fminf(mn, buf[k]); }
        float t1 = rT(mx - mn, dt);
        float t2 = rT(fmaxf(t1, 1e-5f), dt);
        float s = rT(div_scalar(t2, qmax, sem), dt);
        float t3 = rT(mn / s, dt);
        float z = fminf(fmaxf(-rintf(t3), 0.f), qmax);
        double Gs = 0.0;
        int nmax = 0, nmin = 0;
        for (int k = 0; k < G; ++k) {
            float u = rT(buf[k] / s, dt);
            float r = round_half_away(u, dt);
            float q = rT(r + z, dt);
            float c = fminf(fmaxf(q, 0.f), qmax);
            int m = (q >= 0.f && q <= qmax);
            double gk = gy[g * G + k];
            Gs += gk * (((double)c - (double)z) - (m ? (double)buf[k] / (double)s : 0.0));
            gw[g * G + k] = m ? gk : 0.0;
            nmax += (buf[k] == mx);
            nmin += (buf[k] == mn);
        }
        if (t1 >= 1e-5f) {
            double gs = Gs / (double)qmax;
            for (int k = 0; k < G; ++k) {
                if (buf[k] == mx) gw[g * G + k] += gs / nmax;
                if (buf[k] == mn) gw[g * G + k] -= gs / nmin;
            }
        }
    }
    free(buf);
}

/* ---------------------------------------------------------------- NF3 ------------------------------------------ */
/* round_n2f3, quantization/quantizer.py:224-236: level tables and decision thresholds */
static const double NF3_POS_LVL[5] = {0.0, 0.1833375245332718, 0.3819939494132996, 0.6229856610298157, 1.0};
static const double NF3_POS_THR[4] = {0.0916687622666359, 0.2826657369732857, 0.5024898052215576, 0.8114928305149078};
static const double NF3_NEG_LVL[4] = {0.0, -0.2469314038753510, -0.5350227355957031, -1.0};
static const double NF3_NEG_THR[3] = {-0.1234657019376755, -0.39097706973552704, -0.7675113677978516};

static inline float thrT(double t, int dt, int sem) {
    (void)sem; /* measured on B200: CUDA eager, like CPU eager, casts the wrapped python scalar to T */
    return rT((float)t, dt);
}

/* returns level index 0..4 (NaN -> -1: every where() leaves q untouched) */
static inline int nf3_pos_idx(float q, int dt, int sem) {
    if (q != q) return -1;
    if (q >= thrT(NF3_POS_THR[3], dt, sem)) return 4;
    if (q >= thrT(NF3_POS_THR[2], dt, sem)) return 3;
    if (q >= thrT(NF3_POS_THR[1], dt, sem)) return 2;
    if (q >= thrT(NF3_POS_THR[0], dt, sem)) return 1;
    return 0;
}
static inline int nf3_neg_idx(float q, int dt, int sem) {
    if (q != q) return -1;
    if (q >= thrT(NF3_NEG_THR[0], dt, sem)) return 0;
    if (q >= thrT(NF3_NEG_THR[1], dt, sem)) return 1;
    if (q >= thrT(NF3_NEG_THR[2], dt, sem)) return 2;
    return 3;
}

/*
 * SteN2F3Quantizer.forward, quantization/quantizer.py:189-242 (pseudo_quantize_n2f3_tensor 113-116).
 * codes: 0..4 = positive level index, 5..7 = negative level index 1..3 (0 is shared), 255 = NaN.
 */
void orc_nf3_fakequant(const float *w, int64_t n, int G, int dt, int sem,
                       const float *clip_min, const float *clip_max,
                       float *dq, float *scale_pos, float *scale_neg, uint8_t *codes) {
    float *buf = (float *)malloc(sizeof(float) * (size_t)G);
    for (int64_t g = 0; g < n / G; ++g) {
        for (int k = 0; k < G; ++k) {
            float v = w[g * G + k];
            if (clip_min) v = fminf(fmaxf(v, clip_min[g]), clip_max[g]);
            buf[k] = v;
        }
        float mx = buf[0], mn = buf[0];
        for (int k = 1; k < G; ++k) { mx = fmaxf(mx, buf[k]); mn = fminf(mn, buf[k]); }
        float sp = fabsf(mx), sn = fabsf(mn);                      /* quantizer.py:205-206 */
        if (scale_pos) scale_pos[g] = sp;
        if (scale_neg) scale_neg[g] = sn;
        for (int k = 0; k < G; ++k) {
            float x = buf[k];
            float xp = (x >= 0.f) ? x : 0.f, xn = (x < 0.f) ? x : 0.f;   /* :209-212 */
            float qp = rT(xp / sp, dt), qn = rT(xn / sn, dt);           /* :213-214 */
            int ip = nf3_pos_idx(qp, dt, sem), in = nf3_neg_idx(qn, dt, sem);
            float lp = ip < 0 ? qp : rT((float)NF3_POS_LVL[ip], dt);
            float ln = in < 0 ? qn : rT((float)NF3_NEG_LVL[in], dt);
            float yp = rT(rT(lp - qp, dt) + qp, dt);                   /* :242 */
            float yn = rT(rT(ln - qn, dt) + qn, dt);
            if (dq) dq[g * G + k] = rT(rT(yp * sp, dt) + rT(yn * sn, dt), dt); /* :218 */
            if (codes) codes[g * G + k] = (ip < 0 || in < 0) ? 255 : (uint8_t)(in > 0 ? 4 + in : ip);
        }
    }
    free(buf);
}

/* Backward of SteN2F3Quantizer, in double:
 * gW = g + [x==mx]*sign(mx)*Gp/#max + [x==mn]*sign(mn)*Gn/#min, Gp = sum g*(yp-qp), Gn = sum g*(yn-qn). */
void orc_nf3_ste_bwd(const float *w, const float *gy, int64_t n, int G, int dt, int sem, double *gw) {
    for (int64_t g = 0; g < n / G; ++g) {
        const float *x0 = w + g * G;
        float mx = x0[0], mn = x0[0];
        for (int k = 1; k < G; ++k) { mx = fmaxf(mx, x0[k]); mn = fminf(mn, x0[k]); }
        float sp = fabsf(mx), sn = fabsf(mn);
        double Gp = 0, Gn = 0;
        int nmax = 0, nmin = 0;
        for (int k = 0; k < G; ++k) {
            float x = x0[k];
            float xp = (x >= 0.f) ? x : 0.f, xn = (x < 0.f) ? x : 0.f;
            float qp = rT(xp / sp, dt), qn = rT(xn / sn, dt);
            int ip = nf3_pos_idx(qp, dt, sem), in = nf3_neg_idx(qn, dt, sem);
            float lp = ip < 0 ? qp : rT((float)NF3_POS_LVL[ip], dt);
            float ln = in < 0 ? qn : rT((float)NF3_NEG_LVL[in], dt);
            float yp = rT(rT(lp - qp, dt) + qp, dt);
            float yn = rT(rT(ln - qn, dt) + qn, dt);
            double gk = gy[g * G + k];
            Gp += gk * ((double)yp - (double)xp / (double)sp);
            Gn += gk * ((double)yn - (double)xn / (double)sn);
            gw[g * G + k] = gk;
            nmax += (x == mx);
            nmin += (x == mn);
        }
        double sgx = mx > 0 ? 1.0 : (mx < 0 ? -1.0 : 0.0), sgn = mn > 0 ? 1.0 : (mn < 0 ? -1.0 : 0.0);
        for (int k = 0; k < G; ++k) {
            if (x0[k] == mx) gw[g * G + k] += sgx * Gp / nmax;
            if (x0[k] == mn) gw[g * G + k] += sgn * Gn / nmin;
        }
    }
}

/* apply_clip, quantization/clip_utils.py:343-345: w = clamp(w.reshape(co, n_g, G), min_val, max_val) */
void orc_apply_clip(float *w, int64_t n, int G, const float *clip_min, const float *clip_max) {
    for (int64_t i = 0; i < n; ++i) w[i] = fminf(fmaxf(w[i], clip_min[i / G]), clip_max[i / G]);
}

/* ---------------------------------------------------------------- autoclip -------------------------------------- */
/*
 * auto_2clip_layer, quantization/autoclip.py:13-83.  w[co*ci] (values in T), x[T_tok*ci] = the ALREADY
 * sub-sampled activations (autoclip.py:27, done by the caller), quant_type 0 = "int" (:56), 1 = "nf3" (:58).
 * For every (o, g): candidates (i_p outer, i_n inner; n_cand = int(max_shrink*n_grid) each, :49-52),
 *   max' = rT(mx * (float)(1 - i_p/n_grid)), min' = rT(mn * (float)(1 - i_n/n_grid)),
 *   err = mean_t (sum_k x[t,g,k] * (Q(clamp(w, min', max'))[k] - w[k]))^2      (:54-65),
 * strict '<' running minimum from 1e9 (:45, :69-72).  The candidate weights are generated with the exact
 * T roundings of the reference; the error contraction is evaluated in double (the "fp32 oracle" of
 * the reference's own bf16 contraction is noise-limited).
 * Outputs: best_max/best_min [co*n_g] (T values), best_idx [co*n_g] = i_p*n_cand+i_n (or -1), best_err, all_err (opt).
 */
void orc_auto_2clip(const float *w, const float *x, int64_t co, int64_t ci, int64_t n_tok, int G, int bits,
                    int dt, int sem, int quant_type, int n_grid, double max_shrink,
                    float *best_max, float *best_min, int32_t *best_idx, double *best_err, double *all_err) {
    const int nc = (int)(max_shrink * n_grid);
    const int64_t ng = ci / G;
    float *cw = (float *)malloc(sizeof(float) * (size_t)G);
    float *qw = (float *)malloc(sizeof(float) * (size_t)G);
    double *d = (double *)malloc(sizeof(double) * (size_t)G);
    for (int64_t o = 0; o < co; ++o) {
        for (int64_t g = 0; g < ng; ++g) {
            const float *w0 = w + o * ci + g * G;
            float mx = w0[0], mn = w0[0];
            for (int k = 1; k < G; ++k) { mx = fmaxf(mx, w0[k]); mn = fminf(mn, w0[k]); }
            double berr = 1e9;
            float bmax = mx, bmin = mn;
            int bidx = -1;
            for (int ip = 0; ip < nc; ++ip) {
                float hi = rT(mx * (float)(1.0 - (double)ip / n_grid), dt);
                for (int in = 0; in < nc; ++in) {
                    float lo = rT(mn * (float)(1.0 - (double)in / n_grid), dt);
                    for (int k = 0; k < G; ++k) cw[k] = fminf(fmaxf(w0[k], lo), hi);
                    if (quant_type == 0)
                        orc_int_fakequant(cw, G, G, bits, dt, 0, sem, NULL, NULL, qw, NULL, NULL, NULL, NULL);
                    else
                        orc_nf3_fakequant(cw, G, G, dt, sem, NULL, NULL, qw, NULL, NULL, NULL);
                    for (int k = 0; k < G; ++k) d[k] = (double)qw[k] - (double)w0[k];
                    double err = 0;
                    for (int64_t t = 0; t < n_tok; ++t) {
                        const float *xt = x + t * ci + g * G;
                        double acc = 0;
                        for (int k = 0; k < G; ++k) acc += (double)xt[k] * d[k];
                        err += acc * acc;
                    }
                    err /= (double)n_tok;
                    if (all_err) all_err[(o * ng + g) * nc * nc + ip * nc + in] = err;
                    if (err < berr) { berr = err; bmax = hi; bmin = lo; bidx = ip * nc + in; }
                }
            }
            best_max[o * ng + g] = bmax;
            best_min[o * ng + g] = bmin;
            if (best_idx) best_idx[o * ng + g] = bidx;
            if (best_err) best_err[o * ng + g] = berr;
        }
    }
    free(cw); free(qw); free(d);
}

/* ---------------------------------------------------------------- packing --------------------------------------- */
/* calculate_zeros_width, quantization/qmodule.py:16-28 */
int orc_zeros_width(int64_t in_features, int G, int pack) {
    int mult = G >= 128 ? 1 : (G == 64 ? 2 : (G == 32 ? 4 : -1));
    if (mult < 0) return -1;
    int64_t bw = ((in_features / G) + pack - 1) / pack;
    bw = ((bw + mult - 1) / mult) * mult;
    return (int)bw;
}

/*
 * WQLinear.from_linear, quantization/qmodule.py:70-146.  w_dq [rows*cols] (T), scales/zeros [rows*ng] (T).
 * qweight int32 [rows, cols/pack]; qzeros int32 [rows, zw]; qscales fp16 (as float) [rows, zw*pack].
 * code = (int)rint( rP( rT(w + rT(z*s)) / fp16(s) ) ), P = promote(T, fp16) (:94-99); sequential packing (:101-139).
 */
void orc_pack(const float *w_dq, const float *scales, const float *zeros, int64_t rows, int64_t cols, int G,
              int w_bit, int dt, int32_t *qweight, int32_t *qzeros, float *qscales) {
    const int pack = 32 / w_bit;
    const int64_t ng = cols / G;
    const int zw = orc_zeros_width(cols, G, pack);
    const int pdt = (dt == ORC_F16) ? ORC_F16 : ORC_F32; /* bf16/fp16 and fp32/fp16 promote to fp32 */
    memset(qweight, 0, sizeof(int32_t) * (size_t)(rows * (cols / pack)));
    memset(qzeros, 0, sizeof(int32_t) * (size_t)(rows * zw));
    for (int64_t r = 0; r < rows; ++r) {
        for (int64_t j = 0; j < (int64_t)zw * pack; ++j)
            qscales[r * zw * pack + j] = j < ng ? f16_rn(scales[r * ng + j]) : 0.f;
        for (int64_t c = 0; c < cols; ++c) {
            int64_t g = c / G;
            float sz = rT(zeros[r * ng + g] * scales[r * ng + g], dt);
            float a = rT(w_dq[r * cols + c] + sz, dt);
            float q = rT(a / qscales[r * zw * pack + g], pdt);
            int32_t code = (int32_t)rintf(q);
            uint32_t *dst = (uint32_t *)&qweight[r * (cols / pack) + c / pack];
            *dst |= ((uint32_t)code) << ((c % pack) * w_bit);
        }
        for (int64_t g = 0; g < ng; ++g) {
            int32_t z = (int32_t)zeros[r * ng + g];
            uint32_t *dst = (uint32_t *)&qzeros[r * zw + g / pack];
            *dst |= ((uint32_t)z) << ((g % pack) * w_bit);
        }
    }
}

/* ---------------------------------------------------------------- packed GEMV / GEMM ---------------------------- */
/*
 * gemv_kernel_g128_2bit, inference/kernels/csrc/quantization/gemv_cuda.cu:202-272 (and the 4-bit siblings):
 * out[b,oc] = fp16( sum_k (s[oc,k/G] * (code[oc,k] - z[oc,k/G])) * x[b,k] ), fp32 partial sums in the reference;
 * evaluated in double here (tolerance rtol 1e-2 per BASELINE.json).  dequant_fp16 = 1 additionally rounds the
 * dequantised weight to fp16 before the product = quant_matmul_kernel, quantization/triton_kernels.py:89-93.
 * x [B*IC] fp16 values, qweight [OC, IC/pack], qzeros [OC, zw], qscales [OC, zw*pack] (fp16 values).
 */
void orc_wq_matmul(const float *x, const int32_t *qweight, const int32_t *qzeros, const float *qscales,
                   int64_t B, int64_t IC, int64_t OC, int G, int w_bit, int dequant_fp16, float *out) {
    const int pack = 32 / w_bit;
    const uint32_t msk = (1u << w_bit) - 1u;
    const int zw = orc_zeros_width(IC, G, pack);
    double *wrow = (double *)malloc(sizeof(double) * (size_t)IC);
    for (int64_t oc = 0; oc < OC; ++oc) {
        for (int64_t k = 0; k < IC; ++k) {
            uint32_t word = (uint32_t)qweight[oc * (IC / pack) + k / pack];
            float code = (float)((word >> ((k % pack) * w_bit)) & msk);
            int64_t g = k / G;
            uint32_t zword = (uint32_t)qzeros[oc * zw + g / pack];
            float z = (float)((zword >> ((g % pack) * w_bit)) & msk);
            float dqv = qscales[oc * zw * pack + g] * (code - z);
            wrow[k] = dequant_fp16 ? (double)f16_rn(dqv) : (double)dqv;
        }
        for (int64_t b = 0; b < B; ++b) {
            double acc = 0;
            for (int64_t k = 0; k < IC; ++k) acc += wrow[k] * (double)x[b * IC + k];
            out[b * OC + oc] = f16_rn((float)acc);
        }
    }
    free(wrow);
}

/*
 * GPTQ-v2 2-bit export: get_hf_quant_params + convert_layer_to_gptq_format,
 * train/convert_hf_to_gptq_manual.py:11-75 and 85-157.
 * w [OC*IC] holds the weight values (already widened to fp32 as line 28 does); dt is the weight's own dtype, used only
 * for the scales' trip through it (line 64) before the bf16 cast (line 100).  All arithmetic is fp32 host arithmetic.
 * qweight [IC/16, OC]: bits 2i of [r][n] = code[n][16r+i] (113-115); qzeros [IC/G, OC/16]: bits 2i of [g][c] =
 * z[16c+i][g] (127-130); scales [IC/G, OC] (bf16 values).
 */
void orc_gptq_export_w2(const float *w, int64_t OC, int64_t IC, int G, int dt, int32_t *qweight, int32_t *qzeros,
                        float *scales) {
    const int64_t ng = IC / G;
    memset(qweight, 0, sizeof(int32_t) * (size_t)((IC / 16) * OC));
    memset(qzeros, 0, sizeof(int32_t) * (size_t)(ng * (OC / 16)));
    for (int64_t o = 0; o < OC; ++o) {
        for (int64_t g = 0; g < ng; ++g) {
            const float *x0 = w + o * IC + g * G;
            float mx = x0[0], mn = x0[0];
            for (int k = 1; k < G; ++k) { mx = fmaxf(mx, x0[k]); mn = fminf(mn, x0[k]); }
            float s = fmaxf(mx - mn, 1e-5f) / 3.0f;                         /* line 45 */
            float z = fminf(fmaxf(-rintf(mn / s), 0.f), 3.f);               /* line 46 */
            uint32_t zi = (uint32_t)(int)z;
            scales[g * OC + o] = bf16_rn(rT(s, dt));
            ((uint32_t *)qzeros)[g * (OC / 16) + o / 16] |= zi << (2 * (o % 16));
            for (int k = 0; k < G; ++k) {
                float c = fminf(fmaxf(rintf(x0[k] / s) + z, 0.f), 3.f);     /* line 58 */
                int64_t kk = g * G + k;
                ((uint32_t *)qweight)[(kk / 16) * OC + o] |= (uint32_t)(int)c << (2 * (kk % 16));
            }
        }
    }
}
