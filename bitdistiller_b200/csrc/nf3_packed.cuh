// Packed NF3 format shared by nf3_packed.cu (pack / dequant / small-shape GEMV) and gemv_nf3_stream.cu (decode GEMV on the
// INT8 tensor pipe).  The reference has NO packed NF3 format, so the layout is this build's own and
// its parity is pinned on VALUES: dequant(pack(w)) == SteN2F3Quantizer.forward (quantization/quantizer.py:113-116,184-242).
//
//   level index e in 0..7 = position in the ascending staircase -1, neg_lvl[2], neg_lvl[1], 0, pos_lvl[1..4]
//     (e < 3: negative level, scaled by sn = |group min|;  e >= 3: zero or positive level, scaled by sp = |group max|)
//   qweight int32 [OC, IC/16]: bits [2i, 2i+2) of word c = low two bits of e of input channel 16c + i (the int2 layout)
//   qhigh   int32 [OC, IC/32]: bit nf3_hbit_pos(k) of word c = high bit of e of input channel 32c + k:
//     with k = 16w + 2n + p (w = which qweight word, n = nibble 0..7, p = parity) the bit sits at 4n + 2w + p, so that
//         ((qweight_word >> 2p) & 0x33333333) | ((qhigh_word << (2 - 2w - p)) & 0x44444444)
//     is eight 4-bit level indices (channels 16w + 2n + p, n = 0..7) = two PRMT selectors, no per-code extraction
//   scales  fp16  [OC, 2*IC/G]: (sp, sn) per group
#pragma once
#include "bdq_common.cuh"
#include "fq_nf3_elem.cuh"

namespace bdq {

struct Nf3Levels {
    float lvl[8];   // ascending staircase, rounded to fp16
};

inline Nf3Levels nf3_levels_f16() {
    const Nf3Tables t = make_tables(kF16, 0);
    Nf3Levels l;
    l.lvl[0] = t.neg_lvl[3]; l.lvl[1] = t.neg_lvl[2]; l.lvl[2] = t.neg_lvl[1]; l.lvl[3] = 0.f;
    for (int i = 1; i < 5; ++i) l.lvl[3 + i] = t.pos_lvl[i];
    return l;
}

// bit of a qhigh word that holds the high bit of channel k (0..31) of its 32-channel chunk
__host__ __device__ constexpr int nf3_hbit_pos(int k) { return 4 * ((k & 15) >> 1) + 2 * (k >> 4) + (k & 1); }

// Streaming decode GEMV (gemv_nf3_stream.cu): BDQ_ERR_UNSUPPORTED when the shape does not fit (caller falls back)
int nf3_gemv_stream_launch(const void* x, const int32_t* qweight, const int32_t* qhigh, const void* scales, void* out, int64_t B,
                           int64_t IC, int64_t OC, int group, cudaStream_t stream);

}  // namespace bdq
