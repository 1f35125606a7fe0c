/*
 * bdq.h -- C ABI of libbdq.so: B200 (sm_100a) kernels for BitDistiller's low-bit weight-quantisation hot path.
 *
 * Every entry point replaces one function of the reference (file:line under the reference tree cited per
 * function).  Conventions:
 *   - plain pointers and sizes only; all tensor pointers are DEVICE pointers unless the name ends in _host;
 *   - the caller owns every buffer, the library never allocates device memory except inside *_host entries;
 *   - `stream` is a cudaStream_t passed as void*; entries are asynchronous, re-entrant and graph-capturable;
 *   - return value: BDQ_OK or a negative BDQ_ERR_* code; bdq_last_error() gives a thread-local message;
 *   - dtype: BDQ_F32 / BDQ_BF16 / BDQ_F16 = logical dtype T of the weight tensors.  Arithmetic reproduces
 *     eager PyTorch op by op (every intermediate rounded to T);
 *   - sem: BDQ_SEM_CPU reproduces CPU-eager rounding (true division by the python scalar), BDQ_SEM_CUDA
 *     reproduces CUDA-eager (multiply by the fp32 reciprocal of the scalar)  -- SURVEY.md App. D-1.
 */
#ifndef BDQ_H_
#define BDQ_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define BDQ_F32 0
#define BDQ_BF16 1
#define BDQ_F16 2

#define BDQ_SEM_CPU 0
#define BDQ_SEM_CUDA 1

#define BDQ_ROUND_HALF_EVEN 0 /* torch.round: pseudo_quantize_tensor, quantizer.py:60 */
#define BDQ_ROUND_HALF_AWAY 1 /* Round.apply: Ste*AsymQuantizer, quantizer.py:12-17,176 */

#define BDQ_QUANT_INT 0 /* q_config["quant_type"] == "int", autoclip.py:55 */
#define BDQ_QUANT_NF3 1 /* q_config["quant_type"] == "nf3", autoclip.py:57 */

#define BDQ_OK 0
#define BDQ_ERR_SHAPE (-1)
#define BDQ_ERR_DTYPE (-2)
#define BDQ_ERR_ALIGN (-3)
#define BDQ_ERR_CUDA (-4)
#define BDQ_ERR_UNSUPPORTED (-5)

typedef void* bdq_stream_t;

int bdq_version(void);
const char* bdq_last_error(void);

/* Kernels launched by this library since load (monotonic; bench.py reports the delta as gpu_launches). */
int64_t bdq_launch_count(void);

/*
 * Group asymmetric integer fake-quant, one fused pass.
 * Replaces pseudo_quantize_tensor (quantization/quantizer.py:25-69; round_mode HALF_EVEN) and the forward of
 * SteInt2AsymQuantizer / SteInt3AsymQuantizer (quantizer.py:119-182; round_mode HALF_AWAY).
 * w, dq: n_elem values of dtype T (dq may alias w = the inplace=True path, quantizer.py:56-58).
 * group: q_group_size (> 0, divides the innermost dimension; the caller resolves -1 to the row length).
 * Optional outputs (NULL to skip): scales, zeros [n_elem/group] of dtype T (get_scale_zp, quantizer.py:66-67),
 * codes [n_elem] uint8.  Optional clip_min/clip_max [n_elem/group] of dtype T: clamp first, i.e. apply_clip
 * (quantization/clip_utils.py:343-345) fused in front.
 */
int bdq_int_fakequant_fwd(const void* w, void* dq, void* scales, void* zeros, uint8_t* codes,
                          const void* clip_min, const void* clip_max, int64_t n_elem, int group, int bits,
                          int dtype, int round_mode, int sem, bdq_stream_t stream);

/*
 * Straight-through backward of Ste{Int2,Int3}AsymQuantizer = autograd over quantizer.py:166-177 in one pass:
 * gw = gy*mask + d(scale)/d(amax,amin) terms with even tie splitting.  Recomputes the
 * forward from w; nothing is saved.  fp32 internal accumulation.  gw may alias gy.
 */
int bdq_int_ste_bwd(const void* w, const void* gy, void* gw, int64_t n_elem, int group, int bits, int dtype,
                    int sem, bdq_stream_t stream);

/*
 * Fused reduce-scatter + STE backward for FSDP2 shards (no reference counterpart: under ZeRO-3 / FSDP the reference runs the
 * STE backward on the full weight on every rank and leaves the reduce-scatter to the framework, quantization/qlinear.py:63-78).
 * The STE backward is linear in the incoming gradient and group-local, so it commutes with the sum over ranks:
 *   bdq_rows_push        copies up to 64 contiguous segments per launch; with peer-mapped destinations (symmetric memory) every
 *                        rank pushes the rows of its gradient that belong to rank r's shard into r's inbox over NVLink
 *                        (src / dst / nbytes are HOST arrays; 32-byte aligned, multiples of 32 bytes)
 *   bdq_int_ste_bwd_nsrc bdq_int_ste_bwd whose incoming gradient is gy[0] + gy[1] + ... + gy[nsrc-1] (1..8 device pointers in a
 *                        HOST array), accumulated in that order with one rounding to the tensor dtype per addition; 16-bit
 *                        dtypes, bits 2/3, group 32/64/128.  The reduced gradient is never materialised.
 */
/*
 * NF3 payload for quantise-before-all-gather (QLinear's class default quant_type is "ste-n2f3", quantization/qlinear.py:51;
 * no reference counterpart for the wire format).  bdq_nf3_shard_encode quantises a dim-0 shard (whole groups) and stores, into its
 * own slot and into up to 7 peer-mapped slots, per 32 consecutive elements three uint32 words (lo0, lo1, hi: bits [2j, 2j+2) of
 * lo(j / 16) = low two bits of the level index e of element j, bit j of hi = its high bit; e = position in the ascending staircase
 * -1, -0.535, -0.247, 0, 0.183, 0.382, 0.623, 1) and per group one (|max|, |min|) pair of T values (NaN for a group whose max or
 * min is exactly 0: the reference's 0/0).  bdq_nf3_shard_decode rebuilds dq = rT(level * scale), bit-equal to
 * SteN2F3Quantizer.forward (quantization/quantizer.py:184-242) on the same weights.  16-bit dtypes, group 32/64/128;
 * 0.40625 B/weight at group 128.
 */
int bdq_nf3_shard_encode(const void* w, void* codes, void* scales, int64_t n_elem, int group, int dtype, int sem, int n_peer,
                         void* const* peer_codes, void* const* peer_scales, bdq_stream_t stream);
int bdq_nf3_shard_decode(const void* codes, const void* scales, void* dq, int64_t n_elem, int group, int dtype, bdq_stream_t stream);

/*
 * Cross-GPU barrier on peer-mapped flags (one 32-thread kernel): flags[r] = rank r's array of 8 uint32 (symmetric memory, zeroed
 * once), `epoch` = 1, 2, 3, ... identical on all ranks for the same barrier.  Stream-ordered: the rank's earlier work (e.g. its
 * NVLink peer stores) has completed when its flags go out; later work starts after every peer's flag of this epoch has arrived.
 */
int bdq_peer_barrier(void* const* flags, int my_rank, int n_ranks, unsigned int epoch, bdq_stream_t stream);
int bdq_rows_push(int nseg, const void* const* src, void* const* dst, const int64_t* nbytes, bdq_stream_t stream);
int bdq_int_ste_bwd_nsrc(const void* w, int nsrc, const void* const* gy, void* gw, int64_t n_elem, int group, int bits, int dtype,
                         int sem, bdq_stream_t stream);
/* The same for SteN2F3Quantizer (QLinear's class default quant_type, quantization/qlinear.py:51): bdq_nf3_ste_bwd with the incoming
 * gradient summed on the fly from 1..8 sources; 16-bit dtypes, group 32 / 64 / 128. */
int bdq_nf3_ste_bwd_nsrc(const void* w, int nsrc, const void* const* gy, void* gw, int64_t n_elem, int group, int dtype, int sem,
                         bdq_stream_t stream);

/*
 * NF3 (8-level, two scales per group) fake-quant.  Replaces SteN2F3Quantizer.forward
 * (quantization/quantizer.py:189-242) and pseudo_quantize_n2f3_tensor (quantizer.py:113-116).
 * Optional: codes [n_elem] uint8 (0..4 positive level, 5..7 negative level 1..3, 255 NaN),
 * scale_pos/scale_neg [n_elem/group] (T), clip_min/clip_max as above.
 */
int bdq_nf3_fakequant_fwd(const void* w, void* dq, void* scale_pos, void* scale_neg, uint8_t* codes,
                          const void* clip_min, const void* clip_max, int64_t n_elem, int group, int dtype,
                          int sem, bdq_stream_t stream);

/* Backward of SteN2F3Quantizer (identity + scale terms through amax/amin). */
int bdq_nf3_ste_bwd(const void* w, const void* gy, void* gw, int64_t n_elem, int group, int dtype, int sem,
                    bdq_stream_t stream);

/* apply_clip (quantization/clip_utils.py:328-346): w = clamp(w, clip_min[g], clip_max[g]) in place. */
int bdq_apply_clip(void* w, const void* clip_min, const void* clip_max, int64_t n_elem, int group, int dtype,
                   bdq_stream_t stream);

/*
 * calculate_zeros_width (quantization/qmodule.py:16-28): int32 words per row of qzeros; -1 for an unsupported group.
 * Host-only helper, no CUDA call.
 */
int bdq_zeros_width(int64_t in_features, int group, int w_bit);

/*
 * Real quantise + pack = WQLinear.from_linear (quantization/qmodule.py:70-146), one pass instead of
 * in_features + 2*in_features/pack tiny eager kernels.
 * w_dq [rows*cols] (T): the fake-quantised weight (output of pseudo_quantize_tensor, quantizer.py:96);
 * scales, zeros [rows*cols/group] (T).  Outputs (caller-allocated): qweight int32 [rows, cols*w_bit/32],
 * qzeros int32 [rows, zw], qscales fp16 [rows, zw*32/w_bit] (zero-padded), zw = bdq_zeros_width().
 * Codes are packed sequentially: bits [i*w_bit, (i+1)*w_bit) of qweight[r, c] = code of column c*pack+i.
 * w_bit in {2, 4}.
 */
int bdq_pack_wq(const void* w_dq, const void* scales, const void* zeros, int32_t* qweight, int32_t* qzeros,
                void* qscales, int64_t rows, int64_t cols, int group, int w_bit, int dtype, bdq_stream_t stream);

/*
 * Packed 2-bit decode GEMV: out[b, oc] = fp16( sum_k x[b,k] * scales[oc,k/G] * (code[oc,k] - zero[oc,k/G]) ),
 * fp32 accumulation.  Replaces gemv_forward_cuda / gemv_kernel_g128_2bit
 * (inference/kernels/csrc/quantization/gemv_cuda.cu:202-272, 289-346; pybind.cpp:13).
 * x [B, IC] fp16, qweight/qzeros/scales as produced by bdq_pack_wq (w_bit 2), out [B, OC] fp16.
 * group in {64, 128}, IC % 128 == 0, OC % 16 == 0, any B >= 1 (weights are streamed once per 8 rows of x).
 * Bounds are checked against IC (the reference kernel reads past the row end when IC % 2048 != 0).
 * flags: BDQ_GEMV_PDL = launch with programmatic stream serialisation (the weight prefetch of this launch
 * overlaps the tail of the previous kernel on the stream; x is read only after the dependency resolves).
 */
#define BDQ_GEMV_PDL 1
#define BDQ_GEMV_CTAS1 2 /* batch-1 streaming kernel: one CTA per SM instead of two (tuning knob) */
#define BDQ_GEMV_RING 4  /* force the shared-memory ring kernel (the batch 2..8 path) also for batch 1 */
#define BDQ_GEMV_TMA 16  /* batch-1 kernel: weight tiles through 2-D TMA + mbarriers instead of register loads */
#define BDQ_GEMV_SILU_MUL 64 /* bdq_gemv_w2_multi with nmat = 2 (gate, up): out[0][i] = silu(gate_i) * up_i, out[1] unused */
#define BDQ_GEMV_PROF 8  /* dev tool: record %globaltimer phase marks into the buffer given to bdq_gemv_set_prof */
int bdq_gemv_set_prof(void* device_buf_u64, int64_t capacity_u64);
int bdq_gemv_w2(const void* x, const int32_t* qweight, const void* scales, const int32_t* qzeros, void* out,
                int64_t B, int64_t IC, int64_t OC, int group, int flags, bdq_stream_t stream);

/*
 * Batch-1 decode GEMV of up to 4 packed matrices that share one activation vector, in ONE launch: the fused q/k/v
 * projection and the fused gate/up projection of the reference's decode modules
 * (inference/modules/fused_attn.py:304-340 concatenates q,k,v into one WQLinear; inference/modules/fused_mlp.py:44-114
 * issues gate and up back to back).  x [IC] fp16; for i < nmat: qweight[i] / scales[i] / qzeros[i] as produced by
 * bdq_pack_wq for a [OC[i], IC] weight, out[i] [OC[i]] fp16.  The pointer arrays are HOST arrays of device pointers.
 * group 128, OC[i] % 16 == 0 take the fused kernel; anything else falls back to one bdq_gemv_w2 launch per matrix.
 * Prefetch hint (optional, n_prefetch 0..4): prefetch[i] / prefetch_bytes[i] name device buffers -- normally the
 * qweight tensors of the NEXT launch of the decode chain -- that this launch pulls into L2 with
 * cp.async.bulk.prefetch.L2 while it computes, so HBM keeps streaming across the launch boundary.  Purely a hint:
 * results do not depend on it.
 */
int bdq_gemv_w2_multi(const void* x, int nmat, const int32_t* const* qweight, const void* const* scales,
                      const int32_t* const* qzeros, void* const* out, const int64_t* OC, int64_t IC, int group, int flags,
                      int n_prefetch, const void* const* prefetch, const int64_t* prefetch_bytes, bdq_stream_t stream);

/*
 * Whole-token decode chain in ONE persistent launch.  A chain is an ordered list of phases; a phase is what one
 * bdq_gemv_w2_multi call computes (up to 4 packed matrices sharing one activation vector x [IC] fp16, group 128,
 * OC[i] % 16 == 0; BDQ_GEMV_SILU_MUL in `flags` as above).  Running the chain is equivalent -- bit for bit -- to issuing
 * the phases as consecutive bdq_gemv_w2_multi launches on one stream: phase p reads x after every earlier phase has
 * stored its outputs (so x may be the out[] of an earlier phase), unless the phase carries BDQ_GEMV_CHAIN_NOWAIT (its x
 * is an external input that is ready at launch).  This replaces the per-linear gemv_forward_cuda launches of a decoded
 * token (quantization/qmodule.py:166; inference/modules/fused_mlp.py:44-114, fused_attn.py:304-340 ->
 * inference/kernels/csrc/quantization/gemv_cuda.cu:289-346): 2 CTAs per SM stay resident, a grid barrier on an L2 counter
 * replaces the launch boundary between dependent phases, and the weight loads of the following phases are already in
 * flight while a phase is reduced, stored and signalled (the weights do not depend on x).
 *   bdq_gemv_w2_chain_max_phases(max_IC)  phases one run can hold (the plan is kept in shared memory); split longer chains
 *   bdq_gemv_w2_chain_workspace_bytes(n)  device bytes the caller must provide (256-byte aligned)
 *   bdq_gemv_w2_chain_prepare(...)        validates the phases and writes the device-side plan (synchronous, once per model)
 *   bdq_gemv_w2_chain_run(...)            one cooperative launch; max_IC = the largest IC of the plan; `lookahead` is reserved
 * The workspace also holds the barrier counters: two runs of the same workspace must not overlap (one stream).
 */
#define BDQ_GEMV_CHAIN_NOWAIT 256
typedef struct bdq_gemv_phase {
    const void* x;
    int nmat;
    int group;
    int flags;
    int reserved;
    int64_t IC;
    const int32_t* qweight[4];
    const void* scales[4];
    const int32_t* qzeros[4];
    void* out[4];
    int64_t OC[4];
} bdq_gemv_phase_t;
int bdq_gemv_w2_chain_max_phases(int64_t max_IC);
int64_t bdq_gemv_w2_chain_workspace_bytes(int n_phases);
int bdq_gemv_w2_chain_prepare(const bdq_gemv_phase_t* phases, int n_phases, void* workspace, int64_t workspace_bytes,
                              bdq_stream_t stream);
int bdq_gemv_w2_chain_run(const void* workspace, int n_phases, int64_t max_IC, int lookahead, bdq_stream_t stream);
int bdq_gemv_w2_chain_set_prof(void* device_buf_u64, int64_t capacity_u64);   /* dev tool: per-phase %globaltimer marks */

/*
 * bdq_gemv_w2_chain_static: the decode chain for callers that cannot allocate device memory (code under CUDA graph
 * capture, tight serving loops, harnesses that forbid cudaMalloc).  It has NO counterpart in the reference.
 * Same kernel, same arithmetic and bit-equal outputs as _prepare + _run, but the
 * kernel's scratch (barrier counters + device-side plan) lives in a file-scope `static __device__ __align__(256)` array
 * inside the library (16.5 KB, held by every GPU that loads the library until exit) instead of a caller workspace.
 * The call allocates nothing -- no cudaMalloc / cudaFree, no host malloc: it validates the phases into a stack array,
 * clears the counters (cudaMemsetAsync), stages the plan (cudaMemcpyAsync) and launches, all on `stream`.
 *   n_phases > BDQ_GEMV_CHAIN_STATIC_MAX_PHASES  -> BDQ_ERR_SHAPE with a message (split the chain or use _prepare / _run)
 *   plan does not fit the kernel's shared memory -> BDQ_ERR_UNSUPPORTED (same limit as bdq_gemv_w2_chain_max_phases)
 * Every call on a device shares the one array, which keeps the previous call's contents (the counters are cleared at
 * every call): calls in one stream are ordered and safe, calls that OVERLAP on the GPU (different streams) race.
 * The caller still passes all inputs and outputs (x, packed weights, out[]) inside `phases`.
 */
#define BDQ_GEMV_CHAIN_STATIC_MAX_PHASES 64
int bdq_gemv_w2_chain_static(const bdq_gemv_phase_t* phases, int n_phases, int lookahead, bdq_stream_t stream);

/*
 * Packed NF3 (3 bits per weight + two fp16 scales per group).  The reference has no packed NF3 format or kernel
 * (quantization/qmodule.py:77 accepts w_bit 2 and 4 only), so the layout is this library's own; its contract is that
 * bdq_nf3_dequant(pack(w)) equals SteN2F3Quantizer.forward(w) / pseudo_quantize_n2f3_tensor(w)
 * (quantization/quantizer.py:113-116,184-242) bit for bit in fp16.
 *   level index e = position in the ascending staircase -1, -0.535, -0.247, 0, 0.183, 0.382, 0.623, 1
 *   qweight int32 [rows, cols/16]: bits [2i, 2i+2) of word c = low two bits of e of column 16c + i
 *   qhigh   int32 [rows, cols/32]: bit 4n + 2w + p of word c = high bit of e of column 32c + 16w + 2n + p (w, p in 0..1,
 *           n in 0..7): ((qweight_word >> 2p) & 0x33333333) | ((qhigh_word << (2 - 2w - p)) & 0x44444444) is then eight
 *           4-bit level indices = two PRMT selectors (the GEMV looks levels up with PRMT, no per-code extraction)
 *   scales  fp16  [rows, 2*cols/group]: (|group max|, |group min|) per group; e < 3 uses the second one
 * bdq_pack_nf3 takes the uint8 codes written by bdq_nf3_fakequant_fwd (cols % 32 == 0, 16-byte aligned).
 * bdq_gemv_nf3: out[b, oc] = fp16(sum_k x[b,k] * level * scale), x [B, IC] fp16.  group % 32 == 0, IC % group == 0.
 * group 128 and OC % 16 == 0 (32-byte aligned qweight, 16-byte aligned qhigh) take the streaming kernel: PRMT level lookup
 * into 16-bit fixed-point tables, INT8 tensor-pipe MMAs against block-fixed-point x (value error 2^-16 of the group's
 * scale / largest activation, below the fp16 rounding of the dequantised weights); other shapes a CUDA-core pair lookup.
 */
int bdq_pack_nf3(const uint8_t* codes, int32_t* qweight, int32_t* qhigh, int64_t rows, int64_t cols, bdq_stream_t stream);
int bdq_nf3_dequant(const int32_t* qweight, const int32_t* qhigh, const void* scales, void* out_f16, int64_t rows, int64_t cols,
                    int group, bdq_stream_t stream);
int bdq_gemv_nf3(const void* x, const int32_t* qweight, const int32_t* qhigh, const void* scales, void* out, int64_t B,
                 int64_t IC, int64_t OC, int group, bdq_stream_t stream);

/*
 * GPTQ-v2 2-bit export (train/convert_hf_to_gptq_manual.py).
 * bdq_gptq_export_w2 = get_hf_quant_params (lines 11-75) + convert_layer_to_gptq_format (85-157) in one pass:
 *   w [OC, IC] (T) is quantised with the converter's fp32 host arithmetic
 *   (s = clamp(max - min, 1e-5) / 3, z = clamp(-round(min / s), 0, 3), code = clamp(round(w / s) + z, 0, 3)) and
 *   qweight int32 [IC/16, OC]: bits [2i, 2i+2) of qweight[r, n] = code[n, 16r + i]            (lines 113-115)
 *   qzeros  int32 [IC/G, OC/16]: bits [2i, 2i+2) of qzeros[g, c] = z[16c + i, g], NOT minus one (v2, 127-130)
 *   scales  bf16  [IC/G, OC]: s rounded to T and then to bf16                                  (lines 64, 100)
 * g_idx[k] = k / G is an arange the caller builds (lines 143-148).  w_bit must be 2 (line 89); group a power of two in
 * [16, 512] with IC % group == 0; OC % 16 == 0 (line 123).  NaN weights are not supported (the reference asserts).
 * bdq_gptq_pack_w2 is the second half alone, for callers that already hold unpacked parameters:
 *   codes uint8 [OC, IC], zeros uint8 [OC, IC/G] -> qweight, qzeros as above.
 */
int bdq_gptq_export_w2(const void* w, int32_t* qweight, int32_t* qzeros, void* scales_bf16, int64_t OC, int64_t IC,
                       int group, int w_bit, int dtype, bdq_stream_t stream);
int bdq_gptq_pack_w2(const uint8_t* codes, const uint8_t* zeros, int32_t* qweight, int32_t* qzeros, int64_t OC,
                     int64_t IC, int group, int w_bit, bdq_stream_t stream);

/*
 * Packed 2-bit prefill GEMM: out[m, oc] = fp16( sum_k x[m,k] * fp16(scales[oc,k/G] * (code[oc,k] - zero[oc,k/G])) ),
 * fp32 accumulation in tensor memory (tcgen05.mma, TMA-fed activations, in-register dequant of the packed weights).
 * Replaces quant_matmul_v2 (quantization/triton_kernels.py:102-134) and gemm_forward_cuda (pybind.cpp:12; the
 * reference's CUDA GEMM is 4-bit only).  Operands are the row-major tensors WQLinear holds (qmodule.py:62-66) -- NOT the
 * transposed copies the reference builds per call (qmodule.py:154-158).  x [M, IC] fp16, out [M, OC] fp16.
 * Tensor-core path: group in {64, 128}, IC % 128 == 0, OC % 32 == 0 (tiles are 256 channels wide, the last one may be
 * partial); other shapes fall back to bdq_gemv_w2 row blocks.
 */
int bdq_gemm_w2(const void* x, const int32_t* qweight, const void* scales, const int32_t* qzeros, void* out, int64_t M,
                int64_t IC, int64_t OC, int group, bdq_stream_t stream);
/* Plain fp16 GEMM on the same tcgen05 kernel (both operands by TMA): out [M, OC] = x [M, IC] @ w [OC, IC]^T, fp32 accumulate.
 * IC % 64 == 0, OC % 32 == 0.  Used by WQLinearNF3's prefill after bdq_nf3_dequant (no cuBLAS on that path). */
int bdq_gemm_f16(const void* x, const void* w, void* out, int64_t M, int64_t IC, int64_t OC, bdq_stream_t stream);

/*
 * Tensor-parallel variants with the all-gather of the channel slices FUSED into the kernel epilogue (the
 * reference has no tensor parallelism: these are the column-parallel shards of WQLinear, qmodule.py:62-66).  Every rank
 * holds rows [r*OC/n, (r+1)*OC/n) of qweight / scales / qzeros, computes its slice and stores it into its own output
 * AND, with plain stores over NVLink peer memory, into the same position of every peer's output.  `out` and
 * `peer_out[...]` already point at this rank's first output channel inside the full-width rows; the peer pointers are
 * peer-mapped device addresses (e.g. torch symmetric memory `buffer_ptrs`).  The caller puts a cross-rank barrier between
 * these launches and the consumers of the gathered rows.
 *   bdq_gemv_w2_tp: batch-1 decode, up to 4 matrices sharing x (peer_out is [nmat][n_peer], row-major), group 128.
 *   bdq_gemm_w2_tp: prefill, output row pitch `ldo` elements (the full OC), tcgen05 shapes only (OC_local % 32 == 0).
 */
int bdq_gemv_w2_tp(const void* x, int nmat, const int32_t* const* qweight, const void* const* scales,
                   const int32_t* const* qzeros, void* const* out, const int64_t* OC_local, int64_t IC, int group, int flags,
                   int n_peer, void* const* peer_out, bdq_stream_t stream);
int bdq_gemm_w2_tp(const void* x, const int32_t* qweight, const void* scales, const int32_t* qzeros, void* out, int64_t M,
                   int64_t IC, int64_t OC_local, int group, int64_t ldo, int n_peer, void* const* peer_out, bdq_stream_t stream);

/*
 * Quantise-before-all-gather for dim-0 (FSDP2-style) shards of a QLinear weight.  The reference
 * gathers the bf16 shards and fake-quantises the full weight on every rank (QLinear.forward, quantization/qlinear.py:63-78,
 * with SteInt2AsymQuantizer, quantizer.py:153-182); groups lie along the last dimension, so the owner of a shard can
 * quantise it once and ship 2-bit codes + one (scale, zero) pair per group: 0.28 B/weight instead of 2.
 *   bdq_int2_shard_encode: w [n_elem] (bf16 / fp16) -> codes [n_elem/4 bytes], sz [n_elem/group] pairs of T (scale, zero).
 *     The payload is stored to `codes`/`sz` AND to the n_peer (<= 7) peer-mapped destinations peer_codes[i]/peer_sz[i]
 *     (NVLink addresses of this rank's slot in every peer's gather buffer, e.g. torch symmetric memory): the all-gather is
 *     the kernel's store phase.  The caller puts a cross-rank barrier between the encode launches and the decode.
 *     round_mode: BDQ_ROUND_HALF_AWAY = SteInt2AsymQuantizer, BDQ_ROUND_HALF_EVEN = pseudo_quantize_tensor.
 *   bdq_int2_shard_decode: dq = rT((code - zero) * scale), the last two operations of quantizer.py:177 -- bit-equal to
 *     bdq_int_fakequant_fwd(bits 2) of the same weights.
 * Code layout: one uint2 per 32 consecutive elements, one 32-bit word per 16; element j of the 16 sits at bits
 * [16*(j%2) + 2*(j/2), +2).  group in {32, 64, 128}; w and dq 32-byte aligned, codes 8-byte, sz 4-byte.
 */
int bdq_int2_shard_encode(const void* w, void* codes, void* sz, int64_t n_elem, int group, int dtype, int round_mode,
                          int sem, int n_peer, void* const* peer_codes, void* const* peer_sz, bdq_stream_t stream);
int bdq_int2_shard_decode(const void* codes, const void* sz, void* dq, int64_t n_elem, int group, int dtype,
                          bdq_stream_t stream);
/*
 * Up to 8 shards (the tensors of one FSDP2 parameter group, e.g. a decoder layer) in ONE encode launch: segment i =
 * w[i] with n_elem[i] elements, its payload goes to byte offsets code_off[i] / sz_off[i] of `codes` / `sz` and of every
 * peer destination.  With the group's payload laid out as the concatenation of the tensors' payloads, one
 * bdq_int2_shard_decode over the whole buffer rebuilds all tensors (sharded.ShardGatherGroup).
 */
int bdq_int2_shard_encode_multi(int nseg, const void* const* w, const int64_t* n_elem, const int64_t* code_off,
                                const int64_t* sz_off, void* codes, void* sz, int group, int dtype, int round_mode, int sem,
                                int n_peer, void* const* peer_codes, void* const* peer_sz, bdq_stream_t stream);

/*
 * Two-sided clip-ratio grid search = auto_2clip_layer (quantization/autoclip.py:13-83), all candidates of all
 * (output channel, group) pairs in one launch.  w [co, ci] (T, row-major); x = the calibration activations
 * [n_tok, ci] of the same dtype with a row stride of x_row_stride elements (the reference's token sub-sampling
 * x[:, 0::n_tok//512], autoclip.py:27, is a stride here, not a copy).  For every (o, g) and every candidate
 * (i_p, i_n), i_* in [0, int(max_shrink*n_grid)):  max' = max*(1 - i_p/n_grid), min' = min*(1 - i_n/n_grid),
 * err = mean_t (sum_k x[t,g,k] * (Q(clamp(w[o,g,:], min', max'))[k] - w[o,g,k]))^2 with Q = pseudo_quantize_tensor
 * (quant_type BDQ_QUANT_INT, `bits`) or pseudo_quantize_n2f3_tensor (BDQ_QUANT_NF3); the first candidate with the
 * smallest error in (i_p outer, i_n inner) order wins (strict '<' update from 1e9, autoclip.py:45,69-72).
 * Candidate weights carry the exact T roundings of the reference; the contraction over k runs on the tensor cores
 * with fp32 accumulation (the reference contracts in T).  Outputs: best_max, best_min [co * ci/group] (T) = the
 * reference's return values flattened; optional best_idx (i_p*nc + i_n, or -1 if no candidate improved on 1e9 / all NaN)
 * and best_err (fp32 mean squared error of the winner).  group in {32, 64, 128}; nc*nc <= 111.
 */
int bdq_autoclip(const void* w, const void* x, void* best_max, void* best_min, int32_t* best_idx, float* best_err,
                 int64_t co, int64_t ci, int64_t n_tok, int64_t x_row_stride, int group, int bits, int quant_type,
                 int n_grid, double max_shrink, int dtype, int sem, bdq_stream_t stream);
int bdq_autoclip_tc_set_prof(void* device_buf_256_int64);   /* dev tool: per-role wait cycles of the tcgen05 kernel, CTA (0, 0) */

#ifdef __cplusplus
}
#endif
#endif /* BDQ_H_ */
