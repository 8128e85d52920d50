/*
 * b200q.h — C-ABI of the B200-native (sm_100a) group-quantized dequantize+matmul path.
 *
 * This is the drop-in boundary for ONE hot path of mlc-llm (reference mounted at /root/reference;
 * all file:line citations below are relative to it): the kernels TVM generates for
 *
 *   fused_decode*_matmul*        decode-step dequant-GEMV      (one PrimFunc per Linear)
 *   decode*                      prefill dequantize-to-dense   (a6)
 *   NT_matmul* / fused_NT_matmul*  prefill dense GEMM           (a7)
 *   fused_decode*_NT_matmul*     prefill dequant fused into the GEMM (a8)
 *   take_decode                  quantized embedding gather
 *
 * Tensor layouts are the reference's, unchanged (mlc_llm/transform/quantization.py:142-177):
 *   packed  uint32 [n_words(K), N]   row-major, N contiguous; value k of column n lives in word
 *                                    k / npu, bits [(k % npu) * nbit, +nbit);  npu = 32 / nbit
 *   scale   dtype  [n_groups(K), N]  one scale per `group` consecutive k of a column
 *   n_groups = ceil(K / group),  n_words = ceil(group / npu) * n_groups
 *   w[k, n] = (dtype(q) - (2^(nbit-1) - 1)) * scale[k / group, n]          (quantization.py:213-230)
 *
 * Conventions (mirroring the PrimFunc destination-passing signature,
 * mlc_llm/dispatch/dispatch_tir_operator_adreno.py:6,128):
 *   - argument order (packed, scale, x[, epilogue operands], out); outputs never alias inputs
 *   - every pointer is a DEVICE pointer unless the function name ends in _host
 *   - the callee allocates nothing, never synchronizes, launches only on `stream`
 *     (a cudaStream_t passed as void*), and is CUDA-graph capturable
 *   - return 0 on success, a negative b200q_status otherwise; b200q_last_error() returns the
 *     calling thread's last message (TVM's "non-zero return + TVMGetLastError" convention)
 *   - there is NO CPU fallback: without a CUDA device every compute entry point fails.
 */
#ifndef B200Q_H_
#define B200Q_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define B200Q_VERSION 109 /* major*100 + minor */

#if defined(__GNUC__)
#define B200Q_API __attribute__((visibility("default")))
#else
#define B200Q_API
#endif

/* Quantization formats ("_0" = symmetric, scale only, transposed (K/.., N) layout, uint32 storage).
 * Reference flags: build.py:19-25,114; utils.py:31-36 (see SURVEY.md section 0 for the name mapping). */
typedef enum {
  B200Q_Q4F16_0 = 0, /* int4, group 32, f16 scale/activation/output            (offset 7)   */
  B200Q_Q3F16_0 = 1, /* int3, group 40, 10 values per word (2 dead bits), f16  (offset 3)   */
  B200Q_Q4F32_0 = 2, /* int4, group 32, f32 scale/activation/output            (offset 7)   */
  B200Q_Q8F16_0 = 3, /* int8, group 32, 4 values per word, f16                 (offset 127) */
  /* Asymmetric int4 - the reference CLI's default (--quantization-sym off, build.py:19; quantization.py:105-140,
   * 183-211):  w = dtype(q) * scale + min, product and sum each rounded to dtype.  Only the *_asym entry points
   * take these (they carry the extra `zmin` tensor). */
  B200Q_Q4F16_ASYM = 4, /* f16: scale f16[G, N] and min f16[G, N] are separate tensors                      */
  B200Q_Q4F32_ASYM = 5, /* f32: `scale` is uint32[G, N] = bf16(scale) | bf16(min) << 16, zmin is NULL        */
  /* The remaining --quantization-* combinations (symmetric; encode / dequant / take_decode / gemv entry points only;
   * streaming kernels, not the headline path): */
  B200Q_Q3F16_U16 = 6, /* int3 in uint16 words: packed uint16[8 * n_groups, N], 5 values per word, group 40
                          (--quantization-storage-nbit 16, the iOS build; dispatch/llama/main.py:4151-4203, 5815-6363) */
  B200Q_FP4F16_0 = 7,  /* fp4 = sign + 3-bit exponent code (quantization.py:51-101), uint32 words, group 32, f16   */
  B200Q_FP4F32_0 = 8,  /* the same with f32 scale / activations                                                   */
  B200Q_NUM_FMT = 9
} b200q_fmt;

/* Epilogues fused by FuseDecodeMatmulEwise (mlc_llm/transform/decode_matmul_ewise.py:49-75) and by
 * the default Relax FuseOps pipeline for the prefill GEMMs.  `y` is the matmul result rounded to
 * the model dtype; each epilogue op is evaluated in the model dtype, op for op like the PrimFunc. */
typedef enum {
  B200Q_EP_NONE = 0,      /* out = y                                   adreno.py:5-23        */
  B200Q_EP_ADD = 1,       /* out = ep0[n] + y        (residual)        adreno.py:86-91       */
  B200Q_EP_MULTIPLY = 2,  /* out = ep0[n] * y        (SwiGLU up*act)   adreno.py:150-152     */
  B200Q_EP_SILU = 3,      /* out = y * (1 / (1 + exp(-y)))             adreno.py:216-228     */
  B200Q_EP_CAST_F32 = 4,  /* out(f32) = y, kept at f32 accumulate precision  adreno.py:364-369 */
  B200Q_EP_BIAS = 5,      /* out = y + ep0[n]        (ep0 = bias[N])   gpt_neox/dolly_v2_3b_mod.py:41-46 */
  B200Q_EP_BIAS_GELU = 6, /* v=y+b; out = v*(0.5 + erf(v*0.70710678)*0.5)   dolly_v2_3b_mod.py:204-251 */
  B200Q_EP_BIAS_ADD2 = 7, /* out = ((y + bias) + ep1) + ep2  (GEMM only)    dolly_v2_3b_mod.py:88-110 */
  B200Q_NUM_EP = 8
} b200q_epilogue;

typedef enum {
  B200Q_OK = 0,
  B200Q_ERR_BAD_SHAPE = -1,   /* K/N not supported by the layout rules below              */
  B200Q_ERR_BAD_ALIGN = -2,   /* a pointer is not 16-byte aligned                          */
  B200Q_ERR_UNSUPPORTED = -3, /* unknown fmt / epilogue / combination                      */
  B200Q_ERR_CUDA = -4,        /* a CUDA runtime call failed (message has cudaGetErrorString) */
  B200Q_ERR_NULL = -5         /* a required pointer is NULL                                 */
} b200q_status;

/* flags for the launch */
#define B200Q_FLAG_PDL 1u /* launch with programmatic dependent launch: the kernel prefetches its
                             (constant) packed weights before waiting on the previous kernel in the
                             stream.  Only legal when no earlier kernel in flight WRITES packed/scale. */

/* b200q_gemm tile selection (default: chosen per shape): the single-CTA 128 x 256 kernel, or the CTA-pair
 * (cta_group::2, M = 256) kernel with 256 x 128 / 256 x 256 tiles.  For tests and experiments. */
#define B200Q_FLAG_GEMM_SINGLE 2u
#define B200Q_FLAG_GEMM_PAIR128 4u
#define B200Q_FLAG_GEMM_PAIR256 8u

/* ---- introspection ------------------------------------------------------------------------- */
B200Q_API int b200q_version(void);
B200Q_API const char* b200q_last_error(void);
/* number of CUDA kernels this library has launched from the calling process (for bench.py's
 * gpu_launches claim) */
B200Q_API unsigned long long b200q_launch_count(void);
/* layout helpers; return <0 on a bad format */
B200Q_API int b200q_n_groups(int fmt, int K);
B200Q_API int b200q_n_words(int fmt, int K);
/* algorithmic HBM bytes of one GEMV: packed + scale + x + out (+ epilogue operands) */
B200Q_API long long b200q_gemv_bytes(int fmt, int K, int N, int epilogue);

/* ---- a5: decode-step fused dequant-GEMV ------------------------------------------------------
 * Replaces fused_decode*_matmul* (dispatch_tir_operator_adreno.py:5-411, dispatch/llama/main.py:
 * 4235-6363, tuning.py:39-108).
 *   y[n] = sum_{k<K} x[k] * w[k, n],  w bit-identical to the reference decode, products exact,
 *   accumulation in f32 (the reference accumulates in dtype), then the epilogue.
 *   x   dtype[K]         out  dtype[N]  (float[N] for B200Q_EP_CAST_F32 and for q4f32_0)
 *   ep0 dtype[N] or NULL (ADD / MULTIPLY / BIAS / BIAS_GELU)
 * Requirements: N % 8 == 0; packed, scale 16-byte aligned; x, ep0, out aligned to their element.
 *
 * workspace: optional device scratch (b200q_gemv_workspace_bytes() bytes, 16-byte aligned) that lets
 * the kernel split a column panel's K range over several CTAs.  The caller zero-fills it ONCE
 * (cudaMemset) before its first use; every launch leaves its counter area zeroed again.  It must not be shared
 * by launches that may run concurrently (one per stream).  NULL: whole panels are dealt to CTAs
 * (correct, slower for small N).  The callee itself never allocates (SURVEY.md 8b "Ownership"). */
B200Q_API size_t b200q_gemv_workspace_bytes(void);
B200Q_API int b200q_gemv(const void* packed, const void* scale, const void* x, const void* ep0, void* out,
               int K, int N, int fmt, int epilogue, unsigned flags, void* workspace,
               size_t workspace_bytes, void* stream);

/* b200q_gemv for callers that cannot allocate device memory (code under CUDA graph capture, tight serving loops,
 * harnesses that forbid cudaMalloc).  This entry point has NO counterpart in the reference.
 * Same kernel, arithmetic and results as b200q_gemv called with a zero-filled
 * workspace of b200q_gemv_workspace_bytes() bytes, but the split-K scratch (panel counters + partial sums) is a
 * static device array inside the library: B200Q_GEMV_STATIC_SCRATCH_BYTES per GPU that loads the library, held until
 * the process exits.  The call allocates nothing (no cudaMalloc / cudaFree, no allocating library routine); the
 * caller still passes inputs and outputs.
 * Limits: symmetric formats (q4f16_0, q3f16_0, q4f32_0, q8f16_0); N <= B200Q_GEMV_STATIC_MAX_N; a shape whose split
 * needs more scratch than the array holds is refused with B200Q_ERR_BAD_SHAPE (never a silent fall-back).
 * Sharing: every call on a GPU uses the same array, which keeps the last call's contents; the entry point clears
 * the counters it needs zeroed (a cudaMemsetAsync on `stream`) before the kernel.  Calls that may overlap on the
 * GPU (two streams, two host threads) RACE on the array: serialize them on one stream, or use b200q_gemv with a
 * workspace per stream. */
#define B200Q_GEMV_STATIC_SCRATCH_BYTES (4096u * 4u + 4u * 1024u * 1024u)
#define B200Q_GEMV_STATIC_MAX_N (4096 * 128)
B200Q_API int b200q_gemv_static(const void* packed, const void* scale, const void* x, const void* ep0, void* out,
                      int K, int N, int fmt, int epilogue, unsigned flags, void* stream);

/* Several GEMVs that read the SAME activation vector in ONE launch (q|k|v, gate|up: the reference
 * launches them back to back on the same x, relax_model/llama.py:233-236,147-152).  Tensors stay
 * the reference's separate per-Linear arrays.
 * x_mul (optional, dtype[K]): the activation is x[k] * x_mul[k], rounded to dtype - the SwiGLU
 * product that the reference forms with fused_decode5_fused_matmul6_multiply1's epilogue
 * (dispatch_tir_operator_adreno.py:150-152) before down_proj reads it; same single rounding. */
typedef struct {
  const void* packed;
  const void* scale;
  const void* ep0;
  void* out;
  int N;
  int epilogue;
} b200q_gemv_op;
B200Q_API int b200q_gemv_multi(const b200q_gemv_op* ops, int n_ops, const void* x, const void* x_mul, int K,
                     int fmt, unsigned flags, void* workspace, size_t workspace_bytes, void* stream);

/* ---- the whole decode step in ONE persistent launch ----------------------------------------------
 * The reference runs the GEMV chain of one token as one compiled Relax function (`decoding`,
 * mlc_llm/relax_model/llama.py:581-611) whose call_tir sites are the fused_decode*_matmul* PrimFuncs.
 * A "program" is that chain as data: a list of steps, each step = 1..4 GEMVs reading the same
 * activation (x, optionally x * x_mul) exactly like b200q_gemv_multi.  The library derives the
 * dependencies between steps from pointer overlap (a step waits for the steps that write its x /
 * x_mul / ep0, or whose inputs / outputs it overwrites) and runs everything in one grid-resident (one CTA per SM),
 * persistent kernel: weights of later steps stream into shared memory while earlier steps finish.
 * Arithmetic: q3f16_0 / q4f32_0 programs (and q4f16_0 / q8f16_0 with B200Q_PROG_TC=0 / B200Q_PROG_TC8=0 in the environment) use
 * the per-weight arithmetic of b200q_gemv_multi (every weight rounded to dtype exactly like the reference's
 * decode, exact products, f32 accumulation; results differ from the step-by-step launches only by the f32
 * summation order of the K split).  The DEFAULT q4f16_0 and q8f16_0 programs run on tcgen05: they accumulate
 * T[g,n] = sum_k (off_k + q[k,n]) x[k] per quantization group in f32 and apply s[g,n] * (T - C[g]) once per
 * group (C[g] = sum_k (off_k + maxint) x[k], rounded to f32 separately), i.e. the group scale multiplies an f32 group
 * sum instead of every weight being rounded to f16 first.  It is therefore NOT bit-equal to b200q_gemv_multi or
 * to the reference; it is closer to real arithmetic than both (tests/test_program_parity_gpu.py: against the f64
 * sum of the UNROUNDED (q-7)*s it meets the 2^-14 bound that the per-weight kernels meet against the rounded
 * weights), and the T - C cancellation costs accuracy only when |sum_k x[k]| of a group is large against
 * |sum_k (q-7) x[k]| (bounded and tested with shifted / outlier activations).  B200Q_PROG_I8=1 selects an integer
 * formulation for q4f16_0 (tcgen05.mma.kind::i8 on three signed base-256 digits of the group-scaled activations:
 * exact s32 group sums, no cancellation; an activation more than 2^10 below its group's largest loses low bits, a
 * non-finite activation yields NaN for the outputs that depend on its group).  Every program is deterministic:
 * replays give identical bits for a given SM count.
 * Launch: the tcgen05 programs use programmatic dependent launch (B200Q_PROG_PDL): consecutive program launches on
 * a stream overlap their set-up with the predecessor's tail; kernels of other kinds before / after are ordered as
 * usual.  As with B200Q_FLAG_PDL, no kernel in flight may WRITE the packed words or scales a program reads.
 * These launches do not carry the cooperative attribute (it serialises consecutive launches: 689 vs 803 tok/s
 * for a program cut into 97 launches), so two programs launched CONCURRENTLY on different streams of one device
 * could each hold part of the SMs and wait for the rest: keep the program launches of a device on one stream, or
 * set B200Q_PROG_PDL=0 (cooperative launches: the driver then never co-schedules two of them).  The watchdog
 * (b200q_gemv_program_set_watchdog) turns such a wait into an error instead of a hang.
 *
 *   1. b200q_gemv_program_build fills a HOST blob of b200q_gemv_program_bytes(n_steps) bytes
 *      (tensor-map descriptors + step table; needs a CUDA context but launches nothing);
 *   2. the caller copies the blob to a 64-byte aligned DEVICE buffer and zero-fills a 256-byte aligned
 *      device workspace of b200q_gemv_program_workspace_bytes(host_blob) bytes ONCE;
 *   3. b200q_gemv_program_run launches it (CUDA-graph capturable; leaves the workspace clean).
 * All buffers named by the steps must stay alive and in place while the program is in use. */
typedef struct {
  b200q_gemv_op ops[4];
  int n_ops;
  const void* x;
  const void* x_mul; /* optional */
  int K;
} b200q_gemv_step;
B200Q_API size_t b200q_gemv_program_bytes(int n_steps);
B200Q_API int b200q_gemv_program_build(const b200q_gemv_step* steps, int n_steps, int fmt, void* host_blob,
                             size_t blob_bytes);
B200Q_API size_t b200q_gemv_program_workspace_bytes(const void* host_blob);
B200Q_API int b200q_gemv_program_run(const void* host_blob, const void* dev_blob, void* workspace,
                           size_t workspace_bytes, void* stream);

/* Tensor-parallel program (q4f16_0): the reference has no tensor parallelism; the sharding is SURVEY.md 8e
 * (q/k/v/gate/up column-sharded, o/down row-sharded, one reduction per sharded block).  reduce_mask[j] bit i
 * marks op i of step j as ROW-PARALLEL: its matmul result is this rank's partial sum.  The reduction over the
 * ranks happens inside the launch, in ONE NVLink hop: every rank pushes its f32 partial into the slot
 * [panel][its rank] of EVERY peer's workspace (peer-mapped symmetric memory), then adds the slots of its own
 * workspace in rank order - the same order everywhere, so every rank holds the same bits -, applies the epilogue
 * and publishes the finished activation locally, where the dependent steps pick it up: the all-reduce of the
 * sharded block without a collective kernel or a launch boundary.  Requirements: every rank builds the same step
 * list (its own shards), zero-fills its workspace once, launches the same number of times, and passes the
 * peer-mapped workspace pointers of ALL ranks (peer_workspaces[tp_rank] == workspace).  The plain `out` of a
 * row-parallel op is complete on every rank.  A peer that never launches is caught by the watchdog below. */
B200Q_API int b200q_gemv_program_build_tp(const b200q_gemv_step* steps, int n_steps, int fmt,
                                const unsigned char* reduce_mask, int tp_rank, int tp_size, void* host_blob,
                                size_t blob_bytes);
B200Q_API int b200q_gemv_program_run_tp(const void* host_blob, const void* dev_blob, void* workspace,
                              size_t workspace_bytes, void* const* peer_workspaces, void* stream);

/* Watchdog of the program kernels.  Everything that flows between CTAs (and, for a tensor-parallel program,
 * between GPUs) inside a program launch is polled in place; a poll that stays unsatisfied - a peer rank that
 * never launched or replayed a different number of times, a workspace shared by two programs - would otherwise
 * spin forever.  Every such wait is bounded: after timeout_ns the kernel writes
 *     status = 0xDEAD | site << 16 | (uint64)step << 32      (site: 1 activation, 2 K-split partial sum,
 *                                                             3 epilogue operand, 4 peer GPU partial sum,
 *                                                             5 workspace belongs to another program)
 * to *status_word (8 bytes; pass PINNED HOST memory mapped into the device so that it survives the context)
 * and traps: the launch fails with B200Q_ERR_CUDA at the next synchronization instead of hanging.
 * timeout_ns < 0: the default (environment B200Q_WATCHDOG_MS, else 10 s); 0: unbounded.  status_word may be NULL.
 * The bound is kept as a poll count (timeout / 0.35 us, the minimum round trip of one poll): a wait expires no
 * earlier than timeout_ns and at most a few times later.
 * Process-wide; a launch (or a CUDA graph capture of it) takes the settings in force when it is issued. */
B200Q_API void b200q_gemv_program_set_watchdog(void* status_word, long long timeout_ns);

/* Tuning override for experiments (0 = default).  ctas_per_sm: 1 or 2 persistent CTAs per SM;
 * stages: cap of the shared-memory ring depth; dbg_flags bit 0: stream only, skip the math
 * (results are garbage - for bandwidth probes). Thread-local. */
B200Q_API void b200q_gemv_tune(int ctas_per_sm, int stages, int dbg_flags);

/* Same GEMV with HOST activation / output buffers (weights stay resident in HBM as in the
 * reference, cpp/llm_chat.cc:858-861): copies x host->device, runs, copies out device->host and
 * synchronizes `stream`.  x_dev / out_dev are caller-provided device scratch of K / N elements. */
B200Q_API int b200q_gemv_host(const void* packed, const void* scale, const void* x_host, const void* ep0,
                    void* out_host, void* x_dev, void* out_dev, int K, int N, int fmt, int epilogue,
                    void* workspace, size_t workspace_bytes, void* stream);

/* ---- a2 / a6: dequantize ---------------------------------------------------------------------
 * Replaces decode* (dispatch/llama/main.py:3984-4232, tuning.py:111-147, quantization.py:213-230).
 *   transposed == 0: out dtype[K, N]      (decode-step shape)
 *   transposed != 0: out dtype[N, K]      (prefill shape, feeds NT_matmul)
 * Bit-exact against the reference decode. */
B200Q_API int b200q_dequant(const void* packed, const void* scale, void* out, int K, int N, int fmt,
                  int transposed, void* stream);

/* Quantized embedding gather, row-major (V, n_words) layout (quantization.py:235-270,397-420).
 *   packed_rm uint32[V, n_words(K)], scale_rm dtype[V, n_groups(K)], ids int32[n_ids]
 *   out dtype[n_ids, K] */
B200Q_API int b200q_take_decode(const void* packed_rm, const void* scale_rm, const int32_t* ids, void* out,
                      int n_ids, int V, int K, int fmt, void* stream);

/* ---- a1: encode ------------------------------------------------------------------------------
 * Group-quantize W dtype[N, K] (te_encode_sym, quantization.py:142-177; the reference runs it once
 * on the host through LLVM, utils.py:98-110).  Bit-exact against the reference encoder.
 *   transpose != 0: packed uint32[n_words(K), N], scale dtype[n_groups(K), N]   (matmul weights)
 *   transpose == 0: packed uint32[N, n_words(K)], scale dtype[N, n_groups(K)]   (embedding table) */
B200Q_API int b200q_encode(const void* W, void* packed, void* scale, int N, int K, int fmt, int transpose,
                 void* stream);

/* ---- asymmetric scheme: the same operators with the group minimum as an extra tensor --------------------------
 * Replace fused_decode4_matmul5* / fused_decode{3,4,5,6}_*matmul* of the asymmetric builds
 * (dispatch/llama/main.py:4235-5313), decode6/7/8 + decode4/5/6_fp16 (:3985-4094), te_take_decode_asym and
 * te_encode_asym.  Argument order = the PrimFuncs' (packed, scale, min, x[, operand], out) (main.py:4979).
 * fmt: B200Q_Q4F16_ASYM (zmin = f16 tensor shaped like scale) or B200Q_Q4F32_ASYM (zmin = NULL, scale = the packed
 * bf16 pair).  K must be a multiple of the group size (quantization.py:107).  Dequantized weights are bit-exact
 * against the reference's CPU builds (two roundings, no fma contraction); matmuls accumulate in f32. */
B200Q_API int b200q_gemv_asym(const void* packed, const void* scale, const void* zmin, const void* x, const void* ep0,
                    void* out, int K, int N, int fmt, int epilogue, unsigned flags, void* workspace,
                    size_t workspace_bytes, void* stream);
B200Q_API int b200q_dequant_asym(const void* packed, const void* scale, const void* zmin, void* out, int K, int N,
                       int fmt, int transposed, void* stream);
B200Q_API int b200q_take_decode_asym(const void* packed_rm, const void* scale_rm, const void* zmin_rm, const int32_t* ids,
                           void* out, int n_ids, int V, int K, int fmt, void* stream);
B200Q_API int b200q_encode_asym(const void* W, void* packed, void* scale, void* zmin, int N, int K, int fmt, int transpose,
                      void* stream);
/* prefill: f16 scheme only (dequantize inside the tcgen05 GEMM) */
B200Q_API int b200q_gemm_asym(const void* packed, const void* scale, const void* zmin, const void* X, const void* ep0,
                    const void* ep1, const void* ep2, void* Y, int M, int K, int N, int fmt, int epilogue,
                    unsigned flags, void* stream);

/* ---- a7 / a8: prefill GEMM -------------------------------------------------------------------
 * Y[m, n] = sum_k X[m, k] * w[k, n]  (+ epilogue), X dtype[M, K] row-major, Y dtype[M, N].
 * b200q_gemm dequantizes inside the GEMM (replaces fused_decode*_NT_matmul*,
 * dispatch_tir_operator_adreno.py:414-1021, and the decode* + NT_matmul* pair on CUDA):
 * packed tiles are unpacked to f16 in registers, staged through shared memory in the tcgen05
 * canonical layout and multiplied on the 5th-gen tensor cores with f32 accumulators in TMEM.
 * f16 formats only.  ep0/ep1/ep2 as in b200q_epilogue (bias is [N]; residuals are [M, N]).
 * Requirements: K % 64 == 0 for q4/q8 (K % 40 == 0 or zero-padded x for q3), N % 8 == 0. */
B200Q_API int b200q_gemm(const void* packed, const void* scale, const void* X, const void* ep0,
               const void* ep1, const void* ep2, void* Y, int M, int K, int N, int fmt,
               int epilogue, unsigned flags, void* stream);

/* b200q_gemm / b200q_gemm_asym with a device workspace (b200q_gemm_workspace_bytes() bytes, 16-byte aligned, zero-filled
 * ONCE by the caller; every launch leaves its flag area zeroed again; not shared by launches that may overlap on
 * the GPU: one per stream).  With it the CTA-pair kernel may cut the (tile, k) space into equal ranges per pair
 * ("stream-K") when whole 256 x 256 tiles would leave most pairs idle in the last round (80 tiles on 74 pairs: M =
 * 2048, N = 2560): pairs that do not finish a tile park their f32 partial tile in the workspace, the tile's
 * finisher adds the partials in pair order - deterministic, same tolerance as b200q_gemm, not bit-identical to
 * it.  The schedule is chosen from the shape (a measured cost model in csrc/gemm.cu; B200Q_GEMM_STREAMK=0/1 in the
 * environment overrides it).  zmin: NULL for the symmetric formats.  workspace NULL, B200Q_FLAG_GEMM_NO_STREAMK,
 * or a shape that fills its last round of tiles: exactly b200q_gemm. */
#define B200Q_FLAG_GEMM_NO_STREAMK 16u
#define B200Q_FLAG_GEMM_STREAMK 32u /* stream-K whenever the workspace allows it, whatever the shape (tests, experiments) */
B200Q_API size_t b200q_gemm_workspace_bytes(void);
B200Q_API int b200q_gemm_ws(const void* packed, const void* scale, const void* zmin, const void* X, const void* ep0,
                  const void* ep1, const void* ep2, void* Y, int M, int K, int N, int fmt, int epilogue,
                  unsigned flags, void* workspace, size_t workspace_bytes, void* stream);

/* Up to four weights that read the SAME X in one launch (a fourth one: h_to_4h of a GPT-NeoX layer, which reads the
 * layer input like q | k | v - one launch when it carries the bias + GELU epilogue and M > 128, two otherwise) (q | k | v: the reference launches them back to back on the
 * same x, relax_model/llama.py:233-236; gpt_neox.py query/key/value).  Tensors stay the reference's separate
 * per-Linear arrays, every weight has its own epilogue and output.  The column tiles of all weights form ONE tile
 * space (and one stream-K split): three 2560-wide NeoX projections cost 3.24 rounds of tiles instead of 3 x 2.
 * Results are those of b200q_gemm_ws weight by weight (bit-identical under the whole-tile schedule).  No output
 * may alias X or another weight's epilogue operand.  zmin: NULL for the symmetric formats. */
typedef struct {
  const void* packed;
  const void* scale;
  const void* zmin;
  const void* ep0;
  const void* ep1;
  const void* ep2;
  void* Y;
  int N;
  int epilogue;
} b200q_gemm_op;
B200Q_API int b200q_gemm_multi(const b200q_gemm_op* ops, int n_ops, const void* X, int M, int K, int fmt,
                     unsigned flags, void* workspace, size_t workspace_bytes, void* stream);

/* Host-side evaluation of the CTA-pair kernel's work distribution (no GPU needed; for tests): writes the segments of
 * `pair` in processing order as 5 ints each (tile, first 16-k step, end step, kind 0 whole / 1 partial writer /
 * 2 finisher, end of the finisher's contributor pairs) and returns the number of segments (-1: bad arguments). */
B200Q_API int b200q_debug_gemm_schedule(int streamk, int n_tiles, int n_ksteps, int chunk, int n_pairs, int pair,
                              int* out, int max_seg);

/* The same for one step of a decode program (no GPU needed; for tests): a step of n_panels 128-column panels x
 * row_blocks 16-row blocks is divided evenly over n_ctas CTAs; writes the panel segments of `cta` in processing order
 * as 6 ints each (panel, first row block, end row block, kind 0 whole panel / 1 writes a partial sum / 2 finishes the
 * panel, partial-sum slot of the CTA inside the panel, number of contributing CTAs the finisher adds up - 1 otherwise),
 * *slots_per_panel (may be NULL) = the slots the program builder reserves per panel, and returns the number of segments
 * (-1: bad arguments).  finisher_first: 1 = the tcgen05 kernel's rule (the owner of the panel's first block finishes it),
 * 0 = the register-level kernel's (the owner of its last block). */
B200Q_API int b200q_debug_gemv_schedule(int n_panels, int row_blocks, int n_ctas, int cta, int finisher_first, int* out,
                              int max_seg, int* slots_per_panel);

/* Dense NT GEMM on an already dequantized weight  T dtype[N, K]  (replaces NT_matmul*,
 * transform/transpose_matmul.py:34-101, dispatch/llama/main.py:3884-3981). f16 only. */
B200Q_API int b200q_nt_matmul(const void* X, const void* T, const void* ep0, const void* ep1,
                    const void* ep2, void* Y, int M, int K, int N, int epilogue, void* stream);
/* the same with the stream-K workspace of b200q_gemm_ws (and B200Q_FLAG_PDL / B200Q_FLAG_GEMM_[NO_]STREAMK) */
B200Q_API int b200q_nt_matmul_ws(const void* X, const void* T, const void* ep0, const void* ep1,
                       const void* ep2, void* Y, int M, int K, int N, int epilogue, unsigned flags,
                       void* workspace, size_t workspace_bytes, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* B200Q_H_ */
