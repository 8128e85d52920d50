// C-ABI entry points declared in include/b200q.h: argument validation, error plumbing, dispatch to
// the kernels.  No torch types, no allocation, no synchronization (except the *_host helpers).
#include <atomic>
#include <cstdarg>
#include <cstdio>

#include "common.cuh"

namespace b200q {

static thread_local char g_err[512] = "";
static std::atomic<unsigned long long> g_launches{0};

int set_error(int code, const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
  return code;
}
int check_cuda(cudaError_t e, const char* what) {
  if (e == cudaSuccess) return B200Q_OK;
  return set_error(B200Q_ERR_CUDA, "%s: %s", what, cudaGetErrorString(e));
}
void count_launch(int n) { g_launches.fetch_add((unsigned long long)n, std::memory_order_relaxed); }

// kernels (gemv.cu, gemv_f32.cu, dequant.cu, gemm.cu)
int gemv_launch(const b200q_gemv_op* ops, int n_ops, const void* x, const void* x_mul, int K, int fmt,
                unsigned flags, void* workspace, size_t workspace_bytes, cudaStream_t stream);
void gemv_tune(int ctas_per_sm, int stages, int flags);
size_t gemv_workspace_bytes();
int gemv_static_launch(const b200q_gemv_op* ops, int n_ops, const void* x, const void* x_mul, int K, int fmt, unsigned flags,
                       cudaStream_t stream);
size_t gemv_program_bytes(int n_steps);
int gemv_program_build(const b200q_gemv_step* steps, int n_steps, int fmt, void* host_blob, size_t blob_bytes,
                       const unsigned char* reduce_mask, int tp_rank, int tp_size);
size_t gemv_program_workspace_bytes(const void* host_blob);
void gemv_program_set_watchdog(unsigned long long* status, long long timeout_ns);
int gemv_program_run(const void* host_blob, const void* dev_blob, void* workspace, size_t workspace_bytes,
                     void* const* peer_workspaces, cudaStream_t stream);
int dequant_launch(const void* packed, const void* scale, const void* zmin, void* out, int K, int N, int fmt, int transposed,
                   cudaStream_t stream);
int take_decode_launch(const void* packed_rm, const void* scale_rm, const void* zmin_rm, const int32_t* ids, void* out,
                       int n_ids, int V, int K, int fmt, cudaStream_t stream);
bool fmt_is_variant(int fmt);
int variant_encode_launch(const void* W, void* packed, void* scale, int N, int K, int fmt, int transpose, cudaStream_t stream);
int variant_dequant_launch(const void* packed, const void* scale, void* out, int K, int N, int fmt, int transposed, cudaStream_t stream);
int variant_take_launch(const void* packed_rm, const void* scale_rm, const int32_t* ids, void* out, int n_ids, int V, int K, int fmt,
                        cudaStream_t stream);
int variant_gemv_launch(const void* packed, const void* scale, const void* x, const void* ep0, void* out, int K, int N, int fmt,
                        int epilogue, cudaStream_t stream);
int encode_asym_launch(const void* W, void* packed, void* scale, void* zmin, int N, int K, int fmt, int transpose,
                       cudaStream_t stream);
int gemv_asym_launch(const void* packed, const void* scale, const void* zmin, const void* x, const void* ep0, void* out, int K,
                     int N, int fmt, int epilogue, unsigned flags, void* workspace, size_t workspace_bytes, cudaStream_t stream);
int gemm_asym_launch(const void* packed, const void* scale, const void* zmin, const void* X, const void* ep0, const void* ep1,
                     const void* ep2, void* Y, int M, int K, int N, int epilogue, unsigned flags, cudaStream_t stream,
                     void* workspace, size_t workspace_bytes);
size_t gemm_workspace_bytes();
int gemm_schedule_debug(int streamk, int n_tiles, int n_ksteps, int chunk, int n_pairs, int pair, int* out, int max_seg);
int gemv_program_schedule_debug(int n_panels, int RB, int G, int cta, int finisher_first, int* out, int max_seg, int* c_max);
int gemm_multi_launch(const b200q_gemm_op* ops, int n_ops, const void* X, int M, int K, int fmt, unsigned flags,
                      cudaStream_t stream, void* workspace, size_t workspace_bytes);
int encode_launch(const void* W, void* packed, void* scale, int N, int K, int fmt, int transpose,
                  cudaStream_t stream);
int gemm_launch(const void* packed, const void* scale, const void* X, const void* ep0, const void* ep1,
                const void* ep2, void* Y, int M, int K, int N, int fmt, int epilogue, unsigned flags,
                cudaStream_t stream, void* workspace, size_t workspace_bytes);
int nt_matmul_launch(const void* X, const void* T, const void* ep0, const void* ep1, const void* ep2, void* Y,
                     int M, int K, int N, int epilogue, cudaStream_t stream, unsigned flags, void* workspace,
                     size_t workspace_bytes);

static bool aligned(const void* p, size_t a) { return (reinterpret_cast<uintptr_t>(p) & (a - 1)) == 0; }

static int check_fmt(int fmt, const char* who, bool asym_entry = false) {
  if (fmt < 0 || fmt >= B200Q_NUM_FMT) return set_error(B200Q_ERR_UNSUPPORTED, "%s: unknown fmt %d", who, fmt);
  if (fmt_is_asym(fmt) != (asym_entry ? 1 : 0))
    return set_error(B200Q_ERR_UNSUPPORTED, asym_entry ? "%s: fmt %d is a symmetric scheme (use the entry point without _asym)"
                                                       : "%s: fmt %d is an asymmetric scheme (use the *_asym entry point)", who, fmt);
  return B200Q_OK;
}
// asymmetric schemes: the f16 one carries a min tensor shaped like scale, the f32 one packs it into `scale`
static int check_asym(int fmt, const void* zmin, int K, const char* who) {
  if (fmt == B200Q_Q4F16_ASYM && !zmin) return set_error(B200Q_ERR_NULL, "%s: the f16 asymmetric scheme needs the min tensor", who);
  if (fmt == B200Q_Q4F32_ASYM && zmin) return set_error(B200Q_ERR_UNSUPPORTED, "%s: the f32 asymmetric scheme packs min into scale (zmin must be NULL)", who);
  if (zmin && !aligned(zmin, 16)) return set_error(B200Q_ERR_BAD_ALIGN, "%s: zmin must be 16-byte aligned", who);
  if (K <= 0 || K % 32 != 0) return set_error(B200Q_ERR_BAD_SHAPE, "%s: K=%d must be a multiple of the group size 32 (quantization.py:107)", who, K);
  return B200Q_OK;
}

// tir.noalias (dispatch_tir_operator_adreno.py:7): an output never overlaps an input or another output.  With
// stream-K a panel owner writes out[n] while other CTAs (or a PDL-overlapped launch) still read x, so an
// in-place call would race silently: reject it.
static bool overlaps(const void* a, size_t na, const void* b, size_t nb) {
  if (!a || !b || na == 0 || nb == 0) return false;
  const uintptr_t a0 = reinterpret_cast<uintptr_t>(a), b0 = reinterpret_cast<uintptr_t>(b);
  return a0 < b0 + nb && b0 < a0 + na;
}
static int check_gemv_alias(const b200q_gemv_op* ops, int n_ops, const void* x, const void* x_mul, int K, int fmt,
                            const char* who) {
  const size_t el = fmt_is_f32(fmt) ? 4 : 2;
  for (int i = 0; i < n_ops; ++i) {
    const size_t ob = ((ops[i].epilogue == B200Q_EP_CAST_F32) ? 4 : el) * (size_t)ops[i].N;
    if (overlaps(ops[i].out, ob, x, el * K) || overlaps(ops[i].out, ob, x_mul, el * K))
      return set_error(B200Q_ERR_UNSUPPORTED, "%s: out of op %d aliases the activation (outputs never alias inputs)", who, i);
    for (int j = 0; j < n_ops; ++j) {
      if (overlaps(ops[i].out, ob, ops[j].ep0, el * (size_t)ops[j].N))
        return set_error(B200Q_ERR_UNSUPPORTED, "%s: out of op %d aliases the epilogue operand of op %d", who, i, j);
      if (j < i && overlaps(ops[i].out, ob, ops[j].out, ((ops[j].epilogue == B200Q_EP_CAST_F32) ? 4 : el) * (size_t)ops[j].N))
        return set_error(B200Q_ERR_UNSUPPORTED, "%s: the outputs of ops %d and %d overlap", who, j, i);
    }
  }
  return B200Q_OK;
}

static int check_gemv_op(const b200q_gemv_op& op, int fmt, const char* who) {
  if (!op.packed || !op.scale || !op.out) return set_error(B200Q_ERR_NULL, "%s: NULL packed/scale/out", who);
  if (op.N <= 0 || op.N % 8 != 0) return set_error(B200Q_ERR_BAD_SHAPE, "%s: N=%d must be a positive multiple of 8", who, op.N);
  if (!aligned(op.packed, 16) || !aligned(op.scale, 16))
    return set_error(B200Q_ERR_BAD_ALIGN, "%s: packed/scale must be 16-byte aligned", who);
  const int ep = op.epilogue;
  if (ep < 0 || ep >= B200Q_NUM_EP || ep == B200Q_EP_BIAS_ADD2)
    return set_error(B200Q_ERR_UNSUPPORTED, "%s: epilogue %d not available for the GEMV", who, ep);
  const bool needs_ep0 = ep == B200Q_EP_ADD || ep == B200Q_EP_MULTIPLY || ep == B200Q_EP_BIAS || ep == B200Q_EP_BIAS_GELU;
  if (needs_ep0 && !op.ep0) return set_error(B200Q_ERR_NULL, "%s: epilogue %d needs ep0", who, ep);
  const size_t el = fmt_is_f32(fmt) ? 4 : 2;
  if (!aligned(op.out, (ep == B200Q_EP_CAST_F32) ? 4 : el) || (needs_ep0 && !aligned(op.ep0, el)))
    return set_error(B200Q_ERR_BAD_ALIGN, "%s: out/ep0 misaligned", who);
  return B200Q_OK;
}

}  // namespace b200q

using namespace b200q;

extern "C" {

int b200q_version(void) { return B200Q_VERSION; }
const char* b200q_last_error(void) { return g_err; }
unsigned long long b200q_launch_count(void) { return g_launches.load(); }

int b200q_n_groups(int fmt, int K) {
  if (fmt < 0 || fmt >= B200Q_NUM_FMT) return B200Q_ERR_UNSUPPORTED;
  return n_groups(fmt, K);
}
int b200q_n_words(int fmt, int K) {
  if (fmt < 0 || fmt >= B200Q_NUM_FMT) return B200Q_ERR_UNSUPPORTED;
  return n_words(fmt, K);
}
long long b200q_gemv_bytes(int fmt, int K, int N, int epilogue) {
  if (fmt < 0 || fmt >= B200Q_NUM_FMT) return B200Q_ERR_UNSUPPORTED;
  const long long sb = fmt_is_f32(fmt) ? 4 : 2;
  const long long ob = (epilogue == B200Q_EP_CAST_F32 || fmt_is_f32(fmt)) ? 4 : 2;
  long long b = (long long)fmt_word_bytes(fmt) * n_words(fmt, K) * N + sb * n_groups(fmt, K) * N + sb * K + ob * N;
  if (fmt == B200Q_Q4F16_ASYM) b += sb * n_groups(fmt, K) * N;  // the min tensor
  if (epilogue == B200Q_EP_ADD || epilogue == B200Q_EP_MULTIPLY) b += ob * N;
  if (epilogue == B200Q_EP_BIAS || epilogue == B200Q_EP_BIAS_GELU) b += sb * N;
  return b;
}

void b200q_gemv_tune(int ctas_per_sm, int stages, int dbg_flags) { gemv_tune(ctas_per_sm, stages, dbg_flags); }
size_t b200q_gemv_workspace_bytes(void) { return gemv_workspace_bytes(); }

int b200q_gemv_multi(const b200q_gemv_op* ops, int n_ops, const void* x, const void* x_mul, int K, int fmt,
                     unsigned flags, void* workspace, size_t workspace_bytes, void* stream) {
  if (int e = check_fmt(fmt, "b200q_gemv")) return e;
  if (!ops || !x) return set_error(B200Q_ERR_NULL, "b200q_gemv: NULL ops/x");
  if (n_ops < 1 || n_ops > 4) return set_error(B200Q_ERR_BAD_SHAPE, "b200q_gemv_multi: n_ops=%d not in [1,4]", n_ops);
  if (K <= 0) return set_error(B200Q_ERR_BAD_SHAPE, "b200q_gemv: K=%d", K);
  const size_t el = fmt_is_f32(fmt) ? 4 : 2;
  if (!aligned(x, el) || (x_mul && !aligned(x_mul, el))) return set_error(B200Q_ERR_BAD_ALIGN, "b200q_gemv: x misaligned");
  if (workspace && !aligned(workspace, 16)) return set_error(B200Q_ERR_BAD_ALIGN, "b200q_gemv: workspace must be 16-byte aligned");
  for (int i = 0; i < n_ops; ++i)
    if (int e = check_gemv_op(ops[i], fmt, "b200q_gemv")) return e;
  if (int e = check_gemv_alias(ops, n_ops, x, x_mul, K, fmt, "b200q_gemv")) return e;
  if (fmt_is_variant(fmt)) {  // uint16-storage int3 / fp4: the streaming kernel of variants.cu, one Linear per launch
    if (n_ops != 1 || x_mul) return set_error(B200Q_ERR_UNSUPPORTED, "b200q_gemv_multi: fmt %d takes one op and no x_mul", fmt);
    return variant_gemv_launch(ops[0].packed, ops[0].scale, x, ops[0].ep0, ops[0].out, K, ops[0].N, fmt, ops[0].epilogue,
                               reinterpret_cast<cudaStream_t>(stream));
  }
  return gemv_launch(ops, n_ops, x, x_mul, K, fmt, flags, workspace, workspace ? workspace_bytes : 0,
                     reinterpret_cast<cudaStream_t>(stream));
}

int b200q_gemv(const void* packed, const void* scale, const void* x, const void* ep0, void* out, int K,
               int N, int fmt, int epilogue, unsigned flags, void* workspace, size_t workspace_bytes,
               void* stream) {
  b200q_gemv_op op{packed, scale, ep0, out, N, epilogue};
  return b200q_gemv_multi(&op, 1, x, nullptr, K, fmt, flags, workspace, workspace_bytes, stream);
}

int b200q_gemv_static(const void* packed, const void* scale, const void* x, const void* ep0, void* out, int K, int N,
                      int fmt, int epilogue, unsigned flags, void* stream) {
  if (int e = check_fmt(fmt, "b200q_gemv_static")) return e;
  if (fmt_is_variant(fmt)) return set_error(B200Q_ERR_UNSUPPORTED, "b200q_gemv_static: fmt %d has no split-K kernel (no scratch to replace)", fmt);
  if (!x) return set_error(B200Q_ERR_NULL, "b200q_gemv_static: NULL x");
  if (K <= 0) return set_error(B200Q_ERR_BAD_SHAPE, "b200q_gemv_static: K=%d", K);
  const size_t el = fmt_is_f32(fmt) ? 4 : 2;
  if (!aligned(x, el)) return set_error(B200Q_ERR_BAD_ALIGN, "b200q_gemv_static: x misaligned");
  b200q_gemv_op op{packed, scale, ep0, out, N, epilogue};
  if (int e = check_gemv_op(op, fmt, "b200q_gemv_static")) return e;
  if (int e = check_gemv_alias(&op, 1, x, nullptr, K, fmt, "b200q_gemv_static")) return e;
  return gemv_static_launch(&op, 1, x, nullptr, K, fmt, flags, reinterpret_cast<cudaStream_t>(stream));
}

size_t b200q_gemv_program_bytes(int n_steps) { return n_steps > 0 ? gemv_program_bytes(n_steps) : 0; }

int b200q_gemv_program_build(const b200q_gemv_step* steps, int n_steps, int fmt, void* host_blob, size_t blob_bytes) {
  return b200q_gemv_program_build_tp(steps, n_steps, fmt, nullptr, 0, 1, host_blob, blob_bytes);
}

int b200q_gemv_program_build_tp(const b200q_gemv_step* steps, int n_steps, int fmt, const unsigned char* reduce_mask,
                                int tp_rank, int tp_size, void* host_blob, size_t blob_bytes) {
  if (int e = check_fmt(fmt, "b200q_gemv_program_build")) return e;  // (asymmetric schemes: per-operator launches only)
  if (fmt_is_variant(fmt)) return set_error(B200Q_ERR_UNSUPPORTED, "b200q_gemv_program_build: fmt %d has per-operator kernels only", fmt);
  if (!steps || !host_blob) return set_error(B200Q_ERR_NULL, "b200q_gemv_program_build: NULL steps/blob");
  if (n_steps < 1 || n_steps > 65535) return set_error(B200Q_ERR_BAD_SHAPE, "b200q_gemv_program_build: n_steps=%d", n_steps);
  const size_t el = fmt_is_f32(fmt) ? 4 : 2;
  for (int j = 0; j < n_steps; ++j) {
    const b200q_gemv_step& st = steps[j];
    if (st.n_ops < 1 || st.n_ops > 4) return set_error(B200Q_ERR_BAD_SHAPE, "b200q_gemv_program_build: step %d has %d ops", j, st.n_ops);
    if (!st.x || st.K <= 0) return set_error(B200Q_ERR_NULL, "b200q_gemv_program_build: step %d: NULL x or K=%d", j, st.K);
    if (!aligned(st.x, el) || (st.x_mul && !aligned(st.x_mul, el))) return set_error(B200Q_ERR_BAD_ALIGN, "b200q_gemv_program_build: step %d: x misaligned", j);
    for (int i = 0; i < st.n_ops; ++i)
      if (int e = check_gemv_op(st.ops[i], fmt, "b200q_gemv_program_build")) return e;
  }
  return gemv_program_build(steps, n_steps, fmt, host_blob, blob_bytes, reduce_mask, tp_rank, tp_size);
}

size_t b200q_gemv_program_workspace_bytes(const void* host_blob) {
  return host_blob ? gemv_program_workspace_bytes(host_blob) : 0;
}

int b200q_gemv_program_run(const void* host_blob, const void* dev_blob, void* workspace, size_t workspace_bytes,
                           void* stream) {
  return b200q_gemv_program_run_tp(host_blob, dev_blob, workspace, workspace_bytes, nullptr, stream);
}

int b200q_gemv_program_run_tp(const void* host_blob, const void* dev_blob, void* workspace, size_t workspace_bytes,
                              void* const* peer_workspaces, void* stream) {
  if (!host_blob || !dev_blob || !workspace) return set_error(B200Q_ERR_NULL, "b200q_gemv_program_run: NULL blob/workspace");
  return gemv_program_run(host_blob, dev_blob, workspace, workspace_bytes, peer_workspaces, reinterpret_cast<cudaStream_t>(stream));
}

void b200q_gemv_program_set_watchdog(void* status_word, long long timeout_ns) {
  gemv_program_set_watchdog(reinterpret_cast<unsigned long long*>(status_word), timeout_ns);
}

int b200q_gemv_host(const void* packed, const void* scale, const void* x_host, const void* ep0,
                    void* out_host, void* x_dev, void* out_dev, int K, int N, int fmt, int epilogue,
                    void* workspace, size_t workspace_bytes, void* stream) {
  if (int e = check_fmt(fmt, "b200q_gemv_host")) return e;
  if (!x_host || !out_host || !x_dev || !out_dev) return set_error(B200Q_ERR_NULL, "b200q_gemv_host: NULL buffer");
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  const size_t el = fmt_is_f32(fmt) ? 4 : 2;
  const size_t oel = (epilogue == B200Q_EP_CAST_F32) ? 4 : el;
  if (int e = check_cuda(cudaMemcpyAsync(x_dev, x_host, el * K, cudaMemcpyHostToDevice, s), "H2D x")) return e;
  if (int e = b200q_gemv(packed, scale, x_dev, ep0, out_dev, K, N, fmt, epilogue, 0, workspace, workspace_bytes, stream)) return e;
  if (int e = check_cuda(cudaMemcpyAsync(out_host, out_dev, oel * N, cudaMemcpyDeviceToHost, s), "D2H out")) return e;
  return check_cuda(cudaStreamSynchronize(s), "cudaStreamSynchronize");
}

int b200q_dequant(const void* packed, const void* scale, void* out, int K, int N, int fmt, int transposed,
                  void* stream) {
  if (int e = check_fmt(fmt, "b200q_dequant")) return e;
  if (!packed || !scale || !out) return set_error(B200Q_ERR_NULL, "b200q_dequant: NULL pointer");
  if (K <= 0 || N <= 0 || N % 8 != 0) return set_error(B200Q_ERR_BAD_SHAPE, "b200q_dequant: K=%d N=%d (N %% 8 == 0 required)", K, N);
  if (!aligned(packed, 16) || !aligned(scale, 16) || !aligned(out, 16))
    return set_error(B200Q_ERR_BAD_ALIGN, "b200q_dequant: pointers must be 16-byte aligned");
  if (fmt_is_variant(fmt)) return variant_dequant_launch(packed, scale, out, K, N, fmt, transposed, reinterpret_cast<cudaStream_t>(stream));
  return dequant_launch(packed, scale, nullptr, out, K, N, fmt, transposed, reinterpret_cast<cudaStream_t>(stream));
}

int b200q_take_decode(const void* packed_rm, const void* scale_rm, const int32_t* ids, void* out, int n_ids,
                      int V, int K, int fmt, void* stream) {
  if (int e = check_fmt(fmt, "b200q_take_decode")) return e;
  if (!packed_rm || !scale_rm || !ids || !out) return set_error(B200Q_ERR_NULL, "b200q_take_decode: NULL pointer");
  if (K <= 0 || V <= 0 || n_ids < 0) return set_error(B200Q_ERR_BAD_SHAPE, "b200q_take_decode: K=%d V=%d n_ids=%d", K, V, n_ids);
  if (n_ids == 0) return B200Q_OK;
  if (n_ids > 65535) return set_error(B200Q_ERR_BAD_SHAPE, "b200q_take_decode: n_ids=%d > 65535", n_ids);
  if (fmt_is_variant(fmt)) return variant_take_launch(packed_rm, scale_rm, ids, out, n_ids, V, K, fmt, reinterpret_cast<cudaStream_t>(stream));
  return take_decode_launch(packed_rm, scale_rm, nullptr, ids, out, n_ids, V, K, fmt, reinterpret_cast<cudaStream_t>(stream));
}

int b200q_encode(const void* W, void* packed, void* scale, int N, int K, int fmt, int transpose, void* stream) {
  if (int e = check_fmt(fmt, "b200q_encode")) return e;
  if (!W || !packed || !scale) return set_error(B200Q_ERR_NULL, "b200q_encode: NULL pointer");
  if (N <= 0 || K <= 0) return set_error(B200Q_ERR_BAD_SHAPE, "b200q_encode: N=%d K=%d", N, K);
  if (fmt_is_variant(fmt)) return variant_encode_launch(W, packed, scale, N, K, fmt, transpose, reinterpret_cast<cudaStream_t>(stream));
  return encode_launch(W, packed, scale, N, K, fmt, transpose, reinterpret_cast<cudaStream_t>(stream));
}

static int gemm_sym(const void* packed, const void* scale, const void* X, const void* ep0, const void* ep1,
                    const void* ep2, void* Y, int M, int K, int N, int fmt, int epilogue, unsigned flags,
                    void* workspace, size_t workspace_bytes, void* stream) {
  if (int e = check_fmt(fmt, "b200q_gemm")) return e;
  if (fmt_is_variant(fmt)) return set_error(B200Q_ERR_UNSUPPORTED, "b200q_gemm: fmt %d has decode / GEMV kernels only (prefill: b200q_dequant + b200q_nt_matmul)", fmt);
  if (fmt_is_f32(fmt)) return set_error(B200Q_ERR_UNSUPPORTED, "b200q_gemm: f16 formats only (q4f32_0 prefill: dequant + f32 GEMM is out of scope)");
  if (!packed || !scale || !X || !Y) return set_error(B200Q_ERR_NULL, "b200q_gemm: NULL pointer");
  if (M < 0 || K <= 0 || N <= 0 || N % 8 != 0) return set_error(B200Q_ERR_BAD_SHAPE, "b200q_gemm: M=%d K=%d N=%d", M, K, N);
  if (M == 0) return B200Q_OK;
  if (epilogue < 0 || epilogue >= B200Q_NUM_EP) return set_error(B200Q_ERR_UNSUPPORTED, "b200q_gemm: epilogue %d", epilogue);
  if (!aligned(packed, 16) || !aligned(scale, 16) || !aligned(X, 16) || !aligned(Y, 16))
    return set_error(B200Q_ERR_BAD_ALIGN, "b200q_gemm: pointers must be 16-byte aligned");
  return gemm_launch(packed, scale, X, ep0, ep1, ep2, Y, M, K, N, fmt, epilogue, flags,
                     reinterpret_cast<cudaStream_t>(stream), workspace, workspace_bytes);
}

int b200q_gemm(const void* packed, const void* scale, const void* X, const void* ep0, const void* ep1,
               const void* ep2, void* Y, int M, int K, int N, int fmt, int epilogue, unsigned flags,
               void* stream) {
  return gemm_sym(packed, scale, X, ep0, ep1, ep2, Y, M, K, N, fmt, epilogue, flags, nullptr, 0, stream);
}

int b200q_gemv_asym(const void* packed, const void* scale, const void* zmin, const void* x, const void* ep0, void* out, int K,
                    int N, int fmt, int epilogue, unsigned flags, void* workspace, size_t workspace_bytes, void* stream) {
  if (int e = check_fmt(fmt, "b200q_gemv_asym", true)) return e;
  if (!x) return set_error(B200Q_ERR_NULL, "b200q_gemv_asym: NULL x");
  if (int e = check_asym(fmt, zmin, K, "b200q_gemv_asym")) return e;
  const size_t el = fmt_is_f32(fmt) ? 4 : 2;
  if (!aligned(x, el)) return set_error(B200Q_ERR_BAD_ALIGN, "b200q_gemv_asym: x misaligned");
  if (workspace && !aligned(workspace, 16)) return set_error(B200Q_ERR_BAD_ALIGN, "b200q_gemv_asym: workspace must be 16-byte aligned");
  b200q_gemv_op op{packed, scale, ep0, out, N, epilogue};
  if (int e = check_gemv_op(op, fmt, "b200q_gemv_asym")) return e;
  if (int e = check_gemv_alias(&op, 1, x, nullptr, K, fmt, "b200q_gemv_asym")) return e;
  return gemv_asym_launch(packed, scale, zmin, x, ep0, out, K, N, fmt, epilogue, flags, workspace, workspace ? workspace_bytes : 0,
                          reinterpret_cast<cudaStream_t>(stream));
}

int b200q_dequant_asym(const void* packed, const void* scale, const void* zmin, void* out, int K, int N, int fmt,
                       int transposed, void* stream) {
  if (int e = check_fmt(fmt, "b200q_dequant_asym", true)) return e;
  if (!packed || !scale || !out) return set_error(B200Q_ERR_NULL, "b200q_dequant_asym: NULL pointer");
  if (int e = check_asym(fmt, zmin, K, "b200q_dequant_asym")) return e;
  if (N <= 0 || N % 8 != 0) return set_error(B200Q_ERR_BAD_SHAPE, "b200q_dequant_asym: N=%d (N %% 8 == 0 required)", N);
  if (!aligned(packed, 16) || !aligned(scale, 16) || !aligned(out, 16))
    return set_error(B200Q_ERR_BAD_ALIGN, "b200q_dequant_asym: pointers must be 16-byte aligned");
  return dequant_launch(packed, scale, zmin, out, K, N, fmt, transposed, reinterpret_cast<cudaStream_t>(stream));
}

int b200q_take_decode_asym(const void* packed_rm, const void* scale_rm, const void* zmin_rm, const int32_t* ids, void* out,
                           int n_ids, int V, int K, int fmt, void* stream) {
  if (int e = check_fmt(fmt, "b200q_take_decode_asym", true)) return e;
  if (!packed_rm || !scale_rm || !ids || !out) return set_error(B200Q_ERR_NULL, "b200q_take_decode_asym: NULL pointer");
  if (fmt == B200Q_Q4F16_ASYM && !zmin_rm) return set_error(B200Q_ERR_NULL, "b200q_take_decode_asym: the f16 scheme needs the min table");
  if (K <= 0 || K % 32 != 0 || V <= 0 || n_ids < 0) return set_error(B200Q_ERR_BAD_SHAPE, "b200q_take_decode_asym: K=%d V=%d n_ids=%d", K, V, n_ids);
  if (n_ids == 0) return B200Q_OK;
  if (n_ids > 65535) return set_error(B200Q_ERR_BAD_SHAPE, "b200q_take_decode_asym: n_ids=%d > 65535", n_ids);
  return take_decode_launch(packed_rm, scale_rm, zmin_rm, ids, out, n_ids, V, K, fmt, reinterpret_cast<cudaStream_t>(stream));
}

int b200q_encode_asym(const void* W, void* packed, void* scale, void* zmin, int N, int K, int fmt, int transpose, void* stream) {
  if (int e = check_fmt(fmt, "b200q_encode_asym", true)) return e;
  if (!W || !packed || !scale) return set_error(B200Q_ERR_NULL, "b200q_encode_asym: NULL pointer");
  if (fmt == B200Q_Q4F16_ASYM && !zmin) return set_error(B200Q_ERR_NULL, "b200q_encode_asym: the f16 scheme writes a min tensor");
  if (N <= 0 || K <= 0 || K % 32 != 0) return set_error(B200Q_ERR_BAD_SHAPE, "b200q_encode_asym: N=%d K=%d (K %% 32 == 0 required)", N, K);
  return encode_asym_launch(W, packed, scale, zmin, N, K, fmt, transpose, reinterpret_cast<cudaStream_t>(stream));
}

static int gemm_asym(const void* packed, const void* scale, const void* zmin, const void* X, const void* ep0, const void* ep1,
                     const void* ep2, void* Y, int M, int K, int N, int fmt, int epilogue, unsigned flags, void* workspace,
                     size_t workspace_bytes, void* stream) {
  if (int e = check_fmt(fmt, "b200q_gemm_asym", true)) return e;
  if (fmt != B200Q_Q4F16_ASYM) return set_error(B200Q_ERR_UNSUPPORTED, "b200q_gemm_asym: f16 scheme only (f32 prefill is the reference's miss too)");
  if (!packed || !scale || !X || !Y) return set_error(B200Q_ERR_NULL, "b200q_gemm_asym: NULL pointer");
  if (int e = check_asym(fmt, zmin, K, "b200q_gemm_asym")) return e;
  if (M < 0 || N <= 0 || N % 8 != 0) return set_error(B200Q_ERR_BAD_SHAPE, "b200q_gemm_asym: M=%d N=%d", M, N);
  if (M == 0) return B200Q_OK;
  if (epilogue < 0 || epilogue >= B200Q_NUM_EP) return set_error(B200Q_ERR_UNSUPPORTED, "b200q_gemm_asym: epilogue %d", epilogue);
  if (!aligned(packed, 16) || !aligned(scale, 16) || !aligned(X, 16) || !aligned(Y, 16))
    return set_error(B200Q_ERR_BAD_ALIGN, "b200q_gemm_asym: pointers must be 16-byte aligned");
  return gemm_asym_launch(packed, scale, zmin, X, ep0, ep1, ep2, Y, M, K, N, epilogue, flags, reinterpret_cast<cudaStream_t>(stream),
                          workspace, workspace_bytes);
}

int b200q_gemm_asym(const void* packed, const void* scale, const void* zmin, const void* X, const void* ep0, const void* ep1,
                    const void* ep2, void* Y, int M, int K, int N, int fmt, int epilogue, unsigned flags, void* stream) {
  return gemm_asym(packed, scale, zmin, X, ep0, ep1, ep2, Y, M, K, N, fmt, epilogue, flags, nullptr, 0, stream);
}

size_t b200q_gemm_workspace_bytes(void) { return gemm_workspace_bytes(); }

int b200q_debug_gemm_schedule(int streamk, int n_tiles, int n_ksteps, int chunk, int n_pairs, int pair, int* out, int max_seg) {
  if (!out || n_tiles < 1 || n_ksteps < 1 || chunk < 1 || n_pairs < 1 || pair < 0 || pair >= n_pairs || max_seg < 0) return -1;
  return gemm_schedule_debug(streamk, n_tiles, n_ksteps, chunk, n_pairs, pair, out, max_seg);
}

int b200q_debug_gemv_schedule(int n_panels, int row_blocks, int n_ctas, int cta, int finisher_first, int* out, int max_seg,
                              int* slots_per_panel) {
  if (!out || n_panels < 1 || row_blocks < 1 || n_ctas < 1 || cta < 0 || cta >= n_ctas || max_seg < 0) return -1;
  if ((long long)n_panels * row_blocks > 0x3FFFFFFFll) return -1;
  return gemv_program_schedule_debug(n_panels, row_blocks, n_ctas, cta, finisher_first, out, max_seg, slots_per_panel);
}

int b200q_gemm_multi(const b200q_gemm_op* ops, int n_ops, const void* X, int M, int K, int fmt, unsigned flags,
                     void* workspace, size_t workspace_bytes, void* stream) {
  const bool asym = fmt_is_asym(fmt);
  if (int e = check_fmt(fmt, "b200q_gemm_multi", asym)) return e;
  if (fmt_is_variant(fmt) || fmt_is_f32(fmt) || (asym && fmt != B200Q_Q4F16_ASYM))
    return set_error(B200Q_ERR_UNSUPPORTED, "b200q_gemm_multi: fmt %d has no prefill GEMM", fmt);
  if (!ops || !X) return set_error(B200Q_ERR_NULL, "b200q_gemm_multi: NULL ops / X");
  if (n_ops < 1 || n_ops > 4) return set_error(B200Q_ERR_BAD_SHAPE, "b200q_gemm_multi: n_ops=%d not in [1,4]", n_ops);
  if (M < 0 || K <= 0) return set_error(B200Q_ERR_BAD_SHAPE, "b200q_gemm_multi: M=%d K=%d", M, K);
  if (M == 0) return B200Q_OK;
  if (!aligned(X, 16)) return set_error(B200Q_ERR_BAD_ALIGN, "b200q_gemm_multi: X must be 16-byte aligned");
  if (workspace && !aligned(workspace, 16)) return set_error(B200Q_ERR_BAD_ALIGN, "b200q_gemm_multi: workspace must be 16-byte aligned");
  if (!workspace) workspace_bytes = 0;
  const size_t xb = (size_t)M * K * 2;
  for (int i = 0; i < n_ops; ++i) {
    const b200q_gemm_op& o = ops[i];
    if (!o.packed || !o.scale || !o.Y) return set_error(B200Q_ERR_NULL, "b200q_gemm_multi: op %d: NULL packed / scale / Y", i);
    if (o.N <= 0 || o.N % 8 != 0) return set_error(B200Q_ERR_BAD_SHAPE, "b200q_gemm_multi: op %d: N=%d", i, o.N);
    if (o.epilogue < 0 || o.epilogue >= B200Q_NUM_EP) return set_error(B200Q_ERR_UNSUPPORTED, "b200q_gemm_multi: op %d: epilogue %d", i, o.epilogue);
    if (!aligned(o.packed, 16) || !aligned(o.scale, 16) || !aligned(o.Y, 16))
      return set_error(B200Q_ERR_BAD_ALIGN, "b200q_gemm_multi: op %d: pointers must be 16-byte aligned", i);
    if (asym) { if (int e = check_asym(fmt, o.zmin, K, "b200q_gemm_multi")) return e; }
    const size_t yb = (size_t)M * o.N * (o.epilogue == B200Q_EP_CAST_F32 ? 4 : 2);
    if (overlaps(o.Y, yb, X, xb)) return set_error(B200Q_ERR_UNSUPPORTED, "b200q_gemm_multi: output of op %d aliases X", i);
    for (int j = 0; j < n_ops; ++j) {
      if (j != i && overlaps(o.Y, yb, ops[j].Y, (size_t)M * ops[j].N * (ops[j].epilogue == B200Q_EP_CAST_F32 ? 4 : 2)))
        return set_error(B200Q_ERR_UNSUPPORTED, "b200q_gemm_multi: outputs of ops %d and %d overlap", i, j);
      const int ej = ops[j].epilogue;
      const bool bias_like = ej == B200Q_EP_BIAS || ej == B200Q_EP_BIAS_GELU || ej == B200Q_EP_BIAS_ADD2;
      const size_t eb = (bias_like ? (size_t)1 : (size_t)M) * ops[j].N * 2;
      if (overlaps(o.Y, yb, ops[j].ep0, eb) && j != i)
        return set_error(B200Q_ERR_UNSUPPORTED, "b200q_gemm_multi: output of op %d aliases an epilogue operand of op %d", i, j);
    }
  }
  return gemm_multi_launch(ops, n_ops, X, M, K, fmt, flags, reinterpret_cast<cudaStream_t>(stream), workspace, workspace_bytes);
}

int b200q_gemm_ws(const void* packed, const void* scale, const void* zmin, const void* X, const void* ep0, const void* ep1,
                  const void* ep2, void* Y, int M, int K, int N, int fmt, int epilogue, unsigned flags, void* workspace,
                  size_t workspace_bytes, void* stream) {
  if (workspace && !aligned(workspace, 16)) return set_error(B200Q_ERR_BAD_ALIGN, "b200q_gemm_ws: workspace must be 16-byte aligned");
  if (!workspace) workspace_bytes = 0;
  if (fmt_is_asym(fmt)) return gemm_asym(packed, scale, zmin, X, ep0, ep1, ep2, Y, M, K, N, fmt, epilogue, flags, workspace, workspace_bytes, stream);
  return gemm_sym(packed, scale, X, ep0, ep1, ep2, Y, M, K, N, fmt, epilogue, flags, workspace, workspace_bytes, stream);
}

int b200q_nt_matmul(const void* X, const void* T, const void* ep0, const void* ep1, const void* ep2, void* Y,
                    int M, int K, int N, int epilogue, void* stream) {
  return b200q_nt_matmul_ws(X, T, ep0, ep1, ep2, Y, M, K, N, epilogue, 0u, nullptr, 0, stream);
}

int b200q_nt_matmul_ws(const void* X, const void* T, const void* ep0, const void* ep1, const void* ep2, void* Y,
                       int M, int K, int N, int epilogue, unsigned flags, void* workspace, size_t workspace_bytes,
                       void* stream) {
  if (workspace && !aligned(workspace, 16)) return set_error(B200Q_ERR_BAD_ALIGN, "b200q_nt_matmul_ws: workspace must be 16-byte aligned");
  if (!workspace) workspace_bytes = 0;
  if (!X || !T || !Y) return set_error(B200Q_ERR_NULL, "b200q_nt_matmul: NULL pointer");
  if (M < 0 || K <= 0 || N <= 0 || N % 8 != 0 || K % 8 != 0)
    return set_error(B200Q_ERR_BAD_SHAPE, "b200q_nt_matmul: M=%d K=%d N=%d", M, K, N);
  if (M == 0) return B200Q_OK;
  if (!aligned(X, 16) || !aligned(T, 16) || !aligned(Y, 16))
    return set_error(B200Q_ERR_BAD_ALIGN, "b200q_nt_matmul: pointers must be 16-byte aligned");
  return nt_matmul_launch(X, T, ep0, ep1, ep2, Y, M, K, N, epilogue, reinterpret_cast<cudaStream_t>(stream), flags, workspace,
                          workspace_bytes);
}

}  // extern "C"
