// Decode-step fused dequantize + GEMV for all four group-quantized formats, sm_100a.
//
// Replaces the reference's fused_decode*_matmul* PrimFuncs
// (mlc_llm/dispatch/dispatch_tir_operator_adreno.py:5-411, mlc_llm/dispatch/llama/main.py:4235-6363;
// schedule generator mlc_llm/tuning.py:39-108: one output column per thread, sequential K,
// N/256 blocks, 4-byte weight loads, no async copy).  A batch-1 GEMV of these sizes (9-75 MB) lasts
// 1.5-11 us at HBM speed, so the design is about never letting the memory pipe drain:
//
//   * ONE persistent CTA per SM (grid = #SMs); the matrix is cut into blocks of 32 packed-word rows
//     x 128 columns and the flattened block sequence (row blocks fastest inside a 128-column panel)
//     is divided EVENLY over the CTAs ("stream-K"), whatever K and N are;
//   * a producer thread streams its CTA's blocks into a multi-stage shared-memory ring with TMA
//     (cp.async.bulk.tensor.2d ... mbarrier::complete_tx, L2 evict-first): four 32x32-word boxes
//     (128B-swizzled, so the 128-bit reads below are bank-conflict free) + one scale box per stage,
//     ~18 KB per stage, 75-95 KB in flight per SM, no registers tied up, out-of-range rows/columns
//     zero-filled by the hardware;
//   * the weights are constants, so the producer starts BEFORE griddepcontrol.wait: with
//     programmatic dependent launch the next GEMV of the stream fills its ring (the second CTA slot
//     of every SM) while the current one is still computing / reducing;
//   * eight consumer warps unpack the words in registers to the bit-exact f16 weight
//     half((q - maxint) * s) (common.cuh) and contract over K with mma.sync.m16n8k16 (f16 x f16
//     products are exact, f32 accumulate): the packed layout puts 8 consecutive k of ONE column in a
//     register, so weights are the A operand (16 columns x 16 k per instruction) and the activation
//     vector is replicated over the 8 B columns.  A SIMT formulation needs >= 3.5 issue slots per
//     weight against a budget of ~3.1 at HBM speed; this one needs ~1.9.  (q4f32_0 has f32 weights:
//     FFMA, same pipeline.);
//   * panels whose K range is shared by several CTAs are finished by the LAST CTA to arrive
//     (counter in a caller-provided workspace, partial sums added in CTA order => deterministic,
//     no spinning, self-cleaning).  Without a workspace whole panels are dealt out instead;
//   * epilogues (residual add, multiply, silu, bias, bias+gelu, cast) are applied op for op in the
//     model dtype like the reference PrimFuncs; optionally the activation is formed on the fly as
//     x * x_mul (the SwiGLU product feeding down_proj) - bit-identical to the reference's separate
//     `multiply` epilogue.
//
// Lane mapping inside a stage: warp (cq, rh) owns columns [32 cq, +32) x rows [16 rh, +16) of the
// block; lane (c = lane / 4, tig = lane % 4) owns columns 4c..4c+3 of rows {2 tig, 2 tig + 1,
// 8 + 2 tig, 9 + 2 tig} of that half: with the 128-byte swizzle (16-byte chunk index XOR row & 7)
// the eight lanes of a quarter warp hit eight distinct bank groups.
#include <cuda.h>

#include <atomic>

#include "common.cuh"
#include "gemv_device.cuh"
#include "ptx.cuh"

namespace b200q {

namespace {

using namespace ptx;
using namespace gemvdev;

constexpr int kConsumerWarps = 8;
constexpr int kConsumerThreads = kConsumerWarps * 32;
constexpr int kThreads = kConsumerThreads + 32;  // + producer warp
constexpr int kMaxOps = 4;
constexpr int kBlockRows = 32;
constexpr int kBoxBytes = kBlockRows * kBoxCols * 4;           // 4096
constexpr int kWordStageBytes = kBlockRows * kPanelCols * 4;   // 16384
constexpr int kMaxStages = 12;
constexpr int kMaxPanels = 4096;
constexpr int kCounterBytes = kMaxPanels * 4;
constexpr int kSmemHeader = 1280;  // barriers (256) + red[2][128] f32 (1024)
constexpr int kSmemPerSm = 227 * 1024;

template <int FMT>
struct Stream {
  using F = Fmt<FMT>;
  using scale_t = typename F::scale_t;
  static constexpr bool kF32 = fmt_is_f32_c(FMT);
  static constexpr bool kHasMin = fmt_has_min_c(FMT);  // f16 asymmetric: a second box (the group minima) behind the scales
  static constexpr int kGroupsPerBlock = kBlockRows / F::kRowsPerGroup;
  static constexpr int kScalePitch = kPanelCols * (int)sizeof(scale_t);
  static constexpr int kMinOff = kGroupsPerBlock * kScalePitch;
  static constexpr int kScaleStageBytes = kMinOff * (kHasMin ? 2 : 1);
  static constexpr int kStageTx = kWordStageBytes + kScaleStageBytes;       // bytes TMA delivers per stage
  static constexpr int kStageBytes = (kStageTx + 1023) / 1024 * 1024;       // swizzled boxes need 1024B alignment
  static constexpr int kXRowBytes = XLayout<FMT>::kXRowBytes;
};

struct StreamOp {
  const uint32_t* packed;
  const void* scale;
  const void* zmin;  // f16 asymmetric scheme only
  const void* ep0;
  void* out;
  int N;
  int epilogue;
  int panel_begin;
};

struct StreamParams {
  CUtensorMap tm_w[kMaxOps];  // packed  uint32[n_rows, N], box 32 x 32, 128B swizzle
  CUtensorMap tm_s[kMaxOps];  // scale   dtype[n_groups, N], box kGroupsPerBlock x 128
  CUtensorMap tm_m[kMaxOps];  // min     dtype[n_groups, N] (f16 asymmetric scheme), same box
  StreamOp op[kMaxOps];
  int n_ops;
  const void* x;
  const void* x_mul;  // optional: activation = x * x_mul (elementwise, model dtype)
  int K;
  int n_rows;    // packed word rows
  int RB;        // row blocks per panel
  int n_panels;  // over all ops
  int T;         // total blocks
  int stages;
  int xs_bytes;
  int aligned;   // 1: CTAs own whole panels (no workspace)
  int c_max;     // partial slots per panel
  float* ws_partial;
  unsigned* ws_counter;
  int no_compute;  // debug: stream only
};

// ---- PTX helpers: ptx.cuh; shared device code: gemv_device.cuh -------------------------------------
__device__ __forceinline__ void tma_2d(uint32_t dst, const CUtensorMap* tm, int c_inner, int c_outer, uint32_t bar,
                                       uint64_t pol) {
  tma_load_2d_hint(dst, tm, c_inner, c_outer, bar, pol);
}
__device__ __forceinline__ void consumer_sync() { asm volatile("bar.sync 1, %0;" ::"n"(kConsumerThreads) : "memory"); }
__device__ __forceinline__ int find_op(const StreamParams& p, int panel) {
  int oi = 0;
#pragma unroll
  for (int t = 1; t < kMaxOps; ++t)
    if (t < p.n_ops && panel >= p.op[t].panel_begin) oi = t;
  return oi;
}

// first block of CTA i
__device__ __forceinline__ int cta_begin(const StreamParams& p, int i, int G) {
  if (p.aligned) return (int)((long long)i * p.n_panels / G) * p.RB;
  return (int)((long long)i * p.T / G);
}
// CTA that owns block t (non-aligned split)
__device__ __forceinline__ int block_owner(const StreamParams& p, int t, int G) {
  return (int)((((long long)t + 1) * G - 1) / p.T);
}

template <int FMT>
__global__ void __launch_bounds__(kThreads, 2) gemv_stream_kernel(const __grid_constant__ StreamParams p) {
  using S = Stream<FMT>;
  using F = Fmt<FMT>;
  using scale_t = typename S::scale_t;
  extern __shared__ __align__(128) unsigned char smem[];

  pdl_launch_dependents();

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int G = gridDim.x;
  const int t0 = cta_begin(p, blockIdx.x, G), t1 = cta_begin(p, blockIdx.x + 1, G);
  const int stages = p.stages;

  uint64_t* bars = reinterpret_cast<uint64_t*>(smem);  // full[kMaxStages], empty[kMaxStages]
  int* flag = reinterpret_cast<int*>(smem + 240);
  float* red = reinterpret_cast<float*>(smem + 256);  // [2][128]
  unsigned char* xs = smem + kSmemHeader;
  const uint32_t stage_base = (smem_u32(smem + kSmemHeader + p.xs_bytes) + 1023u) & ~1023u;
  const uint32_t full0 = smem_u32(bars), empty0 = smem_u32(bars + kMaxStages);

  if (threadIdx.x == 0) {
    for (int s = 0; s < stages; ++s) {
      mbar_init(full0 + 8 * s, 1);
      mbar_init(empty0 + 8 * s, kConsumerWarps);
    }
    fence_barrier_init();
  }
  __syncthreads();

  if (warp == kConsumerWarps) {
    // ================= producer: stream the CTA's blocks into the ring (weights are constants:
    // legal before the dependency wait) =================================================================
    if (lane == 0) {
      const uint64_t pol = l2_evict_first_policy();
      for (int i = 0; i < p.n_ops; ++i) {
        prefetch_tmap(&p.tm_w[i]);
        prefetch_tmap(&p.tm_s[i]);
      }
      int s = 0;
      uint32_t ph = 0;
      int panel = t0 / p.RB, rb = t0 - panel * p.RB;
      for (int t = t0; t < t1; ++t) {
        const int oi = find_op(p, panel);
        const int col0 = (panel - p.op[oi].panel_begin) * kPanelCols;
        const uint32_t bar = full0 + 8 * s;
        mbar_wait(empty0 + 8 * s, ph ^ 1);
        mbar_expect_tx(bar, (uint32_t)S::kStageTx);
        const uint32_t sw = stage_base + s * S::kStageBytes;
#pragma unroll
        for (int q = 0; q < kPanelCols / kBoxCols; ++q)
          tma_2d(sw + q * kBoxBytes, &p.tm_w[oi], col0 + q * kBoxCols, rb * kBlockRows, bar, pol);
        tma_2d(sw + kWordStageBytes, &p.tm_s[oi], col0, rb * S::kGroupsPerBlock, bar, pol);
        if constexpr (S::kHasMin) tma_2d(sw + kWordStageBytes + S::kMinOff, &p.tm_m[oi], col0, rb * S::kGroupsPerBlock, bar, pol);
        if (++s == stages) { s = 0; ph ^= 1; }
        if (++rb == p.RB) { rb = 0; ++panel; }
      }
    }
    return;
  }

  // ================= consumers ===========================================================================
  const int cq = warp & 3, rh = warp >> 2;  // column quarter (one 32-column box), row half (16 rows)
  const int c = lane >> 2, tig = lane & 3;
  const int rows_alloc = p.RB * kBlockRows;

  pdl_wait();  // x / x_mul / ep0 / out belong to the previous kernel until here
  stage_x<FMT, kConsumerThreads>(xs, p.x, p.x_mul, p.K, rows_alloc);
  consumer_sync();

  // lane rows of the 16-row half: 2 tig + j + 8 m  (j, m in {0,1}); swizzled chunk = c ^ (row & 7)
  const uint32_t xs_base = smem_u32(xs) + (rh * 16 + 2 * tig) * S::kXRowBytes;
  const uint32_t w_off0 = cq * kBoxBytes + (rh * 16 + 2 * tig) * 128 + ((c ^ (2 * tig)) << 4);
  const uint32_t w_off1 = cq * kBoxBytes + (rh * 16 + 2 * tig + 1) * 128 + ((c ^ (2 * tig + 1)) << 4);
  // scale group of rows (16 rh + 8 m + 2 tig): rows-per-group 4 -> 4 rh + 2 m + tig / 2; 8 -> 2 rh + m
  const int g0 = (rh * 16 + 2 * tig) / F::kRowsPerGroup, g1 = (rh * 16 + 8 + 2 * tig) / F::kRowsPerGroup;
  const uint32_t s_col = (cq * 32 + c * 4) * (int)sizeof(scale_t);
  const uint32_t s_off0 = kWordStageBytes + g0 * S::kScalePitch + s_col;
  const uint32_t s_off1 = kWordStageBytes + g1 * S::kScalePitch + s_col;

  int s = 0;
  uint32_t ph = 0;
  int t = t0;
  int panel = t0 / p.RB, rb = t0 - panel * p.RB;
  while (t < t1) {
    const int rb_begin = rb;
    const int seg_end = min(t1, t + (p.RB - rb));
    float acc[2][4];
#pragma unroll
    for (int m = 0; m < 2; ++m)
#pragma unroll
      for (int r = 0; r < 4; ++r) acc[m][r] = 0.f;

    for (; t < seg_end; ++t, ++rb) {
      mbar_wait(full0 + 8 * s, ph);
      const uint32_t sw = stage_base + s * S::kStageBytes;
      if (!p.no_compute) {
        const uint4 w00 = lds128(sw + w_off0), w01 = lds128(sw + w_off1);
        const uint4 w10 = lds128(sw + w_off0 + 8 * 128), w11 = lds128(sw + w_off1 + 8 * 128);
        const uint32_t xr = xs_base + rb * kBlockRows * S::kXRowBytes;
        if constexpr (FMT == B200Q_Q4F32_ASYM) {
          const uint4 sc0 = lds128(sw + s_off0), sc1 = lds128(sw + s_off1);
          compute_row_f32_asym(w00, sc0, xr, acc[0]);
          compute_row_f32_asym(w01, sc0, xr + S::kXRowBytes, acc[0]);
          compute_row_f32_asym(w10, sc1, xr + 8 * S::kXRowBytes, acc[0]);
          compute_row_f32_asym(w11, sc1, xr + 9 * S::kXRowBytes, acc[0]);
        } else if constexpr (S::kF32) {
          const uint4 sc0 = lds128(sw + s_off0), sc1 = lds128(sw + s_off1);
          compute_row_f32(w00, sc0, xr, acc[0]);
          compute_row_f32(w01, sc0, xr + S::kXRowBytes, acc[0]);
          compute_row_f32(w10, sc1, xr + 8 * S::kXRowBytes, acc[0]);
          compute_row_f32(w11, sc1, xr + 9 * S::kXRowBytes, acc[0]);
        } else if constexpr (S::kHasMin) {
          __half2 s2[4], m2[4];
          split_scales(lds64(sw + s_off0), s2);
          split_scales(lds64(sw + s_off0 + S::kMinOff), m2);
          compute_row_f16_asym(w00, s2, m2, xr, acc);
          compute_row_f16_asym(w01, s2, m2, xr + S::kXRowBytes, acc);
          split_scales(lds64(sw + s_off1), s2);
          split_scales(lds64(sw + s_off1 + S::kMinOff), m2);
          compute_row_f16_asym(w10, s2, m2, xr + 8 * S::kXRowBytes, acc);
          compute_row_f16_asym(w11, s2, m2, xr + 9 * S::kXRowBytes, acc);
        } else {
          __half2 s2[4];
          split_scales(lds64(sw + s_off0), s2);
          compute_row_f16<FMT>(w00, s2, xr, acc);
          compute_row_f16<FMT>(w01, s2, xr + S::kXRowBytes, acc);
          split_scales(lds64(sw + s_off1), s2);
          compute_row_f16<FMT>(w10, s2, xr + 8 * S::kXRowBytes, acc);
          compute_row_f16<FMT>(w11, s2, xr + 9 * S::kXRowBytes, acc);
        }
      }
      __syncwarp();
      if (lane == 0) mbar_arrive(empty0 + 8 * s);
      if (++s == stages) { s = 0; ph ^= 1; }
    }

    // ---- reduce the segment: quad lanes (f32 path), the two row halves, then other CTAs -------------
    float4 v4;
    if constexpr (S::kF32) {
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        acc[0][j] += __shfl_xor_sync(0xffffffffu, acc[0][j], 1);
        acc[0][j] += __shfl_xor_sync(0xffffffffu, acc[0][j], 2);
      }
      v4 = make_float4(acc[0][0], acc[0][1], acc[0][2], acc[0][3]);
    } else {
      v4 = make_float4(acc[0][0], acc[0][2], acc[1][0], acc[1][2]);  // columns 4c .. 4c+3
    }
    if (tig == 0) *reinterpret_cast<float4*>(red + rh * kPanelCols + cq * 32 + c * 4) = v4;
    consumer_sync();

    const StreamOp& op = p.op[find_op(p, panel)];
    const int col0 = (panel - op.panel_begin) * kPanelCols;
    const int tc = threadIdx.x;
    const int n = col0 + tc;
    float sum = 0.f;
    if (tc < kPanelCols) sum = red[tc] + red[kPanelCols + tc];
    bool do_epilogue = (tc < kPanelCols);
    const bool complete = (rb_begin == 0) && (rb == p.RB);
    if (!complete) {
      // this panel's K range is shared with other CTAs: publish the partial, last arriver finishes
      const int first = block_owner(p, panel * p.RB, G);
      const int last = block_owner(p, panel * p.RB + p.RB - 1, G);
      const int n_contrib = last - first + 1;
      float* slot0 = p.ws_partial + (size_t)panel * p.c_max * kPanelCols;
      if (tc < kPanelCols) {
        slot0[(size_t)((int)blockIdx.x - first) * kPanelCols + tc] = sum;
        __threadfence();
      }
      consumer_sync();
      if (tc == 0) {
        const unsigned old = atomicAdd(p.ws_counter + panel, 1u);
        const int is_last = (old == (unsigned)(n_contrib - 1));
        if (is_last) p.ws_counter[panel] = 0u;  // self-cleaning: nobody else touches it in this launch
        *flag = is_last;
      }
      consumer_sync();
      do_epilogue = do_epilogue && (*flag != 0);
      if (do_epilogue) {
        __threadfence();
        sum = 0.f;
        for (int o = 0; o < n_contrib; ++o) sum += ld_cg_f32(slot0 + (size_t)o * kPanelCols + tc);
      }
    }
    if (do_epilogue && n < op.N) {
      if constexpr (S::kF32) {
        reinterpret_cast<float*>(op.out)[n] = ep_f32(sum, op.epilogue, reinterpret_cast<const float*>(op.ep0), n);
      } else {
        if (op.epilogue == B200Q_EP_CAST_F32)
          reinterpret_cast<float*>(op.out)[n] = sum;
        else
          reinterpret_cast<__half*>(op.out)[n] = ep_f16(sum, op.epilogue, reinterpret_cast<const __half*>(op.ep0), n);
      }
    }
    consumer_sync();  // red / flag are reused by the next segment
    if (rb == p.RB) { rb = 0; ++panel; }
  }
}

// ---- host side --------------------------------------------------------------------------------------
thread_local int g_tune_ctas_per_sm = 0, g_tune_stages = 0, g_tune_flags = 0;

// per-device caches (one process may drive several GPUs): SM count and the dynamic-smem opt-in, which
// cudaFuncSetAttribute records per device
constexpr int kMaxDevices = 64;
int current_device() {
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= kMaxDevices) return -1;
  return dev;
}
int sm_count() {
  static std::atomic<int> cache[kMaxDevices];
  const int dev = current_device();
  if (dev >= 0) {
    const int c = cache[dev].load(std::memory_order_relaxed);
    if (c > 0) return c;
  }
  int n = 0;
  if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev < 0 ? 0 : dev) != cudaSuccess || n <= 0) n = 148;
  if (dev >= 0) cache[dev].store(n, std::memory_order_relaxed);
  return n;
}

// cuTensorMapEncodeTiled through the runtime's driver entry point (no link-time libcuda dependency:
// the library must load on a machine without a driver for the CPU-side checks).
using EncodeTiledFn = CUresult (*)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                   const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                   CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
EncodeTiledFn encode_tiled_fn() {
  static EncodeTiledFn fn = nullptr;
  static bool tried = false;
  if (!tried) {
    void* f = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &f, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(f);
    tried = true;
  }
  return fn;
}

// 2-D row-major tensor [outer, inner] of `elem_bytes` elements, box [box_outer, box_inner].
int make_tmap(CUtensorMap* tm, CUtensorMapDataType dt, int elem_bytes, const void* ptr, int inner, int outer,
              int box_inner, int box_outer, CUtensorMapSwizzle sw) {
  EncodeTiledFn fn = encode_tiled_fn();
  if (!fn) return set_error(B200Q_ERR_CUDA, "cuTensorMapEncodeTiled is not available from this driver");
  cuuint64_t dims[2] = {(cuuint64_t)inner, (cuuint64_t)outer};
  cuuint64_t strides[1] = {(cuuint64_t)inner * elem_bytes};
  cuuint32_t box[2] = {(cuuint32_t)box_inner, (cuuint32_t)box_outer};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = fn(tm, dt, 2, const_cast<void*>(ptr), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, sw,
                  CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) return set_error(B200Q_ERR_CUDA, "cuTensorMapEncodeTiled failed (%d) for [%d, %d]", (int)r, outer, inner);
  return B200Q_OK;
}

template <int FMT>
int launch_stream(StreamParams& p, unsigned flags, void* workspace, size_t workspace_bytes, cudaStream_t stream) {
  using S = Stream<FMT>;
  using F = Fmt<FMT>;
  p.RB = ceil_div(p.n_rows, kBlockRows);
  p.T = p.n_panels * p.RB;
  p.xs_bytes = ((p.RB * kBlockRows * S::kXRowBytes) + 127) / 128 * 128;
  const int n_grp = p.n_rows / F::kRowsPerGroup;
  for (int i = 0; i < p.n_ops; ++i) {
    if (int e = make_tmap(&p.tm_w[i], CU_TENSOR_MAP_DATA_TYPE_UINT32, 4, p.op[i].packed, p.op[i].N, p.n_rows, kBoxCols,
                          kBlockRows, CU_TENSOR_MAP_SWIZZLE_128B))
      return e;
    if (int e = make_tmap(&p.tm_s[i], S::kF32 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT32 : CU_TENSOR_MAP_DATA_TYPE_FLOAT16,
                          (int)sizeof(typename S::scale_t), p.op[i].scale, p.op[i].N, n_grp, kPanelCols,
                          S::kGroupsPerBlock, CU_TENSOR_MAP_SWIZZLE_NONE))
      return e;
    if constexpr (S::kHasMin) {
      if (int e = make_tmap(&p.tm_m[i], CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, p.op[i].zmin, p.op[i].N, n_grp, kPanelCols,
                            S::kGroupsPerBlock, CU_TENSOR_MAP_SWIZZLE_NONE))
        return e;
    }
  }
  const int per_sm = (g_tune_ctas_per_sm == 2) ? 2 : 1;
  int G = sm_count() * per_sm;
  if (G > p.T) G = p.T;

  // stages: two CTAs must fit per SM (the second slot is the next launch's prefetching CTA)
  const int budget = kSmemPerSm / 2 - 1024 - kSmemHeader - p.xs_bytes - 1024;
  int stages = budget / S::kStageBytes;
  if (stages > kMaxStages) stages = kMaxStages;
  if (g_tune_stages > 0 && g_tune_stages < stages) stages = g_tune_stages;
  if (stages < 2) {
    // very long K: give the CTA the whole SM
    stages = (kSmemPerSm - 1024 - kSmemHeader - p.xs_bytes - 1024) / S::kStageBytes;
    if (stages > kMaxStages) stages = kMaxStages;
    if (stages < 2) return set_error(B200Q_ERR_BAD_SHAPE, "gemv: K=%d too large for the activation stage", p.K);
  }
  p.stages = stages;
  const size_t smem = (size_t)kSmemHeader + p.xs_bytes + 1024 + (size_t)stages * S::kStageBytes;

  // split panels across CTAs only with a workspace big enough for the partial sums
  p.aligned = 1;
  p.c_max = 0;
  p.ws_partial = nullptr;
  p.ws_counter = nullptr;
  if (workspace && p.n_panels <= kMaxPanels && G > 1) {
    const int per_cta = p.T / G;  // >= 1
    const int c_max = ceil_div(p.RB, per_cta) + 1;
    const size_t need = (size_t)kCounterBytes + (size_t)p.n_panels * c_max * kPanelCols * sizeof(float);
    if (need <= workspace_bytes) {
      p.aligned = 0;
      p.c_max = c_max;
      p.ws_counter = reinterpret_cast<unsigned*>(workspace);
      p.ws_partial = reinterpret_cast<float*>(reinterpret_cast<unsigned char*>(workspace) + kCounterBytes);
    }
  }
  if (p.aligned && G > p.n_panels) G = p.n_panels;
  p.no_compute = (g_tune_flags & 1);

  auto kern = gemv_stream_kernel<FMT>;
  static std::atomic<size_t> smem_set[kMaxDevices];  // per instantiation and device
  const int dev = current_device();
  if (dev < 0 || smem > smem_set[dev].load(std::memory_order_relaxed)) {
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return check_cuda(e, "cudaFuncSetAttribute(gemv smem)");
    if (dev >= 0) smem_set[dev].store(smem, std::memory_order_relaxed);
  }
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3(G, 1, 1);
  cfg.blockDim = dim3(kThreads, 1, 1);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = stream;
  cudaLaunchAttribute attrs[1];
  int na = 0;
  if (flags & B200Q_FLAG_PDL) {
    attrs[na].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attrs[na].val.programmaticStreamSerializationAllowed = 1;
    ++na;
  }
  cfg.attrs = attrs;
  cfg.numAttrs = na;
  cudaError_t e = cudaLaunchKernelEx(&cfg, kern, p);
  if (e != cudaSuccess) return check_cuda(e, "cudaLaunchKernelEx(gemv)");
  count_launch();
  return B200Q_OK;
}

}  // namespace

// shared with gemm.cu
int make_tmap_2d(CUtensorMap* tm, CUtensorMapDataType dt, int elem_bytes, const void* ptr, int inner, int outer,
                 int box_inner, int box_outer, CUtensorMapSwizzle sw) {
  return make_tmap(tm, dt, elem_bytes, ptr, inner, outer, box_inner, box_outer, sw);
}
int device_sm_count() { return sm_count(); }

void gemv_tune(int ctas_per_sm, int stages, int flags) {
  g_tune_ctas_per_sm = ctas_per_sm;
  g_tune_stages = stages;
  g_tune_flags = flags;
}

int gemv_asym_launch(const void* packed, const void* scale, const void* zmin, const void* x, const void* ep0, void* out, int K,
                     int N, int fmt, int epilogue, unsigned flags, void* workspace, size_t workspace_bytes, cudaStream_t stream) {
  StreamParams p{};
  p.n_ops = 1;
  p.x = x;
  p.K = K;
  p.n_rows = n_words(fmt, K);
  p.op[0].packed = reinterpret_cast<const uint32_t*>(packed);
  p.op[0].scale = scale;
  p.op[0].zmin = zmin;
  p.op[0].ep0 = ep0;
  p.op[0].out = out;
  p.op[0].N = N;
  p.op[0].epilogue = epilogue;
  p.op[0].panel_begin = 0;
  p.n_panels = ceil_div(N, kPanelCols);
  if (fmt == B200Q_Q4F16_ASYM) return launch_stream<B200Q_Q4F16_ASYM>(p, flags, workspace, workspace_bytes, stream);
  return launch_stream<B200Q_Q4F32_ASYM>(p, flags, workspace, workspace_bytes, stream);
}

size_t gemv_workspace_bytes() { return (size_t)kCounterBytes + (size_t)4 * 1024 * 1024; }

int gemv_launch(const b200q_gemv_op* ops, int n_ops, const void* x, const void* x_mul, int K, int fmt,
                unsigned flags, void* workspace, size_t workspace_bytes, cudaStream_t stream) {
  StreamParams p{};
  p.n_ops = n_ops;
  p.x = x;
  p.x_mul = x_mul;
  p.K = K;
  p.n_rows = n_words(fmt, K);
  int panels = 0;
  for (int i = 0; i < n_ops; ++i) {
    p.op[i].packed = reinterpret_cast<const uint32_t*>(ops[i].packed);
    p.op[i].scale = ops[i].scale;
    p.op[i].ep0 = ops[i].ep0;
    p.op[i].out = ops[i].out;
    p.op[i].N = ops[i].N;
    p.op[i].epilogue = ops[i].epilogue;
    p.op[i].panel_begin = panels;
    panels += ceil_div(ops[i].N, kPanelCols);
  }
  p.n_panels = panels;
  switch (fmt) {
    case B200Q_Q4F16_0: return launch_stream<B200Q_Q4F16_0>(p, flags, workspace, workspace_bytes, stream);
    case B200Q_Q3F16_0: return launch_stream<B200Q_Q3F16_0>(p, flags, workspace, workspace_bytes, stream);
    case B200Q_Q8F16_0: return launch_stream<B200Q_Q8F16_0>(p, flags, workspace, workspace_bytes, stream);
    case B200Q_Q4F32_0: return launch_stream<B200Q_Q4F32_0>(p, flags, workspace, workspace_bytes, stream);
    default: return set_error(B200Q_ERR_UNSUPPORTED, "gemv: fmt %d", fmt);
  }
}

// ---- b200q_gemv_static: the split-K GEMV for callers that cannot allocate device memory -------------------------
// The scratch of gemv_launch (panel counters + partial sums, gemv_workspace_bytes() bytes) lives in this file-scope
// device array instead of a caller-provided buffer.  Every GPU context that loads the library holds the array until
// the process exits.  All calls share it: two calls that overlap on the GPU (different streams, or two host threads)
// race on it.  In host code `g_static_ws` is only a stand-in for the symbol, not a device address: the address is
// looked up once per device with cudaGetSymbolAddress, with that device current.
constexpr size_t kStaticWsBytes = (size_t)kCounterBytes + (size_t)4 * 1024 * 1024;  // == gemv_workspace_bytes()
static __device__ __align__(256) unsigned char g_static_ws[kStaticWsBytes];

int static_ws_ptr(void** out) {
  static std::atomic<void*> cache[kMaxDevices];
  const int dev = current_device();
  if (dev < 0) return set_error(B200Q_ERR_CUDA, "b200q_gemv_static: no current CUDA device (or device id >= %d)", kMaxDevices);
  void* ptr = cache[dev].load(std::memory_order_acquire);
  if (!ptr) {
    cudaError_t e = cudaGetSymbolAddress(&ptr, g_static_ws);
    if (e != cudaSuccess) return check_cuda(e, "cudaGetSymbolAddress(static GEMV scratch)");
    if (!ptr) return set_error(B200Q_ERR_CUDA, "b200q_gemv_static: cudaGetSymbolAddress returned NULL");
    cache[dev].store(ptr, std::memory_order_release);
  }
  *out = ptr;
  return B200Q_OK;
}

int gemv_static_launch(const b200q_gemv_op* ops, int n_ops, const void* x, const void* x_mul, int K, int fmt, unsigned flags,
                       cudaStream_t stream) {
  // the scratch the split needs, as launch_stream sizes it; beyond the array the call is refused (no silent
  // fall-back to the unsplit kernel, no allocation)
  int panels = 0;
  for (int i = 0; i < n_ops; ++i) panels += ceil_div(ops[i].N, kPanelCols);
  if (panels > kMaxPanels)
    return set_error(B200Q_ERR_BAD_SHAPE, "b200q_gemv_static: %d column panels (N = %lld) exceed the static scratch: at most %d panels (N <= %d)",
                     panels, (long long)panels * kPanelCols, kMaxPanels, kMaxPanels * kPanelCols);
  const int RB = ceil_div(n_words(fmt, K), kBlockRows);
  const long long T = (long long)panels * RB;
  long long G = (long long)sm_count() * ((g_tune_ctas_per_sm == 2) ? 2 : 1);
  if (G > T) G = T;
  if (G > 1) {
    const long long per_cta = T / G;
    const long long c_max = (RB + per_cta - 1) / per_cta + 1;
    const size_t need = (size_t)kCounterBytes + (size_t)panels * (size_t)c_max * kPanelCols * sizeof(float);
    if (need > kStaticWsBytes)
      return set_error(B200Q_ERR_BAD_SHAPE, "b200q_gemv_static: K=%d, %d panels need %zu bytes of scratch, the static array holds %zu",
                       K, panels, need, (size_t)kStaticWsBytes);
  }
  void* ws = nullptr;
  if (int e = static_ws_ptr(&ws)) return e;
  // the array keeps the previous call's contents; the panel counters must start at zero (the partial sums are
  // written before they are read).  A memset node, not an allocation: legal under stream capture.
  cudaError_t ce = cudaMemsetAsync(ws, 0, kCounterBytes, stream);
  if (ce != cudaSuccess) return check_cuda(ce, "cudaMemsetAsync(static GEMV scratch)");
  return gemv_launch(ops, n_ops, x, x_mul, K, fmt, flags, ws, kStaticWsBytes, stream);
}

}  // namespace b200q
