// Persistent decode-program kernel: EVERY quantized-Linear GEMV of one decode step in ONE launch.
//
// Why: a batch-1 GEMV of 9-75 MB lasts 1.5-11 us at HBM speed, but a dependent kernel boundary costs
// ~5.5-6.5 us on B200 even with programmatic dependent launch (profiles/r1_gemv_v2_sweep.log), so a
// chain of 129 launches can never pass ~35 % of the HBM roofline.  The reference runs the same chain
// as one compiled Relax function (`decoding`, mlc_llm/relax_model/llama.py:581-611); the B200-native
// shape of that boundary is one persistent kernel:
//
//   * grid = #SMs, one CTA per SM, cooperative launch (co-residency guaranteed);
//   * the "program" is an array of steps in device memory; a step = up to four GEMVs that read one
//     activation vector (q|k|v, gate|up, a single Linear), with TMA descriptors, epilogues and the
//     steps it depends on;
//   * a producer warp streams the CTA's share of every step (blocks of 16 word rows x 128 columns,
//     stream-K split, exactly as gemv.cu) into a ~200 KB shared-memory ring with TMA.  Weights are
//     constants, so the producer NEVER waits for a dependency: while the consumers sit in the
//     dependency bubble of step j (last partial sums, flag, re-read of the activation: ~1-2 us) the
//     ring fills with the weights of step j+1, j+2 ... (29 MB in flight chip-wide = 4.5 us of HBM
//     time), so the memory pipe does not drain at step boundaries;
//   * 16 consumer warps in four groups of four; group g owns ring stages g, g+4, ... (one warp = a
//     16-row x 32-column box, the lane mapping and exact-dequant mma.sync contraction of gemv.cu);
//   * K-split panels: partial sums go to per-step slots, the last CTA to arrive adds them in CTA
//     order (deterministic) and applies the epilogue;
//   * dependencies: every finished panel increments done[step]; before staging its activation a CTA
//     polls (ld.acquire.gpu) until the producer steps of x / x_mul / ep0 are complete.  Activations
//     written inside the launch are read with ld.global.cg (L1 is not coherent across SMs).
//   * the last CTA to leave clears the counters, so the launch is re-playable from a CUDA graph.
#include <cuda.h>

#include <cstdlib>
#include <cstring>

#include "common.cuh"
#include "gemv_device.cuh"
#include "ptx.cuh"

namespace b200q {

// gemv.cu
int make_tmap_2d(CUtensorMap* tm, CUtensorMapDataType dt, int elem_bytes, const void* ptr, int inner, int outer,
                 int box_inner, int box_outer, CUtensorMapSwizzle sw);
int device_sm_count();

namespace {

using namespace ptx;
using namespace gemvdev;

constexpr int kGroups = 4;                    // consumer warp groups
constexpr int kWarpsPerGroup = 4;             // one warp per 32-column box of the 128-column block
constexpr int kConsWarps = kGroups * kWarpsPerGroup;
constexpr int kConsThreads = kConsWarps * 32;  // 512
constexpr int kProgThreads = kConsThreads + 32;  // + producer warp
constexpr int kRows = 16;                     // packed word rows per block
constexpr int kBoxBytes = kRows * kBoxCols * 4;          // 2048
constexpr int kWordBytes = kRows * kPanelCols * 4;       // 8192
constexpr int kMaxStages = 24;
constexpr int kMaxOps = 4;
constexpr int kMaxDeps = 4;
constexpr int kHeader = 4096;                 // barriers (2 x 24 x 8) + flag + red[4][128] f32 (2048)
constexpr int kSmemPerSm = 227 * 1024;

template <int FMT>
struct PCfg {
  using F = Fmt<FMT>;
  using scale_t = typename F::scale_t;
  static constexpr bool kF32 = (FMT == B200Q_Q4F32_0);
  static constexpr int kGroupsPerBlock = kRows / F::kRowsPerGroup;  // 4 / 4 / 2
  static constexpr int kScalePitch = kPanelCols * (int)sizeof(scale_t);
  static constexpr int kScaleBytes = kGroupsPerBlock * kScalePitch;
  static constexpr int kStageTx = kWordBytes + kScaleBytes;
  static constexpr int kStageBytes = (kStageTx + 1023) / 1024 * 1024;
  static constexpr int kXRowBytes = XLayout<FMT>::kXRowBytes;
};

struct ProgOp {
  const void* ep0;
  void* out;
  int N;
  int epilogue;
  int panel_begin;
  int pad_;
};

// one step of the program; lives in DEVICE memory (tensor maps must be 64-byte aligned)
struct alignas(64) ProgStep {
  CUtensorMap tm_w[kMaxOps];
  CUtensorMap tm_s[kMaxOps];
  ProgOp op[kMaxOps];
  const void* x;
  const void* x_mul;
  int n_ops;
  int K;
  int n_rows;    // packed word rows
  int RB;        // row blocks per panel
  int n_panels;  // over all ops of the step
  int T;         // blocks of the step
  int c_max;     // partial slots per panel
  int n_deps;
  int deps[kMaxDeps];          // step indices whose outputs this step reads (or overwrites)
  unsigned counter_off;        // panel counters of this step: ws_counters + counter_off
  unsigned pad_;
  unsigned long long partial_off;  // float offset of this step's partial slots
};

struct ProgHeader {  // first 64 bytes of the blob
  int magic, n_steps, fmt, xs_bytes;
  unsigned n_counters;             // panel counters in total
  unsigned pad_;
  unsigned long long partial_floats;  // partial-slot floats in total
  unsigned long long reserved[4];
};
static_assert(sizeof(ProgHeader) == 64, "header is one cache line");
constexpr int kMagic = 0xB2009E;

// workspace layout: [done: n_steps u32][exit: 1 u32][panel counters][pad to 256][partials f32]
__host__ __device__ inline size_t ws_partial_byte_off(int n_steps, unsigned n_counters) {
  size_t b = ((size_t)n_steps + 1 + n_counters) * 4;
  return (b + 255) / 256 * 256;
}

__device__ __forceinline__ void cons_sync() { asm volatile("bar.sync 1, %0;" ::"n"(kConsThreads) : "memory"); }
__device__ __forceinline__ unsigned ld_acquire_u32(const unsigned* p) {
  unsigned v;
  asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ int find_op(const ProgStep& st, int panel) {
  int oi = 0;
#pragma unroll
  for (int t = 1; t < kMaxOps; ++t)
    if (t < st.n_ops && panel >= st.op[t].panel_begin) oi = t;
  return oi;
}
__device__ __forceinline__ int cta_begin(int T, int i, int G) { return (int)((long long)i * T / G); }
__device__ __forceinline__ int block_owner(int T, int t, int G) { return (int)((((long long)t + 1) * G - 1) / T); }

template <int FMT>
__global__ void __launch_bounds__(kProgThreads, 1)
gemv_program_kernel(const ProgStep* __restrict__ prog, int n_steps, unsigned* __restrict__ ws_u32,
                    float* __restrict__ ws_partial, int stages, int xs_bytes, int debug) {
  using C = PCfg<FMT>;
  using F = Fmt<FMT>;
  using scale_t = typename C::scale_t;
  extern __shared__ __align__(128) unsigned char smem[];

  const int warp = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0), lane = threadIdx.x & 31;
  const int G = gridDim.x, b = blockIdx.x;

  uint64_t* bars = reinterpret_cast<uint64_t*>(smem);            // full[kMaxStages], empty[kMaxStages]
  int* flag = reinterpret_cast<int*>(smem + 2 * kMaxStages * 8);  // byte 384
  float* red = reinterpret_cast<float*>(smem + 2048);             // [kGroups][128]
  unsigned char* xs = smem + kHeader;
  const uint32_t stage_base = (smem_u32(smem + kHeader + xs_bytes) + 1023u) & ~1023u;
  const uint32_t full0 = smem_u32(bars), empty0 = full0 + 8 * kMaxStages;
  unsigned* done = ws_u32;
  unsigned* exit_ctr = ws_u32 + n_steps;
  unsigned* counters = ws_u32 + n_steps + 1;

  if (threadIdx.x == 0) {
    for (int s = 0; s < stages; ++s) {
      mbar_init(full0 + 8 * s, 1);
      mbar_init(empty0 + 8 * s, kWarpsPerGroup);
    }
    fence_barrier_init();
  }
  __syncthreads();

  if (warp == kConsWarps) {
    // ================= producer: never waits for a dependency (weights are constants) ==================
    const uint64_t pol = l2_evict_first_policy();
    int s = 0;
    uint32_t ph = 0;
    for (int j = 0; j < n_steps; ++j) {
      const ProgStep& st = prog[j];
      const int T = st.T, RB = st.RB;
      const int t0 = cta_begin(T, b, G), t1 = cta_begin(T, b + 1, G);
      int panel = t0 / RB, rb = t0 - panel * RB;
      for (int t = t0; t < t1; ++t) {
        const int oi = find_op(st, panel);
        const int col0 = (panel - st.op[oi].panel_begin) * kPanelCols;
        mbar_wait(empty0 + 8 * s, ph ^ 1);
        if (elect_one()) {
          const uint32_t bar = full0 + 8 * s;
          mbar_expect_tx(bar, (uint32_t)C::kStageTx);
          const uint32_t sw = stage_base + s * C::kStageBytes;
#pragma unroll
          for (int qb = 0; qb < kPanelCols / kBoxCols; ++qb)
            tma_load_2d_hint(sw + qb * kBoxBytes, &st.tm_w[oi], col0 + qb * kBoxCols, rb * kRows, bar, pol);
          tma_load_2d_hint(sw + kWordBytes, &st.tm_s[oi], col0, rb * C::kGroupsPerBlock, bar, pol);
        }
        __syncwarp();
        if (++s == stages) { s = 0; ph ^= 1; }
        if (++rb == RB) { rb = 0; ++panel; }
      }
    }
    return;
  }

  // ================= consumers ============================================================================
  const int grp = warp >> 2, cq = warp & 3;  // group (ring stages grp, grp + 4, ...), 32-column box
  const int c = lane >> 2, tig = lane & 3;
  // lane rows of the 16-row block: 2 tig + j + 8 m  (j, m in {0,1}); swizzled chunk = c ^ (row & 7)
  const uint32_t w_off0 = cq * kBoxBytes + (2 * tig) * 128 + ((c ^ (2 * tig)) << 4);
  const uint32_t w_off1 = cq * kBoxBytes + (2 * tig + 1) * 128 + ((c ^ (2 * tig + 1)) << 4);
  const int g0 = (2 * tig) / F::kRowsPerGroup, g1 = (8 + 2 * tig) / F::kRowsPerGroup;
  const uint32_t s_col = (cq * 32 + c * 4) * (int)sizeof(scale_t);
  const uint32_t s_off0 = kWordBytes + g0 * C::kScalePitch + s_col;
  const uint32_t s_off1 = kWordBytes + g1 * C::kScalePitch + s_col;
  const uint32_t xs_lane = smem_u32(xs) + (2 * tig) * C::kXRowBytes;
  const int tc = threadIdx.x;

  int qm = 0;       // (CTA-local block sequence number) & 3
  int sg = grp;     // ring slot of this group's next block
  uint32_t phg = 0;

  for (int j = 0; j < n_steps; ++j) {
    const ProgStep& st = prog[j];
    const int T = st.T, RB = st.RB;
    const int t0 = cta_begin(T, b, G), t1 = cta_begin(T, b + 1, G);
    if (t0 >= t1) continue;  // CTA-uniform

    // ---- dependencies: the steps that produce x / x_mul / ep0 must be complete ------------------------
    if (tc < st.n_deps && !(debug & 1)) {
      const int d = st.deps[tc];
      const unsigned need = (unsigned)prog[d].n_panels;
      while (ld_acquire_u32(done + d) < need) __nanosleep(32);
    }
    cons_sync();
    stage_x<FMT, kConsThreads, true>(xs, st.x, st.x_mul, st.K, RB * kRows);
    cons_sync();

    int panel = t0 / RB, rb = t0 - panel * RB;
    int t = t0;
    while (t < t1) {
      const int rb_begin = rb;
      const int seg_end = min(t1, t + (RB - rb));
      float acc[2][4];
#pragma unroll
      for (int m = 0; m < 2; ++m)
#pragma unroll
        for (int r = 0; r < 4; ++r) acc[m][r] = 0.f;

      for (; t < seg_end; ++t, ++rb) {
        if (qm == grp) {
          mbar_wait(full0 + 8 * sg, phg);
          const uint32_t sw = stage_base + sg * C::kStageBytes;
          const uint4 w00 = lds128(sw + w_off0), w01 = lds128(sw + w_off1);
          const uint4 w10 = lds128(sw + w_off0 + 8 * 128), w11 = lds128(sw + w_off1 + 8 * 128);
          const uint32_t xr = xs_lane + rb * kRows * C::kXRowBytes;
          if constexpr (C::kF32) {
            const uint4 sc0 = lds128(sw + s_off0), sc1 = lds128(sw + s_off1);
            compute_row_f32(w00, sc0, xr, acc[0]);
            compute_row_f32(w01, sc0, xr + C::kXRowBytes, acc[0]);
            compute_row_f32(w10, sc1, xr + 8 * C::kXRowBytes, acc[0]);
            compute_row_f32(w11, sc1, xr + 9 * C::kXRowBytes, acc[0]);
          } else if (debug & 6) {
            // experiments: 2 = unpack only (no MMA), 4 = MMA only (no unpack); results are garbage
            const uint4 ws4[4] = {w00, w01, w10, w11};
            __half2 s2[4];
            split_scales(lds64(sw + s_off0), s2);
#pragma unroll
            for (int r = 0; r < 4; ++r) {
              const uint32_t wv[4] = {ws4[r].x, ws4[r].y, ws4[r].z, ws4[r].w};
              const uint4 xv = lds128(xr + (r & 1) * C::kXRowBytes + (r >> 1) * 8 * C::kXRowBytes);
              if (debug & 2) {
                __half2 h[4];
#pragma unroll
                for (int q4 = 0; q4 < 4; ++q4) {
                  if constexpr (FMT == B200Q_Q4F16_0) unpack_q4(wv[q4], s2[q4], h);
                  acc[0][q4] += __uint_as_float(h2_as_u32(h[0]) ^ h2_as_u32(h[1]) ^ h2_as_u32(h[2]) ^ h2_as_u32(h[3]) ^ xv.x);
                }
              } else {
                mma16816(acc[0], wv[0], wv[1], wv[2], wv[3], xv.x, xv.y);
                mma16816(acc[0], wv[1], wv[2], wv[3], wv[0], xv.z, xv.w);
                mma16816(acc[1], wv[2], wv[3], wv[0], wv[1], xv.x, xv.y);
                mma16816(acc[1], wv[3], wv[0], wv[1], wv[2], xv.z, xv.w);
              }
            }
          } else {
            __half2 s2[4];
            split_scales(lds64(sw + s_off0), s2);
            compute_row_f16<FMT>(w00, s2, xr, acc);
            compute_row_f16<FMT>(w01, s2, xr + C::kXRowBytes, acc);
            split_scales(lds64(sw + s_off1), s2);
            compute_row_f16<FMT>(w10, s2, xr + 8 * C::kXRowBytes, acc);
            compute_row_f16<FMT>(w11, s2, xr + 9 * C::kXRowBytes, acc);
          }
          __syncwarp();
          if (lane == 0) mbar_arrive(empty0 + 8 * sg);
          sg += kGroups;
          if (sg >= stages) { sg -= stages; phg ^= 1; }
        }
        qm = (qm + 1) & (kGroups - 1);
      }

      // ---- reduce the segment over the four groups, then across CTAs ------------------------------------
      float4 v4;
      if constexpr (C::kF32) {
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          acc[0][i] += __shfl_xor_sync(0xffffffffu, acc[0][i], 1);
          acc[0][i] += __shfl_xor_sync(0xffffffffu, acc[0][i], 2);
        }
        v4 = make_float4(acc[0][0], acc[0][1], acc[0][2], acc[0][3]);
      } else {
        v4 = make_float4(acc[0][0], acc[0][2], acc[1][0], acc[1][2]);  // columns 4c .. 4c+3
      }
      if (tig == 0) *reinterpret_cast<float4*>(red + grp * kPanelCols + cq * 32 + c * 4) = v4;
      cons_sync();

      const ProgOp& op = st.op[find_op(st, panel)];
      const int n = (panel - op.panel_begin) * kPanelCols + tc;
      const bool col_thread = (tc < kPanelCols);
      float sum = 0.f;
      if (col_thread) sum = (red[tc] + red[kPanelCols + tc]) + (red[2 * kPanelCols + tc] + red[3 * kPanelCols + tc]);
      const bool complete = (rb_begin == 0) && (rb == RB);
      bool finalize = true;  // CTA-uniform
      if (!complete) {
        // contributors of the panel, in block order.  T >= G: every CTA between the owners of the first
        // and the last block holds at least one block; T < G: a CTA holds at most one block, so every
        // block of the panel is a different contributor
        const int first = block_owner(T, panel * RB, G);
        const int last = block_owner(T, panel * RB + RB - 1, G);
        const bool sparse = (T < G);
        const int n_contrib = sparse ? RB : last - first + 1;
        const int my_slot = sparse ? rb_begin : b - first;
        float* slot0 = ws_partial + st.partial_off + (size_t)panel * st.c_max * kPanelCols;
        if (col_thread) {
          slot0[(size_t)my_slot * kPanelCols + tc] = sum;
          __threadfence();
        }
        cons_sync();
        if (tc == 0) {
          unsigned* ctr = counters + st.counter_off + panel;
          const unsigned old = atomicAdd(ctr, 1u);
          const int is_last = (old == (unsigned)(n_contrib - 1));
          if (is_last) *ctr = 0u;  // self-cleaning
          *flag = is_last;
        }
        cons_sync();
        finalize = (*flag != 0);
        if (finalize && col_thread) {
          __threadfence();
          sum = 0.f;
          for (int o = 0; o < n_contrib; ++o) sum += ld_cg_f32(slot0 + (size_t)o * kPanelCols + tc);
        }
      }
      if (finalize) {
        if (col_thread && n < op.N) {
          if constexpr (C::kF32) {
            float e0 = 0.f;
            if (op.ep0) e0 = __ldcg(reinterpret_cast<const float*>(op.ep0) + n);
            reinterpret_cast<float*>(op.out)[n] = ep_f32(sum, op.epilogue, &e0, 0);
          } else {
            if (op.epilogue == B200Q_EP_CAST_F32) {
              reinterpret_cast<float*>(op.out)[n] = sum;
            } else {
              __half e0 = __ushort_as_half(0);
              if (op.ep0) e0 = __ldcg(reinterpret_cast<const __half*>(op.ep0) + n);
              reinterpret_cast<__half*>(op.out)[n] = ep_f16(sum, op.epilogue, &e0, 0);
            }
          }
        }
        if (col_thread) __threadfence();
        cons_sync();
        if (tc == 0) atomicAdd(done + j, 1u);  // one finished panel
      }
      if (rb == RB) { rb = 0; ++panel; }
    }
  }

  // ---- leave: the last CTA clears the step counters so the launch can be replayed ----------------------
  if (tc == 0) {
    __threadfence();
    const unsigned old = atomicAdd(exit_ctr, 1u);
    if (old == (unsigned)(G - 1)) {
      for (int j = 0; j < n_steps; ++j) done[j] = 0u;
      *exit_ctr = 0u;
      __threadfence();
    }
  }
}

// ---- host side -----------------------------------------------------------------------------------------
struct Range {
  uintptr_t lo, hi;
  bool overlaps(const Range& o) const { return lo < o.hi && o.lo < hi && lo != hi && o.lo != o.hi; }
};
Range range_of(const void* p, size_t bytes) {
  if (!p) return Range{0, 0};
  return Range{reinterpret_cast<uintptr_t>(p), reinterpret_cast<uintptr_t>(p) + bytes};
}

template <int FMT>
int build_step(ProgStep& st, const b200q_gemv_step& in, int G) {
  using C = PCfg<FMT>;
  using F = Fmt<FMT>;
  st.n_ops = in.n_ops;
  st.x = in.x;
  st.x_mul = in.x_mul;
  st.K = in.K;
  st.n_rows = n_words(FMT, in.K);
  st.RB = ceil_div(st.n_rows, kRows);
  const int n_grp = st.n_rows / F::kRowsPerGroup;
  int panels = 0;
  for (int i = 0; i < in.n_ops; ++i) {
    const b200q_gemv_op& o = in.ops[i];
    st.op[i].ep0 = o.ep0;
    st.op[i].out = o.out;
    st.op[i].N = o.N;
    st.op[i].epilogue = o.epilogue;
    st.op[i].panel_begin = panels;
    panels += ceil_div(o.N, kPanelCols);
    if (int e = make_tmap_2d(&st.tm_w[i], CU_TENSOR_MAP_DATA_TYPE_UINT32, 4, o.packed, o.N, st.n_rows, kBoxCols, kRows,
                             CU_TENSOR_MAP_SWIZZLE_128B))
      return e;
    if (int e = make_tmap_2d(&st.tm_s[i], C::kF32 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT32 : CU_TENSOR_MAP_DATA_TYPE_FLOAT16,
                             (int)sizeof(typename C::scale_t), o.scale, o.N, n_grp, kPanelCols, C::kGroupsPerBlock,
                             CU_TENSOR_MAP_SWIZZLE_NONE))
      return e;
  }
  st.n_panels = panels;
  st.T = panels * st.RB;
  const int per_cta = st.T / G > 0 ? st.T / G : 1;
  st.c_max = ceil_div(st.RB, per_cta) + 1;
  return B200Q_OK;
}

int xs_bytes_for(int fmt, int K) {
  const int rows = ceil_div(n_words(fmt, K), kRows) * kRows;
  const int xb = (fmt == B200Q_Q4F32_0) ? 32 : (fmt == B200Q_Q3F16_0 ? 24 : (fmt == B200Q_Q8F16_0 ? 8 : 16));
  return (rows * xb + 127) / 128 * 128;
}
int stage_bytes_for(int fmt) {
  switch (fmt) {
    case B200Q_Q4F16_0: return PCfg<B200Q_Q4F16_0>::kStageBytes;
    case B200Q_Q3F16_0: return PCfg<B200Q_Q3F16_0>::kStageBytes;
    case B200Q_Q8F16_0: return PCfg<B200Q_Q8F16_0>::kStageBytes;
    default: return PCfg<B200Q_Q4F32_0>::kStageBytes;
  }
}

}  // namespace

size_t gemv_program_bytes(int n_steps) { return sizeof(ProgHeader) + (size_t)n_steps * sizeof(ProgStep); }

int gemv_program_build(const b200q_gemv_step* steps, int n_steps, int fmt, void* host_blob, size_t blob_bytes) {
  if (blob_bytes < gemv_program_bytes(n_steps)) return set_error(B200Q_ERR_BAD_SHAPE, "gemv_program_build: blob too small");
  if ((reinterpret_cast<uintptr_t>(host_blob) & 7) != 0) return set_error(B200Q_ERR_BAD_ALIGN, "gemv_program_build: blob must be 8-byte aligned");
  std::memset(host_blob, 0, gemv_program_bytes(n_steps));
  ProgHeader* h = reinterpret_cast<ProgHeader*>(host_blob);
  // the step array is copied to a 64-byte aligned device address by the caller; fill through a bounce
  // copy because the host blob itself need not be 64-byte aligned
  unsigned char* base = reinterpret_cast<unsigned char*>(host_blob) + sizeof(ProgHeader);
  const int G = device_sm_count();
  const size_t el = fmt_is_f32(fmt) ? 4 : 2;
  unsigned n_counters = 0;
  unsigned long long partial_floats = 0;
  int xs_bytes = 0;
  for (int j = 0; j < n_steps; ++j) {
    alignas(64) ProgStep st;
    std::memset(&st, 0, sizeof(st));
    int e;
    switch (fmt) {
      case B200Q_Q4F16_0: e = build_step<B200Q_Q4F16_0>(st, steps[j], G); break;
      case B200Q_Q3F16_0: e = build_step<B200Q_Q3F16_0>(st, steps[j], G); break;
      case B200Q_Q8F16_0: e = build_step<B200Q_Q8F16_0>(st, steps[j], G); break;
      case B200Q_Q4F32_0: e = build_step<B200Q_Q4F32_0>(st, steps[j], G); break;
      default: return set_error(B200Q_ERR_UNSUPPORTED, "gemv_program_build: fmt %d", fmt);
    }
    if (e) return e;
    st.counter_off = n_counters;
    n_counters += (unsigned)st.n_panels;
    st.partial_off = partial_floats;
    partial_floats += (unsigned long long)st.n_panels * st.c_max * kPanelCols;
    const int xb = xs_bytes_for(fmt, st.K);
    if (xb > xs_bytes) xs_bytes = xb;

    // hazards against every earlier step: RAW (we read what it writes), WAW, WAR (we overwrite what it reads)
    Range in_j[2 + kMaxOps], out_j[kMaxOps];
    int n_in = 0;
    in_j[n_in++] = range_of(steps[j].x, el * steps[j].K);
    in_j[n_in++] = range_of(steps[j].x_mul, el * steps[j].K);
    for (int i = 0; i < steps[j].n_ops; ++i) {
      const b200q_gemv_op& o = steps[j].ops[i];
      const size_t oel = (o.epilogue == B200Q_EP_CAST_F32) ? 4 : el;
      in_j[n_in++] = range_of(o.ep0, el * o.N);
      out_j[i] = range_of(o.out, oel * o.N);
    }
    int deps[kMaxDeps];
    int n_deps = 0;
    for (int i = j - 1; i >= 0; --i) {
      bool hazard = false;
      for (int a = 0; a < steps[i].n_ops && !hazard; ++a) {
        const b200q_gemv_op& oi = steps[i].ops[a];
        const Range out_i = range_of(oi.out, ((oi.epilogue == B200Q_EP_CAST_F32) ? 4 : el) * oi.N);
        for (int q = 0; q < n_in; ++q) hazard |= out_i.overlaps(in_j[q]);
        for (int q = 0; q < steps[j].n_ops; ++q) {
          hazard |= out_i.overlaps(out_j[q]);
          hazard |= out_j[q].overlaps(range_of(oi.ep0, el * oi.N));
        }
      }
      for (int q = 0; q < steps[j].n_ops && !hazard; ++q)
        hazard = out_j[q].overlaps(range_of(steps[i].x, el * steps[i].K)) ||
                 out_j[q].overlaps(range_of(steps[i].x_mul, el * steps[i].K));
      if (!hazard) continue;
      // already implied by a dependency we hold?  (i is an ancestor of a held dep d if d depends on it, transitively)
      bool implied = false;
      for (int d = 0; d < n_deps && !implied; ++d) {
        // walk the dependency DAG of deps[d] (steps are few; depth-first over recorded deps)
        int stack[256];
        int sp = 0;
        stack[sp++] = deps[d];
        while (sp > 0 && !implied) {
          const ProgStep* ps = reinterpret_cast<const ProgStep*>(base + (size_t)stack[--sp] * sizeof(ProgStep));
          ProgStep tmp;
          std::memcpy(&tmp, ps, sizeof(tmp));
          for (int q = 0; q < tmp.n_deps; ++q) {
            if (tmp.deps[q] == i) implied = true;
            else if (tmp.deps[q] > i && sp < 256) stack[sp++] = tmp.deps[q];
          }
        }
      }
      if (implied) continue;
      if (n_deps == kMaxDeps)
        return set_error(B200Q_ERR_UNSUPPORTED, "gemv_program_build: step %d has more than %d independent producers", j, kMaxDeps);
      deps[n_deps++] = i;
    }
    st.n_deps = n_deps;
    for (int d = 0; d < n_deps; ++d) st.deps[d] = deps[d];
    std::memcpy(base + (size_t)j * sizeof(ProgStep), &st, sizeof(st));
  }
  h->magic = kMagic;
  h->n_steps = n_steps;
  h->fmt = fmt;
  h->xs_bytes = xs_bytes;
  h->n_counters = n_counters;
  h->partial_floats = partial_floats;
  h->reserved[0] = (unsigned long long)G;
  const int stages = (kSmemPerSm - kHeader - xs_bytes - 1024) / stage_bytes_for(fmt);
  if (stages < 2 * kGroups)
    return set_error(B200Q_ERR_BAD_SHAPE, "gemv_program_build: K too large for the activation stage (%d bytes)", xs_bytes);
  return B200Q_OK;
}

size_t gemv_program_workspace_bytes(const void* host_blob) {
  const ProgHeader* h = reinterpret_cast<const ProgHeader*>(host_blob);
  if (h->magic != kMagic) return 0;
  return ws_partial_byte_off(h->n_steps, h->n_counters) + (size_t)h->partial_floats * 4;
}

int gemv_program_run(const void* host_blob, const void* dev_blob, void* workspace, size_t workspace_bytes,
                     cudaStream_t stream) {
  const ProgHeader* h = reinterpret_cast<const ProgHeader*>(host_blob);
  if (h->magic != kMagic) return set_error(B200Q_ERR_BAD_SHAPE, "gemv_program_run: not a program blob");
  if ((reinterpret_cast<uintptr_t>(dev_blob) & 63) != 0) return set_error(B200Q_ERR_BAD_ALIGN, "gemv_program_run: device blob must be 64-byte aligned");
  if (workspace_bytes < gemv_program_workspace_bytes(host_blob) || (reinterpret_cast<uintptr_t>(workspace) & 255) != 0)
    return set_error(B200Q_ERR_BAD_SHAPE, "gemv_program_run: workspace too small or not 256-byte aligned");
  const int G = (int)h->reserved[0];
  if (G != device_sm_count()) return set_error(B200Q_ERR_BAD_SHAPE, "gemv_program_run: program was built for %d SMs", G);
  int stages = (kSmemPerSm - kHeader - h->xs_bytes - 1024) / stage_bytes_for(h->fmt);
  if (stages > kMaxStages) stages = kMaxStages;
  const size_t smem = (size_t)kHeader + h->xs_bytes + 1024 + (size_t)stages * stage_bytes_for(h->fmt);

  const ProgStep* prog = reinterpret_cast<const ProgStep*>(reinterpret_cast<const unsigned char*>(dev_blob) + sizeof(ProgHeader));
  unsigned* ws_u32 = reinterpret_cast<unsigned*>(workspace);
  float* ws_partial = reinterpret_cast<float*>(reinterpret_cast<unsigned char*>(workspace) + ws_partial_byte_off(h->n_steps, h->n_counters));
  int n_steps = h->n_steps, xs_bytes = h->xs_bytes;
  static const int dbg_env = getenv("B200Q_PROG_DEBUG") ? atoi(getenv("B200Q_PROG_DEBUG")) : 0;  // experiments: 1 = ignore dependencies
  int debug = dbg_env;

  const void* kern = nullptr;
  switch (h->fmt) {
    case B200Q_Q4F16_0: kern = (const void*)gemv_program_kernel<B200Q_Q4F16_0>; break;
    case B200Q_Q3F16_0: kern = (const void*)gemv_program_kernel<B200Q_Q3F16_0>; break;
    case B200Q_Q8F16_0: kern = (const void*)gemv_program_kernel<B200Q_Q8F16_0>; break;
    default: kern = (const void*)gemv_program_kernel<B200Q_Q4F32_0>; break;
  }
  cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  if (e != cudaSuccess) return check_cuda(e, "cudaFuncSetAttribute(gemv program smem)");
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3(G, 1, 1);
  cfg.blockDim = dim3(kProgThreads, 1, 1);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = stream;
  cudaLaunchAttribute attrs[1];
  attrs[0].id = cudaLaunchAttributeCooperative;  // CTAs wait on one another: they must all be resident
  attrs[0].val.cooperative = 1;
  cfg.attrs = attrs;
  cfg.numAttrs = 1;
  void* args[] = {(void*)&prog, (void*)&n_steps, (void*)&ws_u32, (void*)&ws_partial, (void*)&stages, (void*)&xs_bytes, (void*)&debug};
  e = cudaLaunchKernelExC(&cfg, kern, args);
  if (e != cudaSuccess) return check_cuda(e, "cudaLaunchKernelExC(gemv program)");
  count_launch();
  return B200Q_OK;
}

}  // namespace b200q
