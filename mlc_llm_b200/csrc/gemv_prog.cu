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
//   * the compute warps never touch global memory: at the end of a panel segment they hand their sums
//     to four helper warps through shared memory (double-buffered, mbarriers) and go on;
//   * everything that flows between CTAs inside the launch is "LL" data, NCCL-style: every value is
//     stored together with a launch tag ({f16 | tag} in one 32-bit word, {f32, tag} in one 64-bit word)
//     and the reader polls the word itself - no counters, no atomics, no fences, one round trip:
//       - K-split panels: contributors store tagged partial sums into per-step slots; the CTA that
//         owns the panel's last block polls them, adds them in block order (deterministic), applies
//         the epilogue, and stores the result both to the caller's `out` and to the LL activation array;
//       - activations: the helper warps of every CTA poll the LL words of x (x_mul, ep0) of the next
//         step and stage them into the other half of a double-buffered shared-memory activation;
//   * the tag is a launch counter kept in the workspace (bumped by the last CTA to leave), so the same
//     launch can be replayed from a CUDA graph without clearing anything.
#include <cuda.h>

#include <atomic>
#include <cstdlib>
#include <cstring>
#include <cstdio>

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
constexpr int kHelperWarps = 4;               // 128 helper threads = one per panel column
constexpr int kHelperThreads = kHelperWarps * 32;
constexpr int kProducerWarp = kConsWarps;     // warp 16
constexpr int kFirstHelperWarp = kConsWarps + 1;
constexpr int kProgThreads = (kConsWarps + 1 + kHelperWarps) * 32;  // 672
constexpr int kRows = 16;                     // packed word rows per block
constexpr int kBoxBytes = kRows * kBoxCols * 4;          // 2048
constexpr int kWordBytes = kRows * kPanelCols * 4;       // 8192
constexpr int kMaxStages = 24;
constexpr int kMaxOps = 4;
// shared-memory header: ring barriers [2 x 24], hand-off barriers [8], red[2][4][128] f32 at 1024
constexpr int kBarRedFull = 2 * kMaxStages, kBarRedEmpty = kBarRedFull + 2, kBarXFull = kBarRedEmpty + 2,
              kBarXEmpty = kBarXFull + 2;
constexpr int kRedOff = 1024;
constexpr int kHeader = kRedOff + 2 * kGroups * kPanelCols * 4;  // 5120
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
  static constexpr int kLLBytes = kF32 ? 8 : 4;  // one LL activation element
};

struct ProgOp {
  const void* ep0;  // plain epilogue operand (used when ep0_ll < 0)
  void* out;
  int N;
  int epilogue;
  int panel_begin;
  int tp_reduce;  // 1: row-parallel op - the value is summed over the tensor-parallel ranks inside the program
  long long ep0_ll;  // LL element offset of the operand's producer, or -1
  long long out_ll;  // LL element offset of this output, or -1 (f32 logits of an f16 model)
};

// one step of the program; lives in DEVICE memory (tensor maps must be 64-byte aligned)
struct alignas(64) ProgStep {
  CUtensorMap tm_w[kMaxOps];
  CUtensorMap tm_s[kMaxOps];
  ProgOp op[kMaxOps];
  const void* x;      // plain activation (used when x_ll < 0)
  const void* x_mul;  // plain, optional
  long long x_ll;     // LL element offset, or -1
  long long xmul_ll;  // LL element offset, or -1 (x_mul == nullptr: unused)
  int n_ops;
  int K;
  int n_rows;    // packed word rows
  int RB;        // row blocks per panel
  int n_panels;  // over all ops of the step
  int T;         // blocks of the step
  int c_max;     // partial slots per panel
  int has_xmul;
  unsigned long long partial_off;  // uint2 offset of this step's partial slots
  unsigned long long tp_partial_off;  // uint2 offset of the step's [panel][rank][128] cross-GPU partial slots
};

struct ProgHeader {  // first 64 bytes of the blob
  int magic, n_steps, fmt, xs_bytes;
  int G, tc;  // tc: 1 = the tcgen05 kernel of gemv_tc.cuh (q4f16_0)
  unsigned long long ll_elems;       // LL activation elements in total
  unsigned long long partial_slots;  // {f32, tag} partial-sum words in total
  int tp_rank, tp_size;
  unsigned prog_id;  // non-zero hash of the step table: stamped into / checked against ws_ctrl[3] by every launch
  unsigned reserved0;
  unsigned long long reserved1;
};
static_assert(sizeof(ProgHeader) == 64, "header is one cache line");
constexpr int kMagic = 0xB2009F;

// workspace layout: [epoch u32, exit u32, pad to 256][LL activations][pad to 256][partial slots uint2]
inline size_t ws_ll_off() { return 256; }
inline size_t ws_partial_off(unsigned long long ll_elems, int ll_bytes) {
  return (256 + (size_t)ll_elems * ll_bytes + 255) / 256 * 256;
}

// ---- watchdog: every cross-CTA / cross-GPU spin of the program kernels is bounded ---------------------------
// A poll that has not been satisfied after `ns` nanoseconds (a peer rank that never launched or replayed a
// different number of times, a workspace shared between two programs, a corrupted tag) writes a status word
// {0xDEAD | site << 16, step} through `status` (normally pinned host memory, so it survives the dead context)
// and traps: the launch fails with a CUDA error instead of hanging the GPU.  All of it sits on the retry path
// of the polls; a poll that is satisfied at once costs nothing.
struct Wd {
  unsigned long long* status;  // may be NULL
  unsigned max_spins;          // polls one wait may take before it expires; 0: no bound
};
enum WdSite { kWdActivation = 1, kWdPartial = 2, kWdEpOperand = 3, kWdTpPartial = 4, kWdProgId = 5 };
__device__ __noinline__ void wd_expired(const Wd& wd, unsigned site, unsigned step) {
  if (wd.status) {
    *reinterpret_cast<volatile unsigned long long*>(wd.status) =
        ((unsigned long long)step << 32) | ((unsigned long long)site << 16) | 0xDEADull;
    __threadfence_system();
  }
  asm volatile("trap;");
}
// The bound is a poll COUNT (one 32-bit register, live only inside the retry loop, two instructions per retry):
// a time stamp per wait costs two more registers at the kernel's register cap and showed up as -3 % tokens/s.
// One poll is an L2 (or NVLink) round trip of >= ~0.35 us, so the host converts the time-out into polls with that
// figure: a wait expires no earlier than the configured time and at most a few times later.
struct WdSpin {
  unsigned spins = 0;
};
__device__ __forceinline__ void wd_poll(WdSpin& w, const Wd& wd, unsigned site, unsigned step) {
#ifdef B200Q_NO_WATCHDOG  // A/B builds only
  return;
#endif
#ifndef B200Q_WD_ACTIVATION
  // The activation polls sit on the critical path of every dependent step (the activation warps of the tcgen05
  // program); counting there costs 1.5 % tokens/s.  They are left unbounded by default: a wait that can never be
  // satisfied (dead peer, foreign workspace) is always ALSO seen by the epilogue warps' bounded polls of the same
  // step (peer partial sums, K-split partial sums, epilogue operand) or by the workspace id check, and their trap
  // ends the whole launch.  -DB200Q_WD_ACTIVATION bounds these polls as well.
  if (site == kWdActivation) return;
#endif
  if (++w.spins == wd.max_spins) wd_expired(wd, site, step);
}

__device__ __forceinline__ unsigned ld_acquire_u32(const unsigned* p) {
  unsigned v;
  asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ uint32_t ld_vol_u32(const void* p) {
  uint32_t v;
  asm volatile("ld.volatile.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ uint2 ld_vol_v2(const void* p) {
  uint2 v;
  asm volatile("ld.volatile.global.v2.u32 {%0,%1}, [%2];" : "=r"(v.x), "=r"(v.y) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_vol_u32(void* p, uint32_t v) {
  asm volatile("st.volatile.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ void st_vol_v2(void* p, uint32_t a, uint32_t b) {
  asm volatile("st.volatile.global.v2.u32 [%0], {%1,%2};" ::"l"(p), "r"(a), "r"(b) : "memory");
}
__device__ __forceinline__ int find_op(const ProgStep& st, int panel) {
  int oi = 0;
#pragma unroll
  for (int t = 1; t < kMaxOps; ++t)
    if (t < st.n_ops && panel >= st.op[t].panel_begin) oi = t;
  return oi;
}
// (also evaluated on the host: gemv_program_schedule_debug below, tests/test_gemv_schedule_cpu.py)
__host__ __device__ __forceinline__ int cta_begin(int T, int i, int G) { return (int)((long long)i * T / G); }
__host__ __device__ __forceinline__ int block_owner(int T, int t, int G) { return (int)((((long long)t + 1) * G - 1) / T); }
// partial-sum slots reserved per panel: every CTA holds >= T / G blocks, so the RB blocks of a panel span at most
// this many CTAs (T < G: one CTA per block, slot = row block)
inline int partial_slots_per_panel(int T, int RB, int G) {
  const int per_cta = T / G > 0 ? T / G : 1;
  return ceil_div(RB, per_cta) + 1;
}

// A workspace belongs to ONE program: the first launch stamps the program's id into ws_ctrl[3], every later launch
// checks it (a workspace reused by another program would pick up stale tags silently).
__device__ __forceinline__ void check_prog_id(unsigned* ws_ctrl, unsigned prog_id, const Wd& wd, bool elected) {
  if (elected) {
    const unsigned old = atomicCAS(ws_ctrl + 3, 0u, prog_id);
    if (old != 0u && old != prog_id) wd_expired(wd, kWdProgId, 0u);
  }
}

// ---- LL element access ------------------------------------------------------------------------------------
// f16 models: one 32-bit word {tag << 16 | f16 bits}; f32 model (q4f32_0): one 64-bit word {f32 bits, tag}
__device__ __forceinline__ __half ll_poll_f16(const uint32_t* p, uint32_t tag, const Wd& wd, unsigned site, unsigned step) {
  uint32_t w = ld_vol_u32(p);
  if ((w >> 16) != tag) {
    WdSpin sp;
    do {
      wd_poll(sp, wd, site, step);
      w = ld_vol_u32(p);
    } while ((w >> 16) != tag);
  }
  return __ushort_as_half((unsigned short)(w & 0xFFFFu));
}
__device__ __forceinline__ float ll_poll_f32(const uint2* p, uint32_t tag, const Wd& wd, unsigned site, unsigned step) {
  uint2 w = ld_vol_v2(p);
  if (w.y != tag) {
    WdSpin sp;
    do {
      wd_poll(sp, wd, site, step);
      w = ld_vol_v2(p);
    } while (w.y != tag);
  }
  return __uint_as_float(w.x);
}
// kNpu consecutive f16 LL elements (one packed word row of x); elements >= n_valid read as zero
template <int NPU>
__device__ __forceinline__ void ll_load_row_f16(const uint32_t* ll, int n_valid, uint32_t tag, __half (&v)[12], const Wd& wd,
                                                unsigned step) {
  uint32_t w[NPU];
  bool ok;
  WdSpin sp;
  do {
    wd_poll(sp, wd, kWdActivation, step);
    ok = true;
#pragma unroll
    for (int t = 0; t < NPU; t += 2) {
      if (t + 1 < n_valid) {
        const uint2 pr = ld_vol_v2(ll + t);  // rows start 8-byte aligned (NPU even, arrays 16-byte aligned)
        w[t] = pr.x;
        w[t + 1] = pr.y;
      } else {
        w[t] = (t < n_valid) ? ld_vol_u32(ll + t) : (tag << 16);
        w[t + 1] = tag << 16;
      }
    }
#pragma unroll
    for (int t = 0; t < NPU; ++t) ok = ok && ((w[t] >> 16) == tag);
  } while (!ok);
#pragma unroll
  for (int t = 0; t < NPU; ++t) v[t] = __ushort_as_half((unsigned short)(w[t] & 0xFFFFu));
}
template <int NPU>
__device__ __forceinline__ void plain_load_row_f16(const __half* x, int k0, int K, __half (&v)[12]) {
#pragma unroll
  for (int t = 0; t < NPU; ++t) v[t] = (k0 + t < K) ? __ldcg(x + k0 + t) : __ushort_as_half(0);
}

// helper warps: stage the activation of a step (x, optionally x * x_mul) into `xs`
template <int FMT>
__device__ __forceinline__ void helper_stage_x(unsigned char* xs, const ProgStep& st, const void* ws_ll, uint32_t tag,
                                               int ht, const Wd& wd, unsigned step) {
  using C = PCfg<FMT>;
  using F = Fmt<FMT>;
  const int K = st.K, rows = st.RB * kRows;
  if constexpr (C::kF32) {
    const uint2* ll = reinterpret_cast<const uint2*>(ws_ll);
    const float* x = reinterpret_cast<const float*>(st.x);
    const float* m = reinterpret_cast<const float*>(st.x_mul);
    float* dst = reinterpret_cast<float*>(xs);
    const int n = rows * 8;
    for (int k = ht; k < n; k += kHelperThreads) {
      float v = 0.f;
      if (k < K) {
        v = (st.x_ll >= 0) ? ll_poll_f32(ll + st.x_ll + k, tag, wd, kWdActivation, step) : __ldcg(x + k);
        if (st.has_xmul)
          v = __fmul_rn(v, (st.xmul_ll >= 0) ? ll_poll_f32(ll + st.xmul_ll + k, tag, wd, kWdActivation, step) : __ldcg(m + k));
      }
      dst[k] = v;
    }
  } else {
    const uint32_t* ll = reinterpret_cast<const uint32_t*>(ws_ll);
    const __half* x = reinterpret_cast<const __half*>(st.x);
    const __half* m = reinterpret_cast<const __half*>(st.x_mul);
    constexpr int XB = C::kXRowBytes;
    for (int lr = ht; lr < rows; lr += kHelperThreads) {
      const int k0 = lr * F::kNpu;
      const int n_valid = min(F::kNpu, max(0, K - k0));
      __align__(16) __half v[12];
#pragma unroll
      for (int t = 0; t < 12; ++t) v[t] = __ushort_as_half(0);
      if (n_valid > 0) {
        if (st.x_ll >= 0) ll_load_row_f16<F::kNpu>(ll + st.x_ll + k0, n_valid, tag, v, wd, step);
        else plain_load_row_f16<F::kNpu>(x, k0, K, v);
        if (st.has_xmul) {
          __align__(16) __half mv[12];
          if (st.xmul_ll >= 0) ll_load_row_f16<F::kNpu>(ll + st.xmul_ll + k0, n_valid, tag, mv, wd, step);
          else plain_load_row_f16<F::kNpu>(m, k0, K, mv);
#pragma unroll
          for (int t = 0; t < F::kNpu; ++t) v[t] = __hmul(v[t], mv[t]);
        }
      }
      __align__(8) __half o[12];
      permute_x_row<FMT>(v, o);
      unsigned char* d = xs + (size_t)lr * XB;
#pragma unroll
      for (int t = 0; t < XB / 8; ++t) reinterpret_cast<uint2*>(d)[t] = reinterpret_cast<const uint2*>(o)[t];
    }
  }
}

template <int FMT>
__global__ void __launch_bounds__(kProgThreads, 1)
gemv_program_kernel(const ProgStep* __restrict__ prog, int n_steps, unsigned* ws_ctrl, void* ws_ll, uint2* ws_partial,
                    int stages, int xs_bytes, unsigned prog_id, const __grid_constant__ Wd wd) {
  using C = PCfg<FMT>;
  using F = Fmt<FMT>;
  using scale_t = typename C::scale_t;
  extern __shared__ __align__(128) unsigned char smem[];

  const int warp = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0), lane = threadIdx.x & 31;
  const int G = gridDim.x, b = blockIdx.x;

  uint64_t* bars = reinterpret_cast<uint64_t*>(smem);
  float* red = reinterpret_cast<float*>(smem + kRedOff);  // [2][kGroups][128]
  unsigned char* xs0 = smem + kHeader;                    // two activation buffers of xs_bytes each
  const uint32_t stage_base = (smem_u32(smem + kHeader + 2 * xs_bytes) + 1023u) & ~1023u;
  const uint32_t bar0 = smem_u32(bars);
  const uint32_t full0 = bar0, empty0 = bar0 + 8 * kMaxStages;
  const uint32_t red_full = bar0 + 8 * kBarRedFull, red_empty = bar0 + 8 * kBarRedEmpty;
  const uint32_t x_full = bar0 + 8 * kBarXFull, x_empty = bar0 + 8 * kBarXEmpty;

  if (threadIdx.x == 0) {
    for (int s = 0; s < stages; ++s) {
      mbar_init(full0 + 8 * s, 1);
      mbar_init(empty0 + 8 * s, kWarpsPerGroup);
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(red_full + 8 * i, kConsWarps);
      mbar_init(red_empty + 8 * i, kHelperWarps);
      mbar_init(x_full + 8 * i, kHelperWarps);
      mbar_init(x_empty + 8 * i, kConsWarps);
    }
    fence_barrier_init();
  }
  __syncthreads();

  if (warp == kProducerWarp) {
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

  if (warp >= kFirstHelperWarp) {
    // ================= helpers: activation staging, cross-CTA reduction, epilogue (all global traffic) =====
    const int ht = threadIdx.x - kFirstHelperWarp * 32;  // 0..127 = panel column
    const uint32_t epoch = ld_acquire_u32(ws_ctrl);
    const uint32_t tag = (epoch % 65535u) + 1u;  // 1..65535: never the zero of a fresh workspace
    check_prog_id(ws_ctrl, prog_id, wd, b == 0 && ht == 0);
    int u = 0, sgi = 0;  // steps with work / segments handed over so far
    for (int j = 0; j < n_steps; ++j) {
      const ProgStep& st = prog[j];
      const int T = st.T, RB = st.RB;
      const int t0 = cta_begin(T, b, G), t1 = cta_begin(T, b + 1, G);
      if (t0 >= t1) continue;
      // ---- stage x of this step into buffer u & 1 (polls the LL words of its producers) -------------------
      const int xb = u & 1;
      mbar_wait(x_empty + 8 * xb, ((u >> 1) & 1) ^ 1);
      helper_stage_x<FMT>(xs0 + (size_t)xb * xs_bytes, st, ws_ll, tag, ht, wd, (unsigned)j);
      __syncwarp();
      if (lane == 0) mbar_arrive(x_full + 8 * xb);
      ++u;
      // ---- the CTA's panel segments of this step, in the compute warps' order --------------------------------
      int panel = t0 / RB, rb = t0 - panel * RB;
      int t = t0;
      while (t < t1) {
        const int rb_begin = rb;
        const int n_blk = min(t1 - t, RB - rb);
        t += n_blk;
        rb += n_blk;
        const int rbuf = sgi & 1;
        mbar_wait(red_full + 8 * rbuf, (sgi >> 1) & 1);
        const float* r4 = red + rbuf * (kGroups * kPanelCols) + ht;
        float sum = (r4[0] + r4[kPanelCols]) + (r4[2 * kPanelCols] + r4[3 * kPanelCols]);
        __syncwarp();
        if (lane == 0) mbar_arrive(red_empty + 8 * rbuf);
        ++sgi;

        const ProgOp& op = st.op[find_op(st, panel)];
        const int n = (panel - op.panel_begin) * kPanelCols + ht;
        const bool complete = (rb_begin == 0) && (rb == RB);
        bool finalize = true;
        if (!complete) {
          // contributors of the panel in block order (T >= G: every CTA between the owners of the first and
          // last block holds >= 1 block; T < G: every block is a different CTA); the owner of the LAST block
          // finishes the panel
          const int first = block_owner(T, panel * RB, G);
          const bool sparse = (T < G);
          uint2* slot0 = ws_partial + st.partial_off + (size_t)panel * st.c_max * kPanelCols;
          if (rb < RB) {
            const int my_slot = sparse ? rb_begin : b - first;
            st_vol_v2(slot0 + (size_t)my_slot * kPanelCols + ht, __float_as_uint(sum), tag);
            finalize = false;
          } else {
            const int n_before = sparse ? rb_begin : b - first;  // contributors ahead of this CTA
            float tot = 0.f;
            for (int o = 0; o < n_before; ++o) tot += ll_poll_f32(slot0 + (size_t)o * kPanelCols + ht, tag, wd, kWdPartial, (unsigned)j);
            sum = tot + sum;
          }
        }
        if (finalize && n < op.N) {
          if constexpr (C::kF32) {
            float e0 = 0.f;
            if (op.ep0_ll >= 0) e0 = ll_poll_f32(reinterpret_cast<const uint2*>(ws_ll) + op.ep0_ll + n, tag, wd, kWdEpOperand, (unsigned)j);
            else if (op.ep0) e0 = __ldcg(reinterpret_cast<const float*>(op.ep0) + n);
            const float y = ep_f32(sum, op.epilogue, &e0, 0);
            reinterpret_cast<float*>(op.out)[n] = y;
            if (op.out_ll >= 0) st_vol_v2(reinterpret_cast<uint2*>(ws_ll) + op.out_ll + n, __float_as_uint(y), tag);
          } else if (op.epilogue == B200Q_EP_CAST_F32) {
            reinterpret_cast<float*>(op.out)[n] = sum;
          } else {
            __half e0 = __ushort_as_half(0);
            if (op.ep0_ll >= 0) e0 = ll_poll_f16(reinterpret_cast<const uint32_t*>(ws_ll) + op.ep0_ll + n, tag, wd, kWdEpOperand, (unsigned)j);
            else if (op.ep0) e0 = __ldcg(reinterpret_cast<const __half*>(op.ep0) + n);
            const __half y = ep_f16(sum, op.epilogue, &e0, 0);
            reinterpret_cast<__half*>(op.out)[n] = y;
            if (op.out_ll >= 0)
              st_vol_u32(reinterpret_cast<uint32_t*>(ws_ll) + op.out_ll + n, (tag << 16) | (uint32_t)__half_as_ushort(y));
          }
        }
        if (rb == RB) { rb = 0; ++panel; }
      }
    }
    // ---- leave: the last CTA bumps the launch tag ------------------------------------------------------------
    asm volatile("bar.sync 2, %0;" ::"n"(kHelperThreads) : "memory");
    if (ht == 0) {
      __threadfence();
      const unsigned old = atomicAdd(ws_ctrl + 1, 1u);
      if (old == (unsigned)(G - 1)) {
        ws_ctrl[1] = 0u;
        __threadfence();
        ws_ctrl[0] = epoch + 1u;
      }
    }
    return;
  }

  // ================= compute warps ===========================================================================
  const int grp = warp >> 2, cq = warp & 3;  // group (ring stages grp, grp + 4, ...), 32-column box
  const int c = lane >> 2, tig = lane & 3;
  // lane rows of the 16-row block: 2 tig + j + 8 m  (j, m in {0,1}); swizzled chunk = c ^ (row & 7)
  const uint32_t w_off0 = cq * kBoxBytes + (2 * tig) * 128 + ((c ^ (2 * tig)) << 4);
  const uint32_t w_off1 = cq * kBoxBytes + (2 * tig + 1) * 128 + ((c ^ (2 * tig + 1)) << 4);
  const int g0 = (2 * tig) / F::kRowsPerGroup, g1 = (8 + 2 * tig) / F::kRowsPerGroup;
  const uint32_t s_col = (cq * 32 + c * 4) * (int)sizeof(scale_t);
  const uint32_t s_off0 = kWordBytes + g0 * C::kScalePitch + s_col;
  const uint32_t s_off1 = kWordBytes + g1 * C::kScalePitch + s_col;
  const uint32_t xs_lane0 = smem_u32(xs0) + (2 * tig) * C::kXRowBytes;

  int qm = 0;    // (CTA-local block sequence number) & 3
  int sg = grp;  // ring slot of this group's next block
  uint32_t phg = 0;
  int u = 0, sgi = 0;

  for (int j = 0; j < n_steps; ++j) {
    const ProgStep& st = prog[j];
    const int T = st.T, RB = st.RB;
    const int t0 = cta_begin(T, b, G), t1 = cta_begin(T, b + 1, G);
    if (t0 >= t1) continue;  // CTA-uniform

    const int xb = u & 1;
    mbar_wait(x_full + 8 * xb, (u >> 1) & 1);  // the helpers have staged this step's activation
    const uint32_t xs_lane = xs_lane0 + xb * xs_bytes;

    int panel = t0 / RB, rb = t0 - panel * RB;
    int t = t0;
    while (t < t1) {
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

      // ---- hand the segment's sums to the helpers and go on -------------------------------------------------
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
      const int rbuf = sgi & 1;
      mbar_wait(red_empty + 8 * rbuf, ((sgi >> 1) & 1) ^ 1);
      if (tig == 0) *reinterpret_cast<float4*>(red + rbuf * (kGroups * kPanelCols) + grp * kPanelCols + cq * 32 + c * 4) = v4;
      __syncwarp();
      if (lane == 0) mbar_arrive(red_full + 8 * rbuf);
      ++sgi;
      if (rb == RB) { rb = 0; ++panel; }
    }
    __syncwarp();
    if (lane == 0) mbar_arrive(x_empty + 8 * xb);  // this warp is done with the step's activation
    ++u;
  }
}

// tensor-parallel peers of a program launch: every rank's workspace (same layout everywhere), by value
struct TpInfo {
  int rank, size;
  unsigned long long ll_off, partial_off;  // byte offsets of the LL activations / partial slots inside a workspace
  unsigned char* ws[8];
};

#include "gemv_tc.cuh"

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
int build_step(ProgStep& st, const b200q_gemv_step& in, int G, bool use_tc) {
  using C = PCfg<FMT>;
  using F = Fmt<FMT>;
  st.n_ops = in.n_ops;
  st.x = in.x;
  st.x_mul = in.x_mul;
  st.has_xmul = in.x_mul ? 1 : 0;
  st.x_ll = st.xmul_ll = -1;
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
    st.op[i].ep0_ll = st.op[i].out_ll = -1;
    panels += ceil_div(o.N, kPanelCols);
    if (int e = use_tc ? make_tmap_2d(&st.tm_w[i], CU_TENSOR_MAP_DATA_TYPE_UINT32, 4, o.packed, o.N, st.n_rows, kPanelCols,
                                      kRows, CU_TENSOR_MAP_SWIZZLE_NONE)
                       : make_tmap_2d(&st.tm_w[i], CU_TENSOR_MAP_DATA_TYPE_UINT32, 4, o.packed, o.N, st.n_rows, kBoxCols, kRows,
                                      CU_TENSOR_MAP_SWIZZLE_128B))
      return e;
    if (int e = make_tmap_2d(&st.tm_s[i], C::kF32 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT32 : CU_TENSOR_MAP_DATA_TYPE_FLOAT16,
                             (int)sizeof(typename C::scale_t), o.scale, o.N, n_grp, kPanelCols, C::kGroupsPerBlock,
                             CU_TENSOR_MAP_SWIZZLE_NONE))
      return e;
  }
  st.n_panels = panels;
  st.T = panels * st.RB;
  st.c_max = partial_slots_per_panel(st.T, st.RB, G);
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
int stages_for(int fmt, int xs_bytes, bool use_tc = false) {
  if (use_tc) {
    int s = (kSmemPerSm - tc::kHeaderTc - tc::kBSlots * tc::kBSlotBytes - tc::kMiscBytes) / tc::kStageBytesTc;  // no activation staging
    return s > tc::kMaxStagesTc ? tc::kMaxStagesTc : s;
  }
  int s = (kSmemPerSm - kHeader - 2 * xs_bytes - 1024) / stage_bytes_for(fmt);
  return s > kMaxStages ? kMaxStages : s;
}
// environment switches are read ONCE per process (never on the launch path)
struct EnvConfig {
  bool prog_tc = true;          // B200Q_PROG_TC=0: q4f16_0 programs use the exact per-weight mma.sync kernel
  bool prog_tc8 = true;         // B200Q_PROG_TC8=0: the same for q8f16_0 programs
  // B200Q_PROG_PDL: how the tcgen05 program is launched.  2 (default): programmatic dependent launch WITHOUT the
  // cooperative attribute - the CTAs of the next program launch take an SM as soon as this launch's CTA there has left,
  // stream their first weights and wait (griddepcontrol.wait) only before they touch the workspace or an input.  The
  // cooperative attribute makes the driver drain the whole grid first (measured, program cut at the reference's
  // operator boundaries: 689 tok/s cooperative, 719 cooperative + PDL, 803 PDL alone).  Residency: the grid is one
  // CTA per SM and a dependent launch starts only after EVERY CTA of its predecessor has started, so all CTAs of a
  // launch become resident without waiting for one another; the watchdog bounds the wait if foreign work holds SMs.
  // 1 = cooperative + PDL, 0 = cooperative only.
  int prog_pdl = 2;
  bool prog_i8 = false;         // B200Q_PROG_I8=1: q4f16_0 programs use the integer formulation of the tcgen05 kernel (slower: kept for study)
  int tc_dbg = 0;               // B200Q_TC_DBG (only honoured by -DB200Q_TC_DBGFLAGS builds)
  long long watchdog_ms = 10000;  // B200Q_WATCHDOG_MS: bound of one cross-CTA / cross-GPU wait; 0 = unbounded
};
const EnvConfig& env_config() {
  static const EnvConfig cfg = [] {
    EnvConfig c;
    if (const char* e = std::getenv("B200Q_PROG_TC")) c.prog_tc = !(e[0] == '0');
    if (const char* e = std::getenv("B200Q_PROG_TC8")) c.prog_tc8 = !(e[0] == '0');
    if (const char* e = std::getenv("B200Q_PROG_I8")) c.prog_i8 = (e[0] == '1');
    if (const char* e = std::getenv("B200Q_PROG_PDL")) c.prog_pdl = std::atoi(e);
    if (const char* e = std::getenv("B200Q_TC_DBG")) c.tc_dbg = std::atoi(e);
    if (const char* e = std::getenv("B200Q_WATCHDOG_MS")) c.watchdog_ms = std::atoll(e);
    return c;
  }();
  return cfg;
}
bool want_tc(int fmt) {
  return (fmt == B200Q_Q4F16_0 && env_config().prog_tc) || (fmt == B200Q_Q8F16_0 && env_config().prog_tc && env_config().prog_tc8);
}

// watchdog settings of the process (b200q_gemv_program_set_watchdog); a launch captures them by value
std::atomic<unsigned long long*> g_wd_status{nullptr};
std::atomic<long long> g_wd_ns{-1};  // -1: the B200Q_WATCHDOG_MS default

// cudaFuncSetAttribute is a per-device setting: remember the largest opt-in per (device, kernel)
constexpr int kMaxDevices = 64;
int set_max_smem(const void* kern, int kern_id, size_t smem) {
  static std::atomic<size_t> done[kMaxDevices][10];
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= kMaxDevices) dev = -1;
  if (dev >= 0 && done[dev][kern_id].load(std::memory_order_relaxed) >= smem) return B200Q_OK;
  cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  if (e != cudaSuccess) return check_cuda(e, "cudaFuncSetAttribute(gemv program smem)");
  if (dev >= 0) done[dev][kern_id].store(smem, std::memory_order_relaxed);
  return B200Q_OK;
}

}  // namespace

void gemv_program_set_watchdog(unsigned long long* status, long long timeout_ns) {
  g_wd_status.store(status, std::memory_order_relaxed);
  g_wd_ns.store(timeout_ns, std::memory_order_relaxed);
}

size_t gemv_program_bytes(int n_steps) { return sizeof(ProgHeader) + (size_t)n_steps * sizeof(ProgStep); }

int gemv_program_build(const b200q_gemv_step* steps, int n_steps, int fmt, void* host_blob, size_t blob_bytes,
                       const unsigned char* reduce_mask, int tp_rank, int tp_size) {
  if (blob_bytes < gemv_program_bytes(n_steps)) return set_error(B200Q_ERR_BAD_SHAPE, "gemv_program_build: blob too small");
  if ((reinterpret_cast<uintptr_t>(host_blob) & 7) != 0) return set_error(B200Q_ERR_BAD_ALIGN, "gemv_program_build: blob must be 8-byte aligned");
  std::memset(host_blob, 0, gemv_program_bytes(n_steps));
  ProgHeader* h = reinterpret_cast<ProgHeader*>(host_blob);
  unsigned char* base = reinterpret_cast<unsigned char*>(host_blob) + sizeof(ProgHeader);
  const int G = device_sm_count();
  const size_t el = fmt_is_f32(fmt) ? 4 : 2;
  unsigned long long ll_elems = 0, partial_slots = 0;
  int xs_bytes = 0;
  const bool use_tc = want_tc(fmt);
  if (tp_size > 1 && !use_tc)
    return set_error(B200Q_ERR_UNSUPPORTED, "gemv_program_build: in-program tensor-parallel reduction needs the q4f16_0 tcgen05 program");
  if (tp_size < 1 || tp_size > 8 || tp_rank < 0 || tp_rank >= tp_size)
    return set_error(B200Q_ERR_BAD_SHAPE, "gemv_program_build: tp rank %d of %d", tp_rank, tp_size);
  auto step_at = [&](int j, ProgStep& out) { std::memcpy(&out, base + (size_t)j * sizeof(ProgStep), sizeof(ProgStep)); };
  // An input of step j is an LL input when it IS (same address, same length) the output of an earlier
  // step; it is external when it overlaps no output at all.  Anything else (partial overlap, an output
  // that aliases another step's input or output) cannot be ordered by the tags and is rejected.
  auto resolve = [&](int j, const void* p, int len, long long& ll_off) -> int {
    ll_off = -1;
    if (!p) return B200Q_OK;
    const Range r = range_of(p, el * len);
    for (int i = j - 1; i >= 0; --i) {
      ProgStep ps;
      step_at(i, ps);
      for (int a = 0; a < ps.n_ops; ++a) {
        const size_t oel = (!fmt_is_f32(fmt) && ps.op[a].epilogue == B200Q_EP_CAST_F32) ? 4 : el;
        const Range ro = range_of(ps.op[a].out, oel * ps.op[a].N);
        if (!r.overlaps(ro)) continue;
        if (ps.op[a].out == p && ps.op[a].N == len && ps.op[a].out_ll >= 0) {
          ll_off = ps.op[a].out_ll;
          return B200Q_OK;
        }
        return set_error(B200Q_ERR_UNSUPPORTED,
                         "gemv_program_build: an input of step %d partially overlaps (or re-types) the output of step %d", j, i);
      }
    }
    return B200Q_OK;
  };
  for (int j = 0; j < n_steps; ++j) {
    alignas(64) ProgStep st;
    std::memset(&st, 0, sizeof(st));
    int e;
    switch (fmt) {
      case B200Q_Q4F16_0: e = build_step<B200Q_Q4F16_0>(st, steps[j], G, use_tc); break;
      case B200Q_Q3F16_0: e = build_step<B200Q_Q3F16_0>(st, steps[j], G, false); break;
      case B200Q_Q8F16_0: e = build_step<B200Q_Q8F16_0>(st, steps[j], G, use_tc); break;
      case B200Q_Q4F32_0: e = build_step<B200Q_Q4F32_0>(st, steps[j], G, false); break;
      default: return set_error(B200Q_ERR_UNSUPPORTED, "gemv_program_build: fmt %d", fmt);
    }
    if (e) return e;
    st.partial_off = partial_slots;
    partial_slots += (unsigned long long)st.n_panels * st.c_max * kPanelCols;
    st.tp_partial_off = partial_slots;
    if (tp_size > 1 && reduce_mask && reduce_mask[j]) {
      partial_slots += (unsigned long long)st.n_panels * tp_size * kPanelCols;
      for (int i = 0; i < st.n_ops; ++i) st.op[i].tp_reduce = (reduce_mask[j] >> i) & 1;
    }
    const int xb = use_tc ? 0 : xs_bytes_for(fmt, st.K);  // the tcgen05 kernel reads activations per block
    if (xb > xs_bytes) xs_bytes = xb;
    if (int e2 = resolve(j, steps[j].x, st.K, st.x_ll)) return e2;
    if (int e2 = resolve(j, steps[j].x_mul, st.K, st.xmul_ll)) return e2;
    for (int i = 0; i < st.n_ops; ++i) {
      if (int e2 = resolve(j, st.op[i].ep0, st.op[i].N, st.op[i].ep0_ll)) return e2;
      // outputs: must not alias anything an earlier step reads or writes, nor this step's own inputs
      const size_t oel = (!fmt_is_f32(fmt) && st.op[i].epilogue == B200Q_EP_CAST_F32) ? 4 : el;
      const Range ro = range_of(st.op[i].out, oel * st.op[i].N);
      for (int q = 0; q <= j; ++q) {
        const b200q_gemv_step& sq = steps[q];
        bool bad = ro.overlaps(range_of(sq.x, el * sq.K)) || ro.overlaps(range_of(sq.x_mul, el * sq.K));
        for (int a = 0; a < sq.n_ops; ++a) {
          bad = bad || ro.overlaps(range_of(sq.ops[a].ep0, el * sq.ops[a].N));
          if (q < j || a < i) {
            const size_t ael = (!fmt_is_f32(fmt) && sq.ops[a].epilogue == B200Q_EP_CAST_F32) ? 4 : el;
            bad = bad || ro.overlaps(range_of(sq.ops[a].out, ael * sq.ops[a].N));
          }
        }
        if (bad)
          return set_error(B200Q_ERR_UNSUPPORTED,
                           "gemv_program_build: the output %d of step %d aliases a buffer used by step %d (every output "
                           "of a program needs its own buffer)", i, j, q);
      }
      if (fmt_is_f32(fmt) || st.op[i].epilogue != B200Q_EP_CAST_F32) {
        st.op[i].out_ll = (long long)ll_elems;
        ll_elems += (unsigned long long)((st.op[i].N + 3) / 4 * 4);  // keep every array 16-byte aligned
      }
    }
    std::memcpy(base + (size_t)j * sizeof(ProgStep), &st, sizeof(st));
  }
  h->magic = kMagic;
  h->n_steps = n_steps;
  h->fmt = fmt;
  h->xs_bytes = xs_bytes;
  h->G = G;
  h->tc = use_tc ? 1 : 0;
  h->tp_rank = tp_rank;
  h->tp_size = tp_size;
  h->ll_elems = ll_elems;
  h->partial_slots = partial_slots;
  {
    unsigned hsh = 2166136261u;  // FNV-1a over the step table
    const size_t nb = (size_t)n_steps * sizeof(ProgStep);
    for (size_t i = 0; i < nb; ++i) hsh = (hsh ^ base[i]) * 16777619u;
    h->prog_id = hsh ? hsh : 1u;
  }
  if (stages_for(fmt, xs_bytes, use_tc) < 2 * kGroups)
    return set_error(B200Q_ERR_BAD_SHAPE, "gemv_program_build: K too large for the activation stage (%d bytes)", xs_bytes);
  return B200Q_OK;
}

size_t gemv_program_workspace_bytes(const void* host_blob) {
  const ProgHeader* h = reinterpret_cast<const ProgHeader*>(host_blob);
  if (h->magic != kMagic) return 0;
  return ws_partial_off(h->ll_elems, fmt_is_f32(h->fmt) ? 8 : 4) + (size_t)h->partial_slots * 8;
}

int gemv_program_run(const void* host_blob, const void* dev_blob, void* workspace, size_t workspace_bytes,
                     void* const* peer_workspaces, cudaStream_t stream) {
  const ProgHeader* h = reinterpret_cast<const ProgHeader*>(host_blob);
  if (h->magic != kMagic) return set_error(B200Q_ERR_BAD_SHAPE, "gemv_program_run: not a program blob");
  if ((reinterpret_cast<uintptr_t>(dev_blob) & 63) != 0) return set_error(B200Q_ERR_BAD_ALIGN, "gemv_program_run: device blob must be 64-byte aligned");
  if (workspace_bytes < gemv_program_workspace_bytes(host_blob) || (reinterpret_cast<uintptr_t>(workspace) & 255) != 0)
    return set_error(B200Q_ERR_BAD_SHAPE, "gemv_program_run: workspace too small or not 256-byte aligned");
  const int G = h->G;
  if (G != device_sm_count()) return set_error(B200Q_ERR_BAD_SHAPE, "gemv_program_run: program was built for %d SMs", G);
  const bool use_tc = h->tc != 0;
  int stages = stages_for(h->fmt, h->xs_bytes, use_tc);
  const size_t smem = use_tc ? (size_t)tc::kHeaderTc + tc::kBSlots * tc::kBSlotBytes + tc::kMiscBytes + (size_t)stages * tc::kStageBytesTc
                             : (size_t)kHeader + 2 * (size_t)h->xs_bytes + 1024 + (size_t)stages * stage_bytes_for(h->fmt);

  const ProgStep* prog = reinterpret_cast<const ProgStep*>(reinterpret_cast<const unsigned char*>(dev_blob) + sizeof(ProgHeader));
  unsigned char* ws = reinterpret_cast<unsigned char*>(workspace);
  unsigned* ws_ctrl = reinterpret_cast<unsigned*>(ws);
  void* ws_ll = ws + ws_ll_off();
  uint2* ws_partial = reinterpret_cast<uint2*>(ws + ws_partial_off(h->ll_elems, fmt_is_f32(h->fmt) ? 8 : 4));
  int n_steps = h->n_steps, xs_bytes = h->xs_bytes;

  const void* kern = nullptr;
  switch (h->fmt) {
    case B200Q_Q4F16_0: kern = (const void*)gemv_program_kernel<B200Q_Q4F16_0>; break;
    case B200Q_Q3F16_0: kern = (const void*)gemv_program_kernel<B200Q_Q3F16_0>; break;
    case B200Q_Q8F16_0: kern = (const void*)gemv_program_kernel<B200Q_Q8F16_0>; break;
    default: kern = (const void*)gemv_program_kernel<B200Q_Q4F32_0>; break;
  }
  int kern_id = h->fmt;
  if (use_tc) {
    const bool tpk = h->tp_size > 1, q8 = (h->fmt == B200Q_Q8F16_0);
    const bool i8 = !q8 && env_config().prog_i8;
    if (q8) kern = tpk ? (const void*)tc::gemv_program_tc_kernel<B200Q_Q8F16_0, false, true> : (const void*)tc::gemv_program_tc_kernel<B200Q_Q8F16_0, false, false>;
    else if (i8) kern = tpk ? (const void*)tc::gemv_program_tc_kernel<B200Q_Q4F16_0, true, true> : (const void*)tc::gemv_program_tc_kernel<B200Q_Q4F16_0, true, false>;
    else kern = tpk ? (const void*)tc::gemv_program_tc_kernel<B200Q_Q4F16_0, false, true> : (const void*)tc::gemv_program_tc_kernel<B200Q_Q4F16_0, false, false>;
    kern_id = 4 + (tpk ? 1 : 0) + (q8 ? 2 : (i8 ? 4 : 0));
  }
  if (int e2 = set_max_smem(kern, kern_id, smem)) return e2;
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3(G, 1, 1);
  cfg.blockDim = dim3(use_tc ? tc::kThreadsTc : kProgThreads, 1, 1);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = stream;
  cudaLaunchAttribute attrs[2];
  int na = 0;
  const int pdl = use_tc ? env_config().prog_pdl : 0;
  if (pdl != 2) {
    attrs[na].id = cudaLaunchAttributeCooperative;  // CTAs wait on one another: they must all be resident
    attrs[na].val.cooperative = 1;
    ++na;
  }
  if (pdl) {
    // programmatic dependent launch: the next program launch's CTAs take an SM as soon as this launch's CTA there is
    // done, stream their first weights and wait (griddepcontrol.wait) only before they touch the workspace / inputs
    attrs[na].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attrs[na].val.programmaticStreamSerializationAllowed = 1;
    ++na;
  }
  cfg.attrs = attrs;
  cfg.numAttrs = na;
  TpInfo tp{};
  tp.rank = h->tp_rank;
  tp.size = h->tp_size;
  tp.ll_off = ws_ll_off();
  tp.partial_off = ws_partial_off(h->ll_elems, fmt_is_f32(h->fmt) ? 8 : 4);
  for (int r = 0; r < 8; ++r) tp.ws[r] = nullptr;
  if (h->tp_size > 1) {
    if (!peer_workspaces) return set_error(B200Q_ERR_NULL, "gemv_program_run: a tensor-parallel program needs the peers' workspaces");
    for (int r = 0; r < h->tp_size; ++r) {
      tp.ws[r] = reinterpret_cast<unsigned char*>(peer_workspaces[r]);
      if (!tp.ws[r]) return set_error(B200Q_ERR_NULL, "gemv_program_run: NULL workspace of rank %d", r);
    }
    if (tp.ws[h->tp_rank] != ws) return set_error(B200Q_ERR_BAD_SHAPE, "gemv_program_run: peer_workspaces[rank] must be this rank's workspace");
  } else {
    tp.ws[0] = ws;
  }
  int dbg = env_config().tc_dbg;
  unsigned prog_id = h->prog_id;
  Wd wd{};
  wd.status = g_wd_status.load(std::memory_order_relaxed);
  const long long wd_ns0 = g_wd_ns.load(std::memory_order_relaxed);
  const unsigned long long wd_ns = wd_ns0 >= 0 ? (unsigned long long)wd_ns0 : (unsigned long long)env_config().watchdog_ms * 1000000ull;
  const unsigned long long spins = wd_ns / 350ull;  // one poll >= ~0.35 us (an L2 round trip)
  wd.max_spins = wd_ns == 0 ? 0u : (spins >= 0xFFFFFFFFull ? 0xFFFFFFFFu : (spins < 16 ? 16u : (unsigned)spins));
  void* args_tc[] = {(void*)&prog, (void*)&n_steps, (void*)&ws_ctrl, (void*)&ws_ll, (void*)&ws_partial, (void*)&stages,
                     (void*)&xs_bytes, (void*)&dbg, (void*)&prog_id, (void*)&wd, (void*)&tp};
  void* args_gen[] = {(void*)&prog, (void*)&n_steps, (void*)&ws_ctrl, (void*)&ws_ll, (void*)&ws_partial, (void*)&stages,
                      (void*)&xs_bytes, (void*)&prog_id, (void*)&wd};
  cudaError_t e = cudaLaunchKernelExC(&cfg, kern, use_tc ? args_tc : args_gen);
  if (e != cudaSuccess) return check_cuda(e, "cudaLaunchKernelExC(gemv program)");
  count_launch();
  return B200Q_OK;
}

// Host-side evaluation of a step's work distribution (b200q_debug_gemv_schedule): the panel segments of one CTA in
// processing order, with the roles the kernels' epilogue warps derive from cta_begin / block_owner.
int gemv_program_schedule_debug(int n_panels, int RB, int G, int cta, int finisher_first, int* out, int max_seg, int* c_max) {
  const int T = n_panels * RB;
  if (c_max) *c_max = partial_slots_per_panel(T, RB, G);
  const int t0 = cta_begin(T, cta, G), t1 = cta_begin(T, cta + 1, G);
  const bool sparse = (T < G);
  int panel = t0 / RB, rb = t0 - panel * RB, n = 0;
  for (int t = t0; t < t1;) {
    const int n_blk = (t1 - t < RB - rb) ? t1 - t : RB - rb;
    const bool complete = (rb == 0) && (n_blk == RB);
    const int first = block_owner(T, panel * RB, G), last = block_owner(T, panel * RB + RB - 1, G);
    // tcgen05 kernel: the owner of the panel's FIRST block finishes it; register-level kernel: the owner of the LAST
    const bool finisher = complete || (finisher_first ? rb == 0 : rb + n_blk == RB);
    const int n_contrib = complete ? 1 : (sparse ? RB : last - first + 1);
    const int slot = complete ? 0 : (sparse ? rb : cta - first);  // the finisher of the tcgen05 kernel keeps its sum (slot 0 unused)
    if (n < max_seg) {
      int* o = out + 6 * n;
      o[0] = panel; o[1] = rb; o[2] = rb + n_blk; o[3] = complete ? 0 : (finisher ? 2 : 1); o[4] = slot; o[5] = finisher ? n_contrib : 1;
    }
    ++n;
    t += n_blk;
    rb += n_blk;
    if (rb == RB) { rb = 0; ++panel; }
  }
  return n;
}

}  // namespace b200q
