// q4f16_0 decode-program kernel on the 5th-generation tensor cores (included by gemv_prog.cu, inside its
// anonymous namespace; uses ProgStep / ProgOp and the LL helpers defined there).
//
// Why: mma.sync.m16n8k16 occupies the issue port of its SM sub-partition for 8 cycles on sm_100a
// (scripts/ubench/pipes2.cu), and the exact per-weight dequant needs 8 FMA-pipe half2 operations per
// packed word at 2 cycles each: 25 issue cycles per lane-word against a budget of 22.9 at HBM speed
// (scripts/ubench/unpack.cu).  A register-level contraction can therefore not keep up with HBM3e.
// Here the contraction runs on tcgen05.mma, which costs no issue slots of the math warps:
//
//   * a "group tile" is one quantization group of 128 output columns: 4 packed-word rows x 128 columns
//     = 32 k x 128 n.  Thread n of a 128-thread dequant slice reads the four words of ITS column from the
//     TMA ring and turns each word into four f16 pairs with five integer instructions (unpack_q4_direct:
//     nibbles become the numbers 1024 + q or 64 + q by OR-ing an exponent over them - no floating-point
//     instruction at all), and stores the 16 registers with tcgen05.st into a TMEM operand buffer:
//     lane n = output column, 16 columns = 32 k;
//   * the MMA lanes issue two  tcgen05.mma.kind::f16  (M128 N16 K16, A from TMEM, B from shared memory)
//     per group tile.  B is the activation of the group, written into row g of an otherwise zero
//     16-row operand, g = position of the group inside the block: column g of the f32 accumulator receives
//     T[g,n] = sum_k (off(k) + q[k,n]) x[k]  of that group and the other columns receive zeros, so ONE
//     accumulator keeps the four group sums of a block apart without being read in between;
//   * four activation warps write those B operands straight from the LL words of the producing step (per
//     block: a block starts as soon as ITS 128 activations exist) and hand the epilogue the correction
//     C[g] = sum_k (off(k) + 7) x[k];
//   * the four epilogue warps (thread = output column) read a finished accumulator with tcgen05.ld and
//     fold it:  y[n] += s[g,n] * (T[g,n] - C[g])  =  s * sum (q - 7) x  exactly in real arithmetic, so the group
//     scale is applied ONCE per group to an f32 sum instead of once per weight in f16: closer to real
//     arithmetic than the reference's  f16((q - 7) * s)  weights (tolerance: oracle/quant_oracle.py::tolerance);
//   * everything else is the persistent program of gemv_prog.cu: stream-K division, TMA producer that
//     never waits for a dependency, LL (value|tag) dataflow between CTAs, replayable from a CUDA graph.
#pragma once

namespace tc {

// suspend-time hints (ns) of the mbarrier waits per role; 0 = spin.  Waiting dequant warps share their issue ports
// with working ones (a spinning waiter costs the workers issue slots: 35 % of all executed instructions in the
// round-1 profile), the single-lane MMA / producer loops and the epilogue warps do not have that problem to the
// same degree, and a suspended warp wakes up later than a spinning one (scripts: B200Q_TC_HINT_* experiments).
#ifndef B200Q_TC_HINT_DQ
#define B200Q_TC_HINT_DQ 10000000
#endif
#ifndef B200Q_TC_HINT_MMA
#define B200Q_TC_HINT_MMA 10000000
#endif
#ifndef B200Q_TC_HINT_HELP
#define B200Q_TC_HINT_HELP 10000000
#endif
#ifndef B200Q_TC_HINT_X
#define B200Q_TC_HINT_X 10000000
#endif
#ifdef B200Q_TC_NOUNROLL
#define TC_UNROLL1 _Pragma("unroll 1")
#else
#define TC_UNROLL1
#endif
constexpr int kDqWarps = 16;                 // 4 slices (group of a stage) x 4 TMEM lane quadrants
constexpr int kHelperWarp0 = 16;             // warps 16..19: epilogue + staging + cross-CTA traffic
constexpr int kProducerWarpTc = 20;
constexpr int kMmaWarp = 21;                 // blocks with even sequence number
constexpr int kMmaWarp2 = 22;                // blocks with odd sequence number
constexpr int kXWarp0 = 23;                  // warps 23..26: one activation warp per dequant slice (B operand of its blocks)
constexpr int kThreadsTc = 27 * 32;          // 864
constexpr int kASets = 4;                    // stages of operand buffers in flight
constexpr int kABufs = 4 * kASets;           // TMEM operand buffers (16 columns each)
constexpr int kBSlots = 2 * kABufs;          // B operand slots: two per (slice, group), so the next block's is written early
#ifndef B200Q_TC_NACC
#define B200Q_TC_NACC 8
#endif
#ifndef B200Q_TC_PBATCH
#define B200Q_TC_PBATCH 5
#endif
// partial-sum slots of a K-split panel polled together by its finisher.  A panel of the 7B shapes spans 5-6 CTAs
// (4-5 foreign slots).  Measured: 4 per batch 1114, 5: 1123, 6: 1114, 8: 1104 tok/s
constexpr int kPBatch = B200Q_TC_PBATCH;
constexpr int kNAcc = B200Q_TC_NACC;         // accumulators in flight (one per block, 16 columns each): 8 or 16
// ring of per-block group corrections: the activation warps run up to 8 blocks ahead of the MMAs, the MMAs up to
// kNAcc blocks ahead of the epilogue warps
constexpr int kCorrRing = (kNAcc == 8) ? 16 : 32;  // >= 8 + kNAcc, a power of two
static_assert(kNAcc == 8 || kNAcc == 16, "accumulators: D columns [256, 256 + 16 * kNAcc)");
constexpr int kMiscBytes = kCorrRing * 32 + 2 * 80 * 4 + 128;  // corrections (8 words per block) + two step tails + ring alignment
constexpr int kTmemCols = 512;               // A: columns [0,256), D: [256,384)
constexpr int kDCol0 = 256;
constexpr int kStageWordBytes = kRows * kPanelCols * 4;             // 8192: [16 rows][128 cols] u32, no swizzle
constexpr int kStageScaleBytes = 4 * kPanelCols * 2;                // 1024: [4 groups][128 cols] f16
constexpr int kStageBytesTc = kStageWordBytes + kStageScaleBytes;   // 9216
constexpr int kBSlotBytes = 1024;            // two K16 operands of 16 rows: [h][kc][j / 8][j % 8][16 B]
constexpr int kMaxStagesTc = 24;
// barriers (8 bytes each) in the first 1024 bytes
constexpr int kBarFull = 0, kBarEmpty = kMaxStagesTc, kBarAReady = 2 * kMaxStagesTc, kBarAFree = kBarAReady + 2 * kASets,
              kBarAccFull = kBarAFree + 2 * kASets, kBarAccFree = kBarAccFull + kNAcc, kBarBReady = kBarAccFree + kNAcc,
              kBarBFree = kBarBReady + 2 * kASets, kNumBars = kBarBFree + 2 * kASets;  // 96
// the scalar tail of a ProgStep (everything behind the tensor maps) is mirrored in shared memory by the epilogue warps
constexpr int kTailOff = (int)offsetof(ProgStep, op);
constexpr int kTailWords = ((int)sizeof(ProgStep) - kTailOff) / 4;  // 80
static_assert(kTailWords == 80, "kMiscBytes assumes 80 tail words");
static_assert(kTailWords <= kHelperThreads, "the step tail is mirrored one word per helper thread");
constexpr int kTmemPtrOff = 8 * kNumBars;    // 768
constexpr int kHeaderTc = 1024;

__device__ __forceinline__ uint4 ld_vol_v4(const void* p) {
  uint4 v;
  asm volatile("ld.volatile.global.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ uint32_t lds32(uint32_t a) {
  uint32_t r;
  asm volatile("ld.shared.u32 %0, [%1];" : "=r"(r) : "r"(a));
  return r;
}
__device__ __forceinline__ uint32_t lds16(uint32_t a) {
  unsigned short r;
  asm volatile("ld.shared.u16 %0, [%1];" : "=h"(r) : "r"(a));
  return r;
}
__device__ __forceinline__ float lds_f32(uint32_t a) {
  float r;
  asm volatile("ld.shared.f32 %0, [%1];" : "=f"(r) : "r"(a));
  return r;
}
__device__ __forceinline__ void sts64(uint32_t a, uint32_t x, uint32_t y) {
  asm volatile("st.shared.v2.u32 [%0], {%1,%2};" ::"r"(a), "r"(x), "r"(y) : "memory");
}
__device__ __forceinline__ void sts128(uint32_t a, uint4 v) {
  asm volatile("st.shared.v4.u32 [%0], {%1,%2,%3,%4};" ::"r"(a), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}

// one packed word row of an input vector: 8 f16 values, zero beyond K.  LL: poll until all tags match.
struct Row8 {
  uint4 lo, hi;  // LL: 8 x {tag << 16 | f16}; plain: lo = 8 halves
};
__device__ __forceinline__ bool ll_row_ready(const Row8& r, uint32_t tag) {
  const uint32_t t = tag << 16;  // one LOP3 per word: the tag halves of all eight words equal `tag`
  uint32_t acc = (r.lo.x ^ t) | (r.lo.y ^ t);
  acc |= (r.lo.z ^ t);
  acc |= (r.lo.w ^ t);
  acc |= (r.hi.x ^ t);
  acc |= (r.hi.y ^ t);
  acc |= (r.hi.z ^ t);
  acc |= (r.hi.w ^ t);
  return (acc & 0xFFFF0000u) == 0u;
}
__device__ __forceinline__ uint4 ll_row_pack(const Row8& r) {  // 8 halves in k order
  return make_uint4(__byte_perm(r.lo.x, r.lo.y, 0x5410), __byte_perm(r.lo.z, r.lo.w, 0x5410),
                    __byte_perm(r.hi.x, r.hi.y, 0x5410), __byte_perm(r.hi.z, r.hi.w, 0x5410));
}

// ---- activation rows: 8 values of one packed word row from an LL array (polled) or a plain f16 vector -----------
__device__ __forceinline__ __half tc_fetch(const __half* plain, const uint32_t* ll, long long ll_off, int k, uint32_t tag,
                                           const Wd& wd, unsigned step) {
  return (ll_off >= 0) ? ll_poll_f16(ll + ll_off + k, tag, wd, kWdActivation, step) : __ldcg(plain + k);
}
__device__ __forceinline__ uint4 tc_finish_row(Row8& r, const __half* plain, const uint32_t* ll, long long ll_off, int k0,
                                               int n_valid, bool vec, uint32_t tag, const Wd& wd, unsigned step) {
  if (n_valid == 8 && ll_off >= 0) {
    WdSpin sp;
    while (!ll_row_ready(r, tag)) {
      wd_poll(sp, wd, kWdActivation, step);
      r.lo = ld_vol_v4(ll + ll_off + k0);
      r.hi = ld_vol_v4(ll + ll_off + k0 + 4);
    }
    return ll_row_pack(r);
  }
  if (n_valid == 8 && vec) return __ldcg(reinterpret_cast<const uint4*>(plain + k0));
  __align__(16) __half v[8];
#pragma unroll
  for (int t = 0; t < 8; ++t) v[t] = (t < n_valid) ? tc_fetch(plain, ll, ll_off, k0 + t, tag, wd, step) : __ushort_as_half(0);
  return *reinterpret_cast<const uint4*>(v);
}
// One packed word -> four f16 pairs with five integer instructions and no floating-point one: a nibble in mantissa
// bits 0..3 under the exponent of 1024 is the f16 number 1024 + q, a nibble in bits 4..7 under the exponent of 64 is
// 64 + q.  Even nibbles (0,4 | 2,6) take the first form, odd ones (1,5 | 3,7) the second; the offsets leave with
// the correction term  sum_k off(k) x[k]  (kOffEven / kOffOdd below, which also carry the -7 of the format).
__device__ __forceinline__ void unpack_q4_direct(uint32_t w, uint32_t* out) {
  const uint32_t t = w >> 8;
  out[0] = lop3_and_or(w, 0x000F000Fu, 0x64006400u);  // (1024 + v0, 1024 + v4)
  out[1] = lop3_and_or(w, 0x00F000F0u, 0x54005400u);  // (64 + v1, 64 + v5)
  out[2] = lop3_and_or(t, 0x000F000Fu, 0x64006400u);  // (1024 + v2, 1024 + v6)
  out[3] = lop3_and_or(t, 0x00F000F0u, 0x54005400u);  // (64 + v3, 64 + v7)
}
constexpr float kOffEven = 1024.0f + 7.0f, kOffOdd = 64.0f + 7.0f;
constexpr float kOffQ8 = 1024.0f + 127.0f;

// event trace of a few steps (scripts/trace_steps.py reads the printed lines): build with EXTRA=-DB200Q_TC_TRACE
#ifdef B200Q_TC_TRACE
constexpr int kTrStep0 = 41, kTrSteps = 8;
__device__ __forceinline__ unsigned long long tc_globaltimer() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}
#define TC_EV(j, e) do { if ((j) >= kTrStep0 && (j) < kTrStep0 + kTrSteps) tc_trace[(j) - kTrStep0][e] = tc_globaltimer(); } while (0)
// per-block events (SM clock) of the CTA-local blocks [kTrBlk0, kTrBlk0 + kTrBlks)
constexpr int kTrBlk0 = 870, kTrBlks = 64;
#define TC_BEV(m, e) do { if ((m) >= kTrBlk0 && (m) < kTrBlk0 + kTrBlks) tc_btrace[(m) - kTrBlk0][e] = (unsigned)clock(); } while (0)
#else
#define TC_EV(j, e) do { } while (0)
#define TC_BEV(m, e) do { } while (0)
#endif
// B200Q_TC_DBG=<bits> (environment) switches pipeline stages off for bisecting, only in -DB200Q_TC_DBGFLAGS builds:
// the tests cost the dequant loop two instructions per half block
#ifdef B200Q_TC_DBGFLAGS
#define TC_DBG(bit) (dbg & (bit))
#else
#define TC_DBG(bit) false
#endif
#ifdef B200Q_TC_TIMING
#define TC_T0() long long _t0 = clock64()
#define TC_ACC(v) do { long long _t1 = clock64(); (v) += _t1 - _t0; _t0 = _t1; } while (0)
#define TC_DECL(...) long long __VA_ARGS__
#define TC_PRINT(cond, ...) do { if (cond) printf(__VA_ARGS__); } while (0)
#else
#define TC_DECL(...) do { } while (0)
#define TC_PRINT(cond, ...) do { } while (0)
#define TC_T0() do { } while (0)
#define TC_ACC(v) do { } while (0)
#endif

// shared-memory map: [0,1024) barriers + TMEM pointer | B slots 16 x 1 KB | group-sum ring 256 B | weight ring
//
// Block m (CTA-local sequence number of a 16-row x 128-column stage) is handled by
//   dequant slice  m % 4  (warps 4 sl .. 4 sl + 3; thread = output column): its four groups -> operand set m % 4,
//   activation warp m % 4 (B slot set (m / 4) % 2 of the slice),  MMA lane m % 2,  accumulator m % 8 (group g in column g),
// so the four warps that share an SM sub-partition (same lane quadrant, different slice) work on different
// blocks and are never in the same phase of their loop.
// kTp: the tensor-parallel instantiation (row-parallel ops are reduced over the ranks inside the launch); the
// single-GPU instantiation carries none of that code (the epilogue warps' finalize path is on the critical path of
// every dependent step: 4 % tokens/s between the two)
// FMT: B200Q_Q4F16_0 (a block = 16 word rows = 128 k = four groups, two per half block) or B200Q_Q8F16_0 (16 word
// rows = 64 k = two groups, one per half block; a byte under the exponent of 1024 is the f16 number 1024 + q, two
// LOP3 and one shift per word, offset 1024 + 127 for every k).
// kI8 (q4f16_0 only): the integer formulation.  The weights go to tensor memory as the bytes 0 .. 15 themselves (two
// LOP3 and one shift per word, a group = 8 columns = ONE tcgen05.mma.kind::i8 of K = 32: a block costs what a q8
// block costs for the same 8 KB).  The activations of a group are scaled by a power of two (from the group's
// largest magnitude) to integers |X| < 2^21 and split into three signed base-256 digits - three s8 rows of the B
// operand, three s32 accumulator columns per group; T = D0 + 256 D1 + 65536 D2 - 7 sum X is EXACT, and
// y += s * 2^-shift * T.  An activation more than 2^10 below its group's largest loses low bits (error <=
// 2^-21 of the group's largest per element: quant_oracle.tolerance_group_scaled); a non-finite activation makes
// the group's factor NaN.
template <int FMT, bool kI8, bool kTp>
__global__ void __launch_bounds__(kThreadsTc, 1)
gemv_program_tc_kernel(const ProgStep* __restrict__ prog, int n_steps, unsigned* ws_ctrl, void* ws_ll, uint2* ws_partial,
                       int stages, int xs_bytes, int dbg, unsigned prog_id, const __grid_constant__ Wd wd,
                       const __grid_constant__ TpInfo tp) {
  extern __shared__ __align__(128) unsigned char smem[];
#ifdef B200Q_TC_TRACE
  __shared__ unsigned long long tc_trace[kTrSteps][6];
  __shared__ unsigned tc_btrace[kTrBlks][8];
#endif
  static_assert(FMT == B200Q_Q4F16_0 || FMT == B200Q_Q8F16_0, "tcgen05 decode program: q4f16_0 and q8f16_0");
  static_assert(!kI8 || FMT == B200Q_Q4F16_0, "the integer formulation is the q4f16_0 one");
  constexpr bool kQ8 = (FMT == B200Q_Q8F16_0);
  constexpr int kG = kQ8 ? 2 : 4;            // groups per block
  constexpr int kGH = kG / 2;                // groups per half block
  constexpr int kKBlk = kQ8 ? 64 : 128;      // k per block
  constexpr uint32_t kStageTx = kStageWordBytes + kG * kPanelCols * 2;  // bytes TMA delivers per stage
  const int warp = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0), lane = threadIdx.x & 31;
  const int G = gridDim.x, b = blockIdx.x;
  (void)dbg;
  pdl_launch_dependents();  // (no-op unless launched with the programmatic attribute)

  const uint32_t bar0 = smem_u32(smem);
  const uint32_t full0 = bar0 + 8 * kBarFull, empty0 = bar0 + 8 * kBarEmpty;
  const uint32_t aready0 = bar0 + 8 * kBarAReady, afree0 = bar0 + 8 * kBarAFree;
  const uint32_t accfull0 = bar0 + 8 * kBarAccFull, accfree0 = bar0 + 8 * kBarAccFree;
  const uint32_t bready0 = bar0 + 8 * kBarBReady, bfree0 = bar0 + 8 * kBarBFree;  // [slice][slot set of the slice]
  volatile uint32_t* tmem_ptr = reinterpret_cast<volatile uint32_t*>(smem + kTmemPtrOff);
  unsigned char* bslots = smem + kHeaderTc;
  // [16 blocks][8 words]: f16 modes: words 0..3 = offset * sum of the group's x; int8 mode: words 0..3 = 7 * sum of the
  // group's integer activations, words 4..7 = the group's 2^-shift
  float* corr_s = reinterpret_cast<float*>(bslots + kBSlots * kBSlotBytes);
  uint32_t* stepc = reinterpret_cast<uint32_t*>(corr_s + kCorrRing * 8);  // [2][kTailWords]: step tails, see the epilogue warps
  const uint32_t stage_base = (smem_u32(stepc + 2 * kTailWords) + 127u) & ~127u;
  (void)xs_bytes;

  // Start-up is on the critical path of every launch (a program cut at the reference's operator boundaries is ~97
  // launches per token): the producer warp initialises the ring barriers ITSELF and starts streaming weights at
  // once - it only ARRIVES at the start barrier below, it never waits for the TMEM allocation, the zeroing of the B
  // slots or the other roles' step-table reads; the remaining barriers are initialised one per thread.
  if (warp == kProducerWarpTc) {
    for (int s = lane; s < stages; s += 32) {
      mbar_init(full0 + 8 * s, 1);
      mbar_init(empty0 + 8 * s, 4 + kHelperWarps);  // the block's dequant slice + the epilogue warps
    }
    fence_barrier_init();
    __syncwarp();
  } else if (threadIdx.x < 2 * kASets) {  // one pair per half block (two groups)
    mbar_init(aready0 + 8 * threadIdx.x, 4);  // the four warps of the dequant slice
    mbar_init(afree0 + 8 * threadIdx.x, 1);
    mbar_init(bready0 + 8 * threadIdx.x, 1);
    mbar_init(bfree0 + 8 * threadIdx.x, 1);
    fence_barrier_init();
  } else if (threadIdx.x >= 32 && threadIdx.x < 32 + kNAcc) {
    mbar_init(accfull0 + 8 * (threadIdx.x - 32), 1);
    mbar_init(accfree0 + 8 * (threadIdx.x - 32), kHelperWarps);
    fence_barrier_init();
  }
#ifdef B200Q_TC_TRACE
  for (int i = threadIdx.x; i < kTrBlks * 8; i += kThreadsTc) (&tc_btrace[0][0])[i] = 0u;
#endif
  // the B slots start as zeros: slot (set, g) only ever gets activation rows written into operand row g
  for (int i = threadIdx.x; i < kBSlots * kBSlotBytes / 16; i += kThreadsTc)
    reinterpret_cast<uint4*>(bslots)[i] = make_uint4(0, 0, 0, 0);
  if (warp == kMmaWarp) {
    tmem_alloc(smem_u32(const_cast<uint32_t*>(tmem_ptr)), kTmemCols);
    tmem_relinquish();
  }
  fence_proxy_async_smem();
  tc_fence_before();
  if (warp == kProducerWarpTc) {
    asm volatile("bar.arrive 3, %0;" ::"n"(kThreadsTc) : "memory");  // (its barriers are initialised: the consumers may go)
  } else {
    asm volatile("bar.sync 3, %0;" ::"n"(kThreadsTc) : "memory");
  }
  tc_fence_after();
  const uint32_t tmem_base = (warp == kProducerWarpTc) ? 0u : *tmem_ptr;

  if (warp == kProducerWarpTc) {
    // ================= producer: never waits for a dependency (weights are constants) ==================
    // (the whole loop runs in ONE elected lane: an elect + branch + __syncwarp per iteration costs ~100 cycles,
    // scripts/ubench/umma_small.cu)
    if (elect_one()) {
      const uint64_t pol = l2_evict_first_policy();
      int s = 0;
      uint32_t ph = 0;
      for (int j = 0; j < n_steps; ++j) {
        const ProgStep& st = prog[j];
        const int T = st.T, RB = st.RB;
        const int t0 = cta_begin(T, b, G), t1 = cta_begin(T, b + 1, G);
        int panel = t0 / RB, rb = t0 - panel * RB;
        for (int t = t0; t < t1;) {
          // one panel segment: the tensor maps and the column are fixed, the inner loop is the bare minimum
          const int oi = find_op(st, panel);
          const int col0 = (panel - st.op[oi].panel_begin) * kPanelCols;
          const CUtensorMap* tmw = &st.tm_w[oi];
          const CUtensorMap* tms = &st.tm_s[oi];
          const int rb_end = min(RB, rb + (t1 - t));
          t += rb_end - rb;
          for (; rb < rb_end; ++rb) {
            mbar_wait(empty0 + 8 * s, ph ^ 1);
            const uint32_t bar = full0 + 8 * s;
            mbar_expect_tx(bar, kStageTx);
            const uint32_t sw = stage_base + s * kStageBytesTc;
            tma_load_2d_hint(sw, tmw, col0, rb * kRows, bar, pol);
            tma_load_2d_hint(sw + kStageWordBytes, tms, col0, rb * kG, bar, pol);
            if (++s == stages) { s = 0; ph ^= 1; }
          }
          if (rb == RB) { rb = 0; ++panel; }
        }
      }
    }
    __syncwarp();
  } else if (warp < kDqWarps) {
    // ================= dequant warps: ring -> registers -> TMEM operand buffers ============================
    const int q = warp & 3, sl = warp >> 2;  // TMEM lane quadrant (= warp id % 4), slice = operand set
    const int n = q * 32 + lane;
    const uint32_t a_taddr0 = tmem_base + ((uint32_t)(q * 32) << 16) + sl * 64;
    // total number of blocks of this CTA (the slice takes every fourth)
    int n_total = 0;
    for (int j = 0; j < n_steps; ++j) n_total += cta_begin(prog[j].T, b + 1, G) - cta_begin(prog[j].T, b, G);
    uint32_t use = 0;
    TC_DECL(tq_full = 0, tq_unp = 0, tq_afree = 0, tq_st = 0);
    TC_T0();
    int s = sl;        // ring stage of block m = m % stages, phase = (m / stages) & 1 (kept incrementally: a division
    uint32_t ph = 0;   // by a run-time number costs ~20 of the ~120 instructions of a block)
    TC_UNROLL1
    for (int m = sl; m < n_total; m += 4, ++use) {
      mbar_wait_hint<B200Q_TC_HINT_DQ>(full0 + 8 * s, ph);
      TC_ACC(tq_full);
      if (q == 0 && lane == 0) TC_BEV(m, 0);
      const uint32_t sw = stage_base + s * kStageBytesTc + (uint32_t)n * 4;
#pragma unroll
      for (int h = 0; h < 2; ++h) {
        // half a block = two groups = one x32 tcgen05.st (a store costs its warp ~150 cycles whatever its width).
        // The halves are handed to the MMA warp separately: the tensor core works on one while the slice
        // unpacks the other.
        uint32_t a[kI8 ? 16 : 16 * kGH];
        if constexpr (kI8) {
#pragma unroll
          for (int gg = 0; gg < 2; ++gg) {
            const uint32_t sg = sw + (2 * h + gg) * (4 * kPanelCols * 4);
#pragma unroll
            for (int r = 0; r < 4; ++r) {
              const uint32_t w = lds32(sg + r * (kPanelCols * 4));
              a[8 * gg + 2 * r] = w & 0x0F0F0F0Fu;             // nibbles 0, 2, 4, 6 of the word
              a[8 * gg + 2 * r + 1] = (w >> 4) & 0x0F0F0F0Fu;  // nibbles 1, 3, 5, 7
            }
          }
        } else if constexpr (kQ8) {
          const uint32_t sg = sw + h * (8 * kPanelCols * 4);  // group h: word rows 8 h .. 8 h + 7
#pragma unroll
          for (int r = 0; r < 8; ++r) {
            const uint32_t w = lds32(sg + r * (kPanelCols * 4));
            a[2 * r] = lop3_and_or(w, 0x00FF00FFu, 0x64006400u);           // (1024 + b0, 1024 + b2)
            a[2 * r + 1] = lop3_and_or(w >> 8, 0x00FF00FFu, 0x64006400u);  // (1024 + b1, 1024 + b3)
          }
        } else {
#pragma unroll
          for (int gg = 0; gg < 2; ++gg) {
            const uint32_t sg = sw + (2 * h + gg) * (4 * kPanelCols * 4);
            const uint32_t w0 = lds32(sg), w1 = lds32(sg + 512), w2 = lds32(sg + 1024), w3 = lds32(sg + 1536);
            unpack_q4_direct(w0, a + 16 * gg);
            unpack_q4_direct(w1, a + 16 * gg + 4);
            unpack_q4_direct(w2, a + 16 * gg + 8);
            unpack_q4_direct(w3, a + 16 * gg + 12);
          }
        }
        TC_ACC(tq_unp);
        mbar_wait_hint<B200Q_TC_HINT_DQ>(afree0 + 8 * (2 * sl + h), (use & 1u) ^ 1u);  // the MMAs of this half set's previous block are done
        TC_ACC(tq_afree);
        if (q == 0 && lane == 0 && h == 0) TC_BEV(m, 1);
        tc_fence_after();
        if (!TC_DBG(2)) {
          if constexpr (kI8 || kQ8) tmem_st_32x16(a_taddr0 + h * 16, a);
          else tmem_st_32x32(a_taddr0 + h * 32, a);
        }
        tmem_st_wait();  // (.sync.aligned: the whole warp's stores are done - no __syncwarp needed before lane 0 signals)
        tc_fence_before();
        if (lane == 0) {
          mbar_arrive(aready0 + 8 * (2 * sl + h));
          if (h == 1) mbar_arrive(empty0 + 8 * s);  // the words have left shared memory
          if (q == 0 && h == 1) TC_BEV(m, 2);
        }
        TC_ACC(tq_st);
      }
      s += 4;
      if (s >= stages) { s -= stages; ph ^= 1u; }
    }
    TC_PRINT(b == 0 && warp == 0 && lane == 0, "dq: n=%d full %lld unpack %lld afree %lld st %lld\n", (int)use, tq_full / (use + 1),
             tq_unp / (use + 1), tq_afree / (use + 1), tq_st / (use + 1));
  } else if (warp >= kXWarp0) {
    // ================= activation warps: the B operand of the blocks of one dequant slice =====================
    // Lane 4 g + c of the LOWER half warp owns packed word row c of group g: its 8 activations are read straight
    // from their producer's LL words (polled here, per block: a block starts as soon as ITS 128 activations
    // exist) or from a plain input; the UPPER half warp fetches the matching x_mul rows, if the step has one, and
    // hands them down with four shuffles.  The row goes, in the unpack's pair order, into operand row g of B slot
    // (slot set, slice, g); the other rows of the slot stay zero.  The group's correction  sum_k off(k) x[k]
    // goes to the epilogue warps through a small ring.  One generic->async proxy fence per block.
    // Every slice has two slot sets and the loads of the next block are in the air during this one, so the warp
    // runs up to eight blocks ahead of the tensor core.  These warps share their SM sub-partitions with the
    // dequant warps (an instruction costs ~5 cycles here): the common case is kept short.
    const int sl = warp - kXWarp0;
    const int upper = lane >> 4, g = (lane >> 2) & 3, c = lane & 3;
    const uint32_t slot_lane = smem_u32(bslots) + (4 * sl + g) * kBSlotBytes + (c >> 1) * 512 + (c & 1) * 256 + g * 16;
    // int8 mode: rows 3 g + d (digit d) of the group's K = 32 operand, this lane's 8 bytes at K positions 8 c .. 8 c + 7
    const uint32_t slot_lane_i8 = smem_u32(bslots) + (4 * sl + g) * kBSlotBytes + (c >> 1) * 256 + (3 * g) * 16 + (c & 1) * 8;
    (void)slot_lane_i8;
    const uint32_t* ll = reinterpret_cast<const uint32_t*>(ws_ll);
    pdl_wait();  // (programmatic dependent launch) the workspace and the plain inputs belong to the previous launch until here
    const uint32_t tag = (ld_acquire_u32(ws_ctrl) % 65535u) + 1u;
    // iterator: the slice's next block (CTA-local number it_m, it_m % 4 == sl) = block it_t of step it_j.
    // (The step table is in global memory and there is no L1 to speak of next to 227 KB of shared memory: every
    // field read is an L2 round trip, so the fields live in registers while the iterator is inside a step.)
    int it_j = -1, it_t = sl, it_t1 = 0, it_t0 = 0, it_rb = 0, it_RB = 1, it_m = sl, it_K = 0;
    long long it_ll = -1;            // this lane's source: LL element offset, or -1: plain vector it_ptr
    const __half* it_ptr = nullptr;
    bool it_hasm = false, it_valid = true;
    auto settle = [&]() {  // it_t may have run past the end of the step: carry the excess into the following steps
      while (it_t >= it_t1) {
        const int carry = it_t - it_t1;
        if (++it_j >= n_steps) { it_valid = false; return; }
        const ProgStep& st = prog[it_j];
        const int T = st.T;
        it_RB = st.RB;
        it_K = st.K;
        it_hasm = st.has_xmul != 0;
        it_ll = upper ? (it_hasm ? st.xmul_ll : -1) : st.x_ll;
        it_ptr = reinterpret_cast<const __half*>(upper ? st.x_mul : st.x);
        it_t0 = cta_begin(T, b, G);
        it_t1 = cta_begin(T, b + 1, G);
        it_t = it_t0 + carry;
        it_rb = it_t % it_RB;
      }
    };
    // this lane's row of the block under the iterator: loads issued, not yet checked; then the iterator moves on
    Row8 row;
    int p_k0 = 0, p_nvalid = 0, p_m = 0, p_j = 0;
    long long p_ll = -1;
    const __half* p_ptr = nullptr;
    bool p_hasm = false, p_first = false;
    auto preload = [&]() {
      p_first = (it_t == it_t0);
      p_j = it_j;
      p_m = it_m;
      p_ll = it_ll; p_ptr = it_ptr; p_hasm = it_hasm;
      p_k0 = it_rb * kKBlk + 32 * g + 8 * c;  // (q8: the lane's two word rows 2 c, 2 c + 1 of group g)
      p_nvalid = (g < kG && (upper == 0 || it_hasm)) ? min(8, max(0, it_K - p_k0)) : 0;
      row.lo = row.hi = make_uint4(0, 0, 0, 0);
      if (p_nvalid == 8) {
        if (p_ll >= 0) {
          row.lo = ld_vol_v4(ll + p_ll + p_k0);
          row.hi = ld_vol_v4(ll + p_ll + p_k0 + 4);
        } else if ((reinterpret_cast<uintptr_t>(p_ptr) & 15) == 0) {
          row.lo = __ldcg(reinterpret_cast<const uint4*>(p_ptr + p_k0));
        }
      }
      it_t += 4;
      it_m += 4;
      it_rb += 4;
      while (it_rb >= it_RB) it_rb -= it_RB;
      settle();
    };
    // B operand of the preloaded block, into slot set `par` of this slice
    auto bpart = [&](uint32_t par) {
      const bool vec = (reinterpret_cast<uintptr_t>(p_ptr) & 15) == 0;
      uint4 v = make_uint4(0, 0, 0, 0);
      if (p_nvalid == 8 && p_ll >= 0) {  // the common case: a whole row of LL words
        if (!ll_row_ready(row, tag)) {
          WdSpin sp;
          do {
            wd_poll(sp, wd, kWdActivation, (unsigned)p_j);
            row.lo = ld_vol_v4(ll + p_ll + p_k0);
            row.hi = ld_vol_v4(ll + p_ll + p_k0 + 4);
          } while (!ll_row_ready(row, tag));
        }
        v = ll_row_pack(row);
      } else if (p_nvalid == 8 && vec) {
        v = row.lo;  // plain row, loaded by preload()
      } else if (p_nvalid > 0) {
        v = tc_finish_row(row, p_ptr, ll, p_ll, p_k0, p_nvalid, vec, tag, wd, (unsigned)p_j);
      }
      if (p_hasm) {  // (warp-uniform)
        uint4 mv;
        mv.x = __shfl_down_sync(0xffffffffu, v.x, 16);
        mv.y = __shfl_down_sync(0xffffffffu, v.y, 16);
        mv.z = __shfl_down_sync(0xffffffffu, v.z, 16);
        mv.w = __shfl_down_sync(0xffffffffu, v.w, 16);
        v.x = h2_as_u32(__hmul2(u32_as_h2(v.x), u32_as_h2(mv.x)));
        v.y = h2_as_u32(__hmul2(u32_as_h2(v.y), u32_as_h2(mv.y)));
        v.z = h2_as_u32(__hmul2(u32_as_h2(v.z), u32_as_h2(mv.z)));
        v.w = h2_as_u32(__hmul2(u32_as_h2(v.w), u32_as_h2(mv.w)));
      }
      // v = (x0,x1)(x2,x3)(x4,x5)(x6,x7); q4: even k carry the offset 1024 + 7, odd k 64 + 7; q8: every k 1024 + 127
      if (lane < 16) {  // (the upper half warp only supplies x_mul rows: it takes no part in the reduction below)
        const float2 f0 = __half22float2(u32_as_h2(v.x)), f1 = __half22float2(u32_as_h2(v.y));
        const float2 f2 = __half22float2(u32_as_h2(v.z)), f3 = __half22float2(u32_as_h2(v.w));
        float sm;
        if constexpr (kI8) {
          // the group's scale: 2^shift with  largest |x| * 2^shift < 2^21  (from the f16 exponent of the largest)
          const uint32_t m2 = __vmaxu2(__vmaxu2(v.x & 0x7FFF7FFFu, v.y & 0x7FFF7FFFu), __vmaxu2(v.z & 0x7FFF7FFFu, v.w & 0x7FFF7FFFu));
          uint32_t am = max(m2 & 0xFFFFu, m2 >> 16);
          am = max(am, __shfl_xor_sync(0x0000ffffu, am, 1));
          am = max(am, __shfl_xor_sync(0x0000ffffu, am, 2));
          const uint32_t e16 = max(am >> 10, 1u);
          const bool finite = e16 < 31u;
          const float up = finite ? __uint_as_float((162u - e16) << 23) : 0.f;                   // 2^(35 - e16)
          const float fg = finite ? __uint_as_float((92u + e16) << 23) : __uint_as_float(0x7FC00000u);  // its inverse | NaN
          const int X0 = __float2int_rn(f0.x * up), X1 = __float2int_rn(f0.y * up), X2 = __float2int_rn(f1.x * up),
                    X3 = __float2int_rn(f1.y * up), X4 = __float2int_rn(f2.x * up), X5 = __float2int_rn(f2.y * up),
                    X6 = __float2int_rn(f3.x * up), X7 = __float2int_rn(f3.y * up);
          int sx = ((X0 + X1) + (X2 + X3)) + ((X4 + X5) + (X6 + X7));
          sx += __shfl_xor_sync(0x0000ffffu, sx, 1);
          sx += __shfl_xor_sync(0x0000ffffu, sx, 2);
          // three signed base-256 digits of X: the bytes of X + 0x808080 with their top bits flipped.  Byte order of
          // the operand row = the unpack's: nibbles 0, 2, 4, 6 | 1, 3, 5, 7 of the word
          const uint32_t kB = 0x00808080u;
          const uint32_t U0 = (uint32_t)X0 + kB, U1 = (uint32_t)X1 + kB, U2 = (uint32_t)X2 + kB, U3 = (uint32_t)X3 + kB,
                         U4 = (uint32_t)X4 + kB, U5 = (uint32_t)X5 + kB, U6 = (uint32_t)X6 + kB, U7 = (uint32_t)X7 + kB;
          const uint32_t te0 = __byte_perm(U0, U2, 0x5140), te1 = __byte_perm(U4, U6, 0x5140);  // (a0 b0 a1 b1)
          const uint32_t to0 = __byte_perm(U1, U3, 0x5140), to1 = __byte_perm(U5, U7, 0x5140);
          const uint32_t he0 = __byte_perm(U0, U2, 0x0062), he1 = __byte_perm(U4, U6, 0x0062);  // (a2 b2 . .)
          const uint32_t ho0 = __byte_perm(U1, U3, 0x0062), ho1 = __byte_perm(U5, U7, 0x0062);
          const uint32_t dst = slot_lane_i8 + par * (kABufs * kBSlotBytes);
          sts64(dst, __byte_perm(te0, te1, 0x5410) ^ 0x80808080u, __byte_perm(to0, to1, 0x5410) ^ 0x80808080u);
          sts64(dst + 16, __byte_perm(te0, te1, 0x7632) ^ 0x80808080u, __byte_perm(to0, to1, 0x7632) ^ 0x80808080u);
          sts64(dst + 32, __byte_perm(he0, he1, 0x5410) ^ 0x80808080u, __byte_perm(ho0, ho1, 0x5410) ^ 0x80808080u);
          if (c == 0) {
            corr_s[(p_m & (kCorrRing - 1)) * 8 + g] = __int_as_float(7 * sx);
            corr_s[(p_m & (kCorrRing - 1)) * 8 + 4 + g] = fg;
          }
          sm = 0.f;
        } else if constexpr (kQ8) {
          // the unpack's pair order, two words of four bytes: (x0,x2)(x1,x3)(x4,x6)(x5,x7)
          sts128(slot_lane + par * (kABufs * kBSlotBytes),
                 make_uint4(__byte_perm(v.x, v.y, 0x5410), __byte_perm(v.x, v.y, 0x7632), __byte_perm(v.z, v.w, 0x5410),
                            __byte_perm(v.z, v.w, 0x7632)));
          sm = kOffQ8 * (((f0.x + f1.x) + (f2.x + f3.x)) + ((f0.y + f1.y) + (f2.y + f3.y)));
        } else {
          // the unpack's pair order: (x0,x4)(x1,x5)(x2,x6)(x3,x7)
          sts128(slot_lane + par * (kABufs * kBSlotBytes),
                 make_uint4(__byte_perm(v.x, v.z, 0x5410), __byte_perm(v.x, v.z, 0x7632), __byte_perm(v.y, v.w, 0x5410),
                            __byte_perm(v.y, v.w, 0x7632)));
          sm = fmaf(kOffEven, (f0.x + f1.x) + (f2.x + f3.x), kOffOdd * ((f0.y + f1.y) + (f2.y + f3.y)));
        }
        if constexpr (!kI8) {
          sm += __shfl_xor_sync(0x0000ffffu, sm, 1);
          sm += __shfl_xor_sync(0x0000ffffu, sm, 2);
          if (c == 0) corr_s[(p_m & (kCorrRing - 1)) * 8 + g] = sm;
        }
      }
      fence_proxy_async_smem();
      __syncwarp();
    };
    settle();
    if (it_valid) {
      preload();  // p_*: the block to serve; the iterator is already one block further
      TC_UNROLL1
      for (uint32_t use = 0;; ++use) {
        const uint32_t par = use & 1u;
        // the MMAs of the block that used this slot set two uses ago are done (first two uses: passes at once)
        mbar_wait_hint<B200Q_TC_HINT_X>(bfree0 + 8 * (2 * sl + par), ((use >> 1) & 1u) ^ 1u);
        if (p_first && lane == 0) TC_EV(p_j, 5);
        bpart(par);
        if (lane == 0) mbar_arrive(bready0 + 8 * (2 * sl + par));
        if (p_first && lane == 0) TC_EV(p_j, 0);
        if (lane == 0) TC_BEV(p_m, 3);
        if (!it_valid) break;
        preload();
      }
    }
  } else if (warp == kMmaWarp || warp == kMmaWarp2) {
    // ================= MMA warps: eight tcgen05.mma per block (two per group), whole loop in one lane ========
    const uint32_t idesc = kI8 ? make_idesc_i8_u8s8(128, 16) : make_idesc_f16(128, 16, 0, 0);
    const uint64_t bdesc0 = make_smem_desc_nosw(smem_u32(bslots), 256, 128);
    if (elect_one()) {
      int n_total = 0;
      for (int j = 0; j < n_steps; ++j) n_total += cta_begin(prog[j].T, b + 1, G) - cta_begin(prog[j].T, b, G);
      TC_DECL(tm_accfree = 0, tm_bready = 0, tm_a0 = 0, tm_a1 = 0, tm_issue = 0);
      TC_T0();
      int cnt = 0;
      TC_UNROLL1
      for (int m = (warp == kMmaWarp) ? 0 : 1; m < n_total; m += 2, ++cnt) {
        const int set = m & (kASets - 1), ai = m & (kNAcc - 1);
        const uint32_t use = (uint32_t)m / kASets, par = use & 1u;
        const uint32_t d_taddr = tmem_base + kDCol0 + ai * 16;
        const uint32_t a_taddr = tmem_base + set * 64;
        const uint64_t bd = bdesc0 + (uint64_t)((par * kABufs + set * 4) * (kBSlotBytes / 16));
        mbar_wait_hint<B200Q_TC_HINT_MMA>(accfree0 + 8 * ai, ((m / kNAcc) & 1) ^ 1);  // the epilogue has read this accumulator
        TC_ACC(tm_accfree);
        mbar_wait_hint<B200Q_TC_HINT_MMA>(bready0 + 8 * (2 * set + par), (use >> 1) & 1u);
        TC_ACC(tm_bready);
#pragma unroll
        for (int h = 0; h < 2; ++h) {
          mbar_wait_hint<B200Q_TC_HINT_MMA>(aready0 + 8 * (2 * set + h), use & 1u);
          if (h == 0) TC_ACC(tm_a0); else TC_ACC(tm_a1);
          if (h == 1) TC_BEV(m, 4);
          tc_fence_after();
          if (!TC_DBG(1)) {
#pragma unroll
            for (int g = kGH * h; g < kGH * h + kGH; ++g) {
              if constexpr (kI8) {  // one K = 32 MMA per group: 8 columns of weight bytes x three digit rows
                umma_i8_ts(d_taddr, a_taddr + g * 8, bd + (uint64_t)(g * (kBSlotBytes / 16)), idesc, g == 0 ? 0u : 1u);
              } else {
                umma_f16_ts(d_taddr, a_taddr + g * 16, bd + (uint64_t)(g * (kBSlotBytes / 16)), idesc, g == 0 ? 0u : 1u);
                umma_f16_ts(d_taddr, a_taddr + g * 16 + 8, bd + (uint64_t)(g * (kBSlotBytes / 16) + 512 / 16), idesc, 1u);
              }
            }
          }
          umma_commit(afree0 + 8 * (2 * set + h));
        }
        umma_commit(bfree0 + 8 * (2 * set + par));
        umma_commit(accfull0 + 8 * ai);
        TC_ACC(tm_issue);
      }
      TC_PRINT(b == 0 && warp == kMmaWarp, "mma: n=%d accfree %lld bready %lld aready0 %lld aready1 %lld issue %lld\n", cnt,
               tm_accfree / (cnt + 1), tm_bready / (cnt + 1), tm_a0 / (cnt + 1), tm_a1 / (cnt + 1), tm_issue / (cnt + 1));
    }
    __syncwarp();
  } else if (warp >= kHelperWarp0 && warp < kHelperWarp0 + kHelperWarps) {
    // ================= helper warps: staging, accumulator read-out, cross-CTA reduction, epilogue ============
    const int hq = warp - kHelperWarp0;  // TMEM lane quadrant (warp 16 + hq, (16 + hq) % 4 == hq)
    const int ht = hq * 32 + lane;       // panel column
    const uint32_t d_taddr0 = tmem_base + ((uint32_t)(hq * 32) << 16) + kDCol0;
    pdl_wait();  // (programmatic dependent launch) as the activation warps
    const uint32_t epoch = ld_acquire_u32(ws_ctrl);
    const uint32_t tag = (epoch % 65535u) + 1u;
    check_prog_id(ws_ctrl, prog_id, wd, b == 0 && ht == 0);
    int s = 0;
    int m = 0;
    TC_DECL(th_accfull = 0, th_blk = 0, th_fin = 0);
    TC_T0();
    // The step table is in global memory and there is no L1 to speak of next to 227 KB of shared memory: every
    // field read is an L2 round trip (~0.4 us), and a segment needs several dependent ones.  The 128 epilogue
    // threads therefore mirror the scalar tail of the CURRENT step in shared memory (one word per thread) and
    // fetch the next step's while the current one is processed.
    auto tail_word = [&](int j) -> uint32_t {
      return (ht < kTailWords && j < n_steps)
                 ? __ldg(reinterpret_cast<const uint32_t*>(reinterpret_cast<const unsigned char*>(prog + j) + kTailOff) + ht) : 0u;
    };
    if (ht < kTailWords) stepc[ht] = tail_word(0);
    asm volatile("bar.sync 2, %0;" ::"n"(kHelperThreads) : "memory");
    for (int j = 0; j < n_steps; ++j) {
      // (only the tail of this reference is real - the tensor maps are never read through it)
      const ProgStep& st = *reinterpret_cast<const ProgStep*>(reinterpret_cast<const unsigned char*>(stepc + (j & 1) * kTailWords) - kTailOff);
      const uint32_t next_word = tail_word(j + 1);
      const int T = st.T, RB = st.RB;
      const int t0 = cta_begin(T, b, G), t1 = cta_begin(T, b + 1, G);
      int panel = t0 / RB, rb = t0 - panel * RB;
      int t = t0;
      while (t < t1) {
        const int rb_begin = rb;
        const int n_blk = min(t1 - t, RB - rb);
        // ---- everything the panel's epilogue needs is fetched NOW: the step table, the epilogue operand and the
        // other contributors' partial sums are L2 round trips (no L1 next to 227 KB of shared memory) that would
        // otherwise line up behind the segment's last block, on the critical path of the next step
        const ProgOp& op = st.op[find_op(st, panel)];
        const int op_N = op.N, op_ep = op.epilogue, op_reduce = op.tp_reduce;
        const long long op_ep0_ll = op.ep0_ll, op_out_ll = op.out_ll;
        void* const op_out = op.out;
        const void* const op_ep0 = op.ep0;
        const int n = (panel - op.panel_begin) * kPanelCols + ht;
        const bool complete = (rb_begin == 0) && (rb + n_blk == RB);
        // K-split panel.  The owner of the panel's FIRST block finishes it: that segment is the last of its
        // owner's range, so by then the other contributors (whose ranges end at the same time, or - the owner
        // of the panel's tail - much earlier) have published; a CTA never waits in the middle of its range.
        const int first = complete ? 0 : block_owner(T, panel * RB, G);
        const bool sparse = (T < G);  // every block of the step is a different CTA
        uint2* const slot0 = ws_partial + st.partial_off + (size_t)panel * st.c_max * kPanelCols;
        const int n_contrib = (complete || rb_begin > 0) ? 1 : (sparse ? RB : block_owner(T, panel * RB + RB - 1, G) - first + 1);
        uint32_t e0w = tag << 16;  // {tag | f16 epilogue operand}
        if (rb_begin == 0 && n < op_N) {
          if (op_ep0_ll >= 0) e0w = ld_vol_u32(reinterpret_cast<const uint32_t*>(ws_ll) + op_ep0_ll + n);
          else if (op_ep0) e0w |= (uint32_t)__half_as_ushort(__ldcg(reinterpret_cast<const __half*>(op_ep0) + n));
        }
        float sum = 0.f;
        TC_UNROLL1
        for (int bb = 0; bb < n_blk; ++bb, ++m) {
          const int ai = m & (kNAcc - 1);
          TC_ACC(th_fin);
          mbar_wait_hint<B200Q_TC_HINT_HELP>(accfull0 + 8 * ai, (m / kNAcc) & 1);
          TC_ACC(th_accfull);
          if (ht == 0 && t == t0 && bb == 0) TC_EV(j, 2);
          tc_fence_after();
          uint32_t d[kI8 ? 16 : 4];
          if constexpr (kI8) tmem_ld_32x16(d_taddr0 + ai * 16, d);  // three digit columns per group
          else tmem_ld_32x4(d_taddr0 + ai * 16, d);
          // scales and corrections of the block while the TMEM load is in flight
          const uint32_t sc_addr = stage_base + s * kStageBytesTc + kStageWordBytes + ht * 2;
          const uint32_t c_addr = smem_u32(corr_s) + (uint32_t)((m & (kCorrRing - 1)) * 8) * 4;
          float sc[kG];
#pragma unroll
          for (int g = 0; g < kG; ++g) sc[g] = __half2float(__ushort_as_half((unsigned short)lds16(sc_addr + g * (kPanelCols * 2))));
          const uint4 cg4 = gemvdev::lds128(c_addr);
          const float cg[4] = {__uint_as_float(cg4.x), __uint_as_float(cg4.y), __uint_as_float(cg4.z), __uint_as_float(cg4.w)};
          uint4 cf4 = make_uint4(0, 0, 0, 0);
          if constexpr (kI8) cf4 = gemvdev::lds128(c_addr + 16);  // the groups' 2^-shift
          tmem_ld_wait();  // (.sync.aligned, like the shared-memory loads above: warp-wide)
          tc_fence_before();
          if (lane == 0) {
            mbar_arrive(accfree0 + 8 * ai);
            mbar_arrive(empty0 + 8 * s);
          }
          if (!TC_DBG(8)) {
            if constexpr (kI8) {
              const int ci[4] = {(int)cg4.x, (int)cg4.y, (int)cg4.z, (int)cg4.w};
              const float cf[4] = {__uint_as_float(cf4.x), __uint_as_float(cf4.y), __uint_as_float(cf4.z), __uint_as_float(cf4.w)};
#pragma unroll
              for (int g = 0; g < 4; ++g) {  // exact: |T| < 2^31
                const int T = (int)d[3 * g] + ((int)d[3 * g + 1] << 8) + ((int)d[3 * g + 2] << 16) - ci[g];
                sum = fmaf(sc[g] * cf[g], (float)T, sum);
              }
            } else {
#pragma unroll
              for (int g = 0; g < kG; ++g) sum = fmaf(sc[g], __uint_as_float(d[g]) - cg[g], sum);
            }
          }
          if (++s == stages) s = 0;
          TC_ACC(th_blk);
        }
        if (ht == 0 && t + n_blk == t1) TC_EV(j, 3);
        t += n_blk;
        rb += n_blk;
        // ---- cross-CTA reduction + epilogue of the panel (as gemv_program_kernel's helpers) ---------------------
        bool finalize = true;
        if (!complete) {
          if (rb_begin > 0) {
            const int my_slot = sparse ? rb_begin : b - first;
            st_vol_v2(slot0 + (size_t)my_slot * kPanelCols + ht, __float_as_uint(sum), tag);
            finalize = false;
          } else {
            // the other contributors finish their last block of the panel about now: all their slots are polled
            // at once (one L2 round trip when the sums are there), kPBatch at a time
            constexpr int kPB = kTp ? 4 : kPBatch;  // (the tensor-parallel instantiation keeps the batch it was measured with)
            for (int o0 = 1; o0 < n_contrib; o0 += kPB) {
              uint2 pv[kPB];
#pragma unroll
              for (int o = 0; o < kPB; ++o)
                if (o0 + o < n_contrib) pv[o] = ld_vol_v2(slot0 + (size_t)(o0 + o) * kPanelCols + ht);
#pragma unroll
              for (int o = 0; o < kPB; ++o)
                if (o0 + o < n_contrib) {  // added in block order
                  if (pv[o].y != tag) {
                    WdSpin sp;
                    do {
                      wd_poll(sp, wd, kWdPartial, (unsigned)j);
                      pv[o] = ld_vol_v2(slot0 + (size_t)(o0 + o) * kPanelCols + ht);
                    } while (pv[o].y != tag);
                  }
                  sum += __uint_as_float(pv[o].x);
                }
            }
          }
        }
        if (finalize && n < op_N) {
          if (kTp && op_reduce && tp.size > 1) {
            // row-parallel op: `sum` is this rank's partial.  One-hop all-reduce inside the launch: every rank pushes
            // its {f32, tag} partial over NVLink into the slot [panel][its rank] of EVERY peer's workspace and then
            // adds the slots of its own workspace in rank order - the same order on every rank, so all ranks hold the
            // same bits, apply the epilogue redundantly and publish the finished activation LOCALLY: no owner, no
            // second hop for a broadcast, and the plain `out` is complete on every rank.
            const size_t slot = st.tp_partial_off + ((size_t)panel * tp.size) * kPanelCols + ht;
            const uint32_t sum_bits = __float_as_uint(sum);
#pragma unroll
            for (int r = 0; r < 8; ++r)
              if (r < tp.size && r != tp.rank)
                st_vol_v2(reinterpret_cast<uint2*>(tp.ws[r] + tp.partial_off) + slot + (size_t)tp.rank * kPanelCols, sum_bits, tag);
            const uint2* src = ws_partial + slot;
            float tot = 0.f;
#pragma unroll
            for (int r0 = 0; r0 < 8; r0 += 4) {  // four polls in flight together (register budget), added in rank order
              if (r0 >= tp.size) break;
              uint2 rv[4];
#pragma unroll
              for (int q = 0; q < 4; ++q)
                if (r0 + q < tp.size && r0 + q != tp.rank) rv[q] = ld_vol_v2(src + (size_t)(r0 + q) * kPanelCols);
#pragma unroll
              for (int q = 0; q < 4; ++q) {
                const int r = r0 + q;
                if (r >= tp.size) break;
                if (r == tp.rank) { tot += sum; continue; }
                if (rv[q].y != tag) {
                  WdSpin sp;
                  do {
                    wd_poll(sp, wd, kWdTpPartial, (unsigned)j);
                    rv[q] = ld_vol_v2(src + (size_t)r * kPanelCols);
                  } while (rv[q].y != tag);
                }
                tot += __uint_as_float(rv[q].x);
              }
            }
            sum = tot;
          }
          if (finalize) {
            if (op_ep == B200Q_EP_CAST_F32) {
              reinterpret_cast<float*>(op_out)[n] = sum;
            } else {
              if ((e0w >> 16) != tag) {
                WdSpin sp;
                do {
                  wd_poll(sp, wd, kWdEpOperand, (unsigned)j);
                  e0w = ld_vol_u32(reinterpret_cast<const uint32_t*>(ws_ll) + op_ep0_ll + n);
                } while ((e0w >> 16) != tag);
              }
              const __half e0 = __ushort_as_half((unsigned short)(e0w & 0xFFFFu));
              const __half y = ep_f16(sum, op_ep, &e0, 0);
              if (op_out_ll >= 0) {
                const uint32_t word = (tag << 16) | (uint32_t)__half_as_ushort(y);
                st_vol_u32(reinterpret_cast<uint32_t*>(ws_ll) + op_out_ll + n, word);
              }
              reinterpret_cast<__half*>(op_out)[n] = y;
            }
          }
        }
        if (rb == RB) { rb = 0; ++panel; }
      }
      if (ht == 0 && t0 < t1) TC_EV(j, 4);
      if (ht < kTailWords) stepc[((j + 1) & 1) * kTailWords + ht] = next_word;
      asm volatile("bar.sync 2, %0;" ::"n"(kHelperThreads) : "memory");
    }
    TC_PRINT(b == 0 && ht == 0, "helper: n=%d accfull %lld block %lld finalize+step %lld (totals / blocks)\n", m, th_accfull / (m + 1),
             th_blk / (m + 1), th_fin / (m + 1));
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
    (void)m;
  }

  // every accumulator has been read (the helpers waited for the last commit): TMEM can go back
  tc_fence_before();
  __syncthreads();
#ifdef B200Q_TC_TRACE
  if (threadIdx.x < kTrSteps && n_steps > kTrStep0 + kTrSteps) {
    if (b == 5 || b == 77)
      for (int i = threadIdx.x; i < kTrBlks; i += kTrSteps)
        printf("BT %d %d %u %u %u %u %u %u %u %u\n", b, kTrBlk0 + i, tc_btrace[i][0], tc_btrace[i][1], tc_btrace[i][2], tc_btrace[i][3],
               tc_btrace[i][4], tc_btrace[i][5], tc_btrace[i][6], tc_btrace[i][7]);
    printf("TR %d %d %llu %llu %llu %llu %llu %llu\n", b, kTrStep0 + (int)threadIdx.x, tc_trace[threadIdx.x][0], tc_trace[threadIdx.x][1],
           tc_trace[threadIdx.x][2], tc_trace[threadIdx.x][3], tc_trace[threadIdx.x][4], tc_trace[threadIdx.x][5]);
  }
#endif
  if (warp == kMmaWarp) {
    tc_fence_after();
    tmem_dealloc(tmem_base, kTmemCols);
  }
}

}  // namespace tc
