// Prefill GEMM on the 5th-generation tensor cores (tcgen05 / TMEM), sm_100a.
//
//   Y[m, n] = sum_k X[m, k] * w[k, n]  (+ epilogue)
//
// Replaces three reference operators at once:
//   decode* + NT_matmul*           (mlc_llm/dispatch/llama/main.py:3984-4232 + 3884-3981; the reference
//                                   re-materialises the whole f16 weight in HBM on every call and then
//                                   runs a SIMT f16-accumulating GEMM on it)
//   fused_decode*_NT_matmul*       (dispatch_tir_operator_adreno.py:414-1021: dequant in registers, SIMT)
//   NT_matmul* on a dense weight   (transform/transpose_matmul.py:34-101)   -> b200q_nt_matmul
//
// Design (one persistent CTA per SM, 128 x 256 output tiles, m fastest so CTAs that run together
// share one column panel of weights in L2):
//   warp 0      TMA producer: X tiles [128 rows x 64 k] f16, K-major, 128-byte swizzle (and, for the
//               dense-weight variant, T tiles [256 rows x 64 k] the same way)
//   warp 1      allocates TMEM (2 accumulators x 256 columns = all 512) and issues tcgen05.mma
//               (M=128, N=256, K=16 per instruction; one elected thread)
//   warps 2-5   epilogue: tcgen05.ld the f32 accumulator (each warp its 32-lane quarter), round to
//               f16, apply the reference's epilogue op for op, 16-byte global stores; runs on tile i
//               while the MMA warp already accumulates tile i+1 in the other TMEM half
//   warps 6-13  dequantizers: read packed words + scales from global (L2-resident: a panel is
//               shared by all m tiles), unpack to the bit-exact f16 weight half((q - maxint) * s)
//               with the same magic-number arithmetic as the GEMV, and store them into shared memory
//               DIRECTLY in the tcgen05 canonical MN-major 128B-swizzled layout (the packed layout
//               has N contiguous, so eight neighbouring columns of one k are one 16-byte chunk:
//               no transpose, conflict-free 16-byte stores)
// The A ring (64-k stages) and the B ring (64-k stages; 80-k for int3 whose groups are 40 wide) are
// independent: the MMA thread walks 16-k steps and waits / releases each ring at its own stage
// boundaries, so the int3 word geometry never has to line up with the 128-byte swizzle atom of X.
#include <cuda.h>

#include <atomic>

#include <cstdlib>

#include "common.cuh"
#include "ptx.cuh"

namespace b200q {
namespace {

using namespace ptx;

constexpr int kBM = 128;                 // tile rows (UMMA M)
constexpr int kBN = 256;                 // tile columns (UMMA N)
constexpr int kAK = 64;                  // k per A stage (one 128-byte swizzle row of f16)
constexpr int kAStageBytes = kBM * kAK * 2;  // 16 KB
constexpr int kAStages = 4;
#ifndef B200Q_GEMM_ASTAGES
#define B200Q_GEMM_ASTAGES 4
#endif
constexpr int kAStagesP = B200Q_GEMM_ASTAGES;  // X stages of the pair kernel
constexpr int kNAtoms = kBN / 64;        // MN-major swizzle atoms along N per B stage
constexpr int kKAtomBytes = kNAtoms * 1024;  // one 8-k slab of a B stage: [n atom][8 k rows][128 B]
constexpr int kDqWarps = 8;
constexpr int kEpWarps = 4;
constexpr int kFirstEpWarp = 2;
constexpr int kFirstDqWarp = kFirstEpWarp + kEpWarps;
constexpr int kGemmThreads = (kFirstDqWarp + kDqWarps) * 32;  // 448
constexpr int kTmemCols = 512;
constexpr int kDense = -1;               // B operand mode: dense f16 T[N, K] through TMA

template <int FMT>
struct BCfg {  // quantized B: stage geometry per format
  using F = Fmt<FMT>;
  static constexpr int kBK = (FMT == B200Q_Q3F16_0) ? 80 : 64;   // k per B stage (whole groups, multiple of 16)
  static constexpr int kRows = kBK / F::kNpu;                    // packed word rows per stage: 8 / 8 / 16
  static constexpr int kItems = kRows / kDqWarps;                // word rows per dequant warp per stage
  static constexpr int kStages = (FMT == B200Q_Q3F16_0) ? 3 : 4;
  static constexpr int kStageBytes = kBN * kBK * 2;              // 32 KB (40 KB int3)
  static constexpr int kKSteps = kBK / 16;
};
template <>
struct BCfg<kDense> {
  static constexpr int kBK = 64, kStages = 4, kStageBytes = kBN * kAK * 2, kKSteps = 4;
};

// one weight of a multi-weight launch of the pair kernel (q | k | v read the same X): b200q_gemm_multi
struct GemmOp {
  const uint32_t* packed;
  const __half* scale;
  const __half* zmin;
  const __half* ep0;
  const __half* ep1;
  const __half* ep2;
  void* Y;
  int N;
  int epilogue;
};
constexpr int kMaxGemmOps = 4;  // (the fourth only in the GELU-table instantiation: q | k | v | h_to_4h of a NeoX layer)

struct GemmParams {
  CUtensorMap tm_a;  // X f16 [M, K], box 128 x 64, 128B swizzle
  CUtensorMap tm_b;  // dense mode: T f16 [N, K], box 256 x 64, 128B swizzle
  const uint32_t* packed;
  const __half* scale;
  const __half* zmin;  // f16 asymmetric scheme: group minima, shaped like scale
  const __half* ep0;
  const __half* ep1;
  const __half* ep2;
  void* Y;
  int M, K, N;
  int n_rows;    // packed word rows
  int epilogue;
  int tiles_m, tiles_n;
  unsigned flags;  // B200Q_FLAG_GEMM_*: tile selection override
  // pair kernel: the weights of the launch (op[0] mirrors the fields above) and the first column tile of each
  int n_ops;
  GemmOp op[kMaxGemmOps];
  int nt_begin[kMaxGemmOps + 1];  // column tiles [nt_begin[i], nt_begin[i + 1]) belong to op i; unused entries: INT_MAX
  int gelu_lut;  // pair kernel: 1 = a weight of the launch has the bias + GELU epilogue: tabulate erf in shared memory
  // stream-K schedule of the pair kernel (see TileSched): 0 = whole tiles dealt round-robin
  int streamk;
  unsigned* ws_flags;  // [pair][CTA rank][epilogue warp]: 1 = that warp's rows of the pair's partial tile are in ws_partial
  float* ws_partial;   // [pair][CTA rank][128 rows][BN] f32
  int debug;  // experiments (B200Q_GEMM_DEBUG): 1 = operands are not loaded / unpacked (protocol intact, data garbage), 2 = epilogue does not store
};

// ---- dequantize one 8-column chunk of one packed word row into the B stage ---------------------------
// w[8]: the words of columns n0..n0+7 of word row r; s[4]: their scales as half2 pairs (n, n+1).
// Writes kNpu k-rows; k row `kl` (stage-local) of chunk `chunk` (8 columns) lives at
//   (kl / 8) * kKAtomBytes + (chunk / 8) * 1024 + (kl % 8) * 128 + ((chunk % 8) ^ (kl % 8)) * 16
template <int KATOM = kKAtomBytes>
__device__ __forceinline__ uint32_t b_chunk_addr(uint32_t stage, int kl, int chunk) {
  return stage + (kl >> 3) * KATOM + (chunk >> 3) * 1024 + (kl & 7) * 128 + ((((chunk & 7) ^ (kl & 7))) << 4);
}
__device__ __forceinline__ void sts128(uint32_t a, uint32_t x, uint32_t y, uint32_t z, uint32_t w) {
  asm volatile("st.shared.v4.u32 [%0], {%1,%2,%3,%4};" ::"r"(a), "r"(x), "r"(y), "r"(z), "r"(w) : "memory");
}

template <int FMT, int KATOM = kKAtomBytes>
__device__ __forceinline__ void dequant_chunk(const uint32_t (&w)[8], const uint32_t (&s)[4], const uint32_t (&mn)[4],
                                              uint32_t stage, int kl0, int chunk) {
  const uint32_t kMagic = 0x64006400u;
  (void)mn;
  if constexpr (FMT == B200Q_Q4F16_ASYM) {
    // w = half(half(q * s) + m): q exact from the magic numbers, then the reference's two roundings (never fused)
    const __half2 kSubLo = u32_as_h2(0xE400E400u);  // -1024
    const __half2 kMul16 = u32_as_h2(0x2C002C00u);  // 1/16
    const __half2 kSubHi = u32_as_h2(0xD400D400u);  // -64
    uint32_t o[8][4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const __half2 s2 = u32_as_h2(s[j]), m2 = u32_as_h2(mn[j]);
#pragma unroll
      for (int h = 0; h < 2; ++h) {
        const uint32_t p = __byte_perm(w[2 * j], w[2 * j + 1], h ? 0x7632 : 0x5410);
        const uint32_t p8 = p >> 8;
        const __half2 q0 = __hadd2(u32_as_h2(lop3_and_or(p, 0x000F000Fu, kMagic)), kSubLo);
        const __half2 q1 = __hfma2(u32_as_h2(lop3_and_or(p, 0x00F000F0u, kMagic)), kMul16, kSubHi);
        const __half2 q2 = __hadd2(u32_as_h2(lop3_and_or(p8, 0x000F000Fu, kMagic)), kSubLo);
        const __half2 q3 = __hfma2(u32_as_h2(lop3_and_or(p8, 0x00F000F0u, kMagic)), kMul16, kSubHi);
        o[4 * h + 0][j] = h2_as_u32(__hadd2_rn(__hmul2_rn(q0, s2), m2));
        o[4 * h + 1][j] = h2_as_u32(__hadd2_rn(__hmul2_rn(q1, s2), m2));
        o[4 * h + 2][j] = h2_as_u32(__hadd2_rn(__hmul2_rn(q2, s2), m2));
        o[4 * h + 3][j] = h2_as_u32(__hadd2_rn(__hmul2_rn(q3, s2), m2));
      }
    }
#pragma unroll
    for (int i = 0; i < 8; ++i) sts128(b_chunk_addr<KATOM>(stage, kl0 + i, chunk), o[i][0], o[i][1], o[i][2], o[i][3]);
  } else if constexpr (FMT == B200Q_Q4F16_0) {
    const __half2 kSubLo = u32_as_h2(0xE407E407u);  // -(1024 + 7)
    const __half2 kMul16 = u32_as_h2(0x2C002C00u);  // 1/16
    const __half2 kSubHi = u32_as_h2(0xD470D470u);  // -(64 + 7)
    uint32_t o[8][4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const __half2 s2 = u32_as_h2(s[j]);
#pragma unroll
      for (int h = 0; h < 2; ++h) {  // nibbles 0-3 / 4-7 of columns (2j, 2j+1)
        const uint32_t p = __byte_perm(w[2 * j], w[2 * j + 1], h ? 0x7632 : 0x5410);
        const uint32_t p8 = p >> 8;
        o[4 * h + 0][j] = h2_as_u32(__hmul2(__hadd2(u32_as_h2(lop3_and_or(p, 0x000F000Fu, kMagic)), kSubLo), s2));
        o[4 * h + 1][j] = h2_as_u32(__hmul2(__hfma2(u32_as_h2(lop3_and_or(p, 0x00F000F0u, kMagic)), kMul16, kSubHi), s2));
        o[4 * h + 2][j] = h2_as_u32(__hmul2(__hadd2(u32_as_h2(lop3_and_or(p8, 0x000F000Fu, kMagic)), kSubLo), s2));
        o[4 * h + 3][j] = h2_as_u32(__hmul2(__hfma2(u32_as_h2(lop3_and_or(p8, 0x00F000F0u, kMagic)), kMul16, kSubHi), s2));
      }
    }
#pragma unroll
    for (int i = 0; i < 8; ++i) sts128(b_chunk_addr<KATOM>(stage, kl0 + i, chunk), o[i][0], o[i][1], o[i][2], o[i][3]);
  } else if constexpr (FMT == B200Q_Q3F16_0) {
    const __half2 kSub1 = u32_as_h2(0xE403E403u);   // -(1024 + 3)
    const __half2 kMul8 = u32_as_h2(0x30003000u);   // 1/8
    const __half2 kSub8 = u32_as_h2(0xD818D818u);   // -(128 + 3)
    const __half2 kMul64 = u32_as_h2(0x24002400u);  // 1/64
    const __half2 kSub64 = u32_as_h2(0xCCC0CCC0u);  // -(16 + 3)
    uint32_t o[10][4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const __half2 s2 = u32_as_h2(s[j]);
#pragma unroll
      for (int h = 0; h < 2; ++h) {  // values 0-4 / 5-9 of columns (2j, 2j+1)
        const uint32_t p = h ? __byte_perm(w[2 * j] >> 15, w[2 * j + 1] >> 15, 0x5410)
                             : __byte_perm(w[2 * j], w[2 * j + 1], 0x5410);
        const uint32_t q = p >> 9;
        o[5 * h + 0][j] = h2_as_u32(__hmul2(__hadd2(u32_as_h2(lop3_and_or(p, 0x00070007u, kMagic)), kSub1), s2));
        o[5 * h + 1][j] = h2_as_u32(__hmul2(__hfma2(u32_as_h2(lop3_and_or(p, 0x00380038u, kMagic)), kMul8, kSub8), s2));
        o[5 * h + 2][j] = h2_as_u32(__hmul2(__hfma2(u32_as_h2(lop3_and_or(p, 0x01C001C0u, kMagic)), kMul64, kSub64), s2));
        o[5 * h + 3][j] = h2_as_u32(__hmul2(__hadd2(u32_as_h2(lop3_and_or(q, 0x00070007u, kMagic)), kSub1), s2));
        o[5 * h + 4][j] = h2_as_u32(__hmul2(__hfma2(u32_as_h2(lop3_and_or(q, 0x00380038u, kMagic)), kMul8, kSub8), s2));
      }
    }
#pragma unroll
    for (int i = 0; i < 10; ++i) sts128(b_chunk_addr<KATOM>(stage, kl0 + i, chunk), o[i][0], o[i][1], o[i][2], o[i][3]);
  } else {  // Q8: 4 values per word
    const __half2 kSub = u32_as_h2(0xE47FE47Fu);  // -(1024 + 127)
    uint32_t o[4][4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const __half2 s2 = u32_as_h2(s[j]);
      // bytes (b_i of column 2j, b_i of column 2j+1) -> low bytes of the two halves, 0x64 above them
      const uint32_t p02 = __byte_perm(w[2 * j], w[2 * j + 1], 0x6240);  // b0 b0' at bytes 0,1 ; b2 b2' at 2,3
      const uint32_t p13 = __byte_perm(w[2 * j], w[2 * j + 1], 0x7351);
      const uint32_t v0 = __byte_perm(p02, 0x64646464u, 0x4140);  // b0, 0x64, b0', 0x64
      const uint32_t v2 = __byte_perm(p02, 0x64646464u, 0x4342);
      const uint32_t v1 = __byte_perm(p13, 0x64646464u, 0x4140);
      const uint32_t v3 = __byte_perm(p13, 0x64646464u, 0x4342);
      o[0][j] = h2_as_u32(__hmul2(__hadd2(u32_as_h2(v0), kSub), s2));
      o[1][j] = h2_as_u32(__hmul2(__hadd2(u32_as_h2(v1), kSub), s2));
      o[2][j] = h2_as_u32(__hmul2(__hadd2(u32_as_h2(v2), kSub), s2));
      o[3][j] = h2_as_u32(__hmul2(__hadd2(u32_as_h2(v3), kSub), s2));
    }
#pragma unroll
    for (int i = 0; i < 4; ++i) sts128(b_chunk_addr<KATOM>(stage, kl0 + i, chunk), o[i][0], o[i][1], o[i][2], o[i][3]);
  }
}

// ---- epilogue: 32 consecutive columns of one output row ------------------------------------------------
// y = f16(acc) first (the reference's matmul result is an f16 tensor), then every epilogue op in f16,
// op for op (ep_f16 in common.cuh).  bias operands are [N]; residual / multiply operands are [M, N].
// GELU through a table: the reference's op sequence is  v = y + b; t = v * f16(0.7071..); e = f16(erf(f32(t))); out =
// v * (f16(0.5) + e * f16(0.5)),  i.e. e is a function of the 16 bits of t.  The pair kernel's epilogue warps tabulate
// e = f16(erff(t)) for every f16 t in [0, 4) (17408 entries, 34 KB of shared memory; erf is odd, |t| >= 4 gives 1)
// while the first tile's main loop runs, and a lookup replaces the ~40 instructions of erff() - bit-identical by
// construction.  2048 x 2560 -> 10240 + bias + GELU: 144 -> 1xx us.
constexpr int kGeluLutEntries = 0x4400;  // f16 bit patterns of [0, 4)
constexpr int kGeluLutBytes = kGeluLutEntries * 2;
__device__ __forceinline__ __half gelu_lut_f16(float acc, __half bias, uint32_t lut) {
  const __half v = __hadd(__float2half_rn(acc), bias);
  const __half t = __hmul(v, __float2half_rn(0.70710678118654757f));
  const uint32_t tb = __half_as_ushort(t), a = tb & 0x7FFFu;
  uint32_t eb;
  if (a < (uint32_t)kGeluLutEntries) {
    unsigned short r;
    asm volatile("ld.shared.u16 %0, [%1];" : "=h"(r) : "r"(lut + 2u * a));
    eb = r;
  } else {
    eb = (a > 0x7C00u) ? 0x7E00u : 0x3C00u;  // NaN stays NaN, |t| >= 4 (and inf): erf = 1
  }
  const __half e = __ushort_as_half((unsigned short)(eb | (tb & 0x8000u)));
  const __half h = __hadd(__float2half_rn(0.5f), __hmul(e, __float2half_rn(0.5f)));
  return __hmul(v, h);
}

template <class P>  // GemmParams (single-CTA kernel) or GemmOp (pair kernel)
__device__ __forceinline__ void epilogue_row(const P& p, const uint32_t (&acc)[32], int m, int n0, uint32_t gelu_lut) {
  // the table variant: bias + GELU only (the launcher picks it for exactly that epilogue)
  const size_t row = (size_t)m * p.N;
  __half* y = reinterpret_cast<__half*>(p.Y) + row + n0;
#pragma unroll
  for (int c = 0; c < 32; c += 8) {
    if (n0 + c >= p.N) break;
    __align__(16) __half o[8], a0[8];
    *reinterpret_cast<uint4*>(a0) = *reinterpret_cast<const uint4*>(p.ep0 + n0 + c);
#pragma unroll
    for (int i = 0; i < 8; ++i) o[i] = gelu_lut_f16(__uint_as_float(acc[c + i]), a0[i], gelu_lut);
    *reinterpret_cast<uint4*>(y + c) = *reinterpret_cast<const uint4*>(o);
  }
}
template <class P>
__device__ __forceinline__ void epilogue_row(const P& p, const uint32_t (&acc)[32], int m, int n0) {
  const int ep = p.epilogue;
  const size_t row = (size_t)m * p.N;
  if (ep == B200Q_EP_CAST_F32) {
    float* y = reinterpret_cast<float*>(p.Y) + row + n0;
#pragma unroll
    for (int c = 0; c < 32; c += 4) {
      if (n0 + c < p.N)
        *reinterpret_cast<float4*>(y + c) = make_float4(__uint_as_float(acc[c]), __uint_as_float(acc[c + 1]),
                                                        __uint_as_float(acc[c + 2]), __uint_as_float(acc[c + 3]));
    }
    return;
  }
  __half* y = reinterpret_cast<__half*>(p.Y) + row + n0;
  const bool bias_like = (ep == B200Q_EP_BIAS || ep == B200Q_EP_BIAS_GELU || ep == B200Q_EP_BIAS_ADD2);
  const __half* e0 = p.ep0 ? (bias_like ? p.ep0 + n0 : p.ep0 + row + n0) : nullptr;
#pragma unroll
  for (int c = 0; c < 32; c += 8) {
    if (n0 + c >= p.N) break;  // N % 8 == 0: a chunk of 8 is all in or all out
    __align__(16) __half o[8];
    __align__(16) __half a0[8], a1[8], a2[8];
    if (e0) *reinterpret_cast<uint4*>(a0) = *reinterpret_cast<const uint4*>(e0 + c);
    if (ep == B200Q_EP_BIAS_ADD2) {
      *reinterpret_cast<uint4*>(a1) = *reinterpret_cast<const uint4*>(p.ep1 + row + n0 + c);
      *reinterpret_cast<uint4*>(a2) = *reinterpret_cast<const uint4*>(p.ep2 + row + n0 + c);
    }
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      const float v = __uint_as_float(acc[c + i]);
      if (ep == B200Q_EP_BIAS_ADD2) {
        o[i] = __hadd(__hadd(__hadd(__float2half_rn(v), a0[i]), a1[i]), a2[i]);
      } else {
        o[i] = ep_f16(v, ep, a0, i);
      }
    }
    *reinterpret_cast<uint4*>(y + c) = *reinterpret_cast<const uint4*>(o);
  }
}

// smem map (dynamic, 1024-aligned): [A ring][B ring][barriers]
template <int MODE>
struct Smem {
  using B = BCfg<MODE>;
  static constexpr int kAOff = 0;
  static constexpr int kBOff = kAStages * kAStageBytes;
  static constexpr int kBarOff = kBOff + B::kStages * B::kStageBytes;
  // barriers: a_full, a_empty [kAStages]; b_full, b_empty [B::kStages]; tmem_full, tmem_empty [2]; tmem ptr
  static constexpr int kNumBars = 2 * kAStages + 2 * B::kStages + 4;
  static constexpr int kBytes = kBarOff + kNumBars * 8 + 16;
};

template <int MODE>
__global__ void __launch_bounds__(kGemmThreads, 1) gemm_tc_kernel(const __grid_constant__ GemmParams p) {
  using B = BCfg<MODE>;
  using S = Smem<MODE>;
  constexpr bool kDenseB = (MODE == kDense);
  extern __shared__ __align__(1024) unsigned char smem_raw[];
  // the runtime only guarantees 16-byte alignment of dynamic smem: align by hand (host adds 1024 slack)
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  unsigned char* base_ptr = smem_raw + (base - smem_u32(smem_raw));
  const uint32_t a_ring = base + S::kAOff, b_ring = base + S::kBOff;
  const uint32_t bars = base + S::kBarOff;
  const uint32_t a_full = bars, a_empty = a_full + 8 * kAStages;
  const uint32_t b_full = a_empty + 8 * kAStages, b_empty = b_full + 8 * B::kStages;
  const uint32_t t_full = b_empty + 8 * B::kStages, t_empty = t_full + 16;
  volatile uint32_t* tmem_ptr_smem = reinterpret_cast<volatile uint32_t*>(base_ptr + S::kBarOff + S::kNumBars * 8);

  const int warp = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0), lane = threadIdx.x & 31;  // provably warp-uniform
  const int n_tiles = p.tiles_m * p.tiles_n;
  const int n_ksteps = (p.K + 15) >> 4;
  const int n_a_stages = (n_ksteps + 3) >> 2;
  const int n_b_stages = (n_ksteps + B::kKSteps - 1) / B::kKSteps;

  if (warp == 0 && lane == 0) {
    prefetch_tmap(&p.tm_a);
    if constexpr (kDenseB) prefetch_tmap(&p.tm_b);
    for (int s = 0; s < kAStages; ++s) {
      mbar_init(a_full + 8 * s, 1);
      mbar_init(a_empty + 8 * s, 1);
    }
    for (int s = 0; s < B::kStages; ++s) {
      mbar_init(b_full + 8 * s, kDenseB ? 1 : kDqWarps);
      mbar_init(b_empty + 8 * s, 1);
    }
    for (int s = 0; s < 2; ++s) {
      mbar_init(t_full + 8 * s, 1);
      mbar_init(t_empty + 8 * s, kEpWarps);
    }
    fence_barrier_init();
  }
  if (warp == 1) {
    tmem_alloc(smem_u32(const_cast<uint32_t*>(tmem_ptr_smem)), kTmemCols);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr_smem;

  if (warp == 0) {
    // ================= TMA producer (warp-uniform loop, one elected lane issues) ==========================
    int sa = 0, sb = 0;
    uint32_t pa = 0, pb = 0;
    for (int tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
      const int m0 = (tile % p.tiles_m) * kBM, n0 = (tile / p.tiles_m) * kBN;
      for (int i = 0; i < n_a_stages; ++i) {
        mbar_wait(a_empty + 8 * sa, pa ^ 1);
        if (elect_one()) {
          if (p.debug & 1) {
            mbar_arrive(a_full + 8 * sa);
          } else {
            mbar_expect_tx(a_full + 8 * sa, kAStageBytes);
            tma_load_2d(a_ring + sa * kAStageBytes, &p.tm_a, i * kAK, m0, a_full + 8 * sa);
          }
        }
        if (++sa == kAStages) { sa = 0; pa ^= 1; }
        if constexpr (kDenseB) {  // same stage cadence for the dense B operand
          mbar_wait(b_empty + 8 * sb, pb ^ 1);
          if (elect_one()) {
            if (p.debug & 1) {
              mbar_arrive(b_full + 8 * sb);
            } else {
              mbar_expect_tx(b_full + 8 * sb, B::kStageBytes);
              tma_load_2d(b_ring + sb * B::kStageBytes, &p.tm_b, i * kAK, n0, b_full + 8 * sb);
            }
          }
          if (++sb == B::kStages) { sb = 0; pb ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    // ================= MMA issuer (warp-uniform loop, one elected lane issues) ===========================
    // experiments: bits 4-5 of debug shrink the UMMA N (garbage results)
    const uint32_t idesc = make_idesc_f16(kBM, kBN >> ((p.debug >> 4) & 3), 0, kDenseB ? 0 : 1);
    // descriptor high words are constants; the low word carries the start address
    const uint64_t da_hi = make_smem_desc_sw128(0, 16, 1024) & 0xFFFFFFFF00000000ull;
    const uint64_t db_hi = (kDenseB ? make_smem_desc_sw128(0, 16, 1024) : make_smem_desc_sw128(0, 1024, kKAtomBytes)) &
                           0xFFFFFFFF00000000ull;
    const uint32_t da_lo0 = (uint32_t)make_smem_desc_sw128(0, 16, 1024);
    const uint32_t db_lo0 = (uint32_t)(kDenseB ? make_smem_desc_sw128(0, 16, 1024) : make_smem_desc_sw128(0, 1024, kKAtomBytes));
    constexpr uint32_t kBStep = kDenseB ? 32u : 2u * kKAtomBytes;  // bytes per 16-k step inside a B stage
    // the whole loop runs in ONE elected lane: an elect + branch + __syncwarp per 16-k step costs ~100 cycles of
    // issue latency, and a 128 x 256 x 16 MMA lasts only ~176 - the tensor pipe would wait for its issuer
    if (elect_one()) {
      int sa = 0, sb = 0, it = 0;
      uint32_t pa = 0, pb = 0;
      for (int tile = blockIdx.x; tile < n_tiles; tile += gridDim.x, ++it) {
        const int acc = it & 1;
        mbar_wait(t_empty + 8 * acc, ((it >> 1) & 1) ^ 1);
        tc_fence_after();
        const uint32_t d = tmem_base + acc * kBN;
        int ka = 0, kb = 0;  // k-step inside the current A / B stage
        for (int ks = 0; ks < n_ksteps; ++ks) {
          if (ka == 0) mbar_wait(a_full + 8 * sa, pa);
          if (kb == 0) mbar_wait(b_full + 8 * sb, pb);
          if (ka == 0 || kb == 0) tc_fence_after();
          // A: K-major SW128, 8-row groups 1024 B apart; +32 B per 16-k step inside the swizzle row
          // B: MN-major SW128: 64-column atoms 1024 B apart (LBO), 8-k slabs kKAtomBytes apart (SBO)
          const uint32_t a_addr = a_ring + sa * kAStageBytes + ka * 32;
          const uint32_t b_addr = b_ring + sb * B::kStageBytes + kb * kBStep;
          const uint64_t da = da_hi | (uint64_t)(da_lo0 | ((a_addr & 0x3FFFFu) >> 4));
          const uint64_t db = db_hi | (uint64_t)(db_lo0 | ((b_addr & 0x3FFFFu) >> 4));
          const bool last = (ks == n_ksteps - 1);
          const bool rel_a = (ka == 3) || last, rel_b = (kb == B::kKSteps - 1) || last;
          umma_f16(d, da, db, idesc, ks > 0 ? 1u : 0u);
          if (rel_a) umma_commit(a_empty + 8 * sa);
          if (rel_b) umma_commit(b_empty + 8 * sb);
          if (last) umma_commit(t_full + 8 * acc);
          if (rel_a) { ka = 0; if (++sa == kAStages) { sa = 0; pa ^= 1; } } else { ++ka; }
          if (rel_b) { kb = 0; if (++sb == B::kStages) { sb = 0; pb ^= 1; } } else { ++kb; }
        }
      }
    }
    __syncwarp();
  } else if (warp < kFirstDqWarp) {
    // ================= epilogue ============================================================================
    const int quarter = warp & 3;  // TMEM lanes [32 * quarter, +32) are the ones this warp may read
    int it = 0;
    for (int tile = blockIdx.x; tile < n_tiles; tile += gridDim.x, ++it) {
      const int acc = it & 1;
      const int m0 = (tile % p.tiles_m) * kBM, n0 = (tile / p.tiles_m) * kBN;
      mbar_wait(t_full + 8 * acc, (it >> 1) & 1);
      tc_fence_after();
      const int m = m0 + quarter * 32 + lane;
#pragma unroll 1
      for (int c = 0; c < kBN; c += 32) {
        uint32_t v[32];
        tmem_ld_32x32(tmem_base + ((uint32_t)(quarter * 32) << 16) + acc * kBN + c, v);
        tmem_ld_wait();
        if (m < p.M && n0 + c < p.N && !(p.debug & 2)) epilogue_row(p, v, m, n0 + c);
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(t_empty + 8 * acc);
    }
  } else {
    // ================= dequantizers =========================================================================
    if constexpr (!kDenseB) {
      using F = Fmt<MODE>;
      const int dw = warp - kFirstDqWarp;
      const int chunk = lane;  // 8 columns each: 32 chunks = 256 columns
      int sb = 0;
      uint32_t pb = 0;
      uint32_t w[B::kItems][8], s[B::kItems][4], mn[B::kItems][4];
      auto load_stage = [&](int n0, int i) {
#pragma unroll
        for (int t = 0; t < B::kItems; ++t) {
          const int r = i * B::kRows + dw + t * kDqWarps;  // global packed word row
          const int n = n0 + chunk * 8;
          if (r < p.n_rows && n < p.N) {
            const uint4* src = reinterpret_cast<const uint4*>(p.packed + (size_t)r * p.N + n);
            const uint4 lo = __ldg(src), hi = __ldg(src + 1);
            w[t][0] = lo.x; w[t][1] = lo.y; w[t][2] = lo.z; w[t][3] = lo.w;
            w[t][4] = hi.x; w[t][5] = hi.y; w[t][6] = hi.z; w[t][7] = hi.w;
            const uint4 sv = __ldg(reinterpret_cast<const uint4*>(p.scale + (size_t)(r / F::kRowsPerGroup) * p.N + n));
            s[t][0] = sv.x; s[t][1] = sv.y; s[t][2] = sv.z; s[t][3] = sv.w;
            if constexpr (fmt_has_min_c(MODE)) {
              const uint4 mv = __ldg(reinterpret_cast<const uint4*>(p.zmin + (size_t)(r / F::kRowsPerGroup) * p.N + n));
              mn[t][0] = mv.x; mn[t][1] = mv.y; mn[t][2] = mv.z; mn[t][3] = mv.w;
            }
          } else {
#pragma unroll
            for (int j = 0; j < 8; ++j) w[t][j] = 0u;
#pragma unroll
            for (int j = 0; j < 4; ++j) s[t][j] = 0u;
#pragma unroll
            for (int j = 0; j < 4; ++j) mn[t][j] = 0u;
          }
        }
      };
      for (int tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
        const int n0 = (tile / p.tiles_m) * kBN;
        load_stage(n0, 0);
        for (int i = 0; i < n_b_stages; ++i) {
          uint32_t wc[B::kItems][8], sc[B::kItems][4], mc[B::kItems][4];
#pragma unroll
          for (int t = 0; t < B::kItems; ++t) {
#pragma unroll
            for (int j = 0; j < 8; ++j) wc[t][j] = w[t][j];
#pragma unroll
            for (int j = 0; j < 4; ++j) sc[t][j] = s[t][j];
#pragma unroll
            for (int j = 0; j < 4; ++j) mc[t][j] = fmt_has_min_c(MODE) ? mn[t][j] : 0u;
          }
          if (i + 1 < n_b_stages) load_stage(n0, i + 1);  // in flight while this stage is unpacked
          mbar_wait(b_empty + 8 * sb, pb ^ 1);
          const uint32_t stage = b_ring + sb * B::kStageBytes;
          if (!(p.debug & 1)) {
#pragma unroll
            for (int t = 0; t < B::kItems; ++t)
              dequant_chunk<MODE>(wc[t], sc[t], mc[t], stage, (dw + t * kDqWarps) * F::kNpu, chunk);
          }
          fence_proxy_async_smem();
          __syncwarp();
          if (lane == 0) mbar_arrive(b_full + 8 * sb);
          if (++sb == B::kStages) { sb = 0; pb ^= 1; }
        }
      }
    }
  }

  // ---- teardown ---------------------------------------------------------------------------------------------
  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, kTmemCols);
  }
}


// =====================================================================================================================
// CTA-pair variant (cta_group::2): one cluster = the two SMs of a TPC works on a 256 x BN output tile.
//   * every CTA keeps ITS 128 rows of X (TMA) and dequantizes ITS HALF of the weight tile's columns: the unpack work
//     and the shared-memory B traffic per SM are half of the single-CTA kernel's (128 x 256 tiles are shared-memory
//     bound: 139 B/cycle of operand traffic against 128 B/cycle of bandwidth; the pair needs 91);
//   * ONE tcgen05.mma.cta_group::2 (M = 256) issued by the leader's MMA lane reads A from both CTAs and the two B
//     halves; each CTA's TMEM receives the accumulator of its own 128 rows; each CTA's epilogue warps drain it;
//   * barriers: "full" barriers and the accumulator-free barrier live in the LEADER (the peer's TMA counts its bytes
//     there through the pair form of cp.async.bulk.tensor, the peer's dequant / epilogue warps arrive remotely);
//     "empty" barriers and accumulator-full barriers exist in both CTAs and are signalled by multicast commits.
// Warp roles per CTA: 0 TMA producer, 1 MMA issue (leader) + TMEM alloc (both), 2-5 epilogue, 6.. dequant.
// ---- work distribution of the pair kernel ------------------------------------------------------------------------
// classic: pair p computes the whole tiles p, p + P, p + 2P, ..  With 80 tiles on 74 pairs (M = 2048, N = 2560) the
// second round is 8 % full and the kernel takes two tile times for 1.08 tile times of work.
// stream-K: the (tile, k chunk) space - tiles in order, `C` chunks of `chunk` 16-k steps each - is cut into P equal
// ranges.  A range is: the tail of a tile (chunks [c0, C), c0 > 0), whole tiles, the head of a tile ([0, c1), c1 < C).
// The head's owner finishes that tile: the pairs behind it write their f32 partial tile (at most ONE per pair: its
// first segment) into the workspace slot [pair][rank] and raise a flag per epilogue warp; the finisher adds the slots
// in pair order (deterministic) and applies the epilogue.  A writer never waits, so there is no deadlock whatever the
// residency; a finisher processes its head FIRST (right after its own tail), so the partials it waits for - the
// tails the following pairs start with - are on their way, and its fix-up overlaps the whole tiles that follow.
__device__ __forceinline__ unsigned ld_acquire_u32(const unsigned* q) {
  unsigned v;
  asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(q) : "memory");
  return v;
}
__device__ __forceinline__ void st_release_u32(unsigned* q, unsigned v) {
  asm volatile("st.release.gpu.global.u32 [%0], %1;" ::"l"(q), "r"(v) : "memory");
}
struct Seg {
  int tile, ks0, ks1;  // 16-k steps [ks0, ks1) of the tile
  int kind;            // 0 whole tile, 1 partial writer, 2 finisher
};
struct TileSched {
  int streamk, pair, n_pairs, n_ksteps, chunk, C;
  long long W;
  int n_seg, first_tile, a_off, b_off;
  bool has_tail, has_head;
  __host__ __device__ void init(int streamk_, int pair_, int n_pairs_, int n_tiles, int n_ksteps_, int chunk_) {
    streamk = streamk_; pair = pair_; n_pairs = n_pairs_; n_ksteps = n_ksteps_; chunk = chunk_;
    C = (n_ksteps + chunk - 1) / chunk;
    W = (long long)n_tiles * C;
    has_tail = has_head = false;
    first_tile = a_off = b_off = 0;
    if (!streamk) {
      n_seg = pair < n_tiles ? (n_tiles - pair + n_pairs - 1) / n_pairs : 0;
      return;
    }
    const long long a = W * pair / n_pairs, b = W * (pair + 1) / n_pairs;
    if (a == b) { n_seg = 0; return; }
    first_tile = (int)(a / C);
    a_off = (int)(a - (long long)first_tile * C);
    const int last_tile = (int)((b - 1) / C);
    b_off = (int)(b - (long long)last_tile * C);
    n_seg = last_tile - first_tile + 1;
    has_tail = a_off > 0;
    has_head = b_off < C && (n_seg > 1 || !has_tail);
  }
  __host__ __device__ long long range_begin(int q) const { return W * q / n_pairs; }
  // i-th segment in PROCESSING order: the tail, the head, then the whole tiles
  __host__ __device__ Seg get(int i) const {
    Seg sg;
    if (!streamk) {
      sg.tile = pair + i * n_pairs; sg.ks0 = 0; sg.ks1 = n_ksteps; sg.kind = 0;
      return sg;
    }
    int j = i;
    if (has_tail) { if (has_head && i >= 1) j = (i == 1) ? n_seg - 1 : i - 1; }
    else if (has_head) j = (i == 0) ? n_seg - 1 : i - 1;
    const int c0 = (j == 0) ? a_off : 0, c1 = (j == n_seg - 1) ? b_off : C;
    sg.tile = first_tile + j;
    sg.ks0 = c0 * chunk;
    sg.ks1 = (c1 * chunk < n_ksteps) ? c1 * chunk : n_ksteps;
    sg.kind = (c0 > 0) ? 1 : (c1 < C ? 2 : 0);
    return sg;
  }
  // pairs behind this one that hold a partial of `tile` (the finisher's contributors): (pair, end)
  __host__ __device__ int contributors_end(int tile) const {
    int q = pair + 1;
    while (q < n_pairs && range_begin(q) < (long long)(tile + 1) * C) ++q;
    return q;
  }
};
constexpr int kWsFlagBytes = 16384;  // flags of up to 512 pairs

#ifndef B200Q_GEMM_BSTAGES
#define B200Q_GEMM_BSTAGES 6
#endif
template <int MODE, int BN, bool kGelu = false>
struct Pair {
  using B = BCfg<MODE>;
  static constexpr bool kDenseB = (MODE == kDense);
  static constexpr int kHalfN = BN / 2;                 // columns of B this CTA owns
  static constexpr int kAtoms = kHalfN / 64;            // 64-column swizzle atoms: 2 (BN 256) or 1 (BN 128)
  static constexpr int kKAtom = kAtoms * 1024;          // one 8-k slab of a B stage
  static constexpr int kChunks = kHalfN / 8;            // 16-byte column chunks per k row: 16 or 8
  static constexpr int kRowsPerPass = 32 / kChunks;     // word rows one warp covers per pass: 2 or 4
  static constexpr int kDq = 8 / kRowsPerPass;          // dequant warps per B stage: 4 or 2 (8 word rows per pass)
  static constexpr int kDqGroups = 8 / kDq;             // warp groups: group g fills the stages c with c % groups == g, so
                                                        // a warp has `groups` stage times per stage of its own (its global
                                                        // loads run that far ahead) - with one group the warps sat on the
                                                        // L2 latency of their packed words (issue slots 12 % busy)
  static constexpr int kDqWarpsAll = kDq * kDqGroups;   // 8
  static constexpr int kBK = B::kBK;
  static constexpr int kRows = kDenseB ? 8 : B::kRows;  // packed word rows per stage: 8 / 8 / 16
  static constexpr int kItems = kRows / 8;              // passes per stage: 1 (2 for int8)
  static constexpr int kBStage = kHalfN * kBK * 2;      // 16 KB (20 KB int3) / half of it for BN 128
  // a multiple of kDqGroups; the GELU-table instantiation gives 34 KB of the ring to the table
  static constexpr int kBStages = (BN == 256) ? (kGelu ? (kAStagesP > 4 ? 4 : 6) : B200Q_GEMM_BSTAGES) : 8;
  static constexpr int kThreads = (kFirstDqWarp + (kDenseB ? 0 : kDqWarpsAll)) * 32;
  static constexpr int kAOff = 0;
  static constexpr int kBOff = kAStagesP * kAStageBytes;
  static constexpr int kBarOff = kBOff + kBStages * kBStage;
  static constexpr int kNumBars = 2 * kAStagesP + 2 * kBStages + 4;
  static constexpr int kBytes = kBarOff + kNumBars * 8 + 16;
  static constexpr int kLutOff = (kBytes + 15) / 16 * 16;  // GELU table (only allocated for launches with a GELU epilogue)
  static constexpr int kTmemCols = 2 * BN;              // two accumulators: 512 or 256 columns
  // stream-K cuts a tile's k range where an X stage (4 steps of 16 k) and a weight stage (kKSteps) both begin
  static constexpr int kChunk = (B::kKSteps % 4 == 0) ? B::kKSteps : 4 * B::kKSteps;
};

template <int MODE, int BN, bool kGelu = false>
__global__ void __launch_bounds__(Pair<MODE, BN, kGelu>::kThreads, 1) gemm_tc2_kernel(const __grid_constant__ GemmParams p) {
  using C = Pair<MODE, BN, kGelu>;
  using B = BCfg<MODE>;
  constexpr bool kDenseB = C::kDenseB;
  extern __shared__ __align__(1024) unsigned char smem_raw[];
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  unsigned char* base_ptr = smem_raw + (base - smem_u32(smem_raw));
  const uint32_t a_ring = base + C::kAOff, b_ring = base + C::kBOff;
  const uint32_t bars = base + C::kBarOff;
  const uint32_t a_full = bars, a_empty = a_full + 8 * kAStagesP;
  const uint32_t b_full = a_empty + 8 * kAStagesP, b_empty = b_full + 8 * C::kBStages;
  const uint32_t t_full = b_empty + 8 * C::kBStages, t_empty = t_full + 16;
  volatile uint32_t* tmem_ptr_smem = reinterpret_cast<volatile uint32_t*>(base_ptr + C::kBarOff + C::kNumBars * 8);

  const int warp = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0), lane = threadIdx.x & 31;
  const uint32_t rank = cluster_ctarank();  // 0 = leader
  pdl_launch_dependents();  // (no-op unless launched with B200Q_FLAG_PDL)
  const int pair = blockIdx.x >> 1, n_pairs = gridDim.x >> 1;
  const int tiles_m = (p.M + 255) >> 8;
  const int tiles_n = p.nt_begin[p.n_ops];  // over all weights of the launch
  const int n_tiles = tiles_m * tiles_n;
  // column tile -> (weight, first column inside it)
  auto op_of = [&](int tn) { return (tn >= p.nt_begin[1] ? 1 : 0) + (tn >= p.nt_begin[2] ? 1 : 0) + (tn >= p.nt_begin[3] ? 1 : 0); };
  // (the weights are kernel parameters: they are only ever indexed with constants, so every field stays a
  // constant-bank operand - a run-time index made ptxas copy the table to local memory, 2x slower kernels)
  const int n_ksteps = (p.K + 15) >> 4;
  TileSched sched;
  sched.init(p.streamk, pair, n_pairs, n_tiles, n_ksteps, C::kChunk);

  if (warp == 0 && lane == 0) {
    prefetch_tmap(&p.tm_a);
    if constexpr (kDenseB) prefetch_tmap(&p.tm_b);
    for (int s = 0; s < kAStagesP; ++s) {
      mbar_init(a_full + 8 * s, 1);   // the leader's expect_tx arrive; both CTAs' TMA bytes
      mbar_init(a_empty + 8 * s, 1);  // multicast commit
    }
    for (int s = 0; s < C::kBStages; ++s) {
      mbar_init(b_full + 8 * s, kDenseB ? 1 : 2 * C::kDq);  // the dequant warps of BOTH CTAs
      mbar_init(b_empty + 8 * s, 1);
    }
    for (int s = 0; s < 2; ++s) {
      mbar_init(t_full + 8 * s, 1);
      mbar_init(t_empty + 8 * s, 2 * kEpWarps);  // the epilogue warps of BOTH CTAs
    }
    fence_barrier_init();
  }
  if (warp == 1) {
    tmem_alloc2(smem_u32(const_cast<uint32_t*>(tmem_ptr_smem)), C::kTmemCols);
    tmem_relinquish2();
  }
  tc_fence_before();
  cluster_sync_all();  // both CTAs' barriers exist before anything is signalled across the pair
  tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr_smem;

  if (warp == 0) {
    // ================= TMA producer: this CTA's 128 rows of X (and, dense mode, its BN/2 rows of T) ==============
    pdl_wait();  // (programmatic dependent launch) X / T may be the previous kernel's output; the weights are constants:
                 // the dequant warps start at once
    int sa = 0, sb = 0;
    uint32_t pa = 0, pb = 0;
    for (int si = 0; si < sched.n_seg; ++si) {
      const Seg sg = sched.get(si);
      const int tile = sg.tile;
      const int m0 = (tile % tiles_m) * 256 + (int)rank * kBM, n0 = (tile / tiles_m) * BN + (int)rank * C::kHalfN;
      for (int i = sg.ks0 >> 2; i < (sg.ks1 + 3) >> 2; ++i) {
        mbar_wait(a_empty + 8 * sa, pa ^ 1);
        if (elect_one()) {
          if (rank == 0) mbar_expect_tx(a_full + 8 * sa, 2 * kAStageBytes);
          tma_load_2d_pair(a_ring + sa * kAStageBytes, &p.tm_a, i * kAK, m0, a_full + 8 * sa);
        }
        if (++sa == kAStagesP) { sa = 0; pa ^= 1; }
        if constexpr (kDenseB) {
          mbar_wait(b_empty + 8 * sb, pb ^ 1);
          if (elect_one()) {
            if (rank == 0) mbar_expect_tx(b_full + 8 * sb, 2 * C::kBStage);
            tma_load_2d_pair(b_ring + sb * C::kBStage, &p.tm_b, i * kAK, n0, b_full + 8 * sb);
          }
          if (++sb == C::kBStages) { sb = 0; pb ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    // ================= MMA issue: the leader's elected lane, for the pair ==========================================
    if (rank == 0) {
      const uint32_t idesc = make_idesc_f16(256, BN, 0, kDenseB ? 0 : 1);
      const uint64_t da_hi = make_smem_desc_sw128(0, 16, 1024) & 0xFFFFFFFF00000000ull;
      const uint64_t db_hi = (kDenseB ? make_smem_desc_sw128(0, 16, 1024) : make_smem_desc_sw128(0, 1024, C::kKAtom)) &
                             0xFFFFFFFF00000000ull;
      const uint32_t da_lo0 = (uint32_t)make_smem_desc_sw128(0, 16, 1024);
      const uint32_t db_lo0 = (uint32_t)(kDenseB ? make_smem_desc_sw128(0, 16, 1024) : make_smem_desc_sw128(0, 1024, C::kKAtom));
      constexpr uint32_t kBStep = kDenseB ? 32u : 2u * C::kKAtom;  // bytes per 16-k step inside a B stage
      if (elect_one()) {  // (the whole loop in one lane, see the single-CTA kernel)
        int sa = 0, sb = 0, it = 0;
        uint32_t pa = 0, pb = 0;
        for (; it < sched.n_seg; ++it) {
          const Seg sg = sched.get(it);
          const int acc = it & 1;
          mbar_wait(t_empty + 8 * acc, ((it >> 1) & 1) ^ 1);
          tc_fence_after();
          const uint32_t d = tmem_base + acc * BN;
          int ka = 0, kb = 0;
          for (int ks = sg.ks0; ks < sg.ks1; ++ks) {
            if (ka == 0) mbar_wait(a_full + 8 * sa, pa);
            if (kb == 0) mbar_wait(b_full + 8 * sb, pb);
            if (ka == 0 || kb == 0) tc_fence_after();
            const uint32_t a_addr = a_ring + sa * kAStageBytes + ka * 32;
            const uint32_t b_addr = b_ring + sb * C::kBStage + kb * kBStep;
            const uint64_t da = da_hi | (uint64_t)(da_lo0 | ((a_addr & 0x3FFFFu) >> 4));
            const uint64_t db = db_hi | (uint64_t)(db_lo0 | ((b_addr & 0x3FFFFu) >> 4));
            const bool last = (ks == sg.ks1 - 1);
            const bool rel_a = (ka == 3) || last, rel_b = (kb == B::kKSteps - 1) || last;
            umma_f16_pair(d, da, db, idesc, ks > sg.ks0 ? 1u : 0u);
            if (rel_a) umma_commit_pair(a_empty + 8 * sa);
            if (rel_b) umma_commit_pair(b_empty + 8 * sb);
            if (last) umma_commit_pair(t_full + 8 * acc);
            if (rel_a) { ka = 0; if (++sa == kAStagesP) { sa = 0; pa ^= 1; } } else { ++ka; }
            if (rel_b) { kb = 0; if (++sb == C::kBStages) { sb = 0; pb ^= 1; } } else { ++kb; }
          }
        }
      }
      __syncwarp();
    }
  } else if (warp < kFirstDqWarp) {
    // ================= epilogue: this CTA's 128 rows of the pair's tile ============================================
    const int quarter = warp & 3;
    const int row_local = quarter * 32 + lane;
    // (a separate instantiation: the table path in the common kernel pushed it over its register budget - a stack
    // frame in this kernel halves its speed)
    uint32_t gelu_lut = 0u;
    if constexpr (kGelu) {  // (before the wait: nothing here depends on the previous kernel)
      gelu_lut = base + C::kLutOff;
      for (int i = row_local; i < kGeluLutEntries; i += kEpWarps * 32) {
        const __half e = __float2half_rn(erff(__half2float(__ushort_as_half((unsigned short)i))));
        asm volatile("st.shared.u16 [%0], %1;" ::"r"(gelu_lut + 2u * i), "h"(__half_as_ushort(e)) : "memory");
      }
      asm volatile("bar.sync 1, %0;" ::"n"(kEpWarps * 32) : "memory");
    }
    pdl_wait();  // the workspace, the epilogue operands and Y belong to the previous kernel until here
    // workspace slot / flags of (pair q, this rank): the slot is [BN / 4 column quads][128 rows] float4, so that a warp
    // (32 consecutive rows) moves 512 contiguous bytes per instruction
    auto slot_of = [&](int q) {
      return reinterpret_cast<uint4*>(p.ws_partial + (size_t)(q * 2 + (int)rank) * kBM * BN) + row_local;
    };
    auto flag_of = [&](int q) { return p.ws_flags + (q * 2 + (int)rank) * 4 + quarter; };
    auto wait_flag = [&](int q) {
      if (lane == 0) {
        unsigned spins = 0;
        while (ld_acquire_u32(flag_of(q)) == 0u) {
          __nanosleep(64);
          if (++spins == (1u << 27)) asm volatile("trap;");  // ~10 s: a writer never waits, this cannot happen
        }
      }
      __syncwarp();
    };
    for (int it = 0; it < sched.n_seg; ++it) {
      const Seg sg = sched.get(it);
      const int tile = sg.tile;
      const int acc = it & 1;
      const int tn = tile / tiles_m, oi = op_of(tn);
      const int m0 = (tile % tiles_m) * 256 + (int)rank * kBM, n0 = (tn - (oi == 0 ? 0 : (oi == 1 ? p.nt_begin[1] : (oi == 2 ? p.nt_begin[2] : p.nt_begin[3])))) * BN;
      const int m = m0 + row_local;
      const uint32_t t_acc = tmem_base + ((uint32_t)(quarter * 32) << 16) + acc * BN;
      if (sg.kind == 1) {
        // partial tile of a pair that does not finish it: f32 accumulators -> this pair's slot, then the flag
        mbar_wait(t_full + 8 * acc, (it >> 1) & 1);
        tc_fence_after();
        uint4* dst = slot_of(pair);
#pragma unroll 1
        for (int c = 0; c < BN; c += 32) {
          uint32_t v[32];
          tmem_ld_32x32(t_acc + c, v);
          tmem_ld_wait();
          if (!(p.debug & 4)) {
#pragma unroll
            for (int j = 0; j < 8; ++j)
              __stcg(dst + (size_t)(c / 4 + j) * kBM, make_uint4(v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]));
          }
        }
        __threadfence();
        __syncwarp();
        if (lane == 0) st_release_u32(flag_of(pair), 1u);
      } else {
        // whole tile, or finisher (kind 2): the pairs behind this one hold the rest of the tile's k range.  Their
        // partials are usually there long before this pair's own MMAs are done: the flags are awaited and the first
        // columns fetched meanwhile, and the partial of the next 32 columns is in flight while the current ones
        // are finished.
        const bool fin = sg.kind == 2;
        const int q_end = fin ? sched.contributors_end(tile) : pair + 1;
        const uint4* src1 = slot_of(pair + 1);
        uint4 nx[8];
        if (fin) {
          wait_flag(pair + 1);
#pragma unroll
          for (int j = 0; j < 8; ++j) nx[j] = __ldcg(src1 + (size_t)j * kBM);
          for (int q = pair + 2; q < q_end; ++q) wait_flag(q);
        }
        mbar_wait(t_full + 8 * acc, (it >> 1) & 1);
        tc_fence_after();
#pragma unroll 1
        for (int c = 0; c < BN; c += 32) {
          uint32_t v[32];
          tmem_ld_32x32(t_acc + c, v);
          tmem_ld_wait();
          if (fin) {
#pragma unroll
            for (int j = 0; j < 8; ++j) {
              v[4 * j + 0] = __float_as_uint(__uint_as_float(v[4 * j + 0]) + __uint_as_float(nx[j].x));
              v[4 * j + 1] = __float_as_uint(__uint_as_float(v[4 * j + 1]) + __uint_as_float(nx[j].y));
              v[4 * j + 2] = __float_as_uint(__uint_as_float(v[4 * j + 2]) + __uint_as_float(nx[j].z));
              v[4 * j + 3] = __float_as_uint(__uint_as_float(v[4 * j + 3]) + __uint_as_float(nx[j].w));
            }
            if (c + 32 < BN && !(p.debug & 8)) {
#pragma unroll
              for (int j = 0; j < 8; ++j) nx[j] = __ldcg(src1 + (size_t)((c + 32) / 4 + j) * kBM);
            }
            for (int q = pair + 2; q < q_end; ++q) {  // (a tile cut over more than two pairs) added in pair order
              const uint4* src = slot_of(q);
#pragma unroll
              for (int j = 0; j < 8; ++j) {
                const uint4 pv = __ldcg(src + (size_t)(c / 4 + j) * kBM);
                v[4 * j + 0] = __float_as_uint(__uint_as_float(v[4 * j + 0]) + __uint_as_float(pv.x));
                v[4 * j + 1] = __float_as_uint(__uint_as_float(v[4 * j + 1]) + __uint_as_float(pv.y));
                v[4 * j + 2] = __float_as_uint(__uint_as_float(v[4 * j + 2]) + __uint_as_float(pv.z));
                v[4 * j + 3] = __float_as_uint(__uint_as_float(v[4 * j + 3]) + __uint_as_float(pv.w));
              }
            }
          }
          if (m < p.M) {  // (weights indexed with constants only, see pick_op)
            if constexpr (kGelu) {  // the LAST weight of the launch (of one, or of four) has the bias + GELU epilogue
              if (p.n_ops == 1) { if (n0 + c < p.op[0].N) epilogue_row(p.op[0], v, m, n0 + c, gelu_lut); }
              else if (oi == 3) { if (n0 + c < p.op[3].N) epilogue_row(p.op[3], v, m, n0 + c, gelu_lut); }
              else if (oi == 0) { if (n0 + c < p.op[0].N) epilogue_row(p.op[0], v, m, n0 + c); }
              else if (oi == 1) { if (n0 + c < p.op[1].N) epilogue_row(p.op[1], v, m, n0 + c); }
              else { if (n0 + c < p.op[2].N) epilogue_row(p.op[2], v, m, n0 + c); }
            } else {
              if (oi == 0) { if (n0 + c < p.op[0].N) epilogue_row(p.op[0], v, m, n0 + c); }
              else if (oi == 1) { if (n0 + c < p.op[1].N) epilogue_row(p.op[1], v, m, n0 + c); }
              else { if (n0 + c < p.op[2].N) epilogue_row(p.op[2], v, m, n0 + c); }
            }
          }
        }
        if (fin) {  // the slots have been read: lower the flags for the next launch
          __syncwarp();
          if (lane == 0)
            for (int q = pair + 1; q < q_end; ++q) *flag_of(q) = 0u;
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) {
        if (rank == 0) mbar_arrive(t_empty + 8 * acc);
        else mbar_arrive_cluster(t_empty + 8 * acc, 0);
      }
    }
  } else {
    // ================= dequantizers: this CTA's BN/2 columns of the weight tile ====================================
    if constexpr (!kDenseB) {
      using F = Fmt<MODE>;
      static_assert(C::kBStages % C::kDqGroups == 0, "a warp group owns a fixed subset of the ring slots");
      const int dwa = warp - kFirstDqWarp;
      const int grp = dwa / C::kDq, dw = dwa % C::kDq;
      const int sub = lane / C::kChunks, chunk = lane % C::kChunks;  // word row of the pass, 8-column chunk
      uint32_t w[C::kItems][8], s[C::kItems][4], mn[C::kItems][4];
      auto row_of = [&](int t) { return dw * C::kRowsPerPass + sub + 8 * t; };  // stage-local word row
      auto load_stage_op = [&](const GemmOp& o, int n0, int i) {
#pragma unroll
        for (int t = 0; t < C::kItems; ++t) {
          const int r = i * C::kRows + row_of(t);
          const int n = n0 + chunk * 8;
          if (r < p.n_rows && n < o.N) {
            const uint4* src = reinterpret_cast<const uint4*>(o.packed + (size_t)r * o.N + n);
            const uint4 lo = __ldg(src), hi = __ldg(src + 1);
            w[t][0] = lo.x; w[t][1] = lo.y; w[t][2] = lo.z; w[t][3] = lo.w;
            w[t][4] = hi.x; w[t][5] = hi.y; w[t][6] = hi.z; w[t][7] = hi.w;
            const uint4 sv = __ldg(reinterpret_cast<const uint4*>(o.scale + (size_t)(r / F::kRowsPerGroup) * o.N + n));
            s[t][0] = sv.x; s[t][1] = sv.y; s[t][2] = sv.z; s[t][3] = sv.w;
            if constexpr (fmt_has_min_c(MODE)) {
              const uint4 mv = __ldg(reinterpret_cast<const uint4*>(o.zmin + (size_t)(r / F::kRowsPerGroup) * o.N + n));
              mn[t][0] = mv.x; mn[t][1] = mv.y; mn[t][2] = mv.z; mn[t][3] = mv.w;
            }
          } else {
#pragma unroll
            for (int j = 0; j < 8; ++j) w[t][j] = 0u;
#pragma unroll
            for (int j = 0; j < 4; ++j) s[t][j] = 0u;
#pragma unroll
            for (int j = 0; j < 4; ++j) mn[t][j] = 0u;
          }
        }
      };
      auto load_stage = [&](int tile, int i) {
        const int tn = tile / tiles_m, oi = op_of(tn);
        const int half0 = (int)rank * C::kHalfN;
        if (oi == 0) load_stage_op(p.op[0], tn * BN + half0, i);
        else if (oi == 1) load_stage_op(p.op[1], (tn - p.nt_begin[1]) * BN + half0, i);
        else if (oi == 2) load_stage_op(p.op[2], (tn - p.nt_begin[2]) * BN + half0, i);
        else load_stage_op(p.op[3], (tn - p.nt_begin[3]) * BN + half0, i);
      };
      // c = running number of B stages of this CTA over all its segments; this group takes those with c % groups == grp
      long long c0 = 0;  // stages before the current segment
      int si = 0, seg_tile = 0, bs0 = 0, nb = 0;  // current segment: tile, first weight stage of its k range, stages
      auto open = [&]() {
        if (si < sched.n_seg) {
          const Seg sg = sched.get(si);
          seg_tile = sg.tile;
          bs0 = sg.ks0 / B::kKSteps;
          nb = (sg.ks1 + B::kKSteps - 1) / B::kKSteps - bs0;
        }
      };
      open();
      int i = grp;  // first own stage, counted from the segment's first
      auto settle = [&]() {  // move (segment, i) on while i runs past the segment's stages
        while (si < sched.n_seg && i >= nb) {
          i -= nb;
          c0 += nb;
          ++si;
          open();
        }
      };
      settle();
      if (si < sched.n_seg) load_stage(seg_tile, bs0 + i);
      while (si < sched.n_seg) {
        const long long c = c0 + i;
        const int sb = (int)(c % C::kBStages);
        const uint32_t pb = (uint32_t)((c / C::kBStages) & 1);
        uint32_t wc[C::kItems][8], sc[C::kItems][4], mc[C::kItems][4];
#pragma unroll
        for (int t = 0; t < C::kItems; ++t) {
#pragma unroll
          for (int j = 0; j < 8; ++j) wc[t][j] = w[t][j];
#pragma unroll
          for (int j = 0; j < 4; ++j) sc[t][j] = s[t][j];
#pragma unroll
          for (int j = 0; j < 4; ++j) mc[t][j] = fmt_has_min_c(MODE) ? mn[t][j] : 0u;
        }
        // this group's next stage (possibly in the next segment): its loads are in flight during this stage's unpack
        i += C::kDqGroups;
        settle();
        if (si < sched.n_seg) load_stage(seg_tile, bs0 + i);
        mbar_wait(b_empty + 8 * sb, pb ^ 1);
        const uint32_t stage = b_ring + sb * C::kBStage;
#pragma unroll
        for (int t = 0; t < C::kItems; ++t) dequant_chunk<MODE, C::kKAtom>(wc[t], sc[t], mc[t], stage, row_of(t) * F::kNpu, chunk);
        fence_proxy_async_smem();
        __syncwarp();
        if (lane == 0) {
          if (rank == 0) mbar_arrive(b_full + 8 * sb);
          else mbar_arrive_cluster(b_full + 8 * sb, 0);
        }
      }
    }
  }

  // ---- teardown: nobody leaves while the other CTA may still signal into this one ------------------------------
  tc_fence_before();
  cluster_sync_all();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc2(tmem_base, C::kTmemCols);
  }
}

}  // namespace

// gemv.cu
int make_tmap_2d(CUtensorMap* tm, CUtensorMapDataType dt, int elem_bytes, const void* ptr, int inner, int outer,
                 int box_inner, int box_outer, CUtensorMapSwizzle sw);
int device_sm_count();

namespace {

// environment switches, read once: B200Q_GEMM_PAIR = 0 (single-CTA kernel only) | 128 | 256 (force the pair tile N)
int gemm_pair_mode() {
  static const int mode = [] {
    const char* e = getenv("B200Q_GEMM_PAIR");
    return e ? atoi(e) : -1;
  }();
  return mode;
}

// B200Q_GEMM_STREAMK = 0 (never) | 1 (whenever a workspace is given); default: when whole tiles fill the pairs badly
int gemm_streamk_mode() {
  static const int mode = [] {
    const char* e = getenv("B200Q_GEMM_STREAMK");
    return e ? atoi(e) : -1;
  }();
  return mode;
}

template <int MODE, int BN>
int launch_gemm_pair(GemmParams& p, int n_pairs, void* workspace, size_t workspace_bytes, cudaStream_t stream) {
  using C = Pair<MODE, BN>;
  if (p.n_ops < 1) {  // a single weight: op[0] mirrors the plain fields
    p.n_ops = 1;
    p.op[0] = GemmOp{p.packed, p.scale, p.zmin, p.ep0, p.ep1, p.ep2, p.Y, p.N, p.epilogue};
  }
  int col_tiles = 0;
  for (int i = 0; i <= kMaxGemmOps; ++i) p.nt_begin[i] = 0x7FFFFFFF;
  for (int i = 0; i < p.n_ops; ++i) {
    p.nt_begin[i] = col_tiles;
    col_tiles += ceil_div(p.op[i].N, BN);
  }
  p.nt_begin[p.n_ops] = col_tiles;
  const int n_tiles = ceil_div(p.M, 256) * col_tiles;
  // stream-K when the caller gave a workspace and whole tiles would leave pairs idle in the last round
  p.streamk = 0;
  {
    const int n_ksteps = ceil_div(p.K, 16);
    const long long W = (long long)n_tiles * ceil_div(n_ksteps, C::kChunk);
    // Measured on B200 (q4f16_0, M = 2048, scripts/gemm_sk_sweep.py): whole tiles take 12 us + 21.5 us x (K / 2560) per
    // round of tiles; stream-K takes ~10 us more than the same with the FRACTIONAL number of rounds x = tiles / pairs
    // (two accumulator hand-offs more per pair, the partial tiles' trip through L2, the finisher's longer epilogue),
    // and much more when a tile is cut over more than two or three pairs (x < 0.45).  It is taken when the idle part
    // of the last round is worth more than that.
    const double x = (double)n_tiles / n_pairs;
    const double idle_us = ((double)ceil_div(n_tiles, n_pairs) - x) * 21.5 * ((double)p.K / 2560.0);
    const int pairs_sk = (int)(W < n_pairs ? W : n_pairs);
    const size_t need = (size_t)kWsFlagBytes + (size_t)pairs_sk * 2 * kBM * BN * sizeof(float);
    const int mode = gemm_streamk_mode();
    const bool want = (mode == 1) || (p.flags & B200Q_FLAG_GEMM_STREAMK) || (mode != 0 && BN == 256 && x >= 0.45 && idle_us > 12.0);
    if (want && !(p.flags & B200Q_FLAG_GEMM_NO_STREAMK) && workspace && workspace_bytes >= need && pairs_sk * 2 * 4 * 4 <= kWsFlagBytes) {
      p.streamk = 1;
      p.ws_flags = reinterpret_cast<unsigned*>(workspace);
      p.ws_partial = reinterpret_cast<float*>(reinterpret_cast<unsigned char*>(workspace) + kWsFlagBytes);
      n_pairs = pairs_sk;
    }
  }
  if (!p.streamk && n_pairs > n_tiles) n_pairs = n_tiles;
  {
    static const int dbg = getenv("B200Q_GEMM_DEBUG") ? atoi(getenv("B200Q_GEMM_DEBUG")) : 0;
    p.debug = dbg;  // experiments: 4 = partial tiles are not stored, 8 = not loaded (results wrong, protocol intact)
  }
  // bias + GELU on a single weight with 256-wide tiles: the instantiation that tabulates erf in shared memory
  const bool gelu = (BN == 256) && (p.n_ops == 1 || p.n_ops == 4) && p.op[p.n_ops - 1].epilogue == B200Q_EP_BIAS_GELU;
  if (p.n_ops == 4 && !gelu) return set_error(B200Q_ERR_UNSUPPORTED, "gemm pair: four weights only as q | k | v | (bias + GELU) on 256-wide tiles");
  p.gelu_lut = gelu ? 1 : 0;
  using CG = Pair<MODE, BN, true>;
  const size_t smem = gelu ? (size_t)CG::kLutOff + kGeluLutBytes + 1024 : (size_t)C::kBytes + 1024;
  static_assert(CG::kLutOff + kGeluLutBytes + 1024 <= 232448, "pair kernel + GELU table exceed 227 KB of shared memory");
  static_assert(C::kBytes + 1024 <= 232448, "pair kernel exceeds 227 KB of shared memory");
  using KernT = void (*)(GemmParams);
  KernT kern = gemm_tc2_kernel<MODE, BN, false>;
  if constexpr (BN == 256) {
    if (gelu) kern = gemm_tc2_kernel<MODE, BN, true>;
  }
  static std::atomic<bool> attr_set[64][2];
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) dev = -1;
  if (dev < 0 || !attr_set[dev][gelu].load(std::memory_order_relaxed)) {
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return check_cuda(e, "cudaFuncSetAttribute(gemm pair smem)");
    if (dev >= 0) attr_set[dev][gelu].store(true, std::memory_order_relaxed);
  }
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3(2 * n_pairs, 1, 1);
  cfg.blockDim = dim3(C::kThreads, 1, 1);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = stream;
  cudaLaunchAttribute attrs[2];
  attrs[0].id = cudaLaunchAttributeClusterDimension;  // the two CTAs of a cluster run on the two SMs of one TPC
  attrs[0].val.clusterDim.x = 2;
  attrs[0].val.clusterDim.y = 1;
  attrs[0].val.clusterDim.z = 1;
  int na = 1;
  if (p.flags & B200Q_FLAG_PDL) {
    // the next launch's CTAs take an SM as soon as this launch's pair there is done: set-up and the first weight
    // stages (constants) overlap this launch's tail; X, the epilogue operands and the workspace wait (griddepcontrol.wait)
    attrs[na].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attrs[na].val.programmaticStreamSerializationAllowed = 1;
    ++na;
  }
  cfg.attrs = attrs;
  cfg.numAttrs = na;
  cudaError_t e = cudaLaunchKernelEx(&cfg, kern, p);
  if (e != cudaSuccess) return check_cuda(e, "cudaLaunchKernelEx(gemm pair)");
  count_launch();
  return B200Q_OK;
}

// tile choice: the CTA-pair kernel with 256 x 256 tiles whenever there is more than one 128-row block (measured on
// B200, M = 2048: 1.05-1.16x the single-CTA kernel on every Llama / NeoX shape); 256 x 128 pair tiles have twice the
// X traffic per FLOP and lose everywhere (kept behind B200Q_FLAG_GEMM_PAIR128 for experiments).  Returns 0: use the
// single-CTA kernel.
int pick_pair_bn(int M, int N, int n_pairs, unsigned flags = 0) {
  int mode = gemm_pair_mode();
  if (flags & B200Q_FLAG_GEMM_SINGLE) mode = 0;
  if (flags & B200Q_FLAG_GEMM_PAIR128) mode = 128;
  if (flags & B200Q_FLAG_GEMM_PAIR256) mode = 256;
  (void)N;
  if (mode == 0 || n_pairs < 1) return 0;
  if (mode == 128 || mode == 256) return mode;
  return M > kBM ? 256 : 0;
}

template <int MODE>
int launch_gemm(GemmParams& p, cudaStream_t stream, void* workspace = nullptr, size_t workspace_bytes = 0) {
  using S = Smem<MODE>;
  {
    const int n_pairs = device_sm_count() / 2;
    const int bn = pick_pair_bn(p.M, p.N, n_pairs, p.flags);
    if (bn == 256) return launch_gemm_pair<MODE, 256>(p, n_pairs, workspace, workspace_bytes, stream);
    if (bn == 128) return launch_gemm_pair<MODE, 128>(p, n_pairs, workspace, workspace_bytes, stream);
  }
  p.tiles_m = ceil_div(p.M, kBM);
  p.tiles_n = ceil_div(p.N, kBN);
  const int n_tiles = p.tiles_m * p.tiles_n;
  {
    static const int dbg = getenv("B200Q_GEMM_DEBUG") ? atoi(getenv("B200Q_GEMM_DEBUG")) : 0;
    p.debug = dbg;
  }
  int grid = device_sm_count();
  if (grid > n_tiles) grid = n_tiles;
  const size_t smem = (size_t)S::kBytes + 1024;  // + alignment slack
  auto kern = gemm_tc_kernel<MODE>;
  static std::atomic<bool> attr_set[64];  // per instantiation and device (the opt-in is a per-device setting)
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) dev = -1;
  if (dev < 0 || !attr_set[dev].load(std::memory_order_relaxed)) {
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return check_cuda(e, "cudaFuncSetAttribute(gemm smem)");
    if (dev >= 0) attr_set[dev].store(true, std::memory_order_relaxed);
  }
  kern<<<grid, kGemmThreads, smem, stream>>>(p);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return check_cuda(e, "gemm launch");
  count_launch();
  return B200Q_OK;
}

int check_epilogue_operands(int epilogue, const void* ep0, const void* ep1, const void* ep2, const char* who) {
  const bool needs0 = epilogue == B200Q_EP_ADD || epilogue == B200Q_EP_MULTIPLY || epilogue == B200Q_EP_BIAS ||
                      epilogue == B200Q_EP_BIAS_GELU || epilogue == B200Q_EP_BIAS_ADD2;
  if (needs0 && !ep0) return set_error(B200Q_ERR_NULL, "%s: epilogue %d needs ep0", who, epilogue);
  if (epilogue == B200Q_EP_BIAS_ADD2 && (!ep1 || !ep2)) return set_error(B200Q_ERR_NULL, "%s: bias_add2 needs ep1 and ep2", who);
  auto mis = [](const void* q) { return q && (reinterpret_cast<uintptr_t>(q) & 15) != 0; };
  if (mis(ep0) || mis(ep1) || mis(ep2)) return set_error(B200Q_ERR_BAD_ALIGN, "%s: epilogue operands must be 16-byte aligned", who);
  return B200Q_OK;
}

}  // namespace

// The pair kernel's work distribution, evaluated on the host (tests/test_cabi_cpu.py checks its invariants without a
// GPU): segments of `pair` in PROCESSING order as (tile, ks0, ks1, kind), plus the finisher's contributor range end.
int gemm_schedule_debug(int streamk, int n_tiles, int n_ksteps, int chunk, int n_pairs, int pair, int* out, int max_seg) {
  TileSched sc;
  sc.init(streamk, pair, n_pairs, n_tiles, n_ksteps, chunk);
  const int n = sc.n_seg < max_seg ? sc.n_seg : max_seg;
  for (int i = 0; i < n; ++i) {
    const Seg sg = sc.get(i);
    out[5 * i + 0] = sg.tile; out[5 * i + 1] = sg.ks0; out[5 * i + 2] = sg.ks1; out[5 * i + 3] = sg.kind;
    out[5 * i + 4] = (sg.kind == 2) ? sc.contributors_end(sg.tile) : pair + 1;
  }
  return sc.n_seg;
}

size_t gemm_workspace_bytes() {  // flags + one 256 x 256 f32 partial tile per CTA pair
  return (size_t)kWsFlagBytes + (size_t)(device_sm_count() / 2) * 2 * kBM * 256 * sizeof(float);
}

int gemm_launch(const void* packed, const void* scale, const void* X, const void* ep0, const void* ep1,
                const void* ep2, void* Y, int M, int K, int N, int fmt, int epilogue, unsigned flags,
                cudaStream_t stream, void* workspace, size_t workspace_bytes) {
  if (K % 8 != 0) return set_error(B200Q_ERR_BAD_SHAPE, "b200q_gemm: K=%d must be a multiple of 8 (16-byte rows of X)", K);
  if (int e = check_epilogue_operands(epilogue, ep0, ep1, ep2, "b200q_gemm")) return e;
  GemmParams p{};
  p.packed = reinterpret_cast<const uint32_t*>(packed);
  p.scale = reinterpret_cast<const __half*>(scale);
  p.ep0 = reinterpret_cast<const __half*>(ep0);
  p.ep1 = reinterpret_cast<const __half*>(ep1);
  p.ep2 = reinterpret_cast<const __half*>(ep2);
  p.Y = Y;
  p.M = M; p.K = K; p.N = N;
  p.n_rows = n_words(fmt, K);
  p.epilogue = epilogue;
  p.flags = flags;
  if (int e = make_tmap_2d(&p.tm_a, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, X, K, M, kAK, kBM, CU_TENSOR_MAP_SWIZZLE_128B)) return e;
  switch (fmt) {
    case B200Q_Q4F16_0: return launch_gemm<B200Q_Q4F16_0>(p, stream, workspace, workspace_bytes);
    case B200Q_Q3F16_0: return launch_gemm<B200Q_Q3F16_0>(p, stream, workspace, workspace_bytes);
    case B200Q_Q8F16_0: return launch_gemm<B200Q_Q8F16_0>(p, stream, workspace, workspace_bytes);
    default: return set_error(B200Q_ERR_UNSUPPORTED, "b200q_gemm: fmt %d", fmt);
  }
}

int gemm_asym_launch(const void* packed, const void* scale, const void* zmin, const void* X, const void* ep0, const void* ep1,
                     const void* ep2, void* Y, int M, int K, int N, int epilogue, unsigned flags, cudaStream_t stream,
                     void* workspace, size_t workspace_bytes) {
  if (int e = check_epilogue_operands(epilogue, ep0, ep1, ep2, "b200q_gemm_asym")) return e;
  GemmParams p{};
  p.packed = reinterpret_cast<const uint32_t*>(packed);
  p.scale = reinterpret_cast<const __half*>(scale);
  p.zmin = reinterpret_cast<const __half*>(zmin);
  p.ep0 = reinterpret_cast<const __half*>(ep0);
  p.ep1 = reinterpret_cast<const __half*>(ep1);
  p.ep2 = reinterpret_cast<const __half*>(ep2);
  p.Y = Y;
  p.M = M; p.K = K; p.N = N;
  p.n_rows = n_words(B200Q_Q4F16_ASYM, K);
  p.epilogue = epilogue;
  p.flags = flags;
  if (int e = make_tmap_2d(&p.tm_a, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, X, K, M, kAK, kBM, CU_TENSOR_MAP_SWIZZLE_128B)) return e;
  return launch_gemm<B200Q_Q4F16_ASYM>(p, stream, workspace, workspace_bytes);
}

// Several weights that read the SAME X in one launch of the pair kernel: their column tiles form one tile space
// (and one stream-K range split), so three 80-tile GEMMs on 74 pairs cost 3.24 rounds instead of three times two.
int gemm_multi_launch(const b200q_gemm_op* ops, int n_ops, const void* X, int M, int K, int fmt, unsigned flags,
                      cudaStream_t stream, void* workspace, size_t workspace_bytes) {
  const bool asym = fmt == B200Q_Q4F16_ASYM;
  const int n_pairs = device_sm_count() / 2;
  const int bn = pick_pair_bn(M, ops[0].N, n_pairs, flags);
  if (n_ops == 4 && !(bn == 256 && ops[3].epilogue == B200Q_EP_BIAS_GELU)) {
    // four weights in one launch only as q | k | v | (bias + GELU) on 256-wide tiles (the GELU-table instantiation)
    if (int e = gemm_multi_launch(ops, 3, X, M, K, fmt, flags, stream, workspace, workspace_bytes)) return e;
    return gemm_multi_launch(ops + 3, 1, X, M, K, fmt, flags, stream, workspace, workspace_bytes);
  }
  if (n_ops == 1 || bn == 0) {  // the single-CTA kernel takes one weight per launch
    for (int i = 0; i < n_ops; ++i) {
      const b200q_gemm_op& o = ops[i];
      const int e = asym ? gemm_asym_launch(o.packed, o.scale, o.zmin, X, o.ep0, o.ep1, o.ep2, o.Y, M, K, o.N, o.epilogue, flags, stream,
                                            workspace, workspace_bytes)
                         : gemm_launch(o.packed, o.scale, X, o.ep0, o.ep1, o.ep2, o.Y, M, K, o.N, fmt, o.epilogue, flags, stream, workspace,
                                       workspace_bytes);
      if (e) return e;
    }
    return B200Q_OK;
  }
  if (K % 8 != 0) return set_error(B200Q_ERR_BAD_SHAPE, "b200q_gemm_multi: K=%d must be a multiple of 8 (16-byte rows of X)", K);
  GemmParams p{};
  p.M = M; p.K = K;
  p.n_rows = n_words(fmt, K);
  p.flags = flags;
  p.n_ops = n_ops;
  for (int i = 0; i < n_ops; ++i) {
    const b200q_gemm_op& o = ops[i];
    if (int e = check_epilogue_operands(o.epilogue, o.ep0, o.ep1, o.ep2, "b200q_gemm_multi")) return e;
    p.op[i] = GemmOp{reinterpret_cast<const uint32_t*>(o.packed), reinterpret_cast<const __half*>(o.scale),
                     reinterpret_cast<const __half*>(o.zmin), reinterpret_cast<const __half*>(o.ep0),
                     reinterpret_cast<const __half*>(o.ep1), reinterpret_cast<const __half*>(o.ep2), o.Y, o.N, o.epilogue};
  }
  p.N = ops[0].N;
  if (int e = make_tmap_2d(&p.tm_a, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, X, K, M, kAK, kBM, CU_TENSOR_MAP_SWIZZLE_128B)) return e;
  switch (fmt) {
    case B200Q_Q4F16_0: return bn == 256 ? launch_gemm_pair<B200Q_Q4F16_0, 256>(p, n_pairs, workspace, workspace_bytes, stream)
                                         : launch_gemm_pair<B200Q_Q4F16_0, 128>(p, n_pairs, workspace, workspace_bytes, stream);
    case B200Q_Q3F16_0: return bn == 256 ? launch_gemm_pair<B200Q_Q3F16_0, 256>(p, n_pairs, workspace, workspace_bytes, stream)
                                         : launch_gemm_pair<B200Q_Q3F16_0, 128>(p, n_pairs, workspace, workspace_bytes, stream);
    case B200Q_Q8F16_0: return bn == 256 ? launch_gemm_pair<B200Q_Q8F16_0, 256>(p, n_pairs, workspace, workspace_bytes, stream)
                                         : launch_gemm_pair<B200Q_Q8F16_0, 128>(p, n_pairs, workspace, workspace_bytes, stream);
    case B200Q_Q4F16_ASYM: return bn == 256 ? launch_gemm_pair<B200Q_Q4F16_ASYM, 256>(p, n_pairs, workspace, workspace_bytes, stream)
                                            : launch_gemm_pair<B200Q_Q4F16_ASYM, 128>(p, n_pairs, workspace, workspace_bytes, stream);
    default: return set_error(B200Q_ERR_UNSUPPORTED, "b200q_gemm_multi: fmt %d", fmt);
  }
}

int nt_matmul_launch(const void* X, const void* T, const void* ep0, const void* ep1, const void* ep2, void* Y, int M,
                     int K, int N, int epilogue, cudaStream_t stream, unsigned flags, void* workspace,
                     size_t workspace_bytes) {
  if (epilogue < 0 || epilogue >= B200Q_NUM_EP) return set_error(B200Q_ERR_UNSUPPORTED, "b200q_nt_matmul: epilogue %d", epilogue);
  if (int e = check_epilogue_operands(epilogue, ep0, ep1, ep2, "b200q_nt_matmul")) return e;
  GemmParams p{};
  p.ep0 = reinterpret_cast<const __half*>(ep0);
  p.ep1 = reinterpret_cast<const __half*>(ep1);
  p.ep2 = reinterpret_cast<const __half*>(ep2);
  p.Y = Y;
  p.M = M; p.K = K; p.N = N;
  p.epilogue = epilogue;
  p.flags = flags & (B200Q_FLAG_PDL | B200Q_FLAG_GEMM_STREAMK | B200Q_FLAG_GEMM_NO_STREAMK);
  if (int e = make_tmap_2d(&p.tm_a, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, X, K, M, kAK, kBM, CU_TENSOR_MAP_SWIZZLE_128B)) return e;
  const int bn = pick_pair_bn(M, N, device_sm_count() / 2);  // the pair kernel loads BN / 2 rows of T per CTA
  if (int e = make_tmap_2d(&p.tm_b, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, T, K, N, kAK, bn ? bn / 2 : kBN, CU_TENSOR_MAP_SWIZZLE_128B)) return e;
  return launch_gemm<kDense>(p, stream, workspace, workspace_bytes);
}

}  // namespace b200q
