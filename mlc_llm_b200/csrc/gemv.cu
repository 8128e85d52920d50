// Decode-step fused dequantize + GEMV for the f16 group-quantized formats (q4f16_0, q3f16_0,
// q8f16_0), sm_100a.
//
// Replaces the reference's fused_decode*_matmul* PrimFuncs
// (mlc_llm/dispatch/dispatch_tir_operator_adreno.py:5-411, mlc_llm/dispatch/llama/main.py:4235-6363;
// schedule generator mlc_llm/tuning.py:39-108: one output column per thread, sequential K,
// N/256 blocks).  This design instead:
//
//   * streams the packed words with 128-bit L1-bypassing loads, 4 consecutive columns per lane,
//     one full 128-byte line per (word row, 32-column warp tile); a lane keeps two "units"
//     (4 word rows x 4 columns + their 4 scales = 72 B) in flight, i.e. ~37 KB per 256-thread CTA;
//   * unpacks in registers to the bit-exact f16 weight  half((q - maxint) * s)  (common.cuh);
//   * contracts over K with mma.sync.m16n8k16 (f16 x f16 products exact, f32 accumulate): the
//     packed layout puts 8 consecutive k of ONE column into a lane's register, so weights are the
//     A operand (16 columns x 16 k per instruction) and the activation vector is replicated over
//     the 8 B columns.  A SIMT formulation needs >= 3.5 issue slots per weight (2 cvt + 2 fma per
//     half2 on top of the unpack) against a budget of 3.1 at HBM speed; this one needs ~1.9;
//   * splits K across the warps of a CTA (shared-memory reduce) and, for small N, across the CTAs
//     of a thread-block cluster (distributed-shared-memory reduce) - deterministic, no atomics,
//     no workspace;
//   * with B200Q_FLAG_PDL, issues the first weight loads BEFORE griddepcontrol.wait: weights are
//     constants, so their fetch overlaps the tail of the previous kernel in the stream.
//
// Layout of the activation vector in shared memory: for word row r (npu consecutive k) the lane
// needs x in the pair order the unpack produces; rows are stored [i = r % 4][unit = r / 4] so the
// four lanes of a quad (which own consecutive units) read consecutive addresses.
#include <cooperative_groups.h>

#include "common.cuh"

namespace cg = cooperative_groups;

namespace b200q {

constexpr int kThreads = 256;
constexpr int kWarps = kThreads / 32;
constexpr int kMaxOps = 4;

struct GemvOp {
  const uint32_t* packed;
  const __half* scale;
  const __half* ep0;
  void* out;
  int N;
  int epilogue;
  int tile_begin;  // first column tile (of 32 * WN columns) of this op in the grid
};

struct GemvParams {
  GemvOp op[kMaxOps];
  int n_ops;
  const __half* x;
  int K;
  int n_rows;   // packed word rows = n_words(K)
  int n_units;  // ceil(n_rows / 4)
  int units_per_cta;  // multiple of 4
};

__device__ __forceinline__ void mma16816(float (&d)[4], uint32_t a0, uint32_t a1, uint32_t a2,
                                         uint32_t a3, uint32_t b0, uint32_t b1) {
  asm volatile(
      "mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, "
      "{%0,%1,%2,%3};"
      : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
      : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}

// bytes of permuted x per word row
template <int FMT>
__host__ __device__ constexpr int x_row_bytes() {
  return Fmt<FMT>::kQuads * 8;
}

// One unit = 4 word rows x 4 columns of packed words + the 4 column scales.
struct Unit {
  uint4 w[4];
  uint2 s;
};

template <int FMT>
__device__ __forceinline__ void load_unit(Unit& u, const uint32_t* __restrict__ packed,
                                          const __half* __restrict__ scale, int unit, int n_rows,
                                          int col, int N, bool valid, uint64_t pol) {
  constexpr int kUnitsPerGroup = Fmt<FMT>::kRowsPerGroup / 4;
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    int row = unit * 4 + i;
    if (valid && row < n_rows) {
      u.w[i] = ldg_stream_v4(packed + (size_t)row * N + col, pol);
    } else {
      u.w[i] = make_uint4(0, 0, 0, 0);
    }
  }
  if (valid) {
    u.s = ldg_stream_v2(scale + (size_t)(unit / kUnitsPerGroup) * N + col, pol);
  } else {
    u.s = make_uint2(0, 0);
  }
}

// Accumulate one unit. acc[m][*]: MMA m covers columns (4c + 2m) [rows 0-7] and (4c + 2m + 1)
// [rows 8-15] of the lane's column quad c.
template <int FMT>
__device__ __forceinline__ void compute_unit(const Unit& u, const unsigned char* xs_unit,
                                             int xs_i_stride, float (&acc)[2][4]) {
  using F = Fmt<FMT>;
  __half2 s2[4];
  {
    uint32_t s01 = u.s.x, s23 = u.s.y;
    s2[0] = u32_as_h2(__byte_perm(s01, 0, 0x1010));
    s2[1] = u32_as_h2(__byte_perm(s01, 0, 0x3232));
    s2[2] = u32_as_h2(__byte_perm(s23, 0, 0x1010));
    s2[3] = u32_as_h2(__byte_perm(s23, 0, 0x3232));
  }
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const uint32_t wv[4] = {u.w[i].x, u.w[i].y, u.w[i].z, u.w[i].w};
    const unsigned char* xr = xs_unit + i * xs_i_stride;
    if constexpr (FMT == B200Q_Q4F16_0) {
      uint4 xv = *reinterpret_cast<const uint4*>(xr);  // (x0,x4)(x1,x5)(x2,x6)(x3,x7)
      __half2 h[4][4];
#pragma unroll
      for (int j = 0; j < 4; ++j) unpack_q4(wv[j], s2[j], h[j]);
#pragma unroll
      for (int m = 0; m < 2; ++m) {
        mma16816(acc[m], h2_as_u32(h[2 * m][0]), h2_as_u32(h[2 * m + 1][0]), h2_as_u32(h[2 * m][1]),
                 h2_as_u32(h[2 * m + 1][1]), xv.x, xv.y);
        mma16816(acc[m], h2_as_u32(h[2 * m][2]), h2_as_u32(h[2 * m + 1][2]), h2_as_u32(h[2 * m][3]),
                 h2_as_u32(h[2 * m + 1][3]), xv.z, xv.w);
      }
    } else if constexpr (FMT == B200Q_Q3F16_0) {
      // (x0,x5)(x1,x6) | (x2,x7)(x3,x8) | (x4,x9)(0,0)
      // rows are 24 bytes apart: only 8-byte alignment is guaranteed
      uint2 x01 = *reinterpret_cast<const uint2*>(xr);
      uint2 x23 = *reinterpret_cast<const uint2*>(xr + 8);
      uint2 xb = *reinterpret_cast<const uint2*>(xr + 16);
      uint4 xa = make_uint4(x01.x, x01.y, x23.x, x23.y);
      __half2 h[4][5];
#pragma unroll
      for (int j = 0; j < 4; ++j) unpack_q3(wv[j], s2[j], h[j]);
#pragma unroll
      for (int m = 0; m < 2; ++m) {
        mma16816(acc[m], h2_as_u32(h[2 * m][0]), h2_as_u32(h[2 * m + 1][0]), h2_as_u32(h[2 * m][1]),
                 h2_as_u32(h[2 * m + 1][1]), xa.x, xa.y);
        mma16816(acc[m], h2_as_u32(h[2 * m][2]), h2_as_u32(h[2 * m + 1][2]), h2_as_u32(h[2 * m][3]),
                 h2_as_u32(h[2 * m + 1][3]), xa.z, xa.w);
        mma16816(acc[m], h2_as_u32(h[2 * m][4]), h2_as_u32(h[2 * m + 1][4]), 0u, 0u, xb.x, xb.y);
      }
    } else {  // Q8: (x0,x1)(x2,x3)
      uint2 xv = *reinterpret_cast<const uint2*>(xr);
      __half2 h[4][2];
#pragma unroll
      for (int j = 0; j < 4; ++j) unpack_q8(wv[j], s2[j], h[j]);
#pragma unroll
      for (int m = 0; m < 2; ++m) {
        mma16816(acc[m], h2_as_u32(h[2 * m][0]), h2_as_u32(h[2 * m + 1][0]), h2_as_u32(h[2 * m][1]),
                 h2_as_u32(h[2 * m + 1][1]), xv.x, xv.y);
      }
    }
  }
}

// Stage x[k0 .. ) for `units` units into shared memory in the permuted pair order, zero beyond K.
template <int FMT>
__device__ __forceinline__ void stage_x(unsigned char* xs, const __half* __restrict__ x, int K,
                                        int unit_begin, int units) {
  using F = Fmt<FMT>;
  constexpr int XB = x_row_bytes<FMT>();
  const int rows = units * 4;
  for (int lr = threadIdx.x; lr < rows; lr += kThreads) {
    int ul = lr >> 2, i = lr & 3;
    int k0 = ((unit_begin + ul) * 4 + i) * F::kNpu;
    __half v[12];
#pragma unroll
    for (int t = 0; t < 12; ++t) v[t] = __ushort_as_half(0);
    if constexpr (F::kNpu == 8) {
      if (k0 + 8 <= K && ((reinterpret_cast<uintptr_t>(x + k0) & 15) == 0)) {
        uint4 q = *reinterpret_cast<const uint4*>(x + k0);
        *reinterpret_cast<uint4*>(v) = q;
      } else {
#pragma unroll
        for (int t = 0; t < 8; ++t)
          if (k0 + t < K) v[t] = x[k0 + t];
      }
    } else {
#pragma unroll
      for (int t = 0; t < F::kNpu; ++t)
        if (k0 + t < K) v[t] = x[k0 + t];
    }
    __half o[12];
    if constexpr (FMT == B200Q_Q4F16_0) {
      o[0] = v[0]; o[1] = v[4]; o[2] = v[1]; o[3] = v[5];
      o[4] = v[2]; o[5] = v[6]; o[6] = v[3]; o[7] = v[7];
    } else if constexpr (FMT == B200Q_Q3F16_0) {
      o[0] = v[0]; o[1] = v[5]; o[2] = v[1]; o[3] = v[6];
      o[4] = v[2]; o[5] = v[7]; o[6] = v[3]; o[7] = v[8];
      o[8] = v[4]; o[9] = v[9]; o[10] = __ushort_as_half(0); o[11] = __ushort_as_half(0);
    } else {
      o[0] = v[0]; o[1] = v[1]; o[2] = v[2]; o[3] = v[3];
    }
    unsigned char* dst = xs + ((size_t)i * units + ul) * XB;
#pragma unroll
    for (int t = 0; t < XB / 8; ++t)
      reinterpret_cast<uint2*>(dst)[t] = reinterpret_cast<const uint2*>(o)[t];
  }
}

// WN warps across N (each 32 columns), WK = 8 / WN warps across K.  gridDim.x = column tiles,
// gridDim.y = K splits (= cluster size along y when > 1).
template <int FMT, int WN, bool CLUSTER>
__global__ void __launch_bounds__(kThreads, 2) gemv_f16_kernel(const GemvParams p) {
  using F = Fmt<FMT>;
  constexpr int WK = kWarps / WN;
  constexpr int XB = x_row_bytes<FMT>();
  constexpr int kCols = 32 * WN;
  extern __shared__ __align__(16) unsigned char smem[];

  pdl_launch_dependents();
  const uint64_t pol = l2_evict_first_policy();

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int wn = warp % WN, wk = warp / WN;
  const int c = lane >> 2, tig = lane & 3;

  // which op does this column tile belong to (ops are laid out back to back along gridDim.x)
  int oi = 0;
#pragma unroll
  for (int t = 1; t < kMaxOps; ++t)
    if (t < p.n_ops && (int)blockIdx.x >= p.op[t].tile_begin) oi = t;
  const GemvOp& op = p.op[oi];
  const int N = op.N;
  const int col_tile0 = ((int)blockIdx.x - op.tile_begin) * kCols;
  const int col = col_tile0 + wn * 32 + c * 4;
  const bool col_ok = col < N;

  const int unit_begin = (int)blockIdx.y * p.units_per_cta;
  const int unit_end = min(unit_begin + p.units_per_cta, p.n_units);
  const int cta_units = p.units_per_cta;          // smem rows are allocated for all of them
  const int n_iter = (unit_end - unit_begin + 3) >> 2;  // warp iterations (4 units each)

  unsigned char* xs = smem;                                     // [4][cta_units][XB]
  float* red = reinterpret_cast<float*>(smem + (size_t)4 * cta_units * XB);  // [WK][kCols]

  float acc[2][4];
#pragma unroll
  for (int m = 0; m < 2; ++m)
#pragma unroll
    for (int r = 0; r < 4; ++r) acc[m][r] = 0.f;

  // prefetch the first unit (constant weights: legal before the dependency wait)
  Unit cur, nxt;
  int it = wk;
  {
    int unit = unit_begin + it * 4 + tig;
    load_unit<FMT>(cur, op.packed, op.scale, unit, p.n_rows, col, N,
                   col_ok && it < n_iter && unit < unit_end, pol);
  }
  {
    int itn = it + WK;
    int unit = unit_begin + itn * 4 + tig;
    load_unit<FMT>(nxt, op.packed, op.scale, unit, p.n_rows, col, N,
                   col_ok && itn < n_iter && unit < unit_end, pol);
  }

  pdl_wait();  // x (and ep0 / out) belong to the previous kernel until here
  stage_x<FMT>(xs, p.x, p.K, unit_begin, cta_units);
  __syncthreads();

  const int xs_i_stride = cta_units * XB;
  while (it < n_iter) {
    compute_unit<FMT>(cur, xs + (size_t)(it * 4 + tig) * XB, xs_i_stride, acc);
    cur = nxt;
    it += WK;
    int itn = it + WK;
    int unit = unit_begin + itn * 4 + tig;
    load_unit<FMT>(nxt, op.packed, op.scale, unit, p.n_rows, col, N,
                   col_ok && itn < n_iter && unit < unit_end, pol);
  }

  // ---- reduce over the K-split warps of the CTA ------------------------------------------------
  if (tig == 0) {
    float4 v = make_float4(acc[0][0], acc[0][2], acc[1][0], acc[1][2]);  // columns 4c .. 4c+3
    *reinterpret_cast<float4*>(red + wk * kCols + wn * 32 + c * 4) = v;
  }
  __syncthreads();

  float sum = 0.f;
  const int tc = threadIdx.x;  // column within the tile handled by this thread in the tail
  if (tc < kCols) {
#pragma unroll
    for (int w = 0; w < WK; ++w) sum += red[w * kCols + tc];
  }

  if constexpr (CLUSTER) {
    cg::cluster_group cluster = cg::this_cluster();
    const unsigned nranks = cluster.num_blocks();
    const unsigned rank = cluster.block_rank();
    __syncthreads();
    if (tc < kCols) red[tc] = sum;  // CTA partial at red[0 .. kCols)
    cluster.sync();
    if (rank == 0 && tc < kCols) {
      for (unsigned r = 1; r < nranks; ++r) {
        const float* remote = cluster.map_shared_rank(red, r);
        sum += remote[tc];
      }
    }
    cluster.sync();  // keep the other CTAs' shared memory alive until rank 0 has read it
    if (rank != 0) return;
  }

  const int n = col_tile0 + tc;
  if (tc < kCols && n < N) {
    if (op.epilogue == B200Q_EP_CAST_F32) {
      reinterpret_cast<float*>(op.out)[n] = sum;
    } else {
      reinterpret_cast<__half*>(op.out)[n] = ep_f16(sum, op.epilogue, op.ep0, n);
    }
  }
}

// ---- host side --------------------------------------------------------------------------------
static thread_local int g_tune_wn = 0, g_tune_ksplit = 0;
void gemv_tune(int wn, int ksplit) {
  g_tune_wn = wn;
  g_tune_ksplit = ksplit;
}
void gemv_get_tune(int* wn, int* ksplit) {
  *wn = g_tune_wn;
  *ksplit = g_tune_ksplit;
}

template <int FMT, int WN, bool CLUSTER>
static int launch_cfg(const GemvParams& p, int tiles, int ksplit, size_t smem, unsigned flags,
                      cudaStream_t stream) {
  auto kern = gemv_f16_kernel<FMT, WN, CLUSTER>;
  static thread_local size_t smem_set = 0;  // per instantiation
  if (smem > 48 * 1024 && smem > smem_set) {
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return check_cuda(e, "cudaFuncSetAttribute(gemv smem)");
    smem_set = smem;
  }
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3(tiles, ksplit, 1);
  cfg.blockDim = dim3(kThreads, 1, 1);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = stream;
  cudaLaunchAttribute attrs[2];
  int na = 0;
  if (CLUSTER) {
    attrs[na].id = cudaLaunchAttributeClusterDimension;
    attrs[na].val.clusterDim.x = 1;
    attrs[na].val.clusterDim.y = ksplit;
    attrs[na].val.clusterDim.z = 1;
    ++na;
  }
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

template <int FMT>
static int launch_fmt(GemvParams& p, const int* Ns, unsigned flags, cudaStream_t stream) {
  constexpr int XB = x_row_bytes<FMT>();
  // ---- heuristic: fill >= 2 CTAs per SM, keep >= 1 iteration per warp ---------------------------
  int total_n = 0;
  for (int i = 0; i < p.n_ops; ++i) total_n += Ns[i];
  int wn = g_tune_wn;
  if (wn != 1 && wn != 2 && wn != 4 && wn != 8) wn = (total_n >= 16384) ? 2 : 1;
  int tiles = 0;
  for (int i = 0; i < p.n_ops; ++i) {
    p.op[i].tile_begin = tiles;
    tiles += ceil_div(Ns[i], 32 * wn);
  }
  const int wk = kWarps / wn;
  const int iters = ceil_div(p.n_units, 4);  // warp iterations over all of K
  int ksplit = g_tune_ksplit;
  if (ksplit != 1 && ksplit != 2 && ksplit != 4 && ksplit != 8) {
    ksplit = 1;
    while (ksplit < 8 && tiles * ksplit < 2 * 148 && iters / (ksplit * 2) >= wk) ksplit *= 2;
  }
  while (ksplit > 1 && iters < ksplit) ksplit /= 2;
  int units_per_cta = ceil_div(ceil_div(p.n_units, ksplit), 4) * 4;
  while (ksplit > 1 && units_per_cta * (ksplit - 1) >= p.n_units) {  // no empty CTAs
    ksplit /= 2;
    units_per_cta = ceil_div(ceil_div(p.n_units, ksplit), 4) * 4;
  }
  p.units_per_cta = units_per_cta;
  size_t smem = (size_t)4 * units_per_cta * XB + (size_t)wk * 32 * wn * sizeof(float);
  if (smem > 200 * 1024) return set_error(B200Q_ERR_BAD_SHAPE, "gemv: K=%d too large for one CTA's x stage", p.K);

#define B200Q_LAUNCH(WN_)                                                                        \
  (ksplit > 1 ? launch_cfg<FMT, WN_, true>(p, tiles, ksplit, smem, flags, stream)                \
              : launch_cfg<FMT, WN_, false>(p, tiles, ksplit, smem, flags, stream))
  switch (wn) {
    case 1: return B200Q_LAUNCH(1);
    case 2: return B200Q_LAUNCH(2);
    case 4: return B200Q_LAUNCH(4);
    default: return B200Q_LAUNCH(8);
  }
#undef B200Q_LAUNCH
}

int gemv_f16_launch(const b200q_gemv_op* ops, int n_ops, const void* x, int K, int fmt,
                    unsigned flags, cudaStream_t stream) {
  GemvParams p{};
  int Ns[kMaxOps];
  p.n_ops = n_ops;
  p.x = reinterpret_cast<const __half*>(x);
  p.K = K;
  p.n_rows = n_words(fmt, K);
  p.n_units = ceil_div(p.n_rows, 4);
  for (int i = 0; i < n_ops; ++i) {
    p.op[i].packed = reinterpret_cast<const uint32_t*>(ops[i].packed);
    p.op[i].scale = reinterpret_cast<const __half*>(ops[i].scale);
    p.op[i].ep0 = reinterpret_cast<const __half*>(ops[i].ep0);
    p.op[i].out = ops[i].out;
    p.op[i].N = ops[i].N;
    p.op[i].epilogue = ops[i].epilogue;
    Ns[i] = ops[i].N;
  }
  switch (fmt) {
    case B200Q_Q4F16_0: return launch_fmt<B200Q_Q4F16_0>(p, Ns, flags, stream);
    case B200Q_Q3F16_0: return launch_fmt<B200Q_Q3F16_0>(p, Ns, flags, stream);
    case B200Q_Q8F16_0: return launch_fmt<B200Q_Q8F16_0>(p, Ns, flags, stream);
    default: return set_error(B200Q_ERR_UNSUPPORTED, "gemv_f16: fmt %d", fmt);
  }
}

}  // namespace b200q
