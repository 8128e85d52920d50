// Decode-step fused dequantize + GEMV for q4f32_0 (f32 scales, activations, weights, outputs).
//
// Replaces the f32-scale int4 GEMVs of the reference (the f32 `embed_out` head of the GPT-NeoX
// build, log_db/RedPajama-INCITE-Chat-3B-v1/database_tuning_record.json:18; decode formula
// mlc_llm/transform/quantization.py:213-230 with dtype="float32").
//
// f32 x f32 products have no tensor-core form at full precision, so this is the SIMT version of
// gemv.cu: same 128-bit streaming loads (4 columns per lane, one 128-byte line per word row and
// warp), in-register nibble unpack to the bit-exact f32 weight  fl32(f32(q - 7) * s), FFMA into
// four column accumulators per lane, K split across the four lanes of a quad (warp shuffles),
// the warps of the CTA (shared memory) and the CTAs of a cluster (distributed shared memory).
#include <cooperative_groups.h>

#include "common.cuh"

namespace cg = cooperative_groups;

namespace b200q {

namespace {
constexpr int kThreads = 256;
constexpr int kWarps = 8;
constexpr int kMaxOps = 4;

struct Op32 {
  const uint32_t* packed;
  const float* scale;
  const float* ep0;
  float* out;
  int N;
  int epilogue;
  int tile_begin;
};
struct Params32 {
  Op32 op[kMaxOps];
  int n_ops;
  const float* x;
  int K;
  int n_rows;
  int n_units;
  int units_per_cta;
};

struct Unit32 {
  uint4 w[4];
  uint4 s;
};

__device__ __forceinline__ void load_unit32(Unit32& u, const uint32_t* __restrict__ packed,
                                            const float* __restrict__ scale, int unit, int n_rows,
                                            int col, int N, bool valid, uint64_t pol) {
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    int row = unit * 4 + i;
    if (valid && row < n_rows)
      u.w[i] = ldg_stream_v4(packed + (size_t)row * N + col, pol);
    else
      u.w[i] = make_uint4(0, 0, 0, 0);
  }
  if (valid)
    u.s = ldg_stream_v4(scale + (size_t)unit * N + col, pol);
  else
    u.s = make_uint4(0, 0, 0, 0);
}

// (q | 0x4B000000) is the f32 2^23 + q; subtracting 2^23 + 7 gives q - 7 exactly.
__device__ __forceinline__ float nib_to_f32(uint32_t w, int t) {
  uint32_t q = (w >> (4 * t)) & 0xFu;
  return __uint_as_float(q | 0x4B000000u) - 8388615.0f;
}

__device__ __forceinline__ void compute_unit32(const Unit32& u, const float* xs_unit, float (&acc)[4]) {
  const float s[4] = {__uint_as_float(u.s.x), __uint_as_float(u.s.y), __uint_as_float(u.s.z),
                      __uint_as_float(u.s.w)};
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const uint32_t wv[4] = {u.w[i].x, u.w[i].y, u.w[i].z, u.w[i].w};
    float4 xa = *reinterpret_cast<const float4*>(xs_unit + i * 8);
    float4 xb = *reinterpret_cast<const float4*>(xs_unit + i * 8 + 4);
    const float xv[8] = {xa.x, xa.y, xa.z, xa.w, xb.x, xb.y, xb.z, xb.w};
#pragma unroll
    for (int j = 0; j < 4; ++j) {
#pragma unroll
      for (int t = 0; t < 8; ++t) {
        float w = __fmul_rn(nib_to_f32(wv[j], t), s[j]);  // the reference's single rounding
        acc[j] = fmaf(xv[t], w, acc[j]);
      }
    }
  }
}

template <int WN, bool CLUSTER>
__global__ void __launch_bounds__(kThreads, 2) gemv_f32_kernel(const Params32 p) {
  constexpr int WK = kWarps / WN;
  constexpr int kCols = 32 * WN;
  extern __shared__ __align__(16) unsigned char smem[];

  pdl_launch_dependents();
  const uint64_t pol = l2_evict_first_policy();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int wn = warp % WN, wk = warp / WN;
  const int c = lane >> 2, tig = lane & 3;

  int oi = 0;
#pragma unroll
  for (int t = 1; t < kMaxOps; ++t)
    if (t < p.n_ops && (int)blockIdx.x >= p.op[t].tile_begin) oi = t;
  const Op32& op = p.op[oi];
  const int N = op.N;
  const int col_tile0 = ((int)blockIdx.x - op.tile_begin) * kCols;
  const int col = col_tile0 + wn * 32 + c * 4;
  const bool col_ok = col < N;

  const int unit_begin = (int)blockIdx.y * p.units_per_cta;
  const int unit_end = min(unit_begin + p.units_per_cta, p.n_units);
  const int cta_units = p.units_per_cta;
  const int n_iter = (unit_end - unit_begin + 3) >> 2;

  float* xs = reinterpret_cast<float*>(smem);  // [cta_units][32] natural k order
  float* red = xs + (size_t)cta_units * 32;    // [WK][kCols]

  float acc[4] = {0.f, 0.f, 0.f, 0.f};
  Unit32 cur, nxt;
  int it = wk;
  {
    int unit = unit_begin + it * 4 + tig;
    load_unit32(cur, op.packed, op.scale, unit, p.n_rows, col, N, col_ok && it < n_iter && unit < unit_end, pol);
    int itn = it + WK;
    unit = unit_begin + itn * 4 + tig;
    load_unit32(nxt, op.packed, op.scale, unit, p.n_rows, col, N, col_ok && itn < n_iter && unit < unit_end, pol);
  }
  pdl_wait();
  for (int k = threadIdx.x; k < cta_units * 32; k += kThreads) {
    int gk = unit_begin * 32 + k;
    xs[k] = gk < p.K ? p.x[gk] : 0.f;
  }
  __syncthreads();

  while (it < n_iter) {
    compute_unit32(cur, xs + (size_t)(it * 4 + tig) * 32, acc);
    cur = nxt;
    it += WK;
    int itn = it + WK;
    int unit = unit_begin + itn * 4 + tig;
    load_unit32(nxt, op.packed, op.scale, unit, p.n_rows, col, N, col_ok && itn < n_iter && unit < unit_end, pol);
  }

  // quad reduce (the 4 lanes of a quad hold different k for the same 4 columns)
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    acc[j] += __shfl_xor_sync(0xffffffffu, acc[j], 1);
    acc[j] += __shfl_xor_sync(0xffffffffu, acc[j], 2);
  }
  if (tig == 0)
    *reinterpret_cast<float4*>(red + wk * kCols + wn * 32 + c * 4) = make_float4(acc[0], acc[1], acc[2], acc[3]);
  __syncthreads();

  float sum = 0.f;
  const int tc = threadIdx.x;
  if (tc < kCols) {
#pragma unroll
    for (int w = 0; w < WK; ++w) sum += red[w * kCols + tc];
  }
  if constexpr (CLUSTER) {
    cg::cluster_group cluster = cg::this_cluster();
    const unsigned nranks = cluster.num_blocks(), rank = cluster.block_rank();
    __syncthreads();
    if (tc < kCols) red[tc] = sum;
    cluster.sync();
    if (rank == 0 && tc < kCols)
      for (unsigned r = 1; r < nranks; ++r) sum += cluster.map_shared_rank(red, r)[tc];
    cluster.sync();
    if (rank != 0) return;
  }
  const int n = col_tile0 + tc;
  if (tc < kCols && n < N) op.out[n] = ep_f32(sum, op.epilogue, op.ep0, n);
}

template <int WN, bool CLUSTER>
int launch32(const Params32& p, int tiles, int ksplit, size_t smem, unsigned flags, cudaStream_t stream) {
  auto kern = gemv_f32_kernel<WN, CLUSTER>;
  static thread_local size_t smem_set = 0;
  if (smem > 48 * 1024 && smem > smem_set) {
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return check_cuda(e, "cudaFuncSetAttribute(gemv_f32 smem)");
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
  if (e != cudaSuccess) return check_cuda(e, "cudaLaunchKernelEx(gemv_f32)");
  count_launch();
  return B200Q_OK;
}
}  // namespace

void gemv_get_tune(int* wn, int* ksplit);

int gemv_f32_launch(const b200q_gemv_op* ops, int n_ops, const void* x, int K, unsigned flags,
                    cudaStream_t stream) {
  Params32 p{};
  p.n_ops = n_ops;
  p.x = reinterpret_cast<const float*>(x);
  p.K = K;
  p.n_rows = n_words(B200Q_Q4F32_0, K);
  p.n_units = ceil_div(p.n_rows, 4);
  int tune_wn, tune_ks;
  gemv_get_tune(&tune_wn, &tune_ks);
  int total_n = 0;
  for (int i = 0; i < n_ops; ++i) total_n += ops[i].N;
  int wn = tune_wn;
  if (wn != 1 && wn != 2 && wn != 4 && wn != 8) wn = (total_n >= 16384) ? 2 : 1;
  int tiles = 0;
  for (int i = 0; i < n_ops; ++i) {
    p.op[i].packed = reinterpret_cast<const uint32_t*>(ops[i].packed);
    p.op[i].scale = reinterpret_cast<const float*>(ops[i].scale);
    p.op[i].ep0 = reinterpret_cast<const float*>(ops[i].ep0);
    p.op[i].out = reinterpret_cast<float*>(ops[i].out);
    p.op[i].N = ops[i].N;
    p.op[i].epilogue = ops[i].epilogue;
    p.op[i].tile_begin = tiles;
    tiles += ceil_div(ops[i].N, 32 * wn);
  }
  const int wk = kWarps / wn;
  const int iters = ceil_div(p.n_units, 4);
  int ksplit = tune_ks;
  if (ksplit != 1 && ksplit != 2 && ksplit != 4 && ksplit != 8) {
    ksplit = 1;
    while (ksplit < 8 && tiles * ksplit < 2 * 148 && iters / (ksplit * 2) >= wk) ksplit *= 2;
  }
  while (ksplit > 1 && iters < ksplit) ksplit /= 2;
  int units_per_cta = ceil_div(ceil_div(p.n_units, ksplit), 4) * 4;
  while (ksplit > 1 && units_per_cta * (ksplit - 1) >= p.n_units) {
    ksplit /= 2;
    units_per_cta = ceil_div(ceil_div(p.n_units, ksplit), 4) * 4;
  }
  p.units_per_cta = units_per_cta;
  size_t smem = (size_t)units_per_cta * 32 * sizeof(float) + (size_t)wk * 32 * wn * sizeof(float);
  if (smem > 200 * 1024) return set_error(B200Q_ERR_BAD_SHAPE, "gemv_f32: K=%d too large for one CTA's x stage", K);
#define B200Q_LAUNCH(WN_)                                                          \
  (ksplit > 1 ? launch32<WN_, true>(p, tiles, ksplit, smem, flags, stream)         \
              : launch32<WN_, false>(p, tiles, ksplit, smem, flags, stream))
  switch (wn) {
    case 1: return B200Q_LAUNCH(1);
    case 2: return B200Q_LAUNCH(2);
    case 4: return B200Q_LAUNCH(4);
    default: return B200Q_LAUNCH(8);
  }
#undef B200Q_LAUNCH
}

}  // namespace b200q
