// Dequantize kernels (bit-exact against the reference decode), sm_100a.
//
//   dequant_kn   out[K, N]  - the decode-step shape of te_decode_sym
//                             (mlc_llm/transform/quantization.py:213-230, transpose_output=False)
//   dequant_nk   out[N, K]  - the prefill shape that feeds NT_matmul (transpose_output=True); replaces
//                             decode4_fp16 ... decode2_int3_int16_fp16 + decode_sch_func
//                             (mlc_llm/dispatch/llama/main.py:3984-4232; generator tuning.py:111-147,
//                             which also stages a transpose through shared memory)
//   take_decode  out[n_ids, K] from the row-major (V, n_words) embedding layout
//                             (quantization.py:235-270)
//
// All three are pure HBM streaming (write-dominated: 2 or 4 bytes out per 0.5-1 byte in).
#include "common.cuh"

namespace b200q {
namespace {

template <int FMT>
struct Dec;  // decode one word into kNpu values of the model dtype

template <>
struct Dec<B200Q_Q4F16_0> {
  using T = __half;
  static constexpr int kNpu = 8;
  __device__ static void run(uint32_t w, __half s, __half, __half (&o)[8]) {
    __half2 h[4];
    unpack_q4(w, __half2half2(s), h);
#pragma unroll
    for (int t = 0; t < 4; ++t) {
      o[t] = __low2half(h[t]);
      o[t + 4] = __high2half(h[t]);
    }
  }
};
template <>
struct Dec<B200Q_Q3F16_0> {
  using T = __half;
  static constexpr int kNpu = 10;
  __device__ static void run(uint32_t w, __half s, __half, __half (&o)[10]) {
    __half2 h[5];
    unpack_q3(w, __half2half2(s), h);
#pragma unroll
    for (int t = 0; t < 5; ++t) {
      o[t] = __low2half(h[t]);
      o[t + 5] = __high2half(h[t]);
    }
  }
};
template <>
struct Dec<B200Q_Q8F16_0> {
  using T = __half;
  static constexpr int kNpu = 4;
  __device__ static void run(uint32_t w, __half s, __half, __half (&o)[4]) {
    __half2 h[2];
    unpack_q8(w, __half2half2(s), h);
    o[0] = __low2half(h[0]);
    o[1] = __high2half(h[0]);
    o[2] = __low2half(h[1]);
    o[3] = __high2half(h[1]);
  }
};
template <>
struct Dec<B200Q_Q4F32_0> {
  using T = float;
  static constexpr int kNpu = 8;
  __device__ static void run(uint32_t w, float s, float, float (&o)[8]) {
#pragma unroll
    for (int t = 0; t < 8; ++t) {
      uint32_t q = (w >> (4 * t)) & 0xFu;
      o[t] = __fmul_rn(__uint_as_float(q | 0x4B000000u) - 8388615.0f, s);
    }
  }
};

template <>
struct Dec<B200Q_Q4F16_ASYM> {  // w = half(half(q * s) + m)   (quantization.py:196)
  using T = __half;
  static constexpr int kNpu = 8;
  __device__ static void run(uint32_t w, __half s, __half m, __half (&o)[8]) {
    __half2 h[4];
    unpack_q4_asym(w, __half2half2(s), __half2half2(m), h);
#pragma unroll
    for (int t = 0; t < 4; ++t) {
      o[t] = __low2half(h[t]);
      o[t + 4] = __high2half(h[t]);
    }
  }
};
template <>
struct Dec<B200Q_Q4F32_ASYM> {  // s carries the bit pattern bf16(scale) | bf16(min) << 16
  using T = float;
  static constexpr int kNpu = 8;
  __device__ static void run(uint32_t w, float s, float, float (&o)[8]) {
    const uint32_t sm = __float_as_uint(s);
#pragma unroll
    for (int t = 0; t < 8; ++t) o[t] = asym_f32_weight(w, t, sm);
  }
};

// out[K, N]: a thread owns 4 consecutive columns of one word row -> writes kNpu rows of 4 values.
template <int FMT>
__global__ void __launch_bounds__(256) dequant_kn_kernel(const uint32_t* __restrict__ packed,
                                                         const typename Dec<FMT>::T* __restrict__ scale,
                                                         const typename Dec<FMT>::T* __restrict__ zmin,
                                                         typename Dec<FMT>::T* __restrict__ out, int K,
                                                         int N, int n_rows) {
  using D = Dec<FMT>;
  using T = typename D::T;
  constexpr int kRowsPerGroup = Fmt<FMT>::kRowsPerGroup;
  const int n4 = N >> 2;
  const long long total = (long long)n_rows * n4;
  for (long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x; idx < total;
       idx += (long long)gridDim.x * blockDim.x) {
    const int row = (int)(idx / n4);
    const int col = (int)(idx % n4) * 4;
    const uint4 w = ldg_stream_v4(packed + (size_t)row * N + col);
    const uint32_t wv[4] = {w.x, w.y, w.z, w.w};
    T s[4], m[4];
    const T* sp = scale + (size_t)(row / kRowsPerGroup) * N + col;
#pragma unroll
    for (int j = 0; j < 4; ++j) s[j] = sp[j];
#pragma unroll
    for (int j = 0; j < 4; ++j) m[j] = fmt_has_min_c(FMT) ? zmin[(size_t)(row / kRowsPerGroup) * N + col + j] : T(0.0f);
    T o[4][D::kNpu];
#pragma unroll
    for (int j = 0; j < 4; ++j) D::run(wv[j], s[j], m[j], o[j]);
#pragma unroll
    for (int t = 0; t < D::kNpu; ++t) {
      const int k = row * D::kNpu + t;
      if (k < K) {
        T* dst = out + (size_t)k * N + col;
        if constexpr (sizeof(T) == 2) {
          __half2 lo = __halves2half2(o[0][t], o[1][t]), hi = __halves2half2(o[2][t], o[3][t]);
          uint2 v = make_uint2(h2_as_u32(lo), h2_as_u32(hi));
          *reinterpret_cast<uint2*>(dst) = v;
        } else {
          *reinterpret_cast<float4*>(dst) = make_float4(o[0][t], o[1][t], o[2][t], o[3][t]);
        }
      }
    }
  }
}

// out[N, K]: tile of 32 word rows x 32 columns transposed through shared memory so that both the
// packed reads (along N) and the dense writes (along K) are contiguous.
template <int FMT>
__global__ void __launch_bounds__(256) dequant_nk_kernel(const uint32_t* __restrict__ packed,
                                                         const typename Dec<FMT>::T* __restrict__ scale,
                                                         const typename Dec<FMT>::T* __restrict__ zmin,
                                                         typename Dec<FMT>::T* __restrict__ out, int K,
                                                         int N, int n_rows) {
  using D = Dec<FMT>;
  using T = typename D::T;
  constexpr int kRowsPerGroup = Fmt<FMT>::kRowsPerGroup;
  constexpr int kGroupsPerTile = 32 / kRowsPerGroup;
  __shared__ uint32_t tw[32][33];
  __shared__ T ts[kGroupsPerTile][33];
  __shared__ T tm[kGroupsPerTile][33];
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;  // 8 warps
  const int row0 = blockIdx.y * 32, col0 = blockIdx.x * 32;
#pragma unroll
  for (int r = ty; r < 32; r += 8) {
    int row = row0 + r, col = col0 + tx;
    tw[r][tx] = (row < n_rows && col < N) ? ldg_stream_u32(packed + (size_t)row * N + col) : 0u;
  }
  if (ty < kGroupsPerTile) {
    int g = row0 / kRowsPerGroup + ty, col = col0 + tx;
    const bool in = (g * kRowsPerGroup < n_rows && col < N);
    ts[ty][tx] = in ? scale[(size_t)g * N + col] : T(0);
    tm[ty][tx] = (in && fmt_has_min_c(FMT)) ? zmin[(size_t)g * N + col] : T(0);
  }
  __syncthreads();
  // now lane <-> word row (contiguous k), warp <-> column
#pragma unroll
  for (int cc = ty; cc < 32; cc += 8) {
    const int col = col0 + cc, row = row0 + tx;
    if (col >= N || row >= n_rows) continue;
    T o[D::kNpu];
    D::run(tw[tx][cc], ts[tx / kRowsPerGroup][cc], tm[tx / kRowsPerGroup][cc], o);
    const int k0 = row * D::kNpu;
    T* dst = out + (size_t)col * K + k0;
    if (k0 + D::kNpu <= K && ((reinterpret_cast<uintptr_t>(dst) & 15) == 0) && (sizeof(T) * D::kNpu) % 16 == 0) {
#pragma unroll
      for (int t = 0; t < (int)(sizeof(T) * D::kNpu / 16); ++t)
        reinterpret_cast<uint4*>(dst)[t] = reinterpret_cast<const uint4*>(o)[t];
    } else {
#pragma unroll
      for (int t = 0; t < D::kNpu; ++t)
        if (k0 + t < K) dst[t] = o[t];
    }
  }
}

// embedding gather: one warp-strided pass over the words of each requested row
template <int FMT>
__global__ void __launch_bounds__(256) take_decode_kernel(const uint32_t* __restrict__ packed_rm,
                                                          const typename Dec<FMT>::T* __restrict__ scale_rm,
                                                          const typename Dec<FMT>::T* __restrict__ zmin_rm,
                                                          const int32_t* __restrict__ ids,
                                                          typename Dec<FMT>::T* __restrict__ out, int V,
                                                          int K, int n_rows, int n_grp) {
  using D = Dec<FMT>;
  using T = typename D::T;
  constexpr int kRowsPerGroup = Fmt<FMT>::kRowsPerGroup;
  const int i = blockIdx.y;
  const int id = ids[i];
  // an out-of-range id is undefined in the reference (TVM take reads out of bounds); here the row is zero-filled
  // so that a caller-allocated (uninitialised) output never leaks stale memory
  const bool bad = (id < 0 || id >= V);
  for (int r = blockIdx.x * blockDim.x + threadIdx.x; r < n_rows; r += gridDim.x * blockDim.x) {
    T o[D::kNpu];
    if (bad) {
#pragma unroll
      for (int t = 0; t < D::kNpu; ++t) o[t] = T(0.0f);
    } else {
      const size_t gi = (size_t)id * n_grp + r / kRowsPerGroup;
      D::run(packed_rm[(size_t)id * n_rows + r], scale_rm[gi], fmt_has_min_c(FMT) ? zmin_rm[gi] : T(0.0f), o);
    }
    const int k0 = r * D::kNpu;
#pragma unroll
    for (int t = 0; t < D::kNpu; ++t)
      if (k0 + t < K) out[(size_t)i * K + k0 + t] = o[t];
  }
}

template <int FMT>
int launch_dequant(const void* packed, const void* scale, const void* zmin, void* out, int K, int N, int transposed,
                   cudaStream_t stream) {
  using T = typename Dec<FMT>::T;
  const int rows = n_words(FMT, K);
  if (!transposed) {
    long long total = (long long)rows * (N / 4);
    int blocks = (int)((total + 255) / 256);
    if (blocks > 148 * 16) blocks = 148 * 16;
    if (blocks < 1) blocks = 1;
    dequant_kn_kernel<FMT><<<blocks, 256, 0, stream>>>(reinterpret_cast<const uint32_t*>(packed),
                                                       reinterpret_cast<const T*>(scale), reinterpret_cast<const T*>(zmin),
                                                       reinterpret_cast<T*>(out), K, N, rows);
  } else {
    dim3 grid(ceil_div(N, 32), ceil_div(rows, 32));
    dequant_nk_kernel<FMT><<<grid, 256, 0, stream>>>(reinterpret_cast<const uint32_t*>(packed),
                                                     reinterpret_cast<const T*>(scale), reinterpret_cast<const T*>(zmin),
                                                     reinterpret_cast<T*>(out), K, N, rows);
  }
  count_launch();
  return check_cuda(cudaGetLastError(), "dequant launch");
}

template <int FMT>
int launch_take(const void* packed_rm, const void* scale_rm, const void* zmin_rm, const int32_t* ids, void* out, int n_ids,
                int V, int K, cudaStream_t stream) {
  using T = typename Dec<FMT>::T;
  const int rows = n_words(FMT, K), grp = n_groups(FMT, K);
  dim3 grid(ceil_div(rows, 256), n_ids);
  take_decode_kernel<FMT><<<grid, 256, 0, stream>>>(reinterpret_cast<const uint32_t*>(packed_rm),
                                                    reinterpret_cast<const T*>(scale_rm), reinterpret_cast<const T*>(zmin_rm), ids,
                                                    reinterpret_cast<T*>(out), V, K, rows, grp);
  count_launch();
  return check_cuda(cudaGetLastError(), "take_decode launch");
}
}  // namespace

int dequant_launch(const void* packed, const void* scale, const void* zmin, void* out, int K, int N, int fmt,
                   int transposed, cudaStream_t stream) {
  switch (fmt) {
    case B200Q_Q4F16_0: return launch_dequant<B200Q_Q4F16_0>(packed, scale, zmin, out, K, N, transposed, stream);
    case B200Q_Q3F16_0: return launch_dequant<B200Q_Q3F16_0>(packed, scale, zmin, out, K, N, transposed, stream);
    case B200Q_Q4F32_0: return launch_dequant<B200Q_Q4F32_0>(packed, scale, zmin, out, K, N, transposed, stream);
    case B200Q_Q8F16_0: return launch_dequant<B200Q_Q8F16_0>(packed, scale, zmin, out, K, N, transposed, stream);
    case B200Q_Q4F16_ASYM: return launch_dequant<B200Q_Q4F16_ASYM>(packed, scale, zmin, out, K, N, transposed, stream);
    case B200Q_Q4F32_ASYM: return launch_dequant<B200Q_Q4F32_ASYM>(packed, scale, zmin, out, K, N, transposed, stream);
    default: return set_error(B200Q_ERR_UNSUPPORTED, "dequant: fmt %d", fmt);
  }
}

int take_decode_launch(const void* packed_rm, const void* scale_rm, const void* zmin_rm, const int32_t* ids, void* out,
                       int n_ids, int V, int K, int fmt, cudaStream_t stream) {
  switch (fmt) {
    case B200Q_Q4F16_0: return launch_take<B200Q_Q4F16_0>(packed_rm, scale_rm, zmin_rm, ids, out, n_ids, V, K, stream);
    case B200Q_Q3F16_0: return launch_take<B200Q_Q3F16_0>(packed_rm, scale_rm, zmin_rm, ids, out, n_ids, V, K, stream);
    case B200Q_Q4F32_0: return launch_take<B200Q_Q4F32_0>(packed_rm, scale_rm, zmin_rm, ids, out, n_ids, V, K, stream);
    case B200Q_Q8F16_0: return launch_take<B200Q_Q8F16_0>(packed_rm, scale_rm, zmin_rm, ids, out, n_ids, V, K, stream);
    case B200Q_Q4F16_ASYM: return launch_take<B200Q_Q4F16_ASYM>(packed_rm, scale_rm, zmin_rm, ids, out, n_ids, V, K, stream);
    case B200Q_Q4F32_ASYM: return launch_take<B200Q_Q4F32_ASYM>(packed_rm, scale_rm, zmin_rm, ids, out, n_ids, V, K, stream);
    default: return set_error(B200Q_ERR_UNSUPPORTED, "take_decode: fmt %d", fmt);
  }
}

}  // namespace b200q
