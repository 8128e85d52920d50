// The remaining --quantization-* combinations of the reference, sm_100a:
//   B200Q_Q3F16_U16   int3 in uint16 words, 5 values per word, group 40 (the iOS build, ios/README.md:39;
//                     decode*_int3_int16_fp16 / fused_decode*_int3_int16_fp16, dispatch/llama/main.py:4151-4203, 5815-6363)
//   B200Q_FP4F16_0 / B200Q_FP4F32_0   fp4 = sign + 3-bit exponent code (quantization.py:51-101), uint32 words, group 32
// Straight streaming kernels (one thread per packed word / output column; HBM-bound shapes, not the headline path):
// encode, dequantize (both layouts), take_decode and the decode-step GEMV, each bit-exact against the reference's
// te_encode_sym / te_decode_sym (oracle + reference-run goldens, tests/test_variants_gpu.py); the GEMV accumulates
// the exact products of the bit-exact weights in f32.
#include "common.cuh"

namespace b200q {
namespace {

template <typename T>
struct VNum;
template <>
struct VNum<__half> {
  __device__ static float ld(const __half* p) { return __half2float(*p); }
  __device__ static float rnd(float v) { return __half2float(__float2half_rn(v)); }
  __device__ static __half cvt(float v) { return __float2half_rn(v); }
  // _tir_f16_to_uint_to_f4 (quantization.py:64-75)
  __device__ static uint32_t to_f4(float v) {
    const uint32_t u = __half_as_ushort(__float2half_rn(v));
    const uint32_t m_h = (u >> 9) & 1u, e = (u >> 10) & 31u, s = u >> 15;
    const uint32_t e4 = e > 8u ? min(e - 8u + m_h, 7u) : (e == 8u ? 1u : 0u);
    return (s << 3) | e4;
  }
  // _tir_u32_to_f4_to_f16 (:88-101): exponent field e4 | 8, i.e. 2^(e4 - 7)
  __device__ static float from_f4(uint32_t c) {
    const uint32_t e4 = c & 7u, s = (c >> 3) & 1u;
    return e4 == 0u ? 0.f : __half2float(__ushort_as_half((unsigned short)(((e4 | 8u) | (s << 5)) << 10)));
  }
};
template <>
struct VNum<float> {
  __device__ static float ld(const float* p) { return *p; }
  __device__ static float rnd(float v) { return v; }
  __device__ static float cvt(float v) { return v; }
  __device__ static uint32_t to_f4(float v) {  // _tir_f32_to_uint_to_f4 (:51-62)
    const uint32_t u = __float_as_uint(v);
    const uint32_t m_h = (u >> 22) & 1u, e = (u >> 23) & 255u, s = u >> 31;
    const uint32_t e4 = e > 120u ? min(e - 120u + m_h, 7u) : (e == 120u ? 1u : 0u);
    return (s << 3) | e4;
  }
  __device__ static float from_f4(uint32_t c) {  // _tir_u32_to_f4_to_f32 (:78-86)
    const uint32_t e4 = c & 7u, s = (c >> 3) & 1u;
    return e4 == 0u ? 0.f : __uint_as_float(((e4 | 120u) | (s << 8)) << 23);
  }
};

struct VFmt {
  int nbit, group, npu, rpg, fp4, u16;  // rpg: storage words per group
};
__host__ __device__ inline VFmt vfmt(int fmt) {
  if (fmt == B200Q_Q3F16_U16) return VFmt{3, 40, 5, 8, 0, 1};
  return VFmt{4, 32, 8, 4, 1, 0};
}
__device__ __forceinline__ uint32_t ld_word(const void* packed, size_t idx, int u16) {
  return u16 ? (uint32_t)reinterpret_cast<const unsigned short*>(packed)[idx] : reinterpret_cast<const uint32_t*>(packed)[idx];
}
// value t of a storage word, times the group scale, rounded once to T (te_decode_sym, quantization.py:213-230)
template <typename T>
__device__ __forceinline__ float dec_val(uint32_t w, int t, const VFmt& f, float s) {
  const uint32_t q = (w >> (t * f.nbit)) & ((1u << f.nbit) - 1u);
  const float v = f.fp4 ? VNum<T>::from_f4(q) : (float)q - 3.0f;
  return VNum<T>::rnd(__fmul_rn(v, s));
}

// ---- encode: one thread per (n, group) ------------------------------------------------------------------------
template <typename T>
__global__ void __launch_bounds__(256) v_encode_kernel(const T* __restrict__ W, void* __restrict__ packed, T* __restrict__ scale,
                                                       int N, int K, int fmt, int transpose) {
  const VFmt f = vfmt(fmt);
  const int G = ceil_div(K, f.group), R = f.rpg * G;
  const long long total = (long long)N * G;
  for (long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x; idx < total; idx += (long long)gridDim.x * blockDim.x) {
    const int n = (int)(idx % N), g = (int)(idx / N);
    const T* w = W + (size_t)n * K + (size_t)g * f.group;
    const int cnt = min(f.group, K - g * f.group);
    float mx = 0.f;
    for (int k = 0; k < cnt; ++k) mx = fmaxf(mx, fabsf(VNum<T>::ld(w + k)));
    mx = fmaxf(mx, VNum<T>::rnd(1e-4f));
    const float s = f.fp4 ? mx : VNum<T>::rnd(__fdiv_rn(mx, 3.0f));
    scale[transpose ? (size_t)g * N + n : (size_t)n * G + g] = VNum<T>::cvt(s);
    for (int r = 0; r < f.rpg; ++r) {
      uint32_t word = 0;
      for (int t = 0; t < f.npu; ++t) {
        const int kk = r * f.npu + t;
        if (kk < cnt) {
          const float q0 = VNum<T>::rnd(__fdiv_rn(VNum<T>::ld(w + kk), s));
          uint32_t code;
          if (f.fp4) {
            code = VNum<T>::to_f4(q0);
          } else {
            float q = roundf(VNum<T>::rnd(__fadd_rn(q0, 3.0f)));
            code = (uint32_t)fminf(fmaxf(q, 0.f), 6.f);
          }
          word |= code << (t * f.nbit);
        }
      }
      const int row = g * f.rpg + r;
      const size_t pi = transpose ? (size_t)row * N + n : (size_t)n * R + row;
      if (f.u16) reinterpret_cast<unsigned short*>(packed)[pi] = (unsigned short)word;
      else reinterpret_cast<uint32_t*>(packed)[pi] = word;
    }
  }
}

// ---- dequantize: one thread per storage word; out[K, N] or out[N, K] ---------------------------------------------
template <typename T>
__global__ void __launch_bounds__(256) v_dequant_kernel(const void* __restrict__ packed, const T* __restrict__ scale,
                                                        T* __restrict__ out, int K, int N, int fmt, int transposed) {
  const VFmt f = vfmt(fmt);
  const int R = f.rpg * ceil_div(K, f.group);
  const long long total = (long long)R * N;
  for (long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x; idx < total; idx += (long long)gridDim.x * blockDim.x) {
    const int n = (int)(idx % N), r = (int)(idx / N);
    const uint32_t w = ld_word(packed, (size_t)r * N + n, f.u16);
    const float s = VNum<T>::ld(scale + (size_t)(r / f.rpg) * N + n);
    for (int t = 0; t < f.npu; ++t) {
      const int k = r * f.npu + t;
      if (k < K) out[transposed ? (size_t)n * K + k : (size_t)k * N + n] = VNum<T>::cvt(dec_val<T>(w, t, f, s));
    }
  }
}

// ---- take_decode: row-major (V, n_words) table ----------------------------------------------------------------------
template <typename T>
__global__ void __launch_bounds__(256) v_take_kernel(const void* __restrict__ packed_rm, const T* __restrict__ scale_rm,
                                                     const int32_t* __restrict__ ids, T* __restrict__ out, int V, int K, int fmt) {
  const VFmt f = vfmt(fmt);
  const int G = ceil_div(K, f.group), R = f.rpg * G;
  const int i = blockIdx.y, id = ids[i];
  const bool bad = id < 0 || id >= V;  // undefined in the reference: zero-filled here
  for (int r = blockIdx.x * blockDim.x + threadIdx.x; r < R; r += gridDim.x * blockDim.x) {
    const uint32_t w = bad ? 0u : ld_word(packed_rm, (size_t)id * R + r, f.u16);
    const float s = bad ? 0.f : VNum<T>::ld(scale_rm + (size_t)id * G + r / f.rpg);
    for (int t = 0; t < f.npu; ++t) {
      const int k = r * f.npu + t;
      if (k < K) out[(size_t)i * K + k] = VNum<T>::cvt(bad ? 0.f : dec_val<T>(w, t, f, s));
    }
  }
}

// ---- decode-step GEMV: y[n] = sum_k x[k] * w[k, n] (+ epilogue) ------------------------------------------------------
// block = 4 k-slices x 64 columns; a thread walks every fourth storage-word row of its column (coalesced along N),
// f32 accumulation of exact products of the bit-exact weights, slices added in order, epilogue op for op.
template <typename T>
__global__ void __launch_bounds__(256) v_gemv_kernel(const void* __restrict__ packed, const T* __restrict__ scale,
                                                     const T* __restrict__ x, const T* __restrict__ ep0, void* __restrict__ out,
                                                     int K, int N, int fmt, int epilogue) {
  __shared__ float red[4][64];
  const VFmt f = vfmt(fmt);
  const int R = f.rpg * ceil_div(K, f.group);
  const int col = blockIdx.x * 64 + (threadIdx.x & 63), slice = threadIdx.x >> 6;
  float acc = 0.f;
  if (col < N) {
    for (int r = slice; r < R; r += 4) {
      const uint32_t w = ld_word(packed, (size_t)r * N + col, f.u16);
      const float s = VNum<T>::ld(scale + (size_t)(r / f.rpg) * N + col);
      for (int t = 0; t < f.npu; ++t) {
        const int k = r * f.npu + t;
        if (k < K) acc = fmaf(VNum<T>::ld(x + k), dec_val<T>(w, t, f, s), acc);
      }
    }
  }
  red[slice][threadIdx.x & 63] = acc;
  __syncthreads();
  if (slice == 0 && col < N) {
    const float sum = ((red[0][threadIdx.x] + red[1][threadIdx.x]) + red[2][threadIdx.x]) + red[3][threadIdx.x];
    if constexpr (sizeof(T) == 4) {
      reinterpret_cast<float*>(out)[col] = ep_f32(sum, epilogue, reinterpret_cast<const float*>(ep0), col);
    } else {
      if (epilogue == B200Q_EP_CAST_F32) reinterpret_cast<float*>(out)[col] = sum;
      else reinterpret_cast<__half*>(out)[col] = ep_f16(sum, epilogue, reinterpret_cast<const __half*>(ep0), col);
    }
  }
}

int grid_for(long long total) {
  long long b = (total + 255) / 256;
  return (int)(b < 1 ? 1 : (b > 148 * 32 ? 148 * 32 : b));
}
}  // namespace

bool fmt_is_variant(int fmt) { return fmt == B200Q_Q3F16_U16 || fmt == B200Q_FP4F16_0 || fmt == B200Q_FP4F32_0; }

int variant_encode_launch(const void* W, void* packed, void* scale, int N, int K, int fmt, int transpose, cudaStream_t stream) {
  const int g = grid_for((long long)N * n_groups(fmt, K));
  if (fmt == B200Q_FP4F32_0)
    v_encode_kernel<float><<<g, 256, 0, stream>>>(reinterpret_cast<const float*>(W), packed, reinterpret_cast<float*>(scale), N, K, fmt, transpose);
  else
    v_encode_kernel<__half><<<g, 256, 0, stream>>>(reinterpret_cast<const __half*>(W), packed, reinterpret_cast<__half*>(scale), N, K, fmt, transpose);
  count_launch();
  return check_cuda(cudaGetLastError(), "variant encode launch");
}

int variant_dequant_launch(const void* packed, const void* scale, void* out, int K, int N, int fmt, int transposed, cudaStream_t stream) {
  const int g = grid_for((long long)n_words(fmt, K) * N);
  if (fmt == B200Q_FP4F32_0)
    v_dequant_kernel<float><<<g, 256, 0, stream>>>(packed, reinterpret_cast<const float*>(scale), reinterpret_cast<float*>(out), K, N, fmt, transposed);
  else
    v_dequant_kernel<__half><<<g, 256, 0, stream>>>(packed, reinterpret_cast<const __half*>(scale), reinterpret_cast<__half*>(out), K, N, fmt, transposed);
  count_launch();
  return check_cuda(cudaGetLastError(), "variant dequant launch");
}

int variant_take_launch(const void* packed_rm, const void* scale_rm, const int32_t* ids, void* out, int n_ids, int V, int K, int fmt,
                        cudaStream_t stream) {
  dim3 grid(ceil_div(n_words(fmt, K), 256), n_ids);
  if (fmt == B200Q_FP4F32_0)
    v_take_kernel<float><<<grid, 256, 0, stream>>>(packed_rm, reinterpret_cast<const float*>(scale_rm), ids, reinterpret_cast<float*>(out), V, K, fmt);
  else
    v_take_kernel<__half><<<grid, 256, 0, stream>>>(packed_rm, reinterpret_cast<const __half*>(scale_rm), ids, reinterpret_cast<__half*>(out), V, K, fmt);
  count_launch();
  return check_cuda(cudaGetLastError(), "variant take_decode launch");
}

int variant_gemv_launch(const void* packed, const void* scale, const void* x, const void* ep0, void* out, int K, int N, int fmt,
                        int epilogue, cudaStream_t stream) {
  const int g = ceil_div(N, 64);
  if (fmt == B200Q_FP4F32_0)
    v_gemv_kernel<float><<<g, 256, 0, stream>>>(packed, reinterpret_cast<const float*>(scale), reinterpret_cast<const float*>(x),
                                                reinterpret_cast<const float*>(ep0), out, K, N, fmt, epilogue);
  else
    v_gemv_kernel<__half><<<g, 256, 0, stream>>>(packed, reinterpret_cast<const __half*>(scale), reinterpret_cast<const __half*>(x),
                                                 reinterpret_cast<const __half*>(ep0), out, K, N, fmt, epilogue);
  count_launch();
  return check_cuda(cudaGetLastError(), "variant gemv launch");
}

}  // namespace b200q
