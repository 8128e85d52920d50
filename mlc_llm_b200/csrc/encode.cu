// Group-quantization encoder (a1), sm_100a: W[N, K] dtype -> packed uint32 + scale dtype.
//
// Bit-exact CUDA restatement of te_encode_sym (mlc_llm/transform/quantization.py:142-177).  The
// reference runs this once, offline, on the host CPU through LLVM (mlc_llm/utils.py:98-110); here it
// is a device kernel so that synthetic full-size models can be packed where they live.  Every op is
// rounded to the model dtype exactly as the TE expression is typed:
//   max_abs = max_{k in group, k < K} |W[n, k]|
//   scale   = max(max_abs, dtype(1e-4)) / dtype(maxint)
//   q       = clamp(round_half_away(dtype(dtype(W / scale) + dtype(maxint))), 0, 2 * maxint)
//   word j  = OR_t q[j * npu + t] << (t * nbit)   (zero for k >= K)
// transpose != 0 writes packed[word, n], scale[group, n] (matmul weights);
// transpose == 0 writes packed[n, word], scale[n, group] (embedding table).
#include "common.cuh"

namespace b200q {
namespace {

template <typename T>
struct Num;
template <>
struct Num<__half> {
  __device__ static float load(const __half* p) { return __half2float(*p); }
  __device__ static float rnd(float v) { return __half2float(__float2half_rn(v)); }
  __device__ static void store(__half* p, float v) { *p = __float2half_rn(v); }
};
template <>
struct Num<float> {
  __device__ static float load(const float* p) { return *p; }
  __device__ static float rnd(float v) { return v; }
  __device__ static void store(float* p, float v) { *p = v; }
};

// one thread per (n, group)
template <typename T>
__global__ void __launch_bounds__(256) encode_kernel(const T* __restrict__ W, uint32_t* __restrict__ packed,
                                                     T* __restrict__ scale, int N, int K, int nbit, int group,
                                                     int transpose) {
  const int npu = 32 / nbit;
  const int maxint = (1 << (nbit - 1)) - 1;
  const int G = ceil_div(K, group);
  const int rpg = ceil_div(group, npu);
  const int R = rpg * G;
  const long long total = (long long)N * G;
  for (long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x; idx < total;
       idx += (long long)gridDim.x * blockDim.x) {
    const int n = (int)(idx % N), g = (int)(idx / N);
    const T* w = W + (size_t)n * K + (size_t)g * group;
    const int cnt = min(group, K - g * group);
    float mx = 0.f;  // |.| >= 0, and at least one k is in range
    for (int k = 0; k < cnt; ++k) mx = fmaxf(mx, fabsf(Num<T>::load(w + k)));
    // all operands are dtype values held in float; f32 division of two f16 values rounds
    // innocuously (24 >= 2 * 11 + 2), so rnd() of the f32 quotient is the correctly rounded f16 one
    const float s = Num<T>::rnd(__fdiv_rn(fmaxf(mx, Num<T>::rnd(1e-4f)), (float)maxint));
    if (transpose)
      Num<T>::store(scale + (size_t)g * N + n, s);
    else
      Num<T>::store(scale + (size_t)n * G + g, s);
    for (int r = 0; r < rpg; ++r) {
      uint32_t word = 0;
      for (int t = 0; t < npu; ++t) {
        const int kk = r * npu + t;
        if (kk < cnt) {
          float q = Num<T>::rnd(__fdiv_rn(Num<T>::load(w + kk), s));
          q = Num<T>::rnd(__fadd_rn(q, (float)maxint));
          q = roundf(q);  // half away from zero (tir.round)
          q = fminf(fmaxf(q, 0.f), (float)(2 * maxint));
          word |= ((uint32_t)q) << (t * nbit);
        }
      }
      const int row = g * rpg + r;
      if (transpose)
        packed[(size_t)row * N + n] = word;
      else
        packed[(size_t)n * R + row] = word;
    }
  }
}
// asymmetric int4 (te_encode_asym, quantization.py:105-140): one thread per (n, group)
//   scale = dtype(max - min) / dtype(15);  q = clamp(int(round(dtype(dtype(W - min) / scale))), 0, 15)
//   f16: scale and min stored as they are; f32: bf16(scale) | bf16(min) << 16 with round-to-nearest-even by the
//   integer bias trick of _tir_f32x2_to_bf16x2_to_u32 (:15-25) - the weights are quantized with the f32 values
template <typename T>
__global__ void __launch_bounds__(256) encode_asym_kernel(const T* __restrict__ W, uint32_t* __restrict__ packed,
                                                          void* __restrict__ scale_out, T* __restrict__ min_out, int N, int K,
                                                          int transpose) {
  constexpr bool kF32 = sizeof(T) == 4;
  const int G = K / 32, R = K / 8;
  const long long total = (long long)N * G;
  for (long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x; idx < total;
       idx += (long long)gridDim.x * blockDim.x) {
    const int n = (int)(idx % N), g = (int)(idx / N);
    const T* w = W + (size_t)n * K + (size_t)g * 32;
    float mn = Num<T>::load(w), mx = mn;
    for (int k = 1; k < 32; ++k) {
      const float v = Num<T>::load(w + k);
      mn = fminf(mn, v);
      mx = fmaxf(mx, v);
    }
    const float s = Num<T>::rnd(__fdiv_rn(Num<T>::rnd(__fsub_rn(mx, mn)), 15.0f));
    const size_t ai = transpose ? (size_t)g * N + n : (size_t)n * G + g;
    if constexpr (kF32) {
      auto bf = [](float v) {
        uint32_t u = __float_as_uint(v);
        u += ((u >> 16) & 1u) + 0x7FFFu;
        return (u >> 16) & 0xFFFFu;
      };
      reinterpret_cast<uint32_t*>(scale_out)[ai] = bf(s) | (bf(mn) << 16);
    } else {
      Num<T>::store(reinterpret_cast<T*>(scale_out) + ai, s);
      Num<T>::store(min_out + ai, mn);
    }
    for (int r = 0; r < 4; ++r) {
      uint32_t word = 0;
      for (int t = 0; t < 8; ++t) {
        float q = Num<T>::rnd(__fdiv_rn(Num<T>::rnd(__fsub_rn(Num<T>::load(w + r * 8 + t), mn)), s));
        q = roundf(q);
        // a constant group divides by zero (undefined in the reference): NaN -> 0, +inf -> 15
        q = (q != q) ? 0.f : fminf(fmaxf(q, 0.f), 15.f);
        word |= ((uint32_t)q) << (t * 4);
      }
      const int row = g * 4 + r;
      if (transpose)
        packed[(size_t)row * N + n] = word;
      else
        packed[(size_t)n * R + row] = word;
    }
  }
}
}  // namespace

int encode_asym_launch(const void* W, void* packed, void* scale, void* zmin, int N, int K, int fmt, int transpose,
                       cudaStream_t stream) {
  const long long total = (long long)N * (K / 32);
  int blocks = (int)((total + 255) / 256);
  if (blocks > 148 * 32) blocks = 148 * 32;
  if (blocks < 1) blocks = 1;
  if (fmt == B200Q_Q4F32_ASYM)
    encode_asym_kernel<float><<<blocks, 256, 0, stream>>>(reinterpret_cast<const float*>(W), reinterpret_cast<uint32_t*>(packed),
                                                         scale, nullptr, N, K, transpose);
  else
    encode_asym_kernel<__half><<<blocks, 256, 0, stream>>>(reinterpret_cast<const __half*>(W), reinterpret_cast<uint32_t*>(packed),
                                                          scale, reinterpret_cast<__half*>(zmin), N, K, transpose);
  count_launch();
  return check_cuda(cudaGetLastError(), "encode_asym launch");
}

int encode_launch(const void* W, void* packed, void* scale, int N, int K, int fmt, int transpose,
                  cudaStream_t stream) {
  const long long total = (long long)N * n_groups(fmt, K);
  int blocks = (int)((total + 255) / 256);
  if (blocks > 148 * 32) blocks = 148 * 32;
  if (blocks < 1) blocks = 1;
  if (fmt_is_f32(fmt))
    encode_kernel<float><<<blocks, 256, 0, stream>>>(reinterpret_cast<const float*>(W),
                                                     reinterpret_cast<uint32_t*>(packed),
                                                     reinterpret_cast<float*>(scale), N, K, fmt_bits(fmt),
                                                     fmt_group(fmt), transpose);
  else
    encode_kernel<__half><<<blocks, 256, 0, stream>>>(reinterpret_cast<const __half*>(W),
                                                      reinterpret_cast<uint32_t*>(packed),
                                                      reinterpret_cast<__half*>(scale), N, K, fmt_bits(fmt),
                                                      fmt_group(fmt), transpose);
  count_launch();
  return check_cuda(cudaGetLastError(), "encode launch");
}
}  // namespace b200q
