// Shared device/host helpers for the b200q kernels (sm_100a only).
#pragma once
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/b200q.h"

#if defined(__CUDA_ARCH__) && (__CUDA_ARCH__ < 1000)
#error "b200q kernels are written for sm_100a (B200) only"
#endif

namespace b200q {

// ---- host-side error plumbing (api.cu) -------------------------------------------------------
int set_error(int code, const char* fmt, ...);
int check_cuda(cudaError_t e, const char* what);
void count_launch(int n = 1);

// ---- format traits ---------------------------------------------------------------------------
// Layout facts of mlc_llm/transform/quantization.py:142-177 (te_encode_sym) per format.
template <int FMT>
struct Fmt;

template <>
struct Fmt<B200Q_Q4F16_0> {
  static constexpr int kBits = 4, kGroup = 32, kNpu = 8, kMaxInt = 7;
  static constexpr int kRowsPerGroup = 4;  // words of one column per group
  static constexpr int kQuads = 2;         // MMA k-quads (4 values) per word
  using scale_t = __half;
};
template <>
struct Fmt<B200Q_Q3F16_0> {
  static constexpr int kBits = 3, kGroup = 40, kNpu = 10, kMaxInt = 3;
  static constexpr int kRowsPerGroup = 4;
  static constexpr int kQuads = 3;  // 10 values -> 5 pairs, padded to 6 (last pair zero)
  using scale_t = __half;
};
template <>
struct Fmt<B200Q_Q8F16_0> {
  static constexpr int kBits = 8, kGroup = 32, kNpu = 4, kMaxInt = 127;
  static constexpr int kRowsPerGroup = 8;
  static constexpr int kQuads = 1;
  using scale_t = __half;
};
template <>
struct Fmt<B200Q_Q4F32_0> {
  static constexpr int kBits = 4, kGroup = 32, kNpu = 8, kMaxInt = 7;
  static constexpr int kRowsPerGroup = 4;
  static constexpr int kQuads = 2;
  using scale_t = float;
};

template <>
struct Fmt<B200Q_Q4F16_ASYM> {
  static constexpr int kBits = 4, kGroup = 32, kNpu = 8, kMaxInt = 0;
  static constexpr int kRowsPerGroup = 4;
  static constexpr int kQuads = 2;
  using scale_t = __half;
};
template <>
struct Fmt<B200Q_Q4F32_ASYM> {
  static constexpr int kBits = 4, kGroup = 32, kNpu = 8, kMaxInt = 0;
  static constexpr int kRowsPerGroup = 4;
  static constexpr int kQuads = 2;
  using scale_t = float;  // bit pattern: bf16(scale) | bf16(min) << 16
};
__host__ __device__ constexpr bool fmt_is_asym_c(int fmt) { return fmt == B200Q_Q4F16_ASYM || fmt == B200Q_Q4F32_ASYM; }
__host__ __device__ constexpr bool fmt_is_f32_c(int fmt) { return fmt == B200Q_Q4F32_0 || fmt == B200Q_Q4F32_ASYM; }
// the f16 asymmetric scheme is the only one with a separate min tensor
__host__ __device__ constexpr bool fmt_has_min_c(int fmt) { return fmt == B200Q_Q4F16_ASYM; }
__host__ __device__ inline int fmt_is_asym(int fmt) { return fmt == B200Q_Q4F16_ASYM || fmt == B200Q_Q4F32_ASYM; }

__host__ __device__ inline int fmt_group(int fmt) { return (fmt == B200Q_Q3F16_0 || fmt == B200Q_Q3F16_U16) ? 40 : 32; }
__host__ __device__ inline int fmt_npu(int fmt) {  // values per storage word
  return fmt == B200Q_Q3F16_0 ? 10 : (fmt == B200Q_Q3F16_U16 ? 5 : (fmt == B200Q_Q8F16_0 ? 4 : 8));
}
__host__ __device__ inline int fmt_bits(int fmt) {
  return (fmt == B200Q_Q3F16_0 || fmt == B200Q_Q3F16_U16) ? 3 : (fmt == B200Q_Q8F16_0 ? 8 : 4);
}
__host__ __device__ inline int fmt_word_bytes(int fmt) { return fmt == B200Q_Q3F16_U16 ? 2 : 4; }
__host__ __device__ inline int fmt_is_f32(int fmt) { return fmt == B200Q_Q4F32_0 || fmt == B200Q_Q4F32_ASYM || fmt == B200Q_FP4F32_0; }
__host__ __device__ inline int ceil_div(int a, int b) { return (a + b - 1) / b; }
__host__ __device__ inline int n_groups(int fmt, int K) { return ceil_div(K, fmt_group(fmt)); }
__host__ __device__ inline int n_words(int fmt, int K) {
  return ceil_div(fmt_group(fmt), fmt_npu(fmt)) * n_groups(fmt, K);
}

#ifdef __CUDACC__
// ---- streaming loads (weights are read exactly once per GEMV: keep them out of L1, and mark
// them evict-first in L2 so the small re-used vectors (x, outputs, scales of the next op) stay) ----
__device__ __forceinline__ uint64_t l2_evict_first_policy() {
  uint64_t pol;
  asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(pol));
  return pol;
}
__device__ __forceinline__ uint4 ldg_stream_v4(const void* p, uint64_t pol) {
  uint4 r;
  asm volatile("ld.global.nc.L1::no_allocate.L2::cache_hint.v4.u32 {%0,%1,%2,%3}, [%4], %5;"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
               : "l"(p), "l"(pol));
  return r;
}
__device__ __forceinline__ uint2 ldg_stream_v2(const void* p, uint64_t pol) {
  uint2 r;
  asm volatile("ld.global.nc.L1::no_allocate.L2::cache_hint.v2.u32 {%0,%1}, [%2], %3;"
               : "=r"(r.x), "=r"(r.y)
               : "l"(p), "l"(pol));
  return r;
}
__device__ __forceinline__ uint4 ldg_stream_v4(const void* p) {
  uint4 r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
               : "l"(p));
  return r;
}
__device__ __forceinline__ uint32_t ldg_stream_u32(const void* p) {
  uint32_t r;
  asm volatile("ld.global.nc.L1::no_allocate.u32 %0, [%1];" : "=r"(r) : "l"(p));
  return r;
}

// ---- programmatic dependent launch ------------------------------------------------------------
__device__ __forceinline__ void pdl_launch_dependents() {
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
}
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }

// ---- bit-exact unpack to f16 ------------------------------------------------------------------
// Each function returns the dequantized weights  half((q - maxint) * s)  of ONE packed word as
// half2 pairs.  (q - maxint) is produced exactly (magic-number trick: 0x6400 | q is the f16
// 1024 + q; subtracting 1024 + maxint is exact), so the single rounding is the multiply by the
// scale, exactly the reference's  (cast(q) - maxint) * scale  (quantization.py:40-48,222-224).
__device__ __forceinline__ uint32_t lop3_and_or(uint32_t a, uint32_t mask, uint32_t orv) {
  uint32_t r;
  asm("lop3.b32 %0, %1, %2, %3, 0xEA;" : "=r"(r) : "r"(a), "r"(mask), "r"(orv));  // (a & b) | c
  return r;
}
__device__ __forceinline__ __half2 u32_as_h2(uint32_t v) { return *reinterpret_cast<__half2*>(&v); }
__device__ __forceinline__ uint32_t h2_as_u32(__half2 v) { return *reinterpret_cast<uint32_t*>(&v); }

// int4: out[0]=(v0,v4) out[1]=(v1,v5) out[2]=(v2,v6) out[3]=(v3,v7), each already times s2
__device__ __forceinline__ void unpack_q4(uint32_t w, __half2 s2, __half2 (&out)[4]) {
  const uint32_t kMagic = 0x64006400u;
  const __half2 kSubLo = u32_as_h2(0xE407E407u);   // -1031 = -(1024 + 7)
  const __half2 kMul16 = u32_as_h2(0x2C002C00u);   // 1/16
  const __half2 kSubHi = u32_as_h2(0xD470D470u);   // -71 = -(64 + 7)
  uint32_t w8 = w >> 8;
  __half2 a = u32_as_h2(lop3_and_or(w, 0x000F000Fu, kMagic));   // 1024 + v0 | 1024 + v4
  __half2 b = u32_as_h2(lop3_and_or(w, 0x00F000F0u, kMagic));   // 1024 + 16 v1 | ...v5
  __half2 c = u32_as_h2(lop3_and_or(w8, 0x000F000Fu, kMagic));  // v2 | v6
  __half2 d = u32_as_h2(lop3_and_or(w8, 0x00F000F0u, kMagic));  // v3 | v7
  out[0] = __hmul2(__hadd2(a, kSubLo), s2);
  out[1] = __hmul2(__hfma2(b, kMul16, kSubHi), s2);
  out[2] = __hmul2(__hadd2(c, kSubLo), s2);
  out[3] = __hmul2(__hfma2(d, kMul16, kSubHi), s2);
}

// asymmetric int4: out[i] = half(half(q * s) + m), the reference's two roundings (no fma contraction:
// __hmul2_rn / __hadd2_rn are never fused).  q itself is exact: (1024 + q) - 1024, (1024 + 16 q) / 16 - 64.
__device__ __forceinline__ void unpack_q4_asym(uint32_t w, __half2 s2, __half2 m2, __half2 (&out)[4]) {
  const uint32_t kMagic = 0x64006400u;
  const __half2 kSubLo = u32_as_h2(0xE400E400u);   // -1024
  const __half2 kMul16 = u32_as_h2(0x2C002C00u);   // 1/16
  const __half2 kSubHi = u32_as_h2(0xD400D400u);   // -64
  uint32_t w8 = w >> 8;
  __half2 a = __hadd2(u32_as_h2(lop3_and_or(w, 0x000F000Fu, kMagic)), kSubLo);
  __half2 b = __hfma2(u32_as_h2(lop3_and_or(w, 0x00F000F0u, kMagic)), kMul16, kSubHi);
  __half2 c = __hadd2(u32_as_h2(lop3_and_or(w8, 0x000F000Fu, kMagic)), kSubLo);
  __half2 d = __hfma2(u32_as_h2(lop3_and_or(w8, 0x00F000F0u, kMagic)), kMul16, kSubHi);
  out[0] = __hadd2_rn(__hmul2_rn(a, s2), m2);
  out[1] = __hadd2_rn(__hmul2_rn(b, s2), m2);
  out[2] = __hadd2_rn(__hmul2_rn(c, s2), m2);
  out[3] = __hadd2_rn(__hmul2_rn(d, s2), m2);
}
// f32 asymmetric: scale / min from the packed bf16 pair; w = fl(fl(q * s) + m)
__device__ __forceinline__ float asym_f32_weight(uint32_t w, int t, uint32_t sm) {
  const float q = (float)((w >> (4 * t)) & 0xFu);
  return __fadd_rn(__fmul_rn(q, __uint_as_float(sm << 16)), __uint_as_float(sm & 0xFFFF0000u));
}

// int3, 10 values in bits [0,30): out[i] = (v_i, v_{i+5}), i = 0..4
__device__ __forceinline__ void unpack_q3(uint32_t w, __half2 s2, __half2 (&out)[5]) {
  const uint32_t kMagic = 0x64006400u;
  uint32_t p = __byte_perm(w, w >> 15, 0x5410);  // lo16 = v0..v4 (+1 stray bit), hi16 = v5..v9
  uint32_t q = p >> 9;                           // v3 | v8 at bits 0-2 of each half, v4 | v9 at 3-5
  const __half2 kSub1 = u32_as_h2(0xE403E403u);   // -(1024 + 3) = -1027
  const __half2 kMul8 = u32_as_h2(0x30003000u);   // 1/8
  const __half2 kSub8 = u32_as_h2(0xD818D818u);   // -(128 + 3) = -131
  const __half2 kMul64 = u32_as_h2(0x24002400u);  // 1/64
  const __half2 kSub64 = u32_as_h2(0xCCC0CCC0u);  // -(16 + 3) = -19
  __half2 a = u32_as_h2(lop3_and_or(p, 0x00070007u, kMagic));
  __half2 b = u32_as_h2(lop3_and_or(p, 0x00380038u, kMagic));
  __half2 c = u32_as_h2(lop3_and_or(p, 0x01C001C0u, kMagic));
  __half2 d = u32_as_h2(lop3_and_or(q, 0x00070007u, kMagic));
  __half2 e = u32_as_h2(lop3_and_or(q, 0x00380038u, kMagic));
  out[0] = __hmul2(__hadd2(a, kSub1), s2);
  out[1] = __hmul2(__hfma2(b, kMul8, kSub8), s2);
  out[2] = __hmul2(__hfma2(c, kMul64, kSub64), s2);
  out[3] = __hmul2(__hadd2(d, kSub1), s2);
  out[4] = __hmul2(__hfma2(e, kMul8, kSub8), s2);
}

// int8, 4 values: out[0] = (v0, v1), out[1] = (v2, v3)
__device__ __forceinline__ void unpack_q8(uint32_t w, __half2 s2, __half2 (&out)[2]) {
  const __half2 kSub = u32_as_h2(0xE47FE47Fu);  // -(1024 + 127) = -1151
  __half2 a = u32_as_h2(__byte_perm(w, 0x64646464u, 0x4140));  // bytes: b0,0x64,b1,0x64
  __half2 b = u32_as_h2(__byte_perm(w, 0x64646464u, 0x4342));
  out[0] = __hmul2(__hadd2(a, kSub), s2);
  out[1] = __hmul2(__hadd2(b, kSub), s2);
}

// ---- epilogues, op for op in f16 like the reference PrimFuncs ----------------------------------
// (dispatch_tir_operator_adreno.py:86-91,150-152,216-228; gpt_neox/dolly_v2_3b_mod.py:41-46,204-251)
__device__ __forceinline__ __half ep_f16(float acc, int ep, const __half* ep0, int n) {
  __half y = __float2half_rn(acc);
  switch (ep) {
    case B200Q_EP_ADD: return __hadd(ep0[n], y);
    case B200Q_EP_MULTIPLY: return __hmul(ep0[n], y);
    case B200Q_EP_SILU: {
      __half e = __float2half_rn(expf(__half2float(__hneg(y))));
      __half d = __hadd(__float2half_rn(1.0f), e);
      __half s = __float2half_rn(1.0f / __half2float(d));
      return __hmul(y, s);
    }
    case B200Q_EP_BIAS: return __hadd(y, ep0[n]);
    case B200Q_EP_BIAS_GELU: {
      __half v = __hadd(y, ep0[n]);
      __half t = __hmul(v, __float2half_rn(0.70710678118654757f));
      __half e = __float2half_rn(erff(__half2float(t)));
      __half h = __hadd(__float2half_rn(0.5f), __hmul(e, __float2half_rn(0.5f)));
      return __hmul(v, h);
    }
    default: return y;
  }
}
__device__ __forceinline__ float ep_f32(float y, int ep, const float* ep0, int n) {
  switch (ep) {
    case B200Q_EP_ADD: return ep0[n] + y;
    case B200Q_EP_MULTIPLY: return ep0[n] * y;
    case B200Q_EP_SILU: return y * (1.0f / (1.0f + expf(-y)));
    case B200Q_EP_BIAS: return y + ep0[n];
    case B200Q_EP_BIAS_GELU: {
      float v = y + ep0[n];
      float t = v * 0.70710678118654757f;
      return v * (0.5f + erff(t) * 0.5f);
    }
    default: return y;
  }
}
#endif  // __CUDACC__

}  // namespace b200q
