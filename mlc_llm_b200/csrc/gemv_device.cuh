// Device-side building blocks shared by the single-op GEMV kernel (gemv.cu) and the persistent
// decode-program kernel (gemv_prog.cu): shared-memory loads, the mma.sync contraction of one packed
// word row, and the activation staging.  See gemv.cu for the design notes and lane mapping.
#pragma once
#include "common.cuh"
#include "ptx.cuh"

namespace b200q {
namespace gemvdev {

constexpr int kPanelCols = 128;  // columns of one panel (= partial-sum slot width)
constexpr int kBoxCols = 32;     // one TMA word box: 32 words = 128-byte rows

template <int FMT>
struct XLayout {
  using F = Fmt<FMT>;
  static constexpr bool kF32 = fmt_is_f32_c(FMT);
  // permuted activation bytes per packed word row
  static constexpr int kXRowBytes = kF32 ? 32 : F::kQuads * 8;
};

__device__ __forceinline__ uint4 lds128(uint32_t a) {
  uint4 r;
  asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "r"(a));
  return r;
}
__device__ __forceinline__ uint2 lds64(uint32_t a) {
  uint2 r;
  asm volatile("ld.shared.v2.u32 {%0,%1}, [%2];" : "=r"(r.x), "=r"(r.y) : "r"(a));
  return r;
}
__device__ __forceinline__ float ld_cg_f32(const float* p) {
  float v;
  asm volatile("ld.global.cg.f32 %0, [%1];" : "=f"(v) : "l"(p));
  return v;
}

__device__ __forceinline__ void mma16816(float (&d)[4], uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3,
                                         uint32_t b0, uint32_t b1) {
  asm volatile(
      "mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, "
      "{%0,%1,%2,%3};"
      : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
      : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}

// ---- f16 formats: one packed row = 4 words (columns 4c..4c+3) of one lane -------------------------
// acc[m][*]: MMA m covers columns (4c + 2m) [A rows 0-7] and (4c + 2m + 1) [A rows 8-15] of the
// lane's column quad c; the four lanes of a quad own different rows (different k).
__device__ __forceinline__ void split_scales(uint2 sc, __half2 (&s2)[4]) {
  s2[0] = u32_as_h2(__byte_perm(sc.x, 0, 0x1010));
  s2[1] = u32_as_h2(__byte_perm(sc.x, 0, 0x3232));
  s2[2] = u32_as_h2(__byte_perm(sc.y, 0, 0x1010));
  s2[3] = u32_as_h2(__byte_perm(sc.y, 0, 0x3232));
}
// asymmetric f16: the same contraction on the weights half(half(q * s) + m)
__device__ __forceinline__ void compute_row_f16_asym(uint4 w, const __half2 (&s2)[4], const __half2 (&m2)[4], uint32_t xr,
                                                     float (&acc)[2][4]) {
  const uint32_t wv[4] = {w.x, w.y, w.z, w.w};
  uint4 xv = lds128(xr);  // (x0,x4)(x1,x5)(x2,x6)(x3,x7)
  __half2 h[4][4];
#pragma unroll
  for (int j = 0; j < 4; ++j) unpack_q4_asym(wv[j], s2[j], m2[j], h[j]);
#pragma unroll
  for (int m = 0; m < 2; ++m) {
    mma16816(acc[m], h2_as_u32(h[2 * m][0]), h2_as_u32(h[2 * m + 1][0]), h2_as_u32(h[2 * m][1]), h2_as_u32(h[2 * m + 1][1]),
             xv.x, xv.y);
    mma16816(acc[m], h2_as_u32(h[2 * m][2]), h2_as_u32(h[2 * m + 1][2]), h2_as_u32(h[2 * m][3]), h2_as_u32(h[2 * m + 1][3]),
             xv.z, xv.w);
  }
}

template <int FMT>
__device__ __forceinline__ void compute_row_f16(uint4 w, const __half2 (&s2)[4], uint32_t xr, float (&acc)[2][4]) {
  const uint32_t wv[4] = {w.x, w.y, w.z, w.w};
  if constexpr (FMT == B200Q_Q4F16_0) {
    uint4 xv = lds128(xr);  // (x0,x4)(x1,x5)(x2,x6)(x3,x7)
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
    // (x0,x5)(x1,x6) | (x2,x7)(x3,x8) | (x4,x9)(0,0); rows are 24 bytes apart: 8-byte alignment
    uint2 x01 = lds64(xr), x23 = lds64(xr + 8), xb = lds64(xr + 16);
    __half2 h[4][5];
#pragma unroll
    for (int j = 0; j < 4; ++j) unpack_q3(wv[j], s2[j], h[j]);
#pragma unroll
    for (int m = 0; m < 2; ++m) {
      mma16816(acc[m], h2_as_u32(h[2 * m][0]), h2_as_u32(h[2 * m + 1][0]), h2_as_u32(h[2 * m][1]),
               h2_as_u32(h[2 * m + 1][1]), x01.x, x01.y);
      mma16816(acc[m], h2_as_u32(h[2 * m][2]), h2_as_u32(h[2 * m + 1][2]), h2_as_u32(h[2 * m][3]),
               h2_as_u32(h[2 * m + 1][3]), x23.x, x23.y);
      mma16816(acc[m], h2_as_u32(h[2 * m][4]), h2_as_u32(h[2 * m + 1][4]), 0u, 0u, xb.x, xb.y);
    }
  } else {  // Q8: (x0,x1)(x2,x3)
    uint2 xv = lds64(xr);
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

// ---- q4f32_0: f32 weights  fl32(f32(q - 7) * s), FFMA into four column accumulators ----------------
// (q | 0x4B000000) is the f32 2^23 + q; subtracting 2^23 + 7 gives q - 7 exactly.
__device__ __forceinline__ float nib_to_f32(uint32_t w, int t) {
  uint32_t q = (w >> (4 * t)) & 0xFu;
  return __uint_as_float(q | 0x4B000000u) - 8388615.0f;
}
__device__ __forceinline__ void compute_row_f32(uint4 w, uint4 sc, uint32_t xr, float (&acc)[4]) {
  const float s[4] = {__uint_as_float(sc.x), __uint_as_float(sc.y), __uint_as_float(sc.z), __uint_as_float(sc.w)};
  const uint32_t wv[4] = {w.x, w.y, w.z, w.w};
  uint4 xa = lds128(xr), xb = lds128(xr + 16);
  const float xv[8] = {__uint_as_float(xa.x), __uint_as_float(xa.y), __uint_as_float(xa.z), __uint_as_float(xa.w),
                       __uint_as_float(xb.x), __uint_as_float(xb.y), __uint_as_float(xb.z), __uint_as_float(xb.w)};
#pragma unroll
  for (int j = 0; j < 4; ++j) {
#pragma unroll
    for (int t = 0; t < 8; ++t) {
      float wt = __fmul_rn(nib_to_f32(wv[j], t), s[j]);  // the reference's single rounding
      acc[j] = fmaf(xv[t], wt, acc[j]);
    }
  }
}

// asymmetric f32: sc carries bf16(scale) | bf16(min) << 16 of the four columns
__device__ __forceinline__ void compute_row_f32_asym(uint4 w, uint4 sc, uint32_t xr, float (&acc)[4]) {
  const uint32_t sm[4] = {sc.x, sc.y, sc.z, sc.w};
  const uint32_t wv[4] = {w.x, w.y, w.z, w.w};
  uint4 xa = lds128(xr), xb = lds128(xr + 16);
  const float xv[8] = {__uint_as_float(xa.x), __uint_as_float(xa.y), __uint_as_float(xa.z), __uint_as_float(xa.w),
                       __uint_as_float(xb.x), __uint_as_float(xb.y), __uint_as_float(xb.z), __uint_as_float(xb.w)};
#pragma unroll
  for (int j = 0; j < 4; ++j) {
#pragma unroll
    for (int t = 0; t < 8; ++t) acc[j] = fmaf(xv[t], asym_f32_weight(wv[j], t, sm[j]), acc[j]);
  }
}

// ---- activation staging: the whole x (optionally x * x_mul) into shared memory ---------------------
// f16 formats: word row r -> xs[r][kXRowBytes] in the pair order the unpack produces.
// one packed word row of activations (kNpu values, zero padded to 12) -> the pair order the unpack produces
template <int FMT>
__device__ __forceinline__ void permute_x_row(const __half (&v)[12], __half (&o)[12]) {
  if constexpr (FMT == B200Q_Q4F16_0 || FMT == B200Q_Q4F16_ASYM) {
    o[0] = v[0]; o[1] = v[4]; o[2] = v[1]; o[3] = v[5];
    o[4] = v[2]; o[5] = v[6]; o[6] = v[3]; o[7] = v[7];
  } else if constexpr (FMT == B200Q_Q3F16_0) {
    o[0] = v[0]; o[1] = v[5]; o[2] = v[1]; o[3] = v[6];
    o[4] = v[2]; o[5] = v[7]; o[6] = v[3]; o[7] = v[8];
    o[8] = v[4]; o[9] = v[9]; o[10] = __ushort_as_half(0); o[11] = __ushort_as_half(0);
  } else {
    o[0] = v[0]; o[1] = v[1]; o[2] = v[2]; o[3] = v[3];
  }
}

// CG: read the activation with ld.global.cg (L2 only): required when another CTA of the SAME launch
// produced it (the persistent decode-program kernel); L1 is not coherent across SMs.
template <bool CG, typename T>
__device__ __forceinline__ T ldx(const T* p) {
  if constexpr (CG) return __ldcg(p);
  else return *p;
}
template <int FMT, int NT, bool CG = false>
__device__ __forceinline__ void stage_x(unsigned char* xs, const void* xg, const void* xmul, int K, int rows,
                                        int tid = threadIdx.x) {
  using S = XLayout<FMT>;
  using F = Fmt<FMT>;
  if constexpr (S::kF32) {
    const float* x = reinterpret_cast<const float*>(xg);
    const float* m = reinterpret_cast<const float*>(xmul);
    float* dst = reinterpret_cast<float*>(xs);
    const int n = rows * 8;
    for (int k = tid; k < n; k += NT) {
      float v = 0.f;
      if (k < K) v = m ? __fmul_rn(ldx<CG>(x + k), ldx<CG>(m + k)) : ldx<CG>(x + k);
      dst[k] = v;
    }
  } else {
    const __half* x = reinterpret_cast<const __half*>(xg);
    const __half* m = reinterpret_cast<const __half*>(xmul);
    constexpr int XB = S::kXRowBytes;
    const bool vec_ok = (F::kNpu == 8) && ((reinterpret_cast<uintptr_t>(x) & 15) == 0) &&
                        (m == nullptr || (reinterpret_cast<uintptr_t>(m) & 15) == 0);
    for (int lr = tid; lr < rows; lr += NT) {
      const int k0 = lr * F::kNpu;
      __align__(16) __half v[12];
#pragma unroll
      for (int t = 0; t < 12; ++t) v[t] = __ushort_as_half(0);
      if (vec_ok && k0 + 8 <= K) {
        *reinterpret_cast<uint4*>(v) = ldx<CG>(reinterpret_cast<const uint4*>(x + k0));
        if (m) {
          __align__(16) __half mv[8];
          *reinterpret_cast<uint4*>(mv) = ldx<CG>(reinterpret_cast<const uint4*>(m + k0));
#pragma unroll
          for (int t = 0; t < 8; ++t) v[t] = __hmul(v[t], mv[t]);
        }
      } else {
#pragma unroll
        for (int t = 0; t < F::kNpu; ++t)
          if (k0 + t < K) v[t] = m ? __hmul(ldx<CG>(x + k0 + t), ldx<CG>(m + k0 + t)) : ldx<CG>(x + k0 + t);
      }
      __align__(8) __half o[12];
      permute_x_row<FMT>(v, o);
      unsigned char* d = xs + (size_t)lr * XB;
#pragma unroll
      for (int t = 0; t < XB / 8; ++t) reinterpret_cast<uint2*>(d)[t] = reinterpret_cast<const uint2*>(o)[t];
    }
  }
}


}  // namespace gemvdev
}  // namespace b200q
