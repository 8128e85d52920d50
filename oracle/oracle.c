/*
 * CPU oracle, C restatement  (TEST INFRASTRUCTURE ONLY - never linked into the product).
 *
 * Same formulas as oracle/quant_oracle.py, fast enough for full Llama-7B shapes and used as the
 * timed CPU baseline ("port") by bench.py.  Pinned bit-for-bit against the reference's own TVM
 * output through tests/golden/ (tests/test_oracle_golden.py).
 *
 * Reference lines followed (relative to /root/reference):
 *   decode      mlc_llm/transform/quantization.py:40-48, 213-230
 *   fused GEMV  mlc_llm/dispatch/dispatch_tir_operator_adreno.py:5-23   (dtype accumulate, init 0,
 *               k ascending); int3 spec  mlc_llm/dispatch/llama/main.py:5316-5343
 *   NT matmul   mlc_llm/transform/transpose_matmul.py:34-101, dispatch/llama/main.py:3884-3898
 *
 * Build: see oracle/Makefile  (gcc -O2 -fopenmp -mf16c -fexcess-precision=16: every half op
 * rounds to half, as LLVM / CUDA codegen of the reference's f16 TIR does).
 *
 * Loop order: the reference's unscheduled nest is  for n: for k  (a strided walk over packed[k/8][n]);
 * here k is the outer loop and n the inner, vectorisable one.  Each n still sees its own strictly
 * sequential k-ascending chain in `half`, so results are bit-identical to the reference nest.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

typedef _Float16 half;

typedef struct {
  int nbit, group, npu, maxint;
} fmt_t;

static fmt_t make_fmt(int nbit, int group) {
  fmt_t f = {nbit, group, 32 / nbit, (1 << (nbit - 1)) - 1};
  return f;
}

int orc_num_threads(void) {
#ifdef _OPENMP
  return omp_get_max_threads();
#else
  return 1;
#endif
}

void orc_set_threads(int n) {
#ifdef _OPENMP
  omp_set_num_threads(n);
#else
  (void)n;
#endif
}

/* ---- a2: decode ------------------------------------------------------------------------- */
#define DEF_DECODE(SUF, T)                                                                       \
  void orc_decode_##SUF(const uint32_t* packed, const T* scale, T* out, int K, int N, int nbit,   \
                        int group, int transpose_output) {                                       \
    fmt_t f = make_fmt(nbit, group);                                                             \
    uint32_t mask = (1u << nbit) - 1u;                                                           \
    _Pragma("omp parallel for schedule(static)") for (int k = 0; k < K; ++k) {                    \
      const uint32_t* prow = packed + (size_t)(k / f.npu) * N;                                   \
      const T* srow = scale + (size_t)(k / group) * N;                                           \
      uint32_t sh = (uint32_t)(k % f.npu) * nbit;                                                \
      for (int n = 0; n < N; ++n) {                                                              \
        T q = (T)((prow[n] >> sh) & mask) - (T)f.maxint;                                         \
        T w = q * srow[n];                                                                       \
        if (transpose_output)                                                                    \
          out[(size_t)n * K + k] = w;                                                            \
        else                                                                                     \
          out[(size_t)k * N + n] = w;                                                            \
      }                                                                                          \
    }                                                                                            \
  }
DEF_DECODE(f16, half)
DEF_DECODE(f32, float)

/* ---- a5: fused GEMV, reference semantics (accumulate in T) --------------------------------- */
/* epilogue ids: 0 none, 1 add, 2 multiply, 3 silu, 4 cast_f32, 5 bias, 6 bias_gelu (include/b200q.h) */
#define DEF_GEMV(SUF, T)                                                                         \
  void orc_gemv_ref_##SUF(const uint32_t* packed, const T* scale, const T* x, const T* ep0,      \
                          void* out, int K, int N, int nbit, int group, int epilogue,            \
                          int reassociate) {                                                     \
    fmt_t f = make_fmt(nbit, group);                                                             \
    uint32_t mask = (1u << nbit) - 1u;                                                           \
    const int BN = 512;                                                                          \
    _Pragma("omp parallel for schedule(dynamic, 1)") for (int n0 = 0; n0 < N; n0 += BN) {         \
      int nn = N - n0 < BN ? N - n0 : BN;                                                        \
      T acc[512];                                                                                \
      for (int j = 0; j < nn; ++j) acc[j] = (T)0;                                                \
      for (int k = 0; k < K; ++k) {                                                              \
        const uint32_t* prow = packed + (size_t)(k / f.npu) * N + n0;                            \
        const T* srow = scale + (size_t)(k / group) * N + n0;                                    \
        uint32_t sh = (uint32_t)(k % f.npu) * nbit;                                              \
        T xk = x[k];                                                                             \
        if (reassociate) {                                                                       \
          for (int j = 0; j < nn; ++j) {                                                         \
            T q = (T)((prow[j] >> sh) & mask) - (T)f.maxint;                                     \
            T t = xk * q;                                                                        \
            t = t * srow[j];                                                                     \
            acc[j] = acc[j] + t;                                                                 \
          }                                                                                      \
        } else {                                                                                 \
          for (int j = 0; j < nn; ++j) {                                                         \
            T q = (T)((prow[j] >> sh) & mask) - (T)f.maxint;                                     \
            T w = q * srow[j];                                                                   \
            T t = xk * w;                                                                        \
            acc[j] = acc[j] + t;                                                                 \
          }                                                                                      \
        }                                                                                        \
      }                                                                                          \
      for (int j = 0; j < nn; ++j) {                                                             \
        T y = acc[j];                                                                            \
        int n = n0 + j;                                                                          \
        switch (epilogue) {                                                                      \
          case 1: ((T*)out)[n] = ep0[n] + y; break;                                              \
          case 2: ((T*)out)[n] = ep0[n] * y; break;                                              \
          case 3: { /* TVM legalises sigmoid(y) to 1/(1+exp(-y)), every op in T */                  \
            T ny = (T)0 - y; T e = (T)expf((float)ny); T d = (T)1 + e; T s = (T)1 / d;           \
            ((T*)out)[n] = y * s; } break;                                                       \
          case 4: ((float*)out)[n] = (float)y; break;                                            \
          case 5: ((T*)out)[n] = y + ep0[n]; break;                                              \
          case 6: {                                                                              \
            T v = y + ep0[n];                                                                    \
            T t = v * (T)0.70710678118654757;                                                    \
            T e = (T)(float)erf((double)(float)t);                                               \
            T h2 = e * (T)0.5;                                                                   \
            T h = (T)0.5 + h2;                                                                   \
            ((T*)out)[n] = v * h;                                                                \
          } break;                                                                               \
          default: ((T*)out)[n] = y;                                                             \
        }                                                                                        \
      }                                                                                          \
    }                                                                                            \
  }
DEF_GEMV(f16, half)
DEF_GEMV(f32, float)

/* ---- f64 truth of the same dequantised weights + sum |x w| --------------------------------- */
#define DEF_GEMV64(SUF, T)                                                                       \
  void orc_gemv_f64_##SUF(const uint32_t* packed, const T* scale, const T* x, double* y,         \
                          double* abs_sum, int K, int N, int nbit, int group) {                  \
    fmt_t f = make_fmt(nbit, group);                                                             \
    uint32_t mask = (1u << nbit) - 1u;                                                           \
    _Pragma("omp parallel for schedule(static)") for (int n0 = 0; n0 < N; n0 += 256) {            \
      int nn = N - n0 < 256 ? N - n0 : 256;                                                      \
      double a[256], b[256];                                                                     \
      for (int j = 0; j < nn; ++j) a[j] = b[j] = 0.0;                                            \
      for (int k = 0; k < K; ++k) {                                                              \
        const uint32_t* prow = packed + (size_t)(k / f.npu) * N + n0;                            \
        const T* srow = scale + (size_t)(k / group) * N + n0;                                    \
        uint32_t sh = (uint32_t)(k % f.npu) * nbit;                                              \
        double xk = (double)x[k];                                                                \
        for (int j = 0; j < nn; ++j) {                                                           \
          T q = (T)((prow[j] >> sh) & mask) - (T)f.maxint;                                       \
          T w = q * srow[j];                                                                     \
          double t = xk * (double)w;                                                             \
          a[j] += t;                                                                             \
          b[j] += fabs(t);                                                                       \
        }                                                                                        \
      }                                                                                          \
      for (int j = 0; j < nn; ++j) { y[n0 + j] = a[j]; abs_sum[n0 + j] = b[j]; }                 \
    }                                                                                            \
  }
DEF_GEMV64(f16, half)
DEF_GEMV64(f32, float)

/* ---- asymmetric int4 (te_decode_asym, quantization.py:183-211; fused_decode4_matmul5*_before, ------------------
 * dispatch/llama/main.py:4397-4415 f32 / 4979-4997 f16):  w = T(q) * scale + min, product and sum each rounded to T
 * (the C / LLVM CPU builds of the reference do not contract them).  f32: scale_min is ONE uint32 per group and
 * column, bf16(scale) in the low half, bf16(min) in the high half (quantization.py:28-32). */
static inline float bf16lo(uint32_t u) { uint32_t v = u << 16; float f; memcpy(&f, &v, 4); return f; }
static inline float bf16hi(uint32_t u) { uint32_t v = u & 0xFFFF0000u; float f; memcpy(&f, &v, 4); return f; }

#define ASYM_SM_f16(G, N0, J) half s_ = scale[(size_t)(G) * N + (N0) + (J)], m_ = mn[(size_t)(G) * N + (N0) + (J)]
#define ASYM_SM_f32(G, N0, J) uint32_t u_ = ((const uint32_t*)scale)[(size_t)(G) * N + (N0) + (J)]; float s_ = bf16lo(u_), m_ = bf16hi(u_)

#define DEF_ASYM(SUF, T)                                                                             \
  void orc_decode_asym_##SUF(const uint32_t* packed, const void* scale_, const void* mn_, T* out, int K, int N,  \
                             int transpose_output) {                                                 \
    const T* scale = (const T*)scale_; const T* mn = (const T*)mn_; (void)mn;                         \
    _Pragma("omp parallel for schedule(static)") for (int k = 0; k < K; ++k) {                        \
      const uint32_t* prow = packed + (size_t)(k / 8) * N;                                            \
      uint32_t sh = (uint32_t)(k % 8) * 4;                                                            \
      for (int n = 0; n < N; ++n) {                                                                   \
        ASYM_SM_##SUF(k / 32, 0, n);                                                                  \
        T q = (T)((prow[n] >> sh) & 15u);                                                             \
        T w = q * s_;                                                                                 \
        w = w + m_;                                                                                   \
        if (transpose_output) out[(size_t)n * K + k] = w; else out[(size_t)k * N + n] = w;            \
      }                                                                                               \
    }                                                                                                 \
  }                                                                                                   \
  void orc_gemv_ref_asym_##SUF(const uint32_t* packed, const void* scale_, const void* mn_, const T* x, const T* ep0, \
                               void* out, int K, int N, int epilogue) {                               \
    const T* scale = (const T*)scale_; const T* mn = (const T*)mn_; (void)mn;                         \
    const int BN = 512;                                                                               \
    _Pragma("omp parallel for schedule(dynamic, 1)") for (int n0 = 0; n0 < N; n0 += BN) {             \
      int nn = N - n0 < BN ? N - n0 : BN;                                                             \
      T acc[512];                                                                                     \
      for (int j = 0; j < nn; ++j) acc[j] = (T)0;                                                     \
      for (int k = 0; k < K; ++k) {                                                                   \
        const uint32_t* prow = packed + (size_t)(k / 8) * N + n0;                                     \
        uint32_t sh = (uint32_t)(k % 8) * 4;                                                          \
        T xk = x[k];                                                                                  \
        for (int j = 0; j < nn; ++j) {                                                                \
          ASYM_SM_##SUF(k / 32, n0, j);                                                               \
          T q = (T)((prow[j] >> sh) & 15u);                                                           \
          T w = q * s_;                                                                               \
          w = w + m_;                                                                                 \
          T t = xk * w;                                                                               \
          acc[j] = acc[j] + t;                                                                        \
        }                                                                                             \
      }                                                                                               \
      for (int j = 0; j < nn; ++j) {                                                                  \
        T y = acc[j]; int n = n0 + j;                                                                 \
        switch (epilogue) {                                                                           \
          case 1: ((T*)out)[n] = ep0[n] + y; break;                                                   \
          case 2: ((T*)out)[n] = ep0[n] * y; break;                                                   \
          case 3: { T ny = (T)0 - y; T e = (T)expf((float)ny); T d = (T)1 + e; T sg = (T)1 / d; ((T*)out)[n] = y * sg; } break; \
          case 4: ((float*)out)[n] = (float)y; break;                                                 \
          default: ((T*)out)[n] = y;                                                                  \
        }                                                                                             \
      }                                                                                               \
    }                                                                                                 \
  }                                                                                                   \
  void orc_gemv_f64_asym_##SUF(const uint32_t* packed, const void* scale_, const void* mn_, const T* x, double* y, \
                               double* abs_sum, int K, int N) {                                       \
    const T* scale = (const T*)scale_; const T* mn = (const T*)mn_; (void)mn;                         \
    _Pragma("omp parallel for schedule(static)") for (int n0 = 0; n0 < N; n0 += 256) {                \
      int nn = N - n0 < 256 ? N - n0 : 256;                                                           \
      double a[256], b[256];                                                                          \
      for (int j = 0; j < nn; ++j) a[j] = b[j] = 0.0;                                                 \
      for (int k = 0; k < K; ++k) {                                                                   \
        const uint32_t* prow = packed + (size_t)(k / 8) * N + n0;                                     \
        uint32_t sh = (uint32_t)(k % 8) * 4;                                                          \
        double xk = (double)x[k];                                                                     \
        for (int j = 0; j < nn; ++j) {                                                                \
          ASYM_SM_##SUF(k / 32, n0, j);                                                               \
          T q = (T)((prow[j] >> sh) & 15u);                                                           \
          T w = q * s_;                                                                               \
          w = w + m_;                                                                                 \
          double t = xk * (double)w;                                                                  \
          a[j] += t; b[j] += fabs(t);                                                                 \
        }                                                                                             \
      }                                                                                               \
      for (int j = 0; j < nn; ++j) { y[n0 + j] = a[j]; abs_sum[n0 + j] = b[j]; }                      \
    }                                                                                                 \
  }
DEF_ASYM(f16, half)
DEF_ASYM(f32, float)

/* ---- f64 truth of the UNROUNDED weights (q - maxint) * s  (f16 formats) ----------------------
 * y, sum |x w| and the cancellation magnitude sum_k off_k |x_k| s of the tcgen05 decode program
 * (4-bit: off_k = 1024 + maxint for even positions of a word, 64 + maxint for odd ones; 8-bit: 1024 + maxint); see
 * quant_oracle.gemv_f64_unrounded / tolerance_group_scaled. */
void orc_gemv_f64u_f16(const uint32_t* packed, const half* scale, const half* x, double* y, double* abs_sum,
                       double* cancel_sum, int K, int N, int nbit, int group) {
  fmt_t f = make_fmt(nbit, group);
  uint32_t mask = (1u << nbit) - 1u;
#pragma omp parallel for schedule(static)
  for (int n0 = 0; n0 < N; n0 += 256) {
    int nn = N - n0 < 256 ? N - n0 : 256;
    double a[256], b[256], c[256];
    for (int j = 0; j < nn; ++j) a[j] = b[j] = c[j] = 0.0;
    for (int k = 0; k < K; ++k) {
      const uint32_t* prow = packed + (size_t)(k / f.npu) * N + n0;
      const half* srow = scale + (size_t)(k / group) * N + n0;
      uint32_t sh = (uint32_t)(k % f.npu) * nbit;
      double xk = (double)x[k];
      double off = ((nbit == 8 || (k % f.npu) % 2 == 0) ? 1024.0 : 64.0) + (double)f.maxint;
      for (int j = 0; j < nn; ++j) {
        double w = ((double)((prow[j] >> sh) & mask) - (double)f.maxint) * (double)srow[j];
        double t = xk * w;
        a[j] += t;
        b[j] += fabs(t);
        c[j] += fabs(xk) * off * fabs((double)srow[j]);
      }
    }
    /* int8 formulation of the 4-bit program (csrc/gemv_tc.cuh, kI8): a group's activations become integers at
     * 2^-21 of the group's largest magnitude -> error <= 2^-21 amax_g sum_{k in g} |w_k| = 2^-22 (2 amax_g sum |w|) */
    if (nbit == 4) {
      for (int k0 = 0; k0 < K; k0 += group) {
        int k1 = k0 + group < K ? k0 + group : K;
        double amax = 0.0;
        for (int k = k0; k < k1; ++k) { double v = fabs((double)x[k]); if (v > amax) amax = v; }
        for (int k = k0; k < k1; ++k) {
          const uint32_t* prow = packed + (size_t)(k / f.npu) * N + n0;
          const half* srow = scale + (size_t)(k / group) * N + n0;
          uint32_t sh = (uint32_t)(k % f.npu) * nbit;
          for (int j = 0; j < nn; ++j)
            c[j] += 2.0 * amax * fabs(((double)((prow[j] >> sh) & mask) - (double)f.maxint) * (double)srow[j]);
        }
      }
    }
    for (int j = 0; j < nn; ++j) { y[n0 + j] = a[j]; abs_sum[n0 + j] = b[j]; cancel_sum[n0 + j] = c[j]; }
  }
}

/* ---- a7: dense NT matmul, reference semantics (accumulate in half) -------------------------- */
void orc_nt_matmul_ref_f16(const half* X, const half* T_, half* Y, int M, int K, int N) {
#pragma omp parallel for collapse(2) schedule(static)
  for (int m = 0; m < M; ++m)
    for (int n = 0; n < N; ++n) {
      half acc = (half)0;
      const half* xr = X + (size_t)m * K;
      const half* tr = T_ + (size_t)n * K;
      for (int k = 0; k < K; ++k) {
        half t = xr[k] * tr[k];
        acc = acc + t;
      }
      Y[(size_t)m * N + n] = acc;
    }
}

/* f64 truth for Y = X * Wdeq with Wdeq given dense as [N,K] half, plus sum|x w| */
void orc_nt_matmul_f64_f16(const half* X, const half* T_, double* Y, double* A, int M, int K, int N) {
#pragma omp parallel for collapse(2) schedule(static)
  for (int m = 0; m < M; ++m)
    for (int n = 0; n < N; ++n) {
      double acc = 0, ab = 0;
      const half* xr = X + (size_t)m * K;
      const half* tr = T_ + (size_t)n * K;
      for (int k = 0; k < K; ++k) {
        double t = (double)xr[k] * (double)tr[k];
        acc += t;
        ab += fabs(t);
      }
      Y[(size_t)m * N + n] = acc;
      A[(size_t)m * N + n] = ab;
    }
}
