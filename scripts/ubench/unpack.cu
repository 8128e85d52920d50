// Microbenchmark 2: (a) do LOP3 and HFMA2 overlap?  (b) cycles per packed word of the exact q4 unpack
// (+ optional HMMA) in registers only, 1..4 warps per SMSP.
#include <cuda_fp16.h>
#include <cstdio>
#include <cstdint>
#define ITERS 2048
__device__ __forceinline__ uint32_t lop3_and_or(uint32_t a, uint32_t mask, uint32_t orv) {
  uint32_t r; asm("lop3.b32 %0, %1, %2, %3, 0xEA;" : "=r"(r) : "r"(a), "r"(mask), "r"(orv)); return r; }
__device__ __forceinline__ __half2 u2h(uint32_t v) { return *reinterpret_cast<__half2*>(&v); }
__device__ __forceinline__ uint32_t h2u(__half2 v) { return *reinterpret_cast<uint32_t*>(&v); }
__device__ __forceinline__ void unpack_q4(uint32_t w, __half2 s2, __half2 (&out)[4]) {
  const uint32_t kMagic = 0x64006400u;
  const __half2 kSubLo = u2h(0xE407E407u), kMul16 = u2h(0x2C002C00u), kSubHi = u2h(0xD470D470u);
  uint32_t w8 = w >> 8;
  out[0] = __hmul2(__hadd2(u2h(lop3_and_or(w, 0x000F000Fu, kMagic)), kSubLo), s2);
  out[1] = __hmul2(__hfma2(u2h(lop3_and_or(w, 0x00F000F0u, kMagic)), kMul16, kSubHi), s2);
  out[2] = __hmul2(__hadd2(u2h(lop3_and_or(w8, 0x000F000Fu, kMagic)), kSubLo), s2);
  out[3] = __hmul2(__hfma2(u2h(lop3_and_or(w8, 0x00F000F0u, kMagic)), kMul16, kSubHi), s2);
}
template <int MODE>
__global__ void k(uint32_t* out, uint32_t seed) {
  uint32_t w[4], r[8];
  for (int i = 0; i < 4; ++i) w[i] = seed * (threadIdx.x + 1) + i;
  for (int i = 0; i < 8; ++i) r[i] = seed + i + threadIdx.x;
  __half2 s2 = u2h(0x20002000u ^ (seed & 0xff));
  float acc[2][4] = {};
  uint32_t x0 = seed, x1 = seed + 1, c1 = seed ^ 0x3c003c00u, c2 = seed ^ 0x2c002c00u;
  long long t0 = clock64();
  for (int it = 0; it < ITERS; ++it) {
    if (MODE == 0) {  // 8 LOP3 + 8 HFMA2, independent
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        asm volatile("lop3.b32 %0, %0, %1, %2, 0xEA;" : "+r"(r[i]) : "r"(c1), "r"(c2));
        asm volatile("fma.rn.f16x2 %0, %0, %1, %2;" : "+r"(r[4 + i]) : "r"(c1), "r"(c2));
        asm volatile("lop3.b32 %0, %0, %1, %2, 0xEA;" : "+r"(r[i]) : "r"(c2), "r"(c1));
        asm volatile("fma.rn.f16x2 %0, %0, %1, %2;" : "+r"(r[4 + i]) : "r"(c2), "r"(c1));
      }
    } else {
      __half2 h[4][4];
#pragma unroll
      for (int j = 0; j < 4; ++j) unpack_q4(w[j], s2, h[j]);
      if (MODE == 1) {  // unpack only: fold results back (2 cheap ops per word)
#pragma unroll
        for (int j = 0; j < 4; ++j) w[j] += h2u(h[j][0]) + h2u(h[j][1]) + h2u(h[j][2]) + h2u(h[j][3]);
      } else {           // + 4 HMMA per 4 words, as in the kernel
#pragma unroll
        for (int m = 0; m < 2; ++m) {
          asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                       : "+f"(acc[m][0]), "+f"(acc[m][1]), "+f"(acc[m][2]), "+f"(acc[m][3])
                       : "r"(h2u(h[2 * m][0])), "r"(h2u(h[2 * m + 1][0])), "r"(h2u(h[2 * m][1])), "r"(h2u(h[2 * m + 1][1])), "r"(x0), "r"(x1));
          asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                       : "+f"(acc[m][0]), "+f"(acc[m][1]), "+f"(acc[m][2]), "+f"(acc[m][3])
                       : "r"(h2u(h[2 * m][2])), "r"(h2u(h[2 * m + 1][2])), "r"(h2u(h[2 * m][3])), "r"(h2u(h[2 * m + 1][3])), "r"(x0), "r"(x1));
        }
#pragma unroll
        for (int j = 0; j < 4; ++j) w[j] = w[j] * 1664525u + 1013904223u;  // new words (1 IMAD per word)
      }
    }
  }
  long long t1 = clock64();
  uint32_t a = 0;
  for (int i = 0; i < 8; ++i) a ^= r[i];
  for (int i = 0; i < 4; ++i) a ^= w[i] ^ __float_as_uint(acc[0][i]) ^ __float_as_uint(acc[1][i]);
  out[blockIdx.x * blockDim.x + threadIdx.x] = a;
  if (threadIdx.x == 0 && blockIdx.x == 0) out[1 << 20] = (uint32_t)(t1 - t0);
}
template <int MODE>
void run(const char* name, uint32_t* d, double units, const char* unit) {
  for (int threads = 128; threads <= 512; threads *= 2) {
    k<MODE><<<148, threads>>>(d, 12345); cudaDeviceSynchronize();
    k<MODE><<<148, threads>>>(d, 12345); cudaDeviceSynchronize();
    uint32_t cyc; cudaMemcpy(&cyc, d + (1 << 20), 4, cudaMemcpyDeviceToHost);
    double n = (double)ITERS * units * (threads / 128);
    printf("%-28s warps/SMSP=%d  %.2f cycles per %s per SMSP (per-warp-instr granularity)\n", name, threads / 128, cyc / n, unit);
  }
}
int main() {
  uint32_t* d; cudaMalloc(&d, ((1 << 20) + 16) * 4);
  run<0>("8 LOP3 + 8 HFMA2 mixed", d, 16, "warp-instr");
  run<1>("q4 unpack only", d, 4, "word(lane)");
  run<2>("q4 unpack + HMMA", d, 4, "word(lane)");
  printf("%s\n", cudaGetErrorString(cudaGetLastError()));
}
