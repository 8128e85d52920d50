// Microbenchmark 4: does HMMA.16816 overlap with ALU (LOP3/SHF), FMA-pipe (HFMA2/FFMA/IMAD) work?
// per iteration: 2 independent HMMA + N independent ops of one kind; cycles per iteration per SMSP.
#include <cuda_fp16.h>
#include <cstdio>
#include <cstdint>
#define ITERS 2048
template <int KIND, int N>
__global__ void k(uint32_t* out, uint32_t seed) {
  uint32_t r[16]; float f[8];
  for (int i = 0; i < 16; ++i) r[i] = seed + i + threadIdx.x;
  for (int i = 0; i < 8; ++i) f[i] = (float)i;
  uint32_t c1 = seed ^ 0x3c003c00u, c2 = seed ^ 0x2c002c00u;
  long long t0 = clock64();
#pragma unroll 1
  for (int it = 0; it < ITERS; ++it) {
    asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+f"(f[0]), "+f"(f[1]), "+f"(f[2]), "+f"(f[3]) : "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(c1), "r"(c2));
#pragma unroll
    for (int i = 0; i < N / 2; ++i) {
      if (KIND == 0) asm volatile("lop3.b32 %0, %0, %1, %2, 0xEA;" : "+r"(r[8 + (i & 7)]) : "r"(c1), "r"(c2));
      if (KIND == 1) asm volatile("fma.rn.f16x2 %0, %0, %1, %2;" : "+r"(r[8 + (i & 7)]) : "r"(c1), "r"(c2));
      if (KIND == 2) asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+r"(r[8 + (i & 7)]) : "r"(c1), "r"(c2));
      if (KIND == 3) asm volatile("shf.r.clamp.b32 %0, %0, %1, 3;" : "+r"(r[8 + (i & 7)]) : "r"(c1));
    }
    asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+f"(f[4]), "+f"(f[5]), "+f"(f[6]), "+f"(f[7]) : "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]), "r"(c1), "r"(c2));
#pragma unroll
    for (int i = N / 2; i < N; ++i) {
      if (KIND == 0) asm volatile("lop3.b32 %0, %0, %1, %2, 0xEA;" : "+r"(r[8 + (i & 7)]) : "r"(c1), "r"(c2));
      if (KIND == 1) asm volatile("fma.rn.f16x2 %0, %0, %1, %2;" : "+r"(r[8 + (i & 7)]) : "r"(c1), "r"(c2));
      if (KIND == 2) asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+r"(r[8 + (i & 7)]) : "r"(c1), "r"(c2));
      if (KIND == 3) asm volatile("shf.r.clamp.b32 %0, %0, %1, 3;" : "+r"(r[8 + (i & 7)]) : "r"(c1));
    }
  }
  long long t1 = clock64();
  uint32_t a = 0;
  for (int i = 0; i < 16; ++i) a ^= r[i];
  for (int i = 0; i < 8; ++i) a ^= __float_as_uint(f[i]);
  out[blockIdx.x * blockDim.x + threadIdx.x] = a;
  if (threadIdx.x == 0 && blockIdx.x == 0) out[1 << 20] = (uint32_t)(t1 - t0);
}
template <int KIND, int N>
void run(const char* name, uint32_t* d) {
  for (int threads = 256; threads <= 512; threads *= 2) {
    k<KIND, N><<<148, threads>>>(d, 12345); cudaDeviceSynchronize();
    k<KIND, N><<<148, threads>>>(d, 12345); cudaDeviceSynchronize();
    uint32_t cyc; cudaMemcpy(&cyc, d + (1 << 20), 4, cudaMemcpyDeviceToHost);
    printf("2 HMMA + %2d %-6s warps/SMSP=%d  %.2f cycles per iteration per warp-slot (=per SMSP per warp-iteration)\n", N, name, threads / 128,
           cyc / ((double)ITERS * (threads / 128)));
  }
}
int main() {
  uint32_t* d; cudaMalloc(&d, ((1 << 20) + 16) * 4);
  run<0, 0>("none", d);
  run<0, 4>("LOP3", d); run<0, 8>("LOP3", d); run<0, 16>("LOP3", d);
  run<1, 4>("HFMA2", d); run<1, 8>("HFMA2", d); run<1, 16>("HFMA2", d);
  run<2, 8>("FFMA", d); run<2, 16>("FFMA", d);
  run<3, 8>("SHF", d); run<3, 16>("SHF", d);
  printf("%s\n", cudaGetErrorString(cudaGetLastError()));
}
