// Microbenchmark: issue rate (warp-instructions / cycle / SMSP) of the instructions the dequant paths use.
// One CTA of 128..512 threads per SM (1..4 warps per SMSP), 8 independent chains per thread.
#include <cuda_fp16.h>
#include <cstdio>
#include <cstdint>
#define ITERS 4096
template <int OP>
__global__ void k(uint32_t* out, uint32_t seed) {
  uint32_t r[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) r[i] = seed + threadIdx.x * 8 + i;
  uint32_t c1 = seed ^ 0x3c003c00u, c2 = seed ^ 0x2c002c00u;
  float f[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) f[i] = __uint_as_float(r[i]);
  long long t0 = clock64();
  for (int it = 0; it < ITERS; ++it) {
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      if (OP == 0) asm volatile("fma.rn.f16x2 %0, %0, %1, %2;" : "+r"(r[i]) : "r"(c1), "r"(c2));
      if (OP == 1) asm volatile("mul.rn.f16x2 %0, %0, %1;" : "+r"(r[i]) : "r"(c1));
      if (OP == 2) asm volatile("add.rn.f16x2 %0, %0, %1;" : "+r"(r[i]) : "r"(c1));
      if (OP == 3) asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+f"(f[i]) : "f"(__uint_as_float(c1)), "f"(__uint_as_float(c2)));
      if (OP == 4) asm volatile("mul.rn.f32 %0, %0, %1;" : "+f"(f[i]) : "f"(__uint_as_float(c1)));
      if (OP == 5) asm volatile("lop3.b32 %0, %0, %1, %2, 0xEA;" : "+r"(r[i]) : "r"(c1), "r"(c2));
      if (OP == 6) asm volatile("prmt.b32 %0, %0, %1, 0x5410;" : "+r"(r[i]) : "r"(c1));
      if (OP == 7) asm volatile("shr.u32 %0, %0, 8; add.u32 %0, %0, %1;" : "+r"(r[i]) : "r"(c1));  // 2 instr
      if (OP == 8) { uint32_t t; asm volatile("cvt.rn.f16x2.f32 %0, %1, %2;" : "=r"(t) : "f"(f[i]), "f"(f[(i + 1) & 7])); f[i] = __uint_as_float(t ^ c1); }
      if (OP == 9) asm volatile("add.rn.f32 %0, %0, %1;" : "+f"(f[i]) : "f"(__uint_as_float(c1)));
      if (OP == 10) asm volatile("mad.lo.u32 %0, %0, %1, %2;" : "+r"(r[i]) : "r"(c1), "r"(c2));
    }
    if (OP == 11) {  // 2 independent HMMA per iteration
      asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                   : "+f"(f[0]), "+f"(f[1]), "+f"(f[2]), "+f"(f[3]) : "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(c1), "r"(c2));
      asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                   : "+f"(f[4]), "+f"(f[5]), "+f"(f[6]), "+f"(f[7]) : "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]), "r"(c1), "r"(c2));
    }
  }
  long long t1 = clock64();
  uint32_t acc = 0;
#pragma unroll
  for (int i = 0; i < 8; ++i) acc ^= r[i] ^ __float_as_uint(f[i]);
  out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
  if (threadIdx.x == 0 && blockIdx.x == 0) out[1 << 20] = (uint32_t)(t1 - t0);
}
template <int OP>
void run(const char* name, uint32_t* d, int n_per_iter) {
  for (int threads = 128; threads <= 512; threads *= 2) {
    k<OP><<<148, threads>>>(d, 12345);
    cudaDeviceSynchronize();
    k<OP><<<148, threads>>>(d, 12345);
    cudaDeviceSynchronize();
    uint32_t cyc;
    cudaMemcpy(&cyc, d + (1 << 20), 4, cudaMemcpyDeviceToHost);
    double wi = (double)ITERS * n_per_iter * (threads / 128);  // warp-instructions per SMSP
    printf("%-22s warps/SMSP=%d  %.3f warp-instr/cycle/SMSP  (%.2f cycles per warp-instr)\n", name, threads / 128, wi / cyc, cyc / wi);
  }
}
int main() {
  uint32_t* d;
  cudaMalloc(&d, ((1 << 20) + 16) * 4);
  run<0>("HFMA2", d, 8); run<1>("HMUL2", d, 8); run<2>("HADD2", d, 8); run<3>("FFMA", d, 8); run<4>("FMUL", d, 8); run<9>("FADD", d, 8);
  run<5>("LOP3", d, 8); run<6>("PRMT", d, 8); run<7>("SHR+IADD (2)", d, 16); run<8>("F2FP.PACK+LOP3 (2)", d, 16); run<10>("IMAD", d, 8);
  run<11>("HMMA.16816.F32", d, 2);
  printf("%s\n", cudaGetErrorString(cudaGetLastError()));
  return 0;
}
