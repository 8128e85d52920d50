// Microbenchmark 6: tcgen05.st throughput (registers -> TMEM), 4..16 warps per SM, x8 / x16 shapes.
#include <cstdio>
#include <cstdint>
#include "../../mlc_llm_b200/csrc/ptx.cuh"
using namespace b200q::ptx;
#define ITERS 2000
__device__ __forceinline__ void st8(uint32_t taddr, const uint32_t (&v)[64]) {
  asm volatile("tcgen05.st.sync.aligned.32x32b.x8.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};" ::"r"(taddr), "r"(v[0]), "r"(v[1]), "r"(v[2]),
               "r"(v[3]), "r"(v[4]), "r"(v[5]), "r"(v[6]), "r"(v[7]) : "memory");
}
__device__ __forceinline__ void st32(uint32_t taddr, const uint32_t (&v)[64]) {
  asm volatile("tcgen05.st.sync.aligned.32x32b.x32.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31,%32};" ::"r"(taddr), "r"(v[0]), "r"(v[1]), "r"(v[2]), "r"(v[3]), "r"(v[4]), "r"(v[5]), "r"(v[6]), "r"(v[7]), "r"(v[8]), "r"(v[9]), "r"(v[10]), "r"(v[11]), "r"(v[12]), "r"(v[13]), "r"(v[14]), "r"(v[15]), "r"(v[16]), "r"(v[17]), "r"(v[18]), "r"(v[19]), "r"(v[20]), "r"(v[21]), "r"(v[22]), "r"(v[23]), "r"(v[24]), "r"(v[25]), "r"(v[26]), "r"(v[27]), "r"(v[28]), "r"(v[29]), "r"(v[30]), "r"(v[31]) : "memory");
}
__device__ __forceinline__ void st64(uint32_t taddr, const uint32_t (&v)[64]) {
  asm volatile("tcgen05.st.sync.aligned.32x32b.x64.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31,%32,%33,%34,%35,%36,%37,%38,%39,%40,%41,%42,%43,%44,%45,%46,%47,%48,%49,%50,%51,%52,%53,%54,%55,%56,%57,%58,%59,%60,%61,%62,%63,%64};" ::"r"(taddr), "r"(v[0]), "r"(v[1]), "r"(v[2]), "r"(v[3]), "r"(v[4]), "r"(v[5]), "r"(v[6]), "r"(v[7]), "r"(v[8]), "r"(v[9]), "r"(v[10]), "r"(v[11]), "r"(v[12]), "r"(v[13]), "r"(v[14]), "r"(v[15]), "r"(v[16]), "r"(v[17]), "r"(v[18]), "r"(v[19]), "r"(v[20]), "r"(v[21]), "r"(v[22]), "r"(v[23]), "r"(v[24]), "r"(v[25]), "r"(v[26]), "r"(v[27]), "r"(v[28]), "r"(v[29]), "r"(v[30]), "r"(v[31]), "r"(v[32]), "r"(v[33]), "r"(v[34]), "r"(v[35]), "r"(v[36]), "r"(v[37]), "r"(v[38]), "r"(v[39]), "r"(v[40]), "r"(v[41]), "r"(v[42]), "r"(v[43]), "r"(v[44]), "r"(v[45]), "r"(v[46]), "r"(v[47]), "r"(v[48]), "r"(v[49]), "r"(v[50]), "r"(v[51]), "r"(v[52]), "r"(v[53]), "r"(v[54]), "r"(v[55]), "r"(v[56]), "r"(v[57]), "r"(v[58]), "r"(v[59]), "r"(v[60]), "r"(v[61]), "r"(v[62]), "r"(v[63]) : "memory");
}
__global__ void k(int x8, int wait_every, long long* out) {
  __shared__ uint32_t tptr;
  const int warp = threadIdx.x >> 5;
  if (warp == 0) { tmem_alloc(smem_u32(&tptr), 512); tmem_relinquish(); }
  tc_fence_before(); __syncthreads(); tc_fence_after();
  const uint32_t tb = tptr + ((uint32_t)((warp & 3) * 32) << 16) + (warp >> 2) * 64;
  uint32_t v[64];
  for (int i = 0; i < 64; ++i) v[i] = threadIdx.x * 16 + i;
  __syncthreads();
  long long t0 = clock64();
  for (int it = 0; it < ITERS; ++it) {
    if (x8 == 1) st8(tb, v); else if (x8 == 2) st32(tb + (it & 1) * 32, v); else if (x8 == 3) st64(tb, v); else { uint32_t w[16]; for (int i = 0; i < 16; ++i) w[i] = v[i]; tmem_st_32x16(tb + (it & 3) * 16, w); }
    if ((it % wait_every) == wait_every - 1) tmem_st_wait();
    v[0] += it; v[7] ^= it;
  }
  tmem_st_wait();
  long long t1 = clock64();
  __syncthreads();
  if (blockIdx.x == 0 && threadIdx.x == 0) out[0] = t1 - t0;
  tc_fence_before(); __syncthreads();
  if (warp == 0) { tc_fence_after(); tmem_dealloc(tptr, 512); }
}
int main() {
  long long* d; cudaMalloc(&d, 64);
  for (int x8 = 0; x8 < 4; ++x8)
    for (int we = 8; we <= 8; we *= 4)
      for (int warps = 4; warps <= 16; warps *= 2) {
        k<<<148, warps * 32>>>(x8, we, d); cudaError_t e = cudaDeviceSynchronize();
        long long h; cudaMemcpy(&h, d, 8, cudaMemcpyDeviceToHost);
        const int nb[4] = {64, 32, 128, 256};
        double bytes = (double)ITERS * warps * 32 * nb[x8];
        printf("%s warps=%2d wait every %d: %.1f clk per st per warp, %.1f B/clk/SM %s\n", (x8 == 0 ? "x16" : x8 == 1 ? "x8 " : x8 == 2 ? "x32" : "x64"), warps, we, h / (double)ITERS, bytes / h,
               e == cudaSuccess ? "" : cudaGetErrorString(e));
      }
}
