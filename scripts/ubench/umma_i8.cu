// Microbenchmark 8: tcgen05.mma.kind::i8 with the A operand in tensor memory (the layout question of the integer
// formulation in csrc/gemv_tc.cuh).  One CTA, M = 128, N = 16, K = 32: A[m][k] u8 is written by thread m as 8 columns of
// 4 K-adjacent bytes (little endian), B[n][k] u8 / s8 sits in shared memory as the no-swizzle K-major operand
// [k / 16][n / 8][n % 8][16 B], D is s32.  Prints the number of mismatches against the host and, for a one-hot B, which k
// each (column, byte) of A feeds; then the issue rate of a stream of small MMAs from one lane (kind::i8 and kind::f16).
// Measured on B200: 0 mismatches (u8 x u8, u8 x s8), k = 4 * column + byte, ~48 cycles per MMA from one lane
// whatever the kind, the accumulator rotation or the number of issuing lanes.
// build: nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -o umma_i8 umma_i8.cu
#include <cuda_runtime.h>

#include <cstdint>
#include <cstdio>
#include <cstdlib>

#include "../../mlc_llm_b200/csrc/ptx.cuh"
using namespace b200q::ptx;

__device__ __forceinline__ void tmem_st_32x8(uint32_t taddr, const uint32_t* v) {
  asm volatile("tcgen05.st.sync.aligned.32x32b.x8.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};" ::"r"(taddr), "r"(v[0]), "r"(v[1]),
               "r"(v[2]), "r"(v[3]), "r"(v[4]), "r"(v[5]), "r"(v[6]), "r"(v[7])
               : "memory");
}

__global__ void __launch_bounds__(128, 1) k(const uint8_t* A, const uint8_t* B, int* D, int b_signed) {
  __shared__ __align__(128) unsigned char bs[512];
  __shared__ __align__(8) uint64_t bar;
  __shared__ uint32_t tptr;
  const int t = threadIdx.x, warp = t >> 5;
  for (int i = t; i < 512; i += 128) {
    const int kc = i >> 8, n = (i & 255) >> 4, kk = i & 15;  // byte i of the operand
    bs[i] = B[n * 32 + kc * 16 + kk];
  }
  if (t == 0) { mbar_init(smem_u32(&bar), 1); fence_barrier_init(); }
  if (warp == 0) { tmem_alloc(smem_u32(&tptr), 32); tmem_relinquish(); }
  fence_proxy_async_smem();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tb = tptr;
  uint32_t a[8];
  for (int c = 0; c < 8; ++c) a[c] = reinterpret_cast<const uint32_t*>(A + t * 32)[c];
  tmem_st_32x8(tb + ((uint32_t)(warp * 32) << 16), a);
  tmem_st_wait();
  tc_fence_before();
  __syncthreads();
  const uint32_t idesc = b_signed ? make_idesc_i8_u8s8(128, 16) : (make_idesc_i8_u8s8(128, 16) & ~(1u << 10));
  const uint64_t bd = make_smem_desc_nosw(smem_u32(bs), 256, 128);
  if (t == 0) {
    tc_fence_after();
    umma_i8_ts(tb + 16, tb, bd, idesc, 0u);
    umma_commit(smem_u32(&bar));
    mbar_wait(smem_u32(&bar), 0);
  }
  __syncthreads();
  tc_fence_after();
  uint32_t d[16];
  tmem_ld_32x16(tb + ((uint32_t)(warp * 32) << 16) + 16, d);
  tmem_ld_wait();
  for (int j = 0; j < 16; ++j) D[t * 16 + j] = (int)d[j];
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tb, 32);
}

// a stream of small MMAs issued by one lane of each of n_warps warps: kind 0 = i8 K32, 1 = f16 K16
template <int kind>
__global__ void __launch_bounds__(128, 1) rate(int n_acc, int n_warps, int iters, long long* out) {
  __shared__ __align__(128) unsigned char bs[1024];
  __shared__ __align__(8) uint64_t bars[4];
  __shared__ uint32_t tptr;
  const int t = threadIdx.x, warp = t >> 5;
  for (int i = t; i < 256; i += 128) reinterpret_cast<uint32_t*>(bs)[i] = 0;
  if (t == 0) { for (int i = 0; i < 4; ++i) mbar_init(smem_u32(&bars[i]), 1); fence_barrier_init(); }
  if (warp == 0) { tmem_alloc(smem_u32(&tptr), 512); tmem_relinquish(); }
  fence_proxy_async_smem();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tb = tptr;
  const uint32_t idesc8 = make_idesc_i8_u8s8(128, 16), idesc16 = make_idesc_f16(128, 16, 0, 0);
  const uint64_t bd = make_smem_desc_nosw(smem_u32(bs), 256, 128);
  if (warp < n_warps && (t & 31) == 0) {
    const long long c0 = clock64();
    uint32_t dd[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) dd[j] = tb + 256 + warp * 64 + ((uint32_t)j & (uint32_t)(n_acc - 1) & 3u) * 16;
    const uint32_t a0 = tb + warp * 64;
    for (int it = 0; it < iters; it += 8) {
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        if (kind == 0) umma_i8_ts(dd[j], a0 + j * 8, bd, idesc8, 1u);
        else umma_f16_ts(dd[j], a0 + j * 8, bd, idesc16, 1u);
      }
    }
    const long long c1 = clock64();
    umma_commit(smem_u32(&bars[warp]));
    mbar_wait(smem_u32(&bars[warp]), 0);
    out[2 * warp] = c1 - c0;
    out[2 * warp + 1] = clock64() - c0;
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) { tc_fence_after(); tmem_dealloc(tb, 512); }
}

int main() {
  setvbuf(stdout, nullptr, _IONBF, 0);
  uint8_t hA[128 * 32], hB[16 * 32];
  int hD[128 * 16];
  uint8_t *dA, *dB;
  int* dD;
  cudaMalloc(&dA, sizeof hA); cudaMalloc(&dB, sizeof hB); cudaMalloc(&dD, sizeof hD);
  for (int pass = 0; pass < 3; ++pass) {
    const int b_signed = (pass == 1);
    srand(1 + pass);
    for (auto& v : hA) v = (uint8_t)(pass == 2 ? (rand() % 251) : rand() % 16);
    for (auto& v : hB) v = (uint8_t)(rand() % 256);
    cudaMemcpy(dA, hA, sizeof hA, cudaMemcpyHostToDevice);
    cudaMemcpy(dB, hB, sizeof hB, cudaMemcpyHostToDevice);
    k<<<1, 128>>>(dA, dB, dD, b_signed);
    cudaError_t e = cudaDeviceSynchronize();
    cudaMemcpy(hD, dD, sizeof hD, cudaMemcpyDeviceToHost);
    int bad = 0;
    for (int m = 0; m < 128; ++m)
      for (int n = 0; n < 16; ++n) {
        int ref = 0;
        for (int kk = 0; kk < 32; ++kk) ref += (int)hA[m * 32 + kk] * (b_signed ? (int)(int8_t)hB[n * 32 + kk] : (int)hB[n * 32 + kk]);
        if (ref != hD[m * 16 + n]) { if (bad < 4) printf("  m %d n %d got %d want %d\n", m, n, hD[m * 16 + n], ref); ++bad; }
      }
    printf("pass %d (B %s): %d mismatches of 2048 [%s]\n", pass, b_signed ? "s8" : "u8", bad, cudaGetErrorString(e));
  }
  // one-hot probe: B[n][k] = (k == n + 16 * half); A[m][k] = k + 1  ->  D[m][n] = 1 + the k that feeds it
  for (int half = 0; half < 2; ++half) {
    for (int m = 0; m < 128; ++m) for (int kk = 0; kk < 32; ++kk) hA[m * 32 + kk] = (uint8_t)(kk + 1);
    for (int n = 0; n < 16; ++n) for (int kk = 0; kk < 32; ++kk) hB[n * 32 + kk] = (kk == n + 16 * half);
    cudaMemcpy(dA, hA, sizeof hA, cudaMemcpyHostToDevice);
    cudaMemcpy(dB, hB, sizeof hB, cudaMemcpyHostToDevice);
    k<<<1, 128>>>(dA, dB, dD, 0);
    cudaDeviceSynchronize();
    cudaMemcpy(hD, dD, sizeof hD, cudaMemcpyDeviceToHost);
    printf("one-hot half %d, row 5:", half);
    for (int n = 0; n < 16; ++n) printf(" %d", hD[5 * 16 + n]);
    printf("\n");
  }
  long long* dR; cudaMalloc(&dR, 64);
  long long h[8];
  for (int kind = 0; kind < 2; ++kind)
    for (int nw : {1, 2, 4})
      for (int n_acc : {1, 4}) {
        cudaMemset(dR, 0, 64);
        if (kind == 0) rate<0><<<1, 128>>>(n_acc, nw, 2000, dR); else rate<1><<<1, 128>>>(n_acc, nw, 2000, dR);
        cudaError_t e = cudaDeviceSynchronize();
        cudaMemcpy(h, dR, 64, cudaMemcpyDeviceToHost);
        printf("%s lanes %d accumulators %d: issue %.1f clk/MMA, drained %.1f clk/MMA (last lane: %.1f / %.1f) %s\n", kind ? "f16 K16" : "i8 K32 ", nw,
               n_acc, h[0] / 2000.0, h[1] / 2000.0, h[2 * nw - 2] / 2000.0, h[2 * nw - 1] / 2000.0, e == cudaSuccess ? "" : cudaGetErrorString(e));
      }
  return 0;
}
