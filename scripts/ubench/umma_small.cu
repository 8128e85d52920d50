// Microbenchmark 5: issue cost of small tcgen05.mma (M128 N16 K16, A from TMEM) and of tcgen05.commit,
// from one or two issuing warps per CTA.
#include <cstdio>
#include <cstdint>
#include "../../mlc_llm_b200/csrc/ptx.cuh"
using namespace b200q::ptx;
#define ITERS 2000
// mode 0: MMAs only (one commit at the end); 1: MMA + commit every 8; 2: 2 MMA + commit; 3: commit only
__global__ void k(int mode, int n_warps, int nsize, long long* out) {
  __shared__ __align__(128) unsigned char bsm[4096];
  __shared__ uint64_t bars[4];
  __shared__ uint32_t tptr;
  const int warp = threadIdx.x >> 5;
  for (int i = threadIdx.x; i < 1024; i += blockDim.x) reinterpret_cast<uint32_t*>(bsm)[i] = 0;
  if (threadIdx.x == 0) { for (int i = 0; i < 4; ++i) mbar_init(smem_u32(&bars[i]), 1); fence_barrier_init(); }
  if (warp == 0) { tmem_alloc(smem_u32(&tptr), 512); tmem_relinquish(); }
  fence_proxy_async_smem(); tc_fence_before(); __syncthreads(); tc_fence_after();
  const uint32_t tb = tptr;
  if (warp < n_warps) {
    const uint32_t idesc = make_idesc_f16(128, nsize, 0, 0);
    const uint64_t bd = make_smem_desc_nosw(smem_u32(bsm), 256, 128);
    const uint32_t bar = smem_u32(&bars[warp]);
    long long t0 = clock64();
    uint32_t ph = 0;
    if (mode >= 5) {
      if (elect_one()) {
        for (int it = 0; it < ITERS; ++it) {
          umma_f16_ts(tb + 256 + warp * 64 + ((mode == 8) ? 0 : (mode == 9) ? (it & 1) * 16 : (it & 3) * 16), tb + (it & 7) * 8 + warp * 64, bd, idesc, 1u);
          if (mode == 6 && (it & 7) == 7) umma_commit(bar);
          if (mode == 7) umma_commit(bar);
        }
      }
      __syncwarp();
    } else
    for (int it = 0; it < ITERS; ++it) {
      if (elect_one()) {
        if (mode != 3) umma_f16_ts(tb + 256 + warp * 64 + (it & 3) * 16, tb + (it & 7) * 8 + warp * 64, bd, idesc, 1u);
        if (mode == 2) umma_f16_ts(tb + 256 + warp * 64 + (it & 3) * 16, tb + (it & 7) * 8 + warp * 64, bd, idesc, 1u);
        if (mode == 3 || mode == 2 || (mode == 1 && (it & 7) == 7)) umma_commit(bar);
      }
      __syncwarp();
      if (mode == 3 || mode == 2 || (mode == 1 && (it & 7) == 7)) { if (mode == 4) { mbar_wait(bar, ph); } ph ^= 1; }
    }
    long long t1 = clock64();
    if (elect_one()) umma_commit(smem_u32(&bars[2 + warp]));
    __syncwarp();
    mbar_wait(smem_u32(&bars[2 + warp]), 0);
    long long t2 = clock64();
    if (blockIdx.x == 0 && (threadIdx.x & 31) == 0) { out[warp * 2] = t1 - t0; out[warp * 2 + 1] = t2 - t0; }
  }
  tc_fence_before(); __syncthreads();
  if (warp == 0) { tc_fence_after(); tmem_dealloc(tb, 512); }
}
int main() {
  long long* d; cudaMalloc(&d, 64);
  const char* names[] = {"MMA only", "MMA, commit every 8", "2 MMA + commit", "commit only", "-", "thread loop: MMA", "thread loop: 8 MMA+commit", "thread loop: MMA+commit", "thread loop: same D", "thread loop: 2 D"};
  for (int nsize = 16; nsize <= 16; nsize *= 2)
  for (int mode = 5; mode < 10; ++mode)
    for (int nw = 1; nw <= 2; ++nw) {
      cudaMemset(d, 0, 64);
      k<<<148, 128>>>(mode, nw, nsize, d); cudaError_t e = cudaDeviceSynchronize();
      long long h[4]; cudaMemcpy(h, d, 32, cudaMemcpyDeviceToHost);
      printf("N=%d %-22s warps=%d: issue %.1f clk/iter, drained %.1f clk/iter (warp1: %.1f / %.1f) %s\n", nsize, names[mode], nw, h[0] / (double)ITERS,
             h[1] / (double)ITERS, h[2] / (double)ITERS, h[3] / (double)ITERS, e == cudaSuccess ? "" : cudaGetErrorString(e));
    }
}
