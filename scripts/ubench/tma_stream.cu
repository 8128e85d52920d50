// Microbenchmark 7: what the DRAM access pattern alone costs.  148 persistent CTAs stream a packed-weight matrix
// u32[R][N] (N contiguous, the reference's layout) into a shared-memory ring with TMA and throw the data away.
// Variables: the box shape (rows x words), the traversal order of a CTA's blocks, the ring depth.
//   order 0: stream-K as gemv_tc.cuh (a CTA walks DOWN one 128-word column stripe: consecutive blocks = consecutive
//            row blocks of one panel; neighbouring CTAs are on far-apart stripes)
//   order 1: panel-major (consecutive blocks of a CTA = neighbouring panels of the same row block: the CTA itself
//            reads contiguous memory)
//   order 2: stream-K over PAIRS of adjacent panels (a CTA alternates between two neighbouring stripes)
//   order 9: no tensor map at all: cp.async.bulk of contiguous chunks of box_rows * box_cols * 4 bytes (upper bound)
// build: nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -o tma_stream tma_stream.cu -lcuda
#include <cuda.h>
#include <cuda_runtime.h>

#include <cstdint>
#include <cstdio>
#include <cstdlib>

#include "../../mlc_llm_b200/csrc/ptx.cuh"
using namespace b200q::ptx;

struct P {
  CUtensorMap tm;
  const unsigned char* base;
  int n_panels, RB, T, stages, stage_bytes, order, box_rows, box_cols;
};

__global__ void __launch_bounds__(64, 1) k(const __grid_constant__ P p, unsigned long long* sink) {
  extern __shared__ __align__(1024) unsigned char smem[];
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem);
  const uint32_t full0 = smem_u32(bars), empty0 = smem_u32(bars + 64);  // up to 64 stages
  const uint32_t stage0 = (smem_u32(smem) + 1024u + 1023u) & ~1023u;
  const int G = gridDim.x, b = blockIdx.x;
  if (threadIdx.x == 0) {
    for (int s = 0; s < p.stages; ++s) { mbar_init(full0 + 8 * s, 1); mbar_init(empty0 + 8 * s, 1); }
    fence_barrier_init();
  }
  __syncthreads();
  const int t0 = (int)((long long)b * p.T / G), t1 = (int)((long long)(b + 1) * p.T / G);
  if (threadIdx.x == 0) {
    int s = 0; uint32_t ph = 0;
    for (int t = t0; t < t1; ++t) {
      int panel, rb;
      if (p.order == 0) { panel = t / p.RB; rb = t % p.RB; }
      else if (p.order == 1) { rb = t / p.n_panels; panel = t % p.n_panels; }
      else { const int pp = t / (2 * p.RB), r2 = t % (2 * p.RB); panel = 2 * pp + (r2 & 1); rb = r2 >> 1; if (panel >= p.n_panels) panel = p.n_panels - 1; }
      mbar_wait(empty0 + 8 * s, ph ^ 1);
      mbar_expect_tx(full0 + 8 * s, (uint32_t)p.stage_bytes);
      if (p.order == 9) {
        const unsigned char* src = p.base + (size_t)t * p.stage_bytes;
        asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                     ::"r"(stage0 + s * p.stage_bytes), "l"(src), "r"(p.stage_bytes), "r"(full0 + 8 * s) : "memory");
      } else {
        tma_load_2d(stage0 + s * p.stage_bytes, &p.tm, panel * p.box_cols, rb * p.box_rows, full0 + 8 * s);
      }
      if (++s == p.stages) { s = 0; ph ^= 1; }
    }
  } else if (threadIdx.x == 32) {
    int s = 0; uint32_t ph = 0; unsigned long long acc = 0;
    for (int t = t0; t < t1; ++t) {
      mbar_wait(full0 + 8 * s, ph);
      acc += *reinterpret_cast<volatile uint32_t*>(smem + (stage0 - smem_u32(smem)) + s * p.stage_bytes);
      mbar_arrive(empty0 + 8 * s);
      if (++s == p.stages) { s = 0; ph ^= 1; }
    }
    if (acc == 0x1234567ull) *sink = acc;
  }
}

int main(int argc, char** argv) {
  setvbuf(stdout, nullptr, _IONBF, 0);
  const size_t total = (size_t)2 << 30;  // 2 GiB matrix >> L2
  unsigned char* d; cudaMalloc(&d, total); cudaMemset(d, 1, total);
  unsigned long long* sink; cudaMalloc(&sink, 8);
  struct Cfg { int N, bc, br, order, stages; };
  const Cfg cfgs[] = {
      {4096, 128, 16, 0, 21}, {11008, 128, 16, 0, 21}, {32000, 128, 16, 0, 21},
      {4096, 128, 16, 1, 21}, {11008, 128, 16, 1, 21},
      {4096, 128, 16, 2, 21}, {11008, 128, 16, 2, 21},
      {4096, 256, 8, 0, 21}, {11008, 256, 8, 0, 21},
      {4096, 64, 32, 0, 21}, {4096, 32, 64, 0, 21},
      {4096, 128, 32, 0, 11}, {4096, 256, 16, 0, 11}, {4096, 128, 8, 0, 42},
      {4096, 128, 16, 0, 10}, {4096, 128, 16, 0, 15},
      {4096, 128, 16, 9, 21}, {4096, 128, 32, 9, 11},
  };
  CUresult (*enc)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                  const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
  cudaDriverEntryPointQueryResult q;
  cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", (void**)&enc, cudaEnableDefault, &q);
  cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
  int sms = 148; cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
  for (const Cfg& c : cfgs) {
    P p{};
    const int R = (int)(total / ((size_t)c.N * 4)) / c.br * c.br;
    cuuint64_t dims[2] = {(cuuint64_t)c.N, (cuuint64_t)R};
    cuuint64_t str[1] = {(cuuint64_t)c.N * 4};
    cuuint32_t box[2] = {(cuuint32_t)c.bc, (cuuint32_t)c.br}, es[2] = {1, 1};
    enc(&p.tm, CU_TENSOR_MAP_DATA_TYPE_UINT32, 2, d, dims, str, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE,
        CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    p.base = d; p.n_panels = c.N / c.bc; p.RB = R / c.br; p.T = p.n_panels * p.RB; p.stages = c.stages;
    p.stage_bytes = c.bc * c.br * 4; p.order = c.order; p.box_rows = c.br; p.box_cols = c.bc;
    const size_t smem = 2048 + (size_t)c.stages * p.stage_bytes + 1024;
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    printf("N %5d box %3d w x %2d rows order %d stages %2d: ", c.N, c.bc, c.br, c.order, c.stages);
    k<<<sms, 64, smem>>>(p, sink);
    cudaDeviceSynchronize();
    cudaEventRecord(e0);
    for (int i = 0; i < 3; ++i) k<<<sms, 64, smem>>>(p, sink);
    cudaEventRecord(e1); cudaDeviceSynchronize();
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    const double bytes = (double)p.T * p.stage_bytes * 3;
    printf("(%3zu KB ring): %7.1f GB/s  [%s]\n", ((size_t)c.stages * p.stage_bytes) >> 10, bytes / ms / 1e6,
           cudaGetErrorString(cudaGetLastError()));
  }
  return 0;
}
