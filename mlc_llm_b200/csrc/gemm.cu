// placeholder until the tcgen05 GEMM lands (next commit): fail loudly, never fall back.
#include "common.cuh"
namespace b200q {
int gemm_launch(const void*, const void*, const void*, const void*, const void*, const void*, void*, int, int,
                int, int, int, unsigned, cudaStream_t) {
  return set_error(B200Q_ERR_UNSUPPORTED, "b200q_gemm: not built yet");
}
int nt_matmul_launch(const void*, const void*, const void*, const void*, const void*, void*, int, int, int, int,
                     cudaStream_t) {
  return set_error(B200Q_ERR_UNSUPPORTED, "b200q_nt_matmul: not built yet");
}
}  // namespace b200q
