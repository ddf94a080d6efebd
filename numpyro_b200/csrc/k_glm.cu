// Translation unit of the streaming (tall-data logistic) kernel and the memory microbenchmark.
#define B200MC_TU_GLM
#include "launchers.h"

namespace b200mc {
namespace kl {

cudaError_t glm_retile(int grid, const float* X, const uint8_t* y, int N, int D, int nt_total,
                       unsigned char* Xt, cudaStream_t st) {
  glm_retile_kernel<<<grid, 256, 0, st>>>(X, y, N, D, nt_total, Xt);
  return cudaGetLastError();
}
cudaError_t glm_prepare(int max_smem, int smem, int* occupancy) {
  cudaError_t e = cudaFuncSetAttribute(glm_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, max_smem);
  if (e != cudaSuccess) return e;
  return cudaOccupancyMaxActiveBlocksPerMultiprocessor(occupancy, glm_kernel, kGlmThreads, smem);
}
cudaError_t glm_launch(const GlmLaunch& L, int ctas, size_t smem, cudaStream_t st) {
  void* args[] = {(void*)&L};
  return cudaLaunchCooperativeKernel((const void*)glm_kernel, dim3(ctas), dim3(kGlmThreads), args, smem, st);
}
cudaError_t membench_prepare(int max_smem) {
  return cudaFuncSetAttribute(membench_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, max_smem);
}
cudaError_t membench_launch(const MemBenchLaunch& L, int ctas, size_t smem, cudaStream_t st) {
  membench_kernel<<<ctas, kGlmThreads, smem, st>>>(L);
  return cudaGetLastError();
}

}  // namespace kl
}  // namespace b200mc
