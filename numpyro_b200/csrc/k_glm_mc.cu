// Translation unit of the multi-chain streaming kernel.
#define B200MC_TU_GLM_MC
#include "launchers.h"

namespace b200mc {
namespace kl {

cudaError_t glm_mc_retile(int grid, const float* X, const uint8_t* y, int N, int D, int nt_total,
                          unsigned char* Xt, cudaStream_t st) {
  glm_mc_retile_kernel<<<grid, 256, 0, st>>>(X, y, N, D, nt_total, Xt);
  return cudaGetLastError();
}
cudaError_t glm_mc_prepare(int max_smem, int smem, int* occupancy) {
  cudaError_t e = cudaFuncSetAttribute(glm_mc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, max_smem);
  if (e != cudaSuccess) return e;
  return cudaOccupancyMaxActiveBlocksPerMultiprocessor(occupancy, glm_mc_kernel, kMcThreads, smem);
}
cudaError_t glm_mc_launch(const GlmMcLaunch& L, int ctas, size_t smem, cudaStream_t st) {
  void* args[] = {(void*)&L};
  return cudaLaunchCooperativeKernel((const void*)glm_mc_kernel, dim3(ctas), dim3(kMcThreads), args, smem, st);
}

}  // namespace kl
}  // namespace b200mc
