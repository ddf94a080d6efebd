// Translation unit of the chain-batched kernels (thread / warp / CTA per chain).
#define B200MC_TU_CHAIN
#include "launchers.h"

namespace b200mc {
namespace kl {

cudaError_t chain_launch(int mode, const ChainLaunch& L, int grid, int block, size_t smem, cudaStream_t st) {
  if (mode == 0) chain_kernel<0><<<grid, block, 0, st>>>(L);
  else if (mode == 1) chain_kernel<1><<<grid, block, 0, st>>>(L);
  else {
    if (smem > 48 * 1024) {
      cudaError_t e = cudaFuncSetAttribute(chain_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
      if (e != cudaSuccess) return e;
    }
    chain_kernel<2><<<grid, block, smem, st>>>(L);
  }
  return cudaGetLastError();
}
cudaError_t chain_potential_launch(int mode, const ModelData& M, const float* z, float* U, float* grad,
                                   int n, cudaStream_t st) {
  if (mode == 0) chain_potential_kernel<0><<<(n + 63) / 64, 64, 0, st>>>(M, z, U, grad, n);
  else if (mode == 1) chain_potential_kernel<1><<<(n + 3) / 4, 128, 0, st>>>(M, z, U, grad, n);
  else chain_potential_kernel<2><<<n, 256, 0, st>>>(M, z, U, grad, n);
  return cudaGetLastError();
}

size_t ecs_chain_static_smem() {
  cudaFuncAttributes a;
  return cudaFuncGetAttributes(&a, ecs_chain_kernel) == cudaSuccess ? a.sharedSizeBytes : (size_t)1 << 20;
}

cudaError_t ecs_chain_launch(const ChainLaunch& L, int grid, size_t smem, cudaStream_t st) {
  cudaError_t e = cudaFuncSetAttribute(ecs_chain_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  if (e != cudaSuccess) return e;
  ecs_chain_kernel<<<grid, 256, smem, st>>>(L);
  return cudaGetLastError();
}

}  // namespace kl
}  // namespace b200mc
