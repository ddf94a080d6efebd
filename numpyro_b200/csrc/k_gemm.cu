// Translation unit of the split-precision tcgen05 contractions.
#define B200MC_TU_GEMM
#include "launchers.h"

namespace b200mc {
namespace kl {

cudaError_t image_build_launch(const float* src, int R_src, int K_src, int transpose, int R, int K,
                               unsigned char* hi, unsigned char* lo, cudaStream_t st) {
  image_build_kernel<<<1184, 256, 0, st>>>(src, R_src, K_src, transpose, R, K, hi, lo);
  return cudaGetLastError();
}
cudaError_t gemm_prepare() {
  cudaError_t e = cudaFuncSetAttribute(gemm3x_kernel<EPI_STORE>, cudaFuncAttributeMaxDynamicSharedMemorySize, kGemmSmem);
  if (e == cudaSuccess) e = cudaFuncSetAttribute(gemm3x_kernel<EPI_RESID>, cudaFuncAttributeMaxDynamicSharedMemorySize, kGemmSmem);
  if (e == cudaSuccess) e = cudaFuncSetAttribute(gemm3x_resid_persistent_kernel<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, kPersistSmem);
  if (e == cudaSuccess) e = cudaFuncSetAttribute(gemm3x_resid_persistent_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, kPersistSmem);
  return e;
}
cudaError_t gemm3x_launch(const Gemm3xLaunch& L, int epi, int grid, cudaStream_t st) {
  if (epi == EPI_STORE) gemm3x_kernel<EPI_STORE><<<grid, kGemmThreads, kGemmSmem, st>>>(L);
  else gemm3x_kernel<EPI_RESID><<<grid, kGemmThreads, kGemmSmem, st>>>(L);
  return cudaGetLastError();
}
cudaError_t gemm3x_persistent_launch(const Gemm3xLaunch& L, int grid, cudaStream_t st) {
  if (L.likelihood == 1) gemm3x_resid_persistent_kernel<1><<<grid, kGemmThreads, kPersistSmem, st>>>(L);
  else gemm3x_resid_persistent_kernel<0><<<grid, kGemmThreads, kPersistSmem, st>>>(L);
  return cudaGetLastError();
}

}  // namespace kl
}  // namespace b200mc
