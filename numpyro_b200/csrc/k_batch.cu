// Translation unit of the lock-step engine's per-chain kernels.
#define B200MC_TU_BATCH
#include "launchers.h"

namespace b200mc {
namespace kl {

cudaError_t batch_step_launch(const BatchLaunch& L, int chains, cudaStream_t st) {
  batch_step_kernel<<<chains, kBatchThreads, 0, st>>>(L);
  return cudaGetLastError();
}
cudaError_t batch_matvec_launch(const BatchLaunch& L, dim3 grid, size_t smem, cudaStream_t st) {
  batch_matvec_kernel<<<grid, kMatvecThreads, smem, st>>>(L);
  return cudaGetLastError();
}
cudaError_t batch_potential_launch(const BatchLaunch& L, int points, int phase, cudaStream_t st) {
  batch_potential_kernel<<<points, kBatchThreads, 0, st>>>(L, phase);
  return cudaGetLastError();
}
cudaError_t factor_cov_launch(float* mats, int D, float n, int regularize, cudaStream_t st) {
  factor_cov_kernel<<<592, 256, 0, st>>>(mats, D, n, regularize);
  return cudaGetLastError();
}
cudaError_t factor_tinv_launch(float* mats, int D, cudaStream_t st) {
  factor_tinv_kernel<<<592, 256, 0, st>>>(mats, D);
  return cudaGetLastError();
}

}  // namespace kl
}  // namespace b200mc
