// Translation unit of the on-device diagnostics.
#define B200MC_TU_DIAG
#include "diag_kernel.cuh"

namespace b200mc {
namespace kl {

cudaError_t diag_prepare(int max_smem) {
  return cudaFuncSetAttribute(diag_stats_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, max_smem);
}
cudaError_t diag_stats_launch(const DiagLaunch& L, size_t smem, cudaStream_t st) {
  const int halves = L.split ? 2 : 1;
  dim3 grid((L.D + L.pb - 1) / L.pb, L.C * halves);
  diag_stats_kernel<<<grid, kDiagThreads, smem, st>>>(L);
  return cudaGetLastError();
}
cudaError_t diag_finish_launch(const DiagFinish& L, cudaStream_t st) {
  diag_finish_kernel<<<(L.D + 127) / 128, 128, 0, st>>>(L);
  return cudaGetLastError();
}

}  // namespace kl
}  // namespace b200mc
