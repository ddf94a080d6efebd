// Translation unit of the observation-site forward kernel (log_likelihood / Predictive).
#define B200MC_TU_OBS
#include "obs_kernel.cuh"

namespace b200mc {
namespace kl {

cudaError_t obs_forward_launch(const ObsLaunch& L, cudaStream_t st) {
  const size_t smem = L.X ? obs_smem(L.K) : 0;
  if (smem > 48 * 1024) {
    cudaError_t e = cudaFuncSetAttribute(obs_forward_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
  }
  dim3 grid((unsigned)((L.N + kObsThreads - 1) / kObsThreads), (unsigned)((L.S + kObsDraws - 1) / kObsDraws));
  obs_forward_kernel<<<grid, kObsThreads, smem, st>>>(L);
  return cudaGetLastError();
}

}  // namespace kl
}  // namespace b200mc
