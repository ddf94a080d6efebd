// Translation unit of the HMCECS kernels (subsample init, Taylor-proxy sums, Gibbs step).
#define B200MC_TU_ECS
#include "ecs_kernel.cuh"

namespace b200mc {
namespace kl {

cudaError_t ecs_randint_launch(const uint32_t* key, int n, int minval, int maxval, int part, int* out,
                               cudaStream_t st) {
  const int cnt = n > 0 ? n : 1;
  randint_kernel<<<(cnt + 255) / 256, 256, 0, st>>>(key, n, minval, maxval, part, out);
  return cudaGetLastError();
}

cudaError_t ecs_subsample_init_launch(const EcsLaunch& L, cudaStream_t st) {
  ecs_subsample_init_kernel<<<(L.C + 31) / 32, 32, 0, st>>>(L);
  return cudaGetLastError();
}

cudaError_t ecs_taylor_sums_launch(const float* X, const uint8_t* y, int N, int D, const float* ref,
                                   int degree, float* partials, int nparts, float* out,
                                   cudaStream_t st) {
  const int nout = 1 + D + (degree == 2 ? D * D : 0), nfull = 1 + D + D * D;
  const size_t smem = sizeof(float) * ((size_t)kEcsSumRows * D + 3 * kEcsSumRows);
  ecs_taylor_sums_kernel<<<nparts, 256, smem, st>>>(X, y, N, D, ref, degree, partials);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return e;
  ecs_taylor_reduce_kernel<<<(nfull + 255) / 256, 256, 0, st>>>(partials, nparts, nout, nfull, out);
  return cudaGetLastError();
}

cudaError_t ecs_gibbs_launch(const EcsLaunch& L, cudaStream_t st) {
  ecs_gibbs_kernel<<<L.C, 256, 0, st>>>(L);
  return cudaGetLastError();
}

}  // namespace kl
}  // namespace b200mc
