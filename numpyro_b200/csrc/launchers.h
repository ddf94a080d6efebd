// Host-side launch wrappers: every heavy kernel family is compiled in its own translation unit
// (k_*.cu, built in parallel and re-built only when its sources change); api.cu sees the launch
// structs and layout helpers of the .cuh files, and these functions.
#pragma once
#include <cuda_runtime.h>

#include "batch_kernel.cuh"
#include "chain_kernel.cuh"
#include "diag_kernel.cuh"
#include "ecs_kernel.cuh"
#include "gemm3x_kernel.cuh"
#include "glm_kernel.cuh"
#include "glm_mc_kernel.cuh"
#include "membench.cuh"
#include "obs_kernel.cuh"

namespace b200mc {
namespace kl {

// k_glm.cu
cudaError_t glm_retile(int grid, const float* X, const uint8_t* y, int N, int D, int nt_total,
                       unsigned char* Xt, cudaStream_t st);
cudaError_t glm_prepare(int max_smem, int smem, int* occupancy);
cudaError_t glm_launch(const GlmLaunch& L, int ctas, size_t smem, cudaStream_t st);  // cooperative
cudaError_t membench_prepare(int max_smem);
cudaError_t membench_launch(const MemBenchLaunch& L, int ctas, size_t smem, cudaStream_t st);
// k_glm_mc.cu
cudaError_t glm_mc_retile(int grid, const float* X, const uint8_t* y, int N, int D, int nt_total,
                          unsigned char* Xt, cudaStream_t st);
cudaError_t glm_mc_prepare(int max_smem, int smem, int* occupancy);
cudaError_t glm_mc_launch(const GlmMcLaunch& L, int ctas, size_t smem, cudaStream_t st);  // cooperative
// k_chain.cu
cudaError_t chain_launch(int mode, const ChainLaunch& L, int grid, int block, size_t smem, cudaStream_t st);
cudaError_t ecs_chain_launch(const ChainLaunch& L, int grid, size_t smem, cudaStream_t st);
size_t ecs_chain_static_smem();
cudaError_t chain_potential_launch(int mode, const ModelData& M, const float* z, float* U, float* grad,
                                   int n, cudaStream_t st);
// k_batch.cu
cudaError_t batch_step_launch(const BatchLaunch& L, int chains, cudaStream_t st);
cudaError_t batch_matvec_launch(const BatchLaunch& L, dim3 grid, size_t smem, cudaStream_t st);
cudaError_t batch_potential_launch(const BatchLaunch& L, int points, int phase, cudaStream_t st);
cudaError_t factor_cov_launch(float* mats, int D, float n, int regularize, cudaStream_t st);
cudaError_t factor_tinv_launch(float* mats, int D, cudaStream_t st);
// k_diag.cu
cudaError_t diag_prepare(int max_smem);
cudaError_t diag_stats_launch(const DiagLaunch& L, size_t smem, cudaStream_t st);
cudaError_t diag_finish_launch(const DiagFinish& L, cudaStream_t st);
// k_obs.cu
cudaError_t obs_forward_launch(const ObsLaunch& L, cudaStream_t st);
// k_ecs.cu
cudaError_t ecs_randint_launch(const uint32_t* key, int n, int minval, int maxval, int part, int* out,
                               cudaStream_t st);
cudaError_t ecs_subsample_init_launch(const EcsLaunch& L, cudaStream_t st);
cudaError_t ecs_taylor_sums_launch(const float* X, const uint8_t* y, int N, int D, const float* ref,
                                   int degree, float* partials, int nparts, float* out,
                                   cudaStream_t st);
cudaError_t ecs_gibbs_launch(const EcsLaunch& L, cudaStream_t st);
// k_gemm.cu
cudaError_t image_build_launch(const float* src, int R_src, int K_src, int transpose, int R, int K,
                               unsigned char* hi, unsigned char* lo, cudaStream_t st);
cudaError_t gemm_prepare();
cudaError_t gemm3x_launch(const Gemm3xLaunch& L, int epi, int grid, cudaStream_t st);
cudaError_t gemm3x_persistent_launch(const Gemm3xLaunch& L, int grid, cudaStream_t st);

}  // namespace kl
}  // namespace b200mc
