// Plain-data structures shared by host code and kernels.
#pragma once
#include <cstdint>

namespace b200mc {

constexpr int kMaxWindows = 40;
constexpr int kMaxSites = 8;
constexpr int kMaxTreeDepth = 20;

enum Algo { ALGO_NUTS = 0, ALGO_HMC = 1 };

// Per-handle constants of the transition kernel: mirror of the arguments of
// hmc.py:193-212 (init_kernel) + the window schedule of hmc_util.py:387-436.
struct SamplerParams {
  int D;
  int algo;
  int max_depth_warmup, max_depth_sample, max_depth;  // hmc.py:299-303, 439
  int num_warmup;
  int adapt_step_size, adapt_mass_matrix, regularize_mass_matrix;
  int partitionable;  // Threefry counter layout
  float target_accept_prob;
  float init_step_size;
  int has_trajectory_length;  // HMC
  float trajectory_length;
  int fixed_num_steps;  // HMC, <=0: none
  float max_delta_energy;  // hmc.py:188
  int dense_mass;          // 1: full D x D mass matrix (hmc_util.py:439-515, 133-239)
  int find_heuristic;      // find_heuristic_step_size (hmc_util.py:314-384)
  int factor_at_init;      // dense M^-1 supplied by the user: factorise it in init_kernel
  int num_windows;
  int window_end[kMaxWindows];
  // init_to_uniform draws, sites in TRACE order (infer/util.py:423-432)
  int nsites;
  int site_off[kMaxSites], site_size[kMaxSites];
  float init_radius;
  // structured mass matrix (hmc_util.py:439-486): num_blocks <= 1 => one block in flat order.
  // Device tables: blk_of[D] block of a flat coordinate, blk_pos[D] its position inside the
  // block, blk_perm[D] flat index of block-order position a, blk_size/blk_dense[num_blocks].
  int num_blocks;
  const int* blk_of;
  const int* blk_pos;
  const int* blk_perm;
  const int* blk_size;
  const int* blk_dense;
};

// Persistent per-chain scalars: the scalar leaves of HMCState (hmc.py:31-48) and
// HMCAdaptState (hmc_util.py:18-30).
struct ChainScalars {
  int i;
  float potential_energy;
  float energy;
  int num_steps;
  float accept_prob;
  float mean_accept_prob;
  int diverging;
  uint32_t rng_key[2];
  float step_size;
  float ss_x_t, ss_x_avg, ss_g_avg;
  int ss_t;
  float ss_prox_center;
  int mm_n;
  int window_idx;
  uint32_t adapt_key[2];
  int status;  // 0 ok, 1 = "Cannot find valid initial parameters"
  int init_tries;
  int64_t grad_evals;  // cumulative potential+gradient evaluations (init included)
};

// Persistent per-chain vectors, each [C][D] row-major.
struct ChainVectors {
  float* z;
  float* z_grad;
  float* inverse_mass_matrix;
  float* mass_matrix_sqrt;
  float* welford_mean;
  float* welford_m2;
  // dense_mass only: [C][4][D][D] = inverse_mass_matrix, mass_matrix_sqrt,
  // mass_matrix_sqrt_inv, welford m2 (row-major); null otherwise
  float* dense;
};

// Where collected draws go: [n_keep][C][...] (draw-major, like the reference's
// vectorized collection before its final swapaxes, mcmc.py:713-715).
struct CollectSpec {
  int num_iters;      // transitions to run in this call
  int collect_start;  // first local iteration that is collected (util.py:330 start_idx)
  int thinning;
  int n_keep;
  float* z;                 // [n_keep][C][D] or null
  float* potential_energy;  // [n_keep][C] or null
  float* energy;
  int* num_steps;
  float* accept_prob;
  float* mean_accept_prob;
  int* diverging;
  float* step_size;
};

}  // namespace b200mc
