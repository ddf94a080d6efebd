// Chain-batched persistent kernels for the per-chain families (eight schools,
// stochastic volatility): every chain is owned by one thread, one warp or one CTA and
// runs ALL of its transitions inside a single launch; chains advance independently
// (no lock-step across chains, unlike the reference's vmap, hmc.py:797).
#pragma once
#include "ecs_kernel.cuh"
#include "models.cuh"
#include "sampler.cuh"

namespace b200mc {

struct ChainLaunch {
  SamplerParams P;
  ModelData M;
  CollectSpec col;
  ChainScalars* cs;
  ChainVectors cv;
  float* work;              // tree / adaptation slots, layout depends on the mode
  const uint32_t* init_keys;  // [C][2] -> run the init search first; null -> continue
  int user_init;            // init: cv.z already holds caller-supplied positions
  int C;
  int c_pad;                // thread-per-chain: chains padded to a multiple of 32
  int end_thresh;           // thread-per-chain: lanes that must wait at a subtree end before
                            // the tail runs (1 = never wait), see chain_kernel<0>
  unsigned hot_mask;        // CTA-per-chain: slots kept in (dynamic) shared memory, see kChainHotMask
  EcsStep ecs;              // HMCECS: the Gibbs step before every transition (enabled = 0: none)
  int cpw;                  // thread-per-chain: chains (live lanes) per warp, 1..32
};

// CTA per chain: the leaf under construction (position, momentum, gradient) and the subtree's
// momentum sum are read and written several times per leapfrog; in the global work array every
// one of those is an L2 / HBM round trip (39 slots x 12 KB per stochastic-volatility chain, 1,024
// chains: 480 MB, past the L2).  They live in shared memory instead when four copies of D floats
// fit four resident CTAs per SM.
constexpr unsigned kChainHotMask = (1u << LZ) | (1u << LR) | (1u << LG) | (1u << S_RSUM);
__host__ __device__ inline size_t chain_hot_bytes(int D, unsigned mask) {
  int n = 0;
  for (unsigned m = mask; m; m &= m - 1) ++n;
  return (size_t)n * D * sizeof(float);
}

// MODE 0: thread per chain (work is [slot][D][c_pad]); 1: warp per chain; 2: CTA per
// chain (work is [chain][slot][D]).
#ifndef B200MC_CTA_MINB
#define B200MC_CTA_MINB 4  // resident CTAs per SM asked of ptxas for the CTA-per-chain kernel (64 registers)
#endif
#ifndef B200MC_CTA_THREADS
#define B200MC_CTA_THREADS 256
#endif
constexpr int kChainCtaThreads = B200MC_CTA_THREADS;

template <int MODE>
__global__ void __launch_bounds__(MODE == 2 ? kChainCtaThreads : 128, MODE == 2 ? B200MC_CTA_MINB : 1)
chain_kernel(const __grid_constant__ ChainLaunch L) {
  __shared__ float red_scratch[64];
  extern __shared__ float hot_slots[];
  const int D = L.P.D;
  const int nslots = num_slots(L.P.max_depth, L.P.dense_mass);
  int chain;
  float* base;
  long sstride;
  int es;
  if (MODE == 0) {
    // cpw chains per warp (lanes 0 .. cpw-1): a GPU that 16,384 chains fill to 3.5 warps per SM at
    // 32 chains per warp is latency-bound with one warp per scheduler (ncu: issue slots 10 %);
    // fewer chains per warp means more warps to hide latency behind and fewer lanes to diverge
    const int gw = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, ln = threadIdx.x & 31;
    chain = (ln < L.cpw) ? gw * L.cpw + ln : L.C;
    base = L.work + chain;
    sstride = (long)D * L.c_pad;
    es = L.c_pad;
  } else {
    chain = (MODE == 1) ? (blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5)) : blockIdx.x;
    base = L.work + (long)chain * nslots * D;
    sstride = D;
    es = 1;
  }
  if (MODE == 0) {
    // Thread per chain.  The 32 chains of a warp run the same leaf step (potential, leapfrog,
    // energy and checkpoint U-turn tests) in every iteration, but each reaches the END of a
    // subtree / of a transition in its own iteration, and that tail is long (the Threefry
    // draws of the next doubling, or split + D normals of the next momentum, plus the
    // adaptation step): run in place it executes with ~3 of 32 lanes active and takes ~78 % of
    // all issue slots (ncu: 9.8 active threads per instruction).  So a lane that reaches a
    // subtree end WAITS (AT_SUBTREE_END) until `end_thresh` lanes of its warp wait too, or no
    // lane has leaf work left; then all of them run the tail together.  Only the schedule
    // changes: every chain computes exactly what it computed before.
    const bool live = chain < L.C;
    ThreadGroup g{};
    Sampler<ThreadGroup> S;
    StepResult r = FINISHED;
    if (live) {
      S.setup(&L.P, g, base, sstride, es, chain, L.C, &L.col);
      S.load(L.cs, L.cv);
      S.end_yield = L.end_thresh > 1;
      if (L.init_keys)
        r = S.begin_init(Key{L.init_keys[2 * chain], L.init_keys[2 * chain + 1]}, L.user_init != 0);
      else
        r = S.begin_sampling();
    }
    // a warp with few live chains (num_chains = 4, the tail warp) scales the threshold down:
    // 7/8 of its live lanes, so one slow chain does not hold the others at every subtree end
    const int nlive = __popc(__ballot_sync(0xffffffffu, live));
    int thresh = (nlive * 7) / 8;
    thresh = thresh < 1 ? 1 : (thresh > L.end_thresh ? L.end_thresh : thresh);
    for (;;) {
      if (r == NEED_GRAD) {
        const float U = model_potential(L.M, g, S.v(LZ), es, S.v(LG), es, chain);
        r = S.advance(U);
      }
      const unsigned waiting = __ballot_sync(0xffffffffu, r == AT_SUBTREE_END);
      const unsigned running = __ballot_sync(0xffffffffu, r == NEED_GRAD);
      if ((waiting | running) == 0u) break;
      if (waiting != 0u && (__popc(waiting) >= thresh || running == 0u)) {
        if (r == AT_SUBTREE_END) r = S.finish_subtree_yielded();
      }
    }
    if (live) S.store(L.cs, L.cv);
    return;
  }
  if (chain >= L.C) return;

  auto run = [&](auto g) {
    const ModelData& M = L.M;
    float* ecs_rows = nullptr;
    if (MODE == 2 && L.ecs.enabled && L.ecs.rows_in_smem) {
      // HMCECS: the subsample is fixed during a transition and changes by one block between
      // transitions -- its rows (m x ncols floats) stay in shared memory behind the hot slots
      float* rows = hot_slots + chain_hot_bytes(D, L.hot_mask) / sizeof(float);
      const int* idx = M.ecs_idx + (long)chain * M.ecs_m;
      for (int t = g.lane(); t < M.ecs_m * M.ncols; t += g.size())
        rows[t] = M.a[(long)idx[t / M.ncols] * M.ncols + t % M.ncols];
      ecs_rows = rows;
      g.sync();
    }
    Sampler<decltype(g)> S;
    S.setup(&L.P, g, base, sstride, es, chain, L.C, &L.col);
    if (MODE == 2 && L.hot_mask) {
      S.hot = hot_slots;
      S.hot_mask = L.hot_mask;
    }
    S.load(L.cs, L.cv);
    S.transition_hook = MODE == 2 && L.ecs.enabled != 0;
    StepResult r;
    if (L.init_keys)
      r = S.begin_init(Key{L.init_keys[2 * chain], L.init_keys[2 * chain + 1]}, L.user_init != 0);
    else
      r = S.begin_sampling();
    for (;;) {
      if (r == NEED_GRAD) {
        const float U = model_potential(M, g, S.v(LZ), es, S.v(LG), es, chain, ecs_rows);
        r = S.advance(U);
      } else if (MODE == 2 && r == AT_TRANSITION_END) {
        // HMCECS (hmc_gibbs.py:619-652): block update of the subsample at the current point
        float pe_new;
        const bool acc = ecs_propose_test(M, L.ecs, g, S.v(Z_PROP), es, S.s.potential_energy, chain,
                                          S.local_iter, L.C, pe_new, ecs_rows);
        if (acc) {
          S.s.potential_energy = pe_new;
          float* gp = S.v(G_PROP);
          const float* gt = L.ecs.gtmp + (long)chain * D;
          for (int d = g.lane(); d < D; d += g.size()) gp[d * es] = gt[d];
        }
        S.s.grad_evals += 1;
        g.sync();
        r = S.resume_transition();
      } else {
        break;
      }
    }
    S.store(L.cs, L.cv);
  };
  if (MODE == 0) run(ThreadGroup{});
  else if (MODE == 1) run(WarpGroup{});
  else run(BlockGroup{red_scratch});
}

// HMCECS, device-resident loop, one CTA per chain (b200mc_ecs_sample): the step machine runs in
// WARP 0 with the whole tree state of the chain in shared memory (Sampler<WarpGroup>: vectors of
// D <= 128 in registers, warp shuffles -- the control warp of glm_kernel, 4.1), the potential on the
// subsample (rows staged in shared memory) and the Gibbs step run on all 256 threads.  The CTA-wide
// step machine of chain_kernel<2> costs ~30 us per leapfrog in block-wide reductions and global
// tree slots, which is what bounded a single HMCECS chain.
__host__ __device__ inline size_t ecs_chain_smem(int D, int max_depth, int dense, int m, int ncols) {
  return sizeof(float) * (((size_t)num_slots(max_depth, dense) * D + 3) / 4 * 4 + (size_t)m * ncols);
}
#ifdef B200MC_TU_CHAIN
__global__ void __launch_bounds__(256, 1) ecs_chain_kernel(const __grid_constant__ ChainLaunch L) {
  extern __shared__ float esm[];
  __shared__ float red_scratch[64];
  __shared__ int s_r, s_iter;
  __shared__ float s_pe;
  __shared__ float ecs_part[8 * kEcsCtaCols];
  const int D = L.P.D, chain = blockIdx.x, warp = threadIdx.x >> 5;
  if (chain >= L.C) return;
  const ModelData& M = L.M;
  float* work = esm;
  float* rows = esm + ((size_t)num_slots(L.P.max_depth, L.P.dense_mass) * D + 3) / 4 * 4;
  const int* idx = M.ecs_idx + (long)chain * M.ecs_m;
  BlockGroup g{red_scratch};
  for (int t = threadIdx.x; t < M.ecs_m * M.ncols; t += blockDim.x)
    rows[t] = M.a[(long)idx[t / M.ncols] * M.ncols + t % M.ncols];
  Sampler<WarpGroup> S;  // warp 0 only
  StepResult r = FINISHED;
  if (warp == 0) {
    S.setup(&L.P, WarpGroup{}, work, D, 1, chain, L.C, &L.col);
    S.load(L.cs, L.cv);
    S.transition_hook = true;
    r = S.begin_sampling();
  }
  for (;;) {
    if (threadIdx.x == 0) {
      s_r = (int)r;
      s_pe = S.s.potential_energy;
      s_iter = S.local_iter;
    }
    __syncthreads();  // also: warp 0's writes to the work array are visible to the CTA
    const int rr = s_r;
    if (rr == (int)NEED_GRAD) {
      const float U = logistic_ecs_potential_cta(M, g, work + (long)LZ * D, work + (long)LG * D, chain, idx,
                                                 rows, 0, 0, ecs_part);
      if (warp == 0) r = S.advance(U);
    } else if (rr == (int)AT_TRANSITION_END) {
      // hmc_gibbs.py:619-652: block update of the subsample at the current point
      float pe_new;
      const bool acc = ecs_propose_test(M, L.ecs, g, work + (long)Z_PROP * D, 1, s_pe, chain, s_iter, L.C,
                                        pe_new, rows, ecs_part);
      if (acc) {
        const float* gt = L.ecs.gtmp + (long)chain * D;
        for (int d = threadIdx.x; d < D; d += blockDim.x) work[(long)G_PROP * D + d] = gt[d];
      }
      __syncthreads();
      if (warp == 0) {
        if (acc) S.s.potential_energy = pe_new;
        S.s.grad_evals += 1;
        r = S.resume_transition();
      }
    } else {
      break;
    }
    __syncthreads();  // everyone has read s_r / s_pe before thread 0 rewrites them
  }
  if (warp == 0) S.store(L.cs, L.cv);
}
#endif  // B200MC_TU_CHAIN

// Tier-2 test hook: U and gradient at caller-supplied points z[C][D].
template <int MODE>
__global__ void chain_potential_kernel(ModelData M, const float* z, float* U, float* grad,
                                       int C) {
  __shared__ float red_scratch[64];
  const int D = M.D;
  int chain;
  if (MODE == 0) chain = blockIdx.x * blockDim.x + threadIdx.x;
  else chain = (MODE == 1) ? (blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5)) : blockIdx.x;
  if (chain >= C) return;
  auto run = [&](auto g) {
    const float u = model_potential(M, g, z + (long)chain * D, 1, grad + (long)chain * D, 1, chain);
    if (g.lane() == 0) U[chain] = u;
  };
  if (MODE == 0) run(ThreadGroup{});
  else if (MODE == 1) run(WarpGroup{});
  else run(BlockGroup{red_scratch});
}

}  // namespace b200mc
