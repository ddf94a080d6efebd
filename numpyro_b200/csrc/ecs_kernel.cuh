// HMC with energy-conserving subsampling (numpyro/infer/hmc_gibbs.py:502-665) around the
// chain-batched NUTS/HMC kernels: everything the wrapper does between two inner transitions,
// on the device, one CTA per chain.
//   ecs_subsample_init_kernel : the plate's initial subsample (primitives.py:443-462, CPU branch:
//                               Fisher-Yates, one jax.random.randint per index) under the seed
//                               handler's key splits (hmc_gibbs.py:547-552)
//   ecs_taylor_sums_kernel    : log-likelihood sum, gradient and Hessian over ALL rows at the
//                               reference point (contrib/ecs_proxies.py:165-169)
//   ecs_gibbs_kernel          : block update of the subsample (contrib/ecs_proxies.py:58-73), the
//                               Metropolis test on exp(pe - pe_new) and the state refresh
//                               (hmc_gibbs.py:619-652)
// jax.random.randint restated from jax/_src/random.py (_randint; third party, unpinned like
// Threefry: SURVEY 8c): k1, k2 = split(key); span = maxval - minval; mult = (2^16 % span)^2 % span;
// offset = ((bits(k1) % span) * mult + bits(k2) % span) % span in uint32 arithmetic.
#pragma once
#include <cuda_runtime.h>

#include "models.cuh"
#include "rng.cuh"

namespace b200mc {

__host__ __device__ __forceinline__ int randint_at(Key key, uint32_t n, uint32_t i, int minval,
                                                   int maxval, bool partitionable) {
  Key k1, k2;
  split2(key, partitionable, k1, k2);
  const uint32_t hi = random_bits_at(k1, n, i, partitionable);
  const uint32_t lo = random_bits_at(k2, n, i, partitionable);
  const uint32_t span = (maxval > minval) ? (uint32_t)(maxval - minval) : 1u;
  uint32_t mult = 65536u % span;
  mult = (mult * mult) % span;
  const uint32_t off = ((hi % span) * mult + (lo % span)) % span;
  return minval + (int)off;
}

// The Gibbs step of HMCECS.sample (hmc_gibbs.py:619-652) for one chain, executed by the chain's
// group g (every lane computes the same keys and the same decision):
//   rng_key = split(key)[0]; _update_block(rng_key, ...) -> proposal; pe_new = U(z; proposal);
//   accept_prob = min(exp(pe - pe_new), 1) (NaN kept); accepted = uniform(rng_key) < accept_prob.
// On acceptance the chain's subsample is replaced; the caller replaces potential energy and
// gradient (left in gtmp[chain]).  keys[chain] <- rng_key, accept_prob[chain] <- accept_prob.
struct EcsStep {
  int enabled;
  int* idx;            // [C][m] current subsamples (== ModelData::ecs_idx)
  int* prop;           // [C][m] proposals
  float* gtmp;         // [C][D] gradient at the proposal
  uint32_t* keys;      // [C][2] wrapper keys, in / out
  float* accept_prob;  // [C] accept_prob of the last step
  float* accept_all;   // [num_iters][C] accept_prob of every step of a device-resident run, or null
  int num_blocks, partitionable;
  int rows_in_smem;    // device-resident loop: stage the subsample's rows behind the hot slots
};

template <class G>
__device__ bool ecs_propose_test(const ModelData& M, const EcsStep& E, const G& g, const float* z,
                                 int zs, float pe, int chain, int local_iter, int C, float& pe_new,
                                 float* rows = nullptr, float* cta_part = nullptr) {
  const bool part = E.partitionable != 0;
  const int m = M.ecs_m, size = M.N, D = M.D;
  int* idx = E.idx + (long)chain * m;
  int* prop = E.prop + (long)chain * m;
  // hmc_gibbs.py:619: rng_key, rng_gibbs = split(state.rng_key); rng_gibbs is not used
  Key rng_key, unused;
  split2(Key{E.keys[2 * chain], E.keys[2 * chain + 1]}, part, rng_key, unused);
  // _update_block (contrib/ecs_proxies.py:58-73)
  Key k_next, subkey, block_key;
  split3(rng_key, part, k_next, subkey, block_key);
  const int block_size = (m - 1) / E.num_blocks + 1;
  const int pad = block_size - (m - 1) % block_size - 1;
  const int chosen = randint_at(block_key, 1u, 0u, 0, E.num_blocks, part);
  int start = chosen * block_size;
  if (start > m + pad - block_size) start = m + pad - block_size;  // dynamic_update_slice clamps
  if (start < 0) start = 0;
  for (int i = g.lane(); i < m; i += g.size()) {
    int v = idx[i];
    if (i >= start && i < start + block_size)
      v = randint_at(subkey, (uint32_t)block_size, (uint32_t)(i - start), 0, size, part);
    prop[i] = v;
  }
  g.sync();
  if constexpr (G::kBlockWide) {
    if (cta_part)  // ecs_chain_kernel: the one-CTA form with the rows in shared memory
      pe_new = logistic_ecs_potential_cta(M, g, z, E.gtmp + (long)chain * D, chain, prop, rows, start,
                                          start + block_size, cta_part);
    else
      pe_new = logistic_ecs_potential(M, g, z, zs, E.gtmp + (long)chain * D, 1, chain, prop, rows, start,
                                      start + block_size);
  } else {
    pe_new = logistic_ecs_potential(M, g, z, zs, E.gtmp + (long)chain * D, 1, chain, prop, rows, start,
                                    start + block_size);
  }
  const float ex = expf(pe - pe_new);
  const float ap = (ex != ex) ? ex : fminf(ex, 1.0f);  // jnp.clip(exp(pe - pe_new), None, 1): NaN stays NaN
  const float u = uniform01(rng_key, part);            // random.bernoulli(rng_key, accept_prob)
  const bool acc = u < ap;
  if (g.lane() == 0) {
    E.accept_prob[chain] = ap;
    if (E.accept_all) E.accept_all[(long)local_iter * C + chain] = ap;
    E.keys[2 * chain] = rng_key.k0;
    E.keys[2 * chain + 1] = rng_key.k1;
  }
  if (acc) {
    for (int i = g.lane(); i < m; i += g.size()) idx[i] = prop[i];
    if (rows) {  // refresh the staged copy of the replaced block
      const int i1 = (start + block_size < m) ? start + block_size : m;
      for (int t = g.lane(); t < (i1 - start) * M.ncols; t += g.size()) {
        const int i = start + t / M.ncols, d = t % M.ncols;
        rows[(long)i * M.ncols + d] = M.a[(long)prop[i] * M.ncols + d];
      }
    }
  }
  g.sync();
  return acc;
}

struct EcsLaunch {
  ModelData M;
  ChainScalars* cs;
  ChainVectors cv;
  int C, partitionable;
  int* idx;            // [C][m] current subsamples (M.ecs_idx points here)
  int* idx_prop;       // [C][m] proposals
  float* grad_tmp;     // [C][D]
  uint32_t* keys;      // [C][2] wrapper keys, in / out
  float* accept_prob;  // [C]
  int num_blocks;
  // subsample init
  const uint32_t* key_u;  // [C][2]
  int sites_before_plate;
  int* hkeys;  // [C][hcap] open-addressing map of displaced positions (Fisher-Yates on arange(N))
  int* hvals;
  int hcap;    // power of two >= 4 m
};

#ifdef B200MC_TU_ECS
__global__ void randint_kernel(const uint32_t* key, int n, int minval, int maxval, int part,
                               int* out) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  const int cnt = n > 0 ? n : 1;
  if (i < cnt) out[i] = randint_at(Key{key[0], key[1]}, (uint32_t)cnt, (uint32_t)i, minval, maxval, part != 0);
}

// one thread per chain: the shuffle is sequential (each swap depends on the previous ones)
__global__ void ecs_subsample_init_kernel(const EcsLaunch L) {
  const int chain = blockIdx.x * blockDim.x + threadIdx.x;
  if (chain >= L.C) return;
  const bool part = L.partitionable != 0;
  const int size = L.M.N, m = L.M.ecs_m, cap = L.hcap;
  int* hk = L.hkeys + (long)chain * cap;
  int* hv = L.hvals + (long)chain * cap;
  for (int i = 0; i < cap; ++i) hk[i] = -1;
  auto slot = [&](int pos) {
    unsigned h = ((unsigned)pos * 2654435761u) & (unsigned)(cap - 1);
    while (hk[h] != -1 && hk[h] != pos) h = (h + 1u) & (unsigned)(cap - 1);
    return (int)h;
  };
  auto get = [&](int pos) {
    const int h = slot(pos);
    return hk[h] == pos ? hv[h] : pos;
  };
  auto set = [&](int pos, int val) {
    const int h = slot(pos);
    hk[h] = pos;
    hv[h] = val;
  };
  // seed handler: one split per unobserved sample site before the plate, one for the plate
  Key k{L.key_u[2 * chain], L.key_u[2 * chain + 1]}, sub{0u, 0u};
  for (int s = 0; s <= L.sites_before_plate; ++s) {
    Key a, b;
    split2(k, part, a, b);
    k = a;
    sub = b;
  }
  for (int t = 0; t < m; ++t) {
    const int i_p1 = size - t, i = i_p1 - 1;
    const Key kt = split_at(sub, (uint32_t)m, (uint32_t)t, part);
    const int j = randint_at(kt, 1u, 0u, 0, i_p1, part);
    const int a = get(i), b = get(j);
    set(i, b);
    set(j, a);
  }
  int* out = L.idx + (long)chain * m;
  for (int t = 0; t < m; ++t) out[t] = get(size - m + t);
}

// All-row sums at the reference point: out[cta][1 + D + D*D] partials, then ecs_taylor_reduce.
// 64 rows per step are staged in shared memory; thread t owns outputs t, t + 256, ...
constexpr int kEcsSumRows = 64;
__global__ void __launch_bounds__(256) ecs_taylor_sums_kernel(const float* __restrict__ X,
                                                              const uint8_t* __restrict__ y, int N,
                                                              int D, const float* __restrict__ ref,
                                                              int degree, float* partials) {
  extern __shared__ float sm[];
  float* xs = sm;                       // [64][D]
  float* gw = xs + kEcsSumRows * D;     // [64] y - sigmoid
  float* hw = gw + kEcsSumRows;         // [64] -sigmoid (1 - sigmoid)
  float* lw = hw + kEcsSumRows;         // [64] log-lik
  const int nout = 1 + D + (degree == 2 ? D * D : 0);
  constexpr int kMaxOwn = 16;  // outputs per thread: D <= 62
  float acc[kMaxOwn];
#pragma unroll
  for (int q = 0; q < kMaxOwn; ++q) acc[q] = 0.f;
  const long rows_per = ((long)N + gridDim.x - 1) / gridDim.x;
  const long r0 = (long)blockIdx.x * rows_per, r1 = (r0 + rows_per < N) ? r0 + rows_per : N;
  for (long base = r0; base < r1; base += kEcsSumRows) {
    const int nr = (int)((r1 - base < kEcsSumRows) ? r1 - base : kEcsSumRows);
    __syncthreads();
    for (int t = threadIdx.x; t < kEcsSumRows * D; t += blockDim.x)
      xs[t] = (t < nr * D) ? X[base * D + t] : 0.f;
    __syncthreads();
    if (threadIdx.x < kEcsSumRows) {
      const int r = threadIdx.x;
      float l = 0.f;
      for (int d = 0; d < D; ++d) l = fmaf(xs[r * D + d], ref[d], l);
      float g = 0.f, h = 0.f, ll = 0.f;
      if (r < nr) {
        const float yv = (float)y[base + r];
        const float e = expf(-fabsf(l));
        const float inv = 1.0f / (1.0f + e);
        const float sig = (l >= 0.f) ? inv : e * inv;
        ll = -((fmaxf(l, 0.f) + log1pf(e)) - l * yv);
        g = yv - sig;
        h = -sig * (1.0f - sig);
      }
      gw[r] = g; hw[r] = h; lw[r] = ll;
    }
    __syncthreads();
#pragma unroll
    for (int q = 0; q < kMaxOwn; ++q) {
      const int o = threadIdx.x + q * 256;
      if (o < nout) {
        float a = acc[q];
        if (o == 0) {
          for (int r = 0; r < kEcsSumRows; ++r) a += lw[r];
        } else if (o <= D) {
          const int d = o - 1;
          for (int r = 0; r < kEcsSumRows; ++r) a = fmaf(gw[r], xs[r * D + d], a);
        } else {
          const int e2 = o - 1 - D, d = e2 / D, k = e2 - d * D;
          for (int r = 0; r < kEcsSumRows; ++r) a = fmaf(hw[r] * xs[r * D + d], xs[r * D + k], a);
        }
        acc[q] = a;
      }
    }
  }
#pragma unroll
  for (int q = 0; q < kMaxOwn; ++q) {
    const int o = threadIdx.x + q * 256;
    if (o < nout) partials[(long)blockIdx.x * nout + o] = acc[q];
  }
}

__global__ void ecs_taylor_reduce_kernel(const float* partials, int nparts, int nout, int nfull,
                                         float* out) {
  const int o = blockIdx.x * blockDim.x + threadIdx.x;
  if (o >= nfull) return;
  double a = 0.0;
  if (o < nout)
    for (int p = 0; p < nparts; ++p) a += (double)partials[(long)p * nout + o];  // fixed order
  out[o] = (float)a;
}

// one CTA per chain
__global__ void __launch_bounds__(256) ecs_gibbs_kernel(const EcsLaunch L) {
  __shared__ float red_scratch[64];
  const int chain = blockIdx.x;
  if (chain >= L.C) return;
  const int D = L.M.D;
  BlockGroup g{red_scratch};
  EcsStep E;
  E.enabled = 1; E.idx = L.idx; E.prop = L.idx_prop; E.gtmp = L.grad_tmp; E.keys = L.keys;
  E.accept_prob = L.accept_prob; E.accept_all = nullptr; E.num_blocks = L.num_blocks;
  E.partitionable = L.partitionable; E.rows_in_smem = 0;
  const float pe = L.cs[chain].potential_energy;
  float pe_new;
  const bool acc = ecs_propose_test(L.M, E, g, L.cv.z + (long)chain * D, 1, pe, chain, 0, L.C, pe_new);
  if (threadIdx.x == 0) {
    if (acc) L.cs[chain].potential_energy = pe_new;
    L.cs[chain].grad_evals += 1;
  }
  if (acc) {
    float* zg = L.cv.z_grad + (long)chain * D;
    const float* gt = L.grad_tmp + (long)chain * D;
    for (int d = threadIdx.x; d < D; d += blockDim.x) zg[d] = gt[d];
  }
}
#endif  // B200MC_TU_ECS

}  // namespace b200mc
