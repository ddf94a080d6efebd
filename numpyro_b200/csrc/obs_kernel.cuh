// Forward evaluation of the registered families' observation sites over ALL posterior draws:
// the device half of log_likelihood / Predictive (numpyro/infer/util.py:803-1148).
//   m[s][i] = X[i] . B[s]   (GLM families: logistic, horseshoe)   or   loc[s][i]   (given)
//   log-likelihood: log_prob(y_i | m, scale, df)      Bernoulli(logits) discrete.py:137-139 +
//                   distributions/util.py:295-298; Normal continuous.py:2203-2207; StudentT :2375-2385
//   sampling:       Bernoulli: uniform(key_s, (N,))[i] < expit(m)  (discrete.py Bernoulli.sample ->
//                   jax.random.bernoulli);  Normal: m + scale * normal(key_s, (N,))[i]
// One thread per observation row, a block of draws per CTA: the X tile goes global -> shared once
// (coalesced), every element feeds kObsDraws FMAs.  FMA- and output-bound (N x S results).
#pragma once
#include <cuda_runtime.h>

#include "rng.cuh"

namespace b200mc {

constexpr int kObsThreads = 256;
constexpr int kObsDraws = 8;  // draws per CTA
enum ObsLikelihood { OBS_BERNOULLI_LOGITS = 0, OBS_NORMAL = 1, OBS_STUDENT_T = 2 };

struct ObsLaunch {
  const float* X;      // [N][K] row-major, or null
  int N, K, S;
  const float* B;      // X != null: coefficients [S][K];  X == null: loc [S][N]
  const float* scale;  // kind 1: [S], 2: [N], 3: [S][N]
  int scale_kind;
  const float* df;     // StudentT: [S]
  const float* y;      // observations [N]: log-likelihood mode
  const uint32_t* keys;  // [S][2] the observation site's key per draw: sampling mode
  int likelihood, partitionable;
  float* out;          // [S][N]; sampled Bernoulli values are written as int32
};

__host__ __device__ inline int obs_stride(int K) { return K | 1; }  // odd: conflict-free rows
__host__ __device__ inline size_t obs_smem(int K) {
  return sizeof(float) * ((size_t)kObsThreads * obs_stride(K) + (size_t)kObsDraws * K);
}

#ifdef B200MC_TU_OBS
__global__ void __launch_bounds__(kObsThreads) obs_forward_kernel(const ObsLaunch L) {
  extern __shared__ float sm[];
  const int K = L.K, st = obs_stride(K);
  float* xs = sm;                                // [256][st]
  float* bs = sm + (size_t)kObsThreads * st;     // [kObsDraws][K]
  const long i0 = (long)blockIdx.x * kObsThreads;
  const int s0 = blockIdx.y * kObsDraws;
  const int ns = (L.S - s0 < kObsDraws) ? L.S - s0 : kObsDraws;
  const long i = i0 + threadIdx.x;
  float m[kObsDraws];
#pragma unroll
  for (int s = 0; s < kObsDraws; ++s) m[s] = 0.f;
  if (L.X) {
    const long rows = (L.N - i0 < kObsThreads) ? L.N - i0 : kObsThreads;
    for (long e = threadIdx.x; e < rows * K; e += kObsThreads) {
      const int r = (int)(e / K), d = (int)(e - (long)r * K);
      xs[r * st + d] = L.X[i0 * K + e];
    }
    for (int e = threadIdx.x; e < ns * K; e += kObsThreads) bs[e] = L.B[(long)s0 * K + e];
    __syncthreads();
    if (i < L.N) {
      const float* xr = xs + threadIdx.x * st;
      for (int d = 0; d < K; ++d) {
        const float x = xr[d];
#pragma unroll
        for (int s = 0; s < kObsDraws; ++s)
          if (s < ns) m[s] = fmaf(x, bs[s * K + d], m[s]);
      }
    }
  } else if (i < L.N) {
#pragma unroll
    for (int s = 0; s < kObsDraws; ++s)
      if (s < ns) m[s] = L.B[(long)(s0 + s) * L.N + i];
  }
  if (i >= L.N) return;
  const float yv = L.y ? L.y[i] : 0.f;
#pragma unroll
  for (int s = 0; s < kObsDraws; ++s) {
    if (s >= ns) break;
    const int sg = s0 + s;
    const float loc = m[s];
    float sc = 1.f;
    if (L.scale_kind == 1) sc = L.scale[sg];
    else if (L.scale_kind == 2) sc = L.scale[i];
    else if (L.scale_kind == 3) sc = L.scale[(long)sg * L.N + i];
    float* o = L.out + (long)sg * L.N + i;
    if (L.keys) {  // ---- Predictive: one draw of the observation site
      const Key k{L.keys[2 * sg], L.keys[2 * sg + 1]};
      if (L.likelihood == OBS_BERNOULLI_LOGITS) {
        const float p = 1.0f / (1.0f + expf(-loc));  // expit
        const float u = uniform_at(k, (uint32_t)L.N, (uint32_t)i, 0.f, 1.f, L.partitionable != 0);
        *reinterpret_cast<int*>(o) = (u < p) ? 1 : 0;
      } else {
        *o = loc + sc * normal_at(k, (uint32_t)L.N, (uint32_t)i, L.partitionable != 0);
      }
      continue;
    }
    // ---- log_likelihood
    if (L.likelihood == OBS_BERNOULLI_LOGITS) {
      *o = -((fmaxf(loc, 0.f) - loc * yv) + log1pf(expf(-fabsf(loc))));
    } else if (L.likelihood == OBS_NORMAL) {
      const float r = yv - loc;
      *o = (-(r * r) / (2.0f * sc * sc) - logf(sc)) - 0.9189385332046727f;
    } else {
      const float df = L.df[sg];
      const float t = (yv - loc) / sc;
      const float z = logf(sc) + 0.5f * logf(df) + 0.5723649429247001f /* log(pi) / 2 */ +
                      lgammaf(0.5f * df) - lgammaf(0.5f * (df + 1.0f));
      *o = -0.5f * (df + 1.0f) * log1pf(t * t / df) - z;
    }
  }
}
#endif  // B200MC_TU_OBS

}  // namespace b200mc
