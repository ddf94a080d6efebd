// Convergence diagnostics on the device, over the draws as the samplers collect them:
// x[draw][chain][D] float32 (the reference's layout before its final swapaxes, mcmc.py:713-715).
//   * split Gelman-Rubin R-hat        numpyro/diagnostics.py:29-80
//   * autocovariance (biased)         numpyro/diagnostics.py:101-155  (the reference goes through
//     an FFT; here the lag products are summed directly in float64, which is exact up to
//     round-off and, for the few thousand draws a run keeps, cheaper than the transposes an FFT
//     batch would need on this layout)
//   * Geyer initial-monotone-sequence effective sample size   numpyro/diagnostics.py:158-203
// so that ESS/s reporting over 16,384 chains x D does not go through host NumPy.
//
// diag_stats_kernel  : per (chain-or-half-chain, parameter) mean and sum of squared deviations;
//                      with max_lag > 0 also adds the chain's autocovariances gamma_t, t < max_lag,
//                      into gamma_sum[t][d] (float64 atomics: the order of the chain sum is not
//                      fixed, the result is to ~1e-16 relative).
// diag_finish_kernel : per parameter: var_within / var_estimator / R-hat from the chain
//                      statistics, and tau / n_eff from the chain-averaged autocovariances.
#pragma once
#include <cuda_runtime.h>

namespace b200mc {

constexpr int kDiagThreads = 256;

struct DiagLaunch {
  const float* x;    // [n_total][C][D]
  int n_total, C, D;
  int first, n;      // draws first .. first + n - 1 of every chain are used
  int split;         // 1: every chain counts as two half chains (first n/2 and last n/2 draws)
  int max_lag;       // autocovariances gamma_0 .. gamma_{max_lag-1} (0: none)
  int pb;            // parameters per CTA (a power of two <= 32)
  double* mean;      // [C * (1 + split)][D]
  double* ssd;       // [C * (1 + split)][D]  sum of squared deviations from the mean
  double* gamma_sum; // [max_lag][D], zeroed by the caller
};

struct DiagFinish {
  const double* mean;       // [M][D] chain (or half-chain) means, M = chains in this statistic
  const double* ssd;        // [M][D]
  int M, n, D;              // n = draws per (half) chain
  const double* gamma_sum;  // [max_lag][D] sum over chains of gamma_t, or null
  int max_lag;
  double* r_hat;            // [D] sqrt(var_estimator / var_within)            (diagnostics.py:45-61)
  double* n_eff;            // [D] M n / tau (null when gamma_sum is null)     (diagnostics.py:158-203)
};

#ifdef B200MC_TU_DIAG
// grid (ceil(D / pb), chains * (1 + split)); dynamic shared memory n_eff_draws * pb floats
__global__ void __launch_bounds__(kDiagThreads) diag_stats_kernel(const DiagLaunch L) {
  extern __shared__ float series[];  // [draw][pb] centred
  __shared__ double red[kDiagThreads];
  const int pb = L.pb, p = threadIdx.x % pb, g = threadIdx.x / pb, ng = kDiagThreads / pb;
  const int d = blockIdx.x * pb + p;
  const int halves = L.split ? 2 : 1;
  const int chain = blockIdx.y / halves, half = blockIdx.y % halves;
  const int n = L.split ? L.n / 2 : L.n;
  const int first = L.first + ((L.split && half) ? (L.n - n) : 0);
  const bool live = d < L.D;
  // mean (float64), then centre into shared memory
  double s = 0.0;
  if (live)
    for (int i = g; i < n; i += ng) s += (double)L.x[((long)(first + i) * L.C + chain) * L.D + d];
  red[threadIdx.x] = s;
  __syncthreads();
  double mean = 0.0;
  for (int k = 0; k < ng; ++k) mean += red[k * pb + p];  // fixed order
  mean /= (double)n;
  __syncthreads();
  double q = 0.0;
  if (live)
    for (int i = g; i < n; i += ng) {
      const double v = (double)L.x[((long)(first + i) * L.C + chain) * L.D + d] - mean;
      series[(long)i * pb + p] = (float)v;
      q += v * v;
    }
  red[threadIdx.x] = q;
  __syncthreads();
  if (g == 0 && live) {
    double ssd = 0.0;
    for (int k = 0; k < ng; ++k) ssd += red[k * pb + p];
    L.mean[(long)blockIdx.y * L.D + d] = mean;
    L.ssd[(long)blockIdx.y * L.D + d] = ssd;
  }
  if (L.max_lag <= 0 || !live) return;
  // biased autocovariance gamma_t = (1/n) sum_i c_i c_{i+t}; lags strided over the thread groups
  for (int t = g; t < L.max_lag && t < n; t += ng) {
    double a0 = 0.0, a1 = 0.0;
    int i = 0;
    for (; i + 1 < n - t; i += 2) {
      a0 += (double)series[(long)i * pb + p] * (double)series[(long)(i + t) * pb + p];
      a1 += (double)series[(long)(i + 1) * pb + p] * (double)series[(long)(i + 1 + t) * pb + p];
    }
    if (i < n - t) a0 += (double)series[(long)i * pb + p] * (double)series[(long)(i + t) * pb + p];
    atomicAdd(&L.gamma_sum[(long)t * L.D + d], (a0 + a1) / (double)n);
  }
}

// one thread per parameter
__global__ void diag_finish_kernel(const DiagFinish L) {
  const int d = blockIdx.x * blockDim.x + threadIdx.x;
  if (d >= L.D) return;
  const double n = (double)L.n, M = (double)L.M;
  // _chain_variance_stats (diagnostics.py:29-42)
  double vw = 0.0, mm = 0.0;
  for (int c = 0; c < L.M; ++c) {
    vw += L.ssd[(long)c * L.D + d] / (n - 1.0);
    mm += L.mean[(long)c * L.D + d];
  }
  vw /= M;
  mm /= M;
  double var_within = vw, var_estimator = vw * (n - 1.0) / n;
  if (L.M > 1) {
    double vb = 0.0;
    for (int c = 0; c < L.M; ++c) {
      const double e = L.mean[(long)c * L.D + d] - mm;
      vb += e * e;
    }
    var_estimator += vb / (M - 1.0);
  } else {
    var_within = var_estimator;
  }
  if (L.r_hat) L.r_hat[d] = sqrt(var_estimator / var_within);
  if (!L.n_eff || !L.gamma_sum) return;
  // Geyer: rho_t = 1 - (var_within - mean_c gamma_t) / var_estimator, rho_0 = 1; pair sums
  // P_k = rho_2k + rho_2k+1; P_0 kept, P_k>=1 clipped at 0 and made non-increasing; tau = -1 + 2 sum P
  const int T = L.max_lag < L.n ? L.max_lag : L.n;
  double tau = -1.0, run_min = 0.0;
  for (int k = 0; 2 * k + 1 < T; ++k) {
    const double r0 = (k == 0) ? 1.0 : 1.0 - (var_within - L.gamma_sum[(long)(2 * k) * L.D + d] / M) / var_estimator;
    const double r1 = 1.0 - (var_within - L.gamma_sum[(long)(2 * k + 1) * L.D + d] / M) / var_estimator;
    double P = r0 + r1;
    if (k > 0) {
      P = P > 0.0 ? P : 0.0;
      run_min = (k == 1) ? P : (P < run_min ? P : run_min);
      P = run_min;
    }
    tau += 2.0 * P;
  }
  L.n_eff[d] = M * n / tau;
}
#endif  // B200MC_TU_DIAG

}  // namespace b200mc
