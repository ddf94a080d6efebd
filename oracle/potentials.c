/* Oracle, C restatement of the tall-data logistic potential and gradient.
 * TEST INFRASTRUCTURE (see oracle/__init__.py): the timed CPU baseline of bench.py and a
 * cross-check of oracle/potentials.py.  Follows examples/covtype.py:66-71 through
 * numpyro/distributions/continuous.py:2203-2207 (Normal.log_prob) and
 * numpyro/distributions/util.py:295-298 (binary_cross_entropy_with_logits), with the
 * gradient theta + X^T (sigmoid(X theta) - y) that reverse-mode AD produces.
 * (XLA's CPU program makes two passes over X; this baseline fuses them, which is faster.)
 *
 *   gcc -O3 -march=native -fopenmp -shared -fPIC potentials.c -o _build/liboracle.so -lm
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#ifdef _OPENMP
#include <omp.h>
#endif

/* torchrun exports OMP_NUM_THREADS=1; the timed CPU baseline asks for all host cores. */
void oracle_set_num_threads(int n) {
#ifdef _OPENMP
  if (n > 0) omp_set_num_threads(n);
#else
  (void)n;
#endif
}

int oracle_num_threads(void) {
#ifdef _OPENMP
  return omp_get_max_threads();
#else
  return 1;
#endif
}

/* returns U; writes grad[D].  scratch[N] receives sigmoid(l) - y (kept for callers that
 * want the residuals).  One fused pass per row block: the logits use a SIMD reduction,
 * the gradient is accumulated in float over blocks of 512 rows and flushed to double. */
float oracle_logistic_potential_and_grad(const float* X, const uint8_t* y, long N, int D,
                                         const float* theta, float* grad, float* scratch) {
  double lik = 0.0;
#ifdef _OPENMP
  const int nt = omp_get_max_threads();
#else
  const int nt = 1;
#endif
  double* part = (double*)calloc((size_t)nt * D, sizeof(double));
  const long BLK = 512;
  const long nblk = (N + BLK - 1) / BLK;
#pragma omp parallel reduction(+ : lik)
  {
#ifdef _OPENMP
    const int t = omp_get_thread_num();
#else
    const int t = 0;
#endif
    double* p = part + (size_t)t * D;
    float* acc = (float*)malloc(sizeof(float) * (size_t)D);
#pragma omp for schedule(static)
    for (long b = 0; b < nblk; ++b) {
      const long i0 = b * BLK, i1 = (i0 + BLK < N) ? i0 + BLK : N;
      for (int d = 0; d < D; ++d) acc[d] = 0.f;
      float blk_lik = 0.f;
      for (long i = i0; i < i1; ++i) {
        const float* x = X + i * D;
        float l = 0.f;
#pragma omp simd reduction(+ : l)
        for (int d = 0; d < D; ++d) l += x[d] * theta[d];
        const float e = expf(-fabsf(l));
        const float yv = (float)y[i];
        blk_lik += fmaxf(l, 0.f) + log1pf(e) - l * yv;
        const float sig = l >= 0.f ? 1.f / (1.f + e) : e / (1.f + e);
        const float g = sig - yv;
        scratch[i] = g;
#pragma omp simd
        for (int d = 0; d < D; ++d) acc[d] += g * x[d];
      }
      lik += (double)blk_lik;
      for (int d = 0; d < D; ++d) p[d] += (double)acc[d];
    }
    free(acc);
  }
  double prior = 0.0;
  for (int d = 0; d < D; ++d) prior += 0.5 * theta[d] * theta[d] + 0.9189385332046727;
  for (int d = 0; d < D; ++d) {
    double s = 0.0;
    for (int t = 0; t < nt; ++t) s += part[(size_t)t * D + d];
    grad[d] = theta[d] + (float)s;
  }
  free(part);
  return (float)(prior + lik);
}
