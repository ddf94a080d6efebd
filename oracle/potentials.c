/* Oracle, C restatement of the tall-data logistic potential and gradient.
 * TEST INFRASTRUCTURE (see oracle/__init__.py): the timed CPU baseline of bench.py and a
 * cross-check of oracle/potentials.py.  Follows examples/covtype.py:66-71 through
 * numpyro/distributions/continuous.py:2203-2207 (Normal.log_prob) and
 * numpyro/distributions/util.py:295-298 (binary_cross_entropy_with_logits), with the
 * gradient theta + X^T (sigmoid(X theta) - y) that reverse-mode AD produces.
 * Like XLA's CPU program it makes two passes over X (forward dot, transposed dot).
 *
 *   gcc -O3 -march=native -fopenmp -shared -fPIC potentials.c -o _build/liboracle.so -lm
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#ifdef _OPENMP
#include <omp.h>
#endif

int oracle_num_threads(void) {
#ifdef _OPENMP
  return omp_get_max_threads();
#else
  return 1;
#endif
}

/* returns U; writes grad[D].  scratch[N] holds sigmoid(l) - y between the two passes. */
float oracle_logistic_potential_and_grad(const float* X, const uint8_t* y, long N, int D,
                                         const float* theta, float* grad, float* scratch) {
  double lik = 0.0;
#pragma omp parallel for reduction(+ : lik) schedule(static)
  for (long i = 0; i < N; ++i) {
    const float* x = X + i * D;
    float l = 0.f;
    for (int d = 0; d < D; ++d) l += x[d] * theta[d];
    const float e = expf(-fabsf(l));
    const float yv = (float)y[i];
    lik += (double)(fmaxf(l, 0.f) + log1pf(e) - l * yv);
    const float sig = l >= 0.f ? 1.f / (1.f + e) : e / (1.f + e);
    scratch[i] = sig - yv;
  }
  double prior = 0.0;
  for (int d = 0; d < D; ++d) prior += 0.5 * theta[d] * theta[d] + 0.9189385332046727;
#ifdef _OPENMP
  const int nt = omp_get_max_threads();
#else
  const int nt = 1;
#endif
  double* part = (double*)calloc((size_t)nt * D, sizeof(double));
#pragma omp parallel
  {
#ifdef _OPENMP
    const int t = omp_get_thread_num();
#else
    const int t = 0;
#endif
    double* p = part + (size_t)t * D;
#pragma omp for schedule(static)
    for (long i = 0; i < N; ++i) {
      const float* x = X + i * D;
      const float g = scratch[i];
      for (int d = 0; d < D; ++d) p[d] += (double)(g * x[d]);
    }
  }
  for (int d = 0; d < D; ++d) {
    double s = 0.0;
    for (int t = 0; t < nt; ++t) s += part[(size_t)t * D + d];
    grad[d] = theta[d] + (float)s;
  }
  free(part);
  return (float)(prior + lik);
}
