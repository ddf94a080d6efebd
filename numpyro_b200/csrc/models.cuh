// Closed-form potential energy U(z) and gradient for the per-chain model families
// (those whose data is small enough that one thread / warp / CTA evaluates a whole
// chain).  The tall-data logistic family has its own streaming kernel (glm_kernel.cu).
//
// Each function restates what the reference obtains by tracing the model and running
// reverse-mode AD (numpyro/infer/util.py:302-327, hmc_util.py:242-252); the log-prob
// formulas come from numpyro/distributions/continuous.py (cited per family).
// Evaluates at z (element stride zs), writes the gradient to grad, returns U in every lane.
#pragma once
#include <cmath>

#include "group.cuh"
#include "types.h"

namespace b200mc {

enum Family { FAM_LOGISTIC = 0, FAM_EIGHT_SCHOOLS = 1, FAM_HORSESHOE = 2, FAM_SV = 3 };

struct ModelData {
  int family;
  int N;           // rows / schools / time steps
  int D;           // flat unconstrained dimension
  const float* a;  // eight schools: y[J];    SV: returns[T];  horseshoe: X[N][ncols] row-major
  const float* b;  // eight schools: sigma[J];                 horseshoe: Y[N] (float)
  int ncols;       // horseshoe: number of covariates
  int likelihood;  // horseshoe: 0 Bernoulli(logits), 1 Normal(mean, prec_obs^-1/2)
  float* scratch;  // horseshoe: [C][N] residuals dU/dmean
};

#define B200MC_LOG_SQRT_2PI 0.9189385332046727f
#define B200MC_SQRT_2PI 2.5066282746310002f

// Eight schools, non-centred (README.md:101-110 + reparam.py:159-185).
// z = [mu, log tau, theta_base[J]].  Normal: continuous.py:2203-2207; HalfCauchy:
// :238-244,721-723; ExpTransform log|J| = z: transforms.py:563-571.
template <class G>
__device__ float eight_schools_potential(const ModelData& M, const G& g, const float* z, int zs,
                                         float* grad, int gs) {
  const int J = M.N;
  g.sync();  // z[0], z[1] were written by their owner lanes
  const float mu = z[0], tu = z[zs];
  const float tau = expf(tu);
  const float q = tau / 5.0f;
  const float qq = q * q;
  float lp0 = -0.5f * (mu / 5.0f) * (mu / 5.0f) - logf(B200MC_SQRT_2PI * 5.0f);
  lp0 += (0.6931471805599453f - 1.1447298858494002f - 1.6094379124341003f - log1pf(qq)) + tu;
  float lp = 0.f, s1 = 0.f, s2 = 0.f;
  for (int j = g.lane(); j < J; j += g.size()) {
    const float b = z[(2 + j) * zs];
    const float sig = M.b[j];
    const float theta = mu + tau * b;
    const float resid = (M.a[j] - theta) / sig;
    lp += (-0.5f * b * b - B200MC_LOG_SQRT_2PI) +
          (-0.5f * resid * resid - logf(B200MC_SQRT_2PI * sig));
    const float dth = -(resid / sig);
    s1 += dth;
    s2 += dth * b;
    grad[(2 + j) * gs] = b + tau * dth;
  }
  g.sum2(s1, s2);
  lp = g.sum(lp);
  if (g.lane() == 0) {
    grad[0] = mu / 25.0f + s1;
    grad[gs] = (2.0f * qq / (1.0f + qq)) - 1.0f + tau * s2;
  }
  g.sync();  // gradient elements are read back by their owner lanes
  return -(lp0 + lp);
}

// digamma(x), x > 0, in double (recurrence up to x >= 6, then the asymptotic series).
__device__ inline double digamma_f64(double x) {
  double r = 0.0;
  while (x < 6.0) {
    r -= 1.0 / x;
    x += 1.0;
  }
  const double f = 1.0 / (x * x);
  return r + log(x) - 0.5 / x -
         f * (1.0 / 12.0 - f * (1.0 / 120.0 - f * (1.0 / 252.0 - f * (1.0 / 240.0 - f / 132.0))));
}

// Stochastic volatility (examples/stochastic_volatility.py:57-65).
// z = [log nu, s[T], log sigma].  Exponential: continuous.py:473-475;
// GaussianRandomWalk: :684-689; StudentT: :2375-2385.
template <class G>
__device__ float sv_potential(const ModelData& M, const G& g, const float* z, int zs, float* grad,
                              int gs) {
  const int T = M.N;
  g.sync();  // neighbours' s[t-1], s[t+1] were written by other lanes
  const float nu_u = z[0], sg_u = z[(1 + T) * zs];
  const float nu = expf(nu_u), sg = expf(sg_u);
  const float nup1 = nu + 1.0f;
  float sum_ds2 = 0.f, sum_t = 0.f, sum_nu = 0.f, sum_l = 0.f;
  for (int e = 1 + g.lane(); e <= T; e += g.size()) {
    const int t = e - 1;
    const float st = z[e * zs];
    const float prev = (t > 0) ? z[(e - 1) * zs] : 0.0f;
    const float ds = (st - prev) / sg;
    float gsv = ds / sg;
    if (t + 1 < T) gsv -= ((z[(e + 1) * zs] - st) / sg) / sg;
    const float y = M.a[t] * expf(-st);
    const float y2n = y * y / nu;
    const float l1p = log1pf(y2n);
    const float wq = y2n / (1.0f + y2n);
    sum_ds2 += ds * ds;
    sum_t += -0.5f * nup1 * l1p - st;
    sum_l += 0.5f * l1p;
    sum_nu += 0.5f * nup1 * wq / nu;
    grad[e * gs] = gsv + (1.0f - nup1 * wq);
  }
  g.sum2(sum_ds2, sum_t);
  g.sum2(sum_l, sum_nu);
  const float fT = (float)T;
  const double hn = 0.5 * (double)nu, hn1 = 0.5 * ((double)nu + 1.0);
  const float zt = 0.5f * nu_u + 0.5723649429247001f + (float)(lgamma(hn) - lgamma(hn1));
  float lp = (3.912023005428146f - 50.0f * sg) + sg_u;    // log 50
  lp += (-2.302585092994046f - 0.1f * nu) + nu_u;         // log 0.1
  lp += -0.5f * sum_ds2 - fT * (B200MC_LOG_SQRT_2PI + sg_u);
  lp += sum_t - fT * zt;
  if (g.lane() == 0) {
    grad[(1 + T) * gs] = 50.0f * sg - 1.0f - sum_ds2 + fT;
    const float dig = (float)(digamma_f64(hn) - digamma_f64(hn1));
    const float dnu = (sum_l - sum_nu) + fT * (0.5f / nu + 0.5f * dig);
    grad[0] = 0.1f * nu - 1.0f + nu * dnu;
  }
  g.sync();
  return -lp;
}

// Horseshoe regression (examples/horseshoe_regression.py:37-78).
// z = [log lambdas[Dh], (log prec_obs,) log tau, unscaled_betas[Dh]] (sorted site names).
// HalfCauchy(1): continuous.py:703-723; Gamma(3, 1): :515-525; Normal: :2203-2207; Bernoulli
// logits: discrete.py:137-139 + distributions/util.py:295-298; exp transform log|J| = z.
// CUDA-core version: X is read twice per evaluation (X beta, then X^T residual) from L2.
template <class G>
__device__ float horseshoe_potential(const ModelData& M, const G& g, const float* z, int zs,
                                     float* grad, int gs, int chain) {
  const int Dh = M.ncols, N = M.N;
  const bool normal = M.likelihood == 1;
  const int o_l = 0, o_p = Dh, o_t = Dh + (normal ? 1 : 0), o_u = o_t + 1;
  const float c_hc = 0.6931471805599453f - 1.1447298858494002f;  // log 2 - log pi
  g.sync();
  const float tu = z[o_t * zs];
  const float tau = expf(tu);
  float pu = 0.f, prec = 1.f;
  if (normal) {
    pu = z[o_p * zs];
    prec = expf(pu);
  }
  float lp = 0.f;
  for (int d = g.lane(); d < Dh; d += g.size()) {  // priors; beta parked in grad[o_u + d]
    const float lu = z[(o_l + d) * zs], ub = z[(o_u + d) * zs];
    const float lam = expf(lu);
    lp += (c_hc - log1pf(lam * lam)) + lu;
    lp += -0.5f * ub * ub - B200MC_LOG_SQRT_2PI;
    grad[(o_u + d) * gs] = (tau * lam) * ub;
  }
  g.sync();
  float* res = M.scratch + (long)chain * N;
  float ll = 0.f, ss = 0.f;
  float sbb = 0.f;
  if constexpr (G::kBlockWide) {
    // CTA per chain: warp per data row (coalesced 128-byte lines of X) for X beta ...
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5, nw = blockDim.x >> 5;
    const float* beta = grad + (long)o_u * gs;
    for (int i = wid; i < N; i += nw) {
      const float* x = M.a + (long)i * Dh;
      float a[4] = {0.f, 0.f, 0.f, 0.f};
      int d = lane;
      for (; d + 96 < Dh; d += 128) {
#pragma unroll
        for (int k = 0; k < 4; ++k) a[k] = fmaf(__ldg(x + d + 32 * k), beta[(d + 32 * k) * gs], a[k]);
      }
      for (; d < Dh; d += 32) a[0] = fmaf(__ldg(x + d), beta[d * gs], a[0]);
      float m = (a[0] + a[1]) + (a[2] + a[3]);
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) m += __shfl_xor_sync(0xffffffffu, m, o);
      if (lane == 0) {
        const float yv = M.b[i];
        if (normal) {
          const float r = yv - m;
          ss += r * r;
          res[i] = -(prec * r);
        } else {
          const float e = expf(-fabsf(m));
          const float inv = 1.0f / (1.0f + e);
          ll += (fmaxf(m, 0.f) + log1pf(e)) - m * yv;
          res[i] = ((m >= 0.f) ? inv : e * inv) - yv;
        }
      }
    }
    g.sync();
    // ... and for dU/dbeta_d = sum_i X[i][d] res_i each thread owns up to 4 columns at a
    // time, rows in pairs => 8 coalesced loads in flight per thread
    const int T = g.size();
    for (int d0 = g.lane(); d0 < Dh; d0 += 4 * T) {
      float a[4][2];
      const float* col[4];
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        a[k][0] = 0.f; a[k][1] = 0.f;
        col[k] = M.a + ((d0 + k * T < Dh) ? d0 + k * T : d0);
      }
      int i = 0;
      for (; i + 1 < N; i += 2) {
        const float r0 = res[i], r1 = res[i + 1];
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          a[k][0] = fmaf(__ldg(col[k] + (long)i * Dh), r0, a[k][0]);
          a[k][1] = fmaf(__ldg(col[k] + (long)(i + 1) * Dh), r1, a[k][1]);
        }
      }
      if (i < N) {
        const float r0 = res[i];
#pragma unroll
        for (int k = 0; k < 4; ++k) a[k][0] = fmaf(__ldg(col[k] + (long)i * Dh), r0, a[k][0]);
      }
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const int d = d0 + k * T;
        if (d < Dh) {
          const float db = a[k][0] + a[k][1];
          const float bt = grad[(o_u + d) * gs];
          const float lu = z[(o_l + d) * zs], ub = z[(o_u + d) * zs];
          const float lam = expf(lu);
          const float l2 = lam * lam;
          grad[(o_l + d) * gs] = (2.0f * l2 / (1.0f + l2) - 1.0f) + db * bt;
          grad[(o_u + d) * gs] = ub + db * (tau * lam);
          sbb += db * bt;
        }
      }
    }
  } else {
    for (int i = g.lane(); i < N; i += g.size()) {  // mean_i = x_i . beta, dU/dmean_i
      const float* x = M.a + (long)i * Dh;
      float m0 = 0.f, m1 = 0.f;
      int d = 0;
      for (; d + 1 < Dh; d += 2) {
        m0 = fmaf(x[d], grad[(o_u + d) * gs], m0);
        m1 = fmaf(x[d + 1], grad[(o_u + d + 1) * gs], m1);
      }
      if (d < Dh) m0 = fmaf(x[d], grad[(o_u + d) * gs], m0);
      const float m = m0 + m1;
      const float yv = M.b[i];
      if (normal) {
        const float r = yv - m;
        ss += r * r;
        res[i] = -(prec * r);
      } else {
        const float e = expf(-fabsf(m));
        const float inv = 1.0f / (1.0f + e);
        ll += (fmaxf(m, 0.f) + log1pf(e)) - m * yv;
        res[i] = ((m >= 0.f) ? inv : e * inv) - yv;
      }
    }
    g.sync();
    for (int d = g.lane(); d < Dh; d += g.size()) {  // dU/dbeta_d = sum_i X[i][d] res_i
      float a0 = 0.f, a1 = 0.f;
      int i = 0;
      for (; i + 1 < N; i += 2) {
        a0 = fmaf(M.a[(long)i * Dh + d], res[i], a0);
        a1 = fmaf(M.a[(long)(i + 1) * Dh + d], res[i + 1], a1);
      }
      if (i < N) a0 = fmaf(M.a[(long)i * Dh + d], res[i], a0);
      const float db = a0 + a1;
      const float beta = grad[(o_u + d) * gs];
      const float lu = z[(o_l + d) * zs], ub = z[(o_u + d) * zs];
      const float lam = expf(lu);
      const float l2 = lam * lam;
      grad[(o_l + d) * gs] = (2.0f * l2 / (1.0f + l2) - 1.0f) + db * beta;
      grad[(o_u + d) * gs] = ub + db * (tau * lam);
      sbb += db * beta;
    }
  }
  g.sum2(lp, sbb);
  g.sum2(ll, ss);
  lp += (c_hc - log1pf(tau * tau)) + tu;
  const float fN = (float)N;
  if (normal) {
    lp += (2.0f * pu - prec - 0.6931471805599453f) + pu;  // Gamma(3,1): lgamma(3) = log 2
    lp += -0.5f * prec * ss - fN * B200MC_LOG_SQRT_2PI + 0.5f * fN * pu;
  } else {
    lp -= ll;
  }
  if (g.lane() == 0) {
    const float t2 = tau * tau;
    grad[o_t * gs] = (2.0f * t2 / (1.0f + t2) - 1.0f) + sbb;
    if (normal) grad[o_p * gs] = -(3.0f - prec) + 0.5f * prec * ss - 0.5f * fN;
  }
  g.sync();
  return -lp;
}

template <class G>
__device__ __forceinline__ float model_potential(const ModelData& M, const G& g, const float* z,
                                                 int zs, float* grad, int gs, int chain) {
  if (M.family == FAM_EIGHT_SCHOOLS) return eight_schools_potential(M, g, z, zs, grad, gs);
  if (M.family == FAM_HORSESHOE) return horseshoe_potential(M, g, z, zs, grad, gs, chain);
  return sv_potential(M, g, z, zs, grad, gs);
}

}  // namespace b200mc
