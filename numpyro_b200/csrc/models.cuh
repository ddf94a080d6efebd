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

enum Family { FAM_LOGISTIC = 0, FAM_EIGHT_SCHOOLS = 1, FAM_HORSESHOE = 2, FAM_SV = 3,
              FAM_LOGISTIC_ECS = 4, FAM_EIGHT_SCHOOLS_LOCSCALE = 5,
              FAM_BETA_BERNOULLI = 6, FAM_DIRICHLET_CATEGORICAL = 7, FAM_GLM = 8 };
constexpr int kEcsSlices = 4;  // row slices of the subsample in the gradient sweep

struct ModelData {
  int family;
  int N;           // rows / schools / time steps
  int D;           // flat unconstrained dimension
  const float* a;  // eight schools: y[J];    SV: returns[T];  horseshoe: X[N][ncols] row-major
  const float* b;  // eight schools: sigma[J];                 horseshoe: Y[N] (float)
  int ncols;       // horseshoe: number of covariates
  int likelihood;  // horseshoe: 0 Bernoulli(logits), 1 Normal(mean, prec_obs^-1/2)
  float* scratch;  // horseshoe: [C][N] residuals dU/dmean; logistic ECS: [C][scratch_stride]
  // logistic with a subsampled plate (HMCECS): a = X[N][ncols], y8 = labels[N]
  const uint8_t* y8;
  const int* ecs_idx;     // [C][ecs_m] current subsample of every chain
  const float* ecs_ref;   // [D] reference point of the Taylor proxy (degree > 0)
  const float* ecs_sums;  // [1 + D + D*D] log-lik sum, its gradient and Hessian at ecs_ref, all rows
  int ecs_m, ecs_degree;
  long scratch_stride;    // 2 * ecs_m + kEcsSlices * D floats per chain
  float centered;         // eight schools under LocScaleReparam(centered)
  // Conditioned sites (HMCGibbs, hmc_gibbs.py:30-34 `condition` + `substitute`): the sampler's
  // vector holds the FREE coordinates only; the family's potential is evaluated on the full vector
  // with the conditioned coordinates held at the values of b200mc_set_conditioned.
  const int* cond_map;       // [cond_nfree] free coordinate -> index in the full vector (null: none)
  float* cond_full;          // [C][2 * cond_dfull]: full position, then full gradient
  const float* cond_offset;  // [C] added to U: the conditioned sites' Jacobian terms the
                             // reference's conditioned density does not have
  int cond_dfull, cond_nfree;
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

// Eight schools under LocScaleReparam(centered) (numpyro/infer/reparam.py:64-145) around the
// centred model theta ~ Normal(mu, tau): z = [mu, log tau, t[J]], t = theta (centered = 1) or
// theta_decentered ~ Normal(c mu, tau^c) (Normal(0, 1) for c = 0, :129-134) with
// theta = mu + tau^(1-c) (t - c mu) (:141-142).
template <class G>
__device__ float eight_schools_locscale_potential(const ModelData& M, const G& g, const float* z,
                                                  int zs, float* grad, int gs) {
  const int J = M.N;
  const float c = M.centered;
  g.sync();
  const float mu = z[0], tu = z[zs];
  const float tau = expf(tu);
  const float q = tau / 5.0f;
  const float qq = q * q;
  float lp0 = -0.5f * (mu / 5.0f) * (mu / 5.0f) - logf(B200MC_SQRT_2PI * 5.0f);
  lp0 += (0.6931471805599453f - 1.1447298858494002f - 1.6094379124341003f - log1pf(qq)) + tu;
  const float loc = (c == 0.f) ? 0.f : mu * c;
  const float s = (c == 0.f) ? 1.f : powf(tau, c);
  const float w = powf(tau, 1.0f - c);
  const float log_s = logf(s);
  float lp = 0.f, smu = 0.f, stu = 0.f;
  for (int j = g.lane(); j < J; j += g.size()) {
    const float t = z[(2 + j) * zs];
    const float sig = M.b[j];
    const float a = (t - loc) / s;
    const float delta = t - c * mu;
    const float theta = mu + w * delta;
    const float resid = (M.a[j] - theta) / sig;
    lp += ((-0.5f * a * a - log_s) - B200MC_LOG_SQRT_2PI) +
          (-0.5f * resid * resid - logf(B200MC_SQRT_2PI * sig));
    const float e = resid / sig;  // d lp / d theta_j
    smu += a * c / s + e * (1.0f - w * c);
    stu += (c * a * a - c) + (1.0f - c) * w * (e * delta);
    grad[(2 + j) * gs] = a / s - e * w;
  }
  g.sum2(smu, stu);
  lp = g.sum(lp);
  if (g.lane() == 0) {
    grad[0] = mu / 25.0f - smu;
    grad[gs] = ((2.0f * qq / (1.0f + qq)) - 1.0f) - stu;
  }
  g.sync();
  return -(lp0 + lp);
}

// Beta-Bernoulli of the reference's own test-suite (test/infer/test_mcmc.py:239-244):
// p ~ Beta(alpha, beta)[K], obs[n][K] ~ Bernoulli(p); z = logit(p).  M.a = column counts of ones
// s[K]; M.b = [K][4]: n, alpha, beta, lgamma(alpha) + lgamma(beta) - lgamma(alpha + beta).
// SigmoidTransform with clipped expit and log|J| = -softplus(z) - softplus(-z)
// (transforms.py:59-61,960-961); Beta via the Dirichlet log-density (continuous.py:193-194,
// 292-299); Bernoulli(probs) with clamped probabilities (discrete.py:90-93).
template <class G>
__device__ float beta_bernoulli_potential(const ModelData& M, const G& g, const float* z, int zs,
                                          float* grad, int gs) {
  const int K = M.N;
  float lp = 0.f;
  for (int k = g.lane(); k < K; k += g.size()) {
    const float zk = z[k * zs];
    float p = 1.0f / (1.0f + expf(-zk));
    p = fminf(fmaxf(p, 1.17549435e-38f), 1.0f - 1.1920929e-07f);
    const float q = 1.0f - p;
    const float s = M.a[k], n = M.b[4 * k], a = M.b[4 * k + 1], b = M.b[4 * k + 2], norm = M.b[4 * k + 3];
    const float sp_pos = fmaxf(zk, 0.f) + log1pf(expf(-fabsf(zk)));   // softplus(z)
    const float sp_neg = fmaxf(-zk, 0.f) + log1pf(expf(-fabsf(zk)));  // softplus(-z)
    const float lgp = logf(p);
    lp += ((lgp * (a - 1.0f) + logf(q) * (b - 1.0f)) - norm) + (-sp_pos - sp_neg) +
          (s * lgp + (n - s) * log1pf(-p));
    grad[k * gs] = -((a - 1.0f) * q - (b - 1.0f) * p + (1.0f - 2.0f * p) + s * q - (n - s) * p);
  }
  g.sync();
  return -g.sum(lp);
}

// Dirichlet-Categorical of the reference's test-suite (test/infer/test_mcmc.py:283-288):
// p ~ Dirichlet(c[K]), obs[n] ~ Categorical(p); z[K-1] through StickBreakingTransform
// (transforms.py:1090-1122).  M.a = class counts[K]; M.b = c[K] followed by the Dirichlet
// normaliser sum lgamma(c) - lgamma(sum c).  K is small: lane 0 walks the stick.
template <class G>
__device__ float dirichlet_categorical_potential(const ModelData& M, const G& g, const float* z,
                                                 int zs, float* grad, int gs) {
  const int K = M.N;
  g.sync();
  float lp = 0.f;
  if (g.lane() == 0) {
    // forward: sticks s_i = clipped expit(z_i - log(K-1-i)), p_i = s_i prod_{j<i}(1 - s_j)
    float rem = 1.0f, tail = 0.f;
    for (int k = 0; k < K; ++k) {
      float s = 1.0f, ls_minus_x = 0.f;
      if (k < K - 1) {
        const float x = z[k * zs] - logf((float)(K - 1 - k));
        s = 1.0f / (1.0f + expf(-x));
        s = fminf(fmaxf(s, 1.17549435e-38f), 1.0f - 1.1920929e-07f);
        ls_minus_x = -(fmaxf(-x, 0.f) + log1pf(expf(-fabsf(x)))) - x;  // log_sigmoid(x) - x
      }
      const float logp = logf(s * rem);
      lp += logp * (M.b[k] - 1.0f) + M.a[k] * fmaxf(logp, -3.40282347e+38f);
      if (k < K - 1) lp += logp + ls_minus_x;
      rem *= 1.0f - s;
    }
    lp -= M.b[K];
    // backward: d lp / d x_i = w_i (1 - s_i) - s_i (sum_{k>i} w_k + 1), w_k = c_k - 1 + count_k + [k < K-1]
    for (int k = K - 1; k >= 0; --k) {
      const float w = M.b[k] - 1.0f + M.a[k] + ((k < K - 1) ? 1.0f : 0.f);
      if (k < K - 1) {
        const float x = z[k * zs] - logf((float)(K - 1 - k));
        float s = 1.0f / (1.0f + expf(-x));
        s = fminf(fmaxf(s, 1.17549435e-38f), 1.0f - 1.1920929e-07f);
        grad[k * gs] = -(w * (1.0f - s) - s * tail - s);
      }
      tail += w;
    }
  }
  g.sync();
  return -g.sum(lp);
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

// Logistic regression with a subsampled plate under HMCECS (examples/covtype.py:66-71 with
// subsample_size; numpyro/infer/hmc_gibbs.py:502-665).  idx[m] is the chain's subsample.
//   degree 0: U = -logN(theta) + (n/m) sum_i bce_i               (plate scale, no proxy)
//   degree 1/2: estimate_likelihood + perturbed_method (contrib/ecs_proxies.py:23-52) with the
//   Taylor proxy (:232-274): U = -logN(theta) - [proxy_sum(theta) + n mean(diff) - n^2/(2m) var(diff)],
//   diff_i = loglik_i(theta) - proxy_i(theta); for Bernoulli(logits = x_i.theta) the per-row proxy is
//   loglik_i* + (y_i - s_i*) d_i - 1/2 s_i*(1 - s_i*) d_i^2 with d_i = x_i.(theta - theta*), and
//   proxy_sum = A + b.dt + 1/2 dt^T H dt with the all-row sums A, b, H of b200mc_create.
template <class G>
__device__ float logistic_ecs_potential(const ModelData& M, const G& g, const float* z, int zs,
                                        float* grad, int gs, int chain, const int* idx,
                                        const float* rows = nullptr, int fresh0 = 0,
                                        int fresh1 = 0) {
  // rows: the subsample's rows staged in shared memory ([m][ncols], row i = X[idx[i]]; the
  // device-resident HMCECS loop) or null (gathered from X through idx); rows [fresh0, fresh1) of
  // the subsample differ from the staged copy (a proposal) and are gathered
  auto rowp = [&](int i) -> const float* {
    return (rows && !(i >= fresh0 && i < fresh1)) ? rows + (long)i * M.ncols
                                                  : M.a + (long)idx[i] * M.ncols;
  };
  const int D = M.D, m = M.ecs_m, deg = M.ecs_degree, lane = g.lane(), size = g.size();
  const float nf = (float)M.N, mf = (float)m, scale = nf / mf;
  float* coef = M.scratch + (long)chain * M.scratch_stride;
  float* dif = coef + m;
  float* part = dif + m;
  const float* ref = M.ecs_ref;
  g.sync();
  float prior = 0.f, s1 = 0.f;
  for (int d = lane; d < D; d += size) {
    const float t = z[d * zs];
    prior += 0.5f * t * t + B200MC_LOG_SQRT_2PI;
  }
  for (int i = lane; i < m; i += size) {
    const int row = idx[i];
    const float* x = rowp(i);
    float l = 0.f, ls = 0.f, dl = 0.f;
    // (the loops of this function are load-latency bound on one CTA: unrolled so that several
    // loads are in flight per thread; the order of every sum is unchanged)
#pragma unroll 11
    for (int d = 0; d < D; ++d) {
      const float xv = x[d], zv = z[d * zs];
      l = fmaf(xv, zv, l);
      if (deg) {
        const float rv = ref[d];
        ls = fmaf(xv, rv, ls);
        dl = fmaf(xv, zv - rv, dl);
      }
    }
    const float yv = (float)M.y8[row];
    const float e = expf(-fabsf(l));
    const float bce = (fmaxf(l, 0.f) + log1pf(e)) - l * yv;  // distributions/util.py:295-298
    const float inv = 1.0f / (1.0f + e);
    const float sig = (l >= 0.f) ? inv : e * inv;
    if (!deg) {
      s1 += bce;
      coef[i] = scale * (yv - sig);
    } else {
      const float es = expf(-fabsf(ls));
      const float bces = (fmaxf(ls, 0.f) + log1pf(es)) - ls * yv;
      const float invs = 1.0f / (1.0f + es);
      const float sigs = (ls >= 0.f) ? invs : es * invs;
      float proxy = -bces + (yv - sigs) * dl;
      float dproxy = yv - sigs;
      if (deg == 2) {
        const float w = sigs * (1.0f - sigs);
        proxy -= 0.5f * w * dl * dl;
        dproxy -= w * dl;
      }
      const float diff = -bce - proxy;
      dif[i] = diff;
      coef[i] = (yv - sig) - dproxy;  // d diff_i / d logit_i
      s1 += diff;
    }
  }
  g.sum2(prior, s1);
  float U;
  if (!deg) {
    U = prior + scale * s1;
  } else {
    const float mean = s1 / mf;
    float s2 = 0.f, ps = 0.f;
    for (int i = lane; i < m; i += size) {  // this lane's own entries
      const float dv = dif[i] - mean;
      s2 += dv * dv;
      coef[i] *= scale - (scale * scale) * dv;
    }
    const float* bsum = M.ecs_sums + 1;
    const float* H = bsum + D;
    for (int d = lane; d < D; d += size) {
      const float dz = z[d * zs] - ref[d];
      float hd = 0.f;
      if (deg == 2) {
#pragma unroll 11
        for (int k = 0; k < D; ++k) hd = fmaf(H[(long)d * D + k], z[k * zs] - ref[k], hd);
      }
      grad[d * gs] = bsum[d] + hd;  // gradient of proxy_sum, folded in below
      ps += bsum[d] * dz + 0.5f * hd * dz;
    }
    g.sum2(s2, ps);
    const float var = s2 / mf;  // jnp.var: population variance
    const float loglik = (M.ecs_sums[0] + ps) + nf * mean - 0.5f * (nf * nf / mf) * var;
    U = prior - loglik;
  }
  g.sync();  // coef[] complete
  for (int t = lane; t < kEcsSlices * D; t += size) {
    const int sl = t / D, d = t - sl * D;
    float a = 0.f;
#pragma unroll 8
    for (int i = sl; i < m; i += kEcsSlices) a = fmaf(coef[i], rowp(i)[d], a);
    part[t] = a;
  }
  g.sync();
  for (int d = lane; d < D; d += size) {
    float a = 0.f;
    for (int sl = 0; sl < kEcsSlices; ++sl) a += part[sl * D + d];  // fixed order
    const float gp = deg ? grad[d * gs] : 0.f;
    grad[d * gs] = z[d * zs] - gp - a;
  }
  g.sync();
  return U;
}

// The same potential for ONE 256-thread CTA with the subsample's rows staged in shared memory
// (ecs_chain_kernel; D <= 64, m <= 2048), written for latency: with 215 KB of rows staged the SM has
// ~24 KB of L1 left, so every global access of the generic form above is an L2 round trip and the
// per-row coefficients it parks in global memory are re-read 55 times.  Here a thread keeps its
// rows' labels, differences and coefficients in registers, accumulates its rows' contribution to ALL
// columns in registers (a second pass over its rows in shared memory) and the CTA combines them with
// two 31-shuffle transpose-reduces per warp and one fixed-order sum over the warps; H (theta - theta*)
// is a warp-per-row sweep with coalesced loads.  part: 8 x 64 floats of shared memory.
constexpr int kEcsCtaRows = 8;   // rows per thread (m <= 8 * 256)
constexpr int kEcsCtaCols = 64;  // columns (D <= 64)

__device__ __forceinline__ float ecs_transpose_reduce32(float (&a)[32]) {
  const int lane = threadIdx.x & 31;
#pragma unroll
  for (int s = 16; s >= 1; s >>= 1) {
    const bool upper = (lane & s) != 0;
#pragma unroll
    for (int i = 0; i < s; ++i) {
      const float send = upper ? a[i] : a[i + s];
      const float keep = upper ? a[i + s] : a[i];
      a[i] = keep + __shfl_xor_sync(0xffffffffu, send, s);
    }
  }
  return a[0];
}

__device__ inline float logistic_ecs_potential_cta(const ModelData& M, const BlockGroup& g, const float* z,
                                            float* grad, int chain, const int* idx, const float* rows,
                                            int fresh0, int fresh1, float* part) {
  const int D = M.D, m = M.ecs_m, deg = M.ecs_degree, tid = threadIdx.x, nthr = blockDim.x;
  const int warp = tid >> 5, lane = tid & 31, nwarp = nthr >> 5;
  const float nf = (float)M.N, mf = (float)m, scale = nf / mf;
  const float* ref = M.ecs_ref;
  auto rowp = [&](int i) -> const float* {
    return (i >= fresh0 && i < fresh1) ? M.a + (long)idx[i] * M.ncols : rows + (long)i * M.ncols;
  };
  __syncthreads();
  // labels of this thread's rows: independent loads, all in flight together
  float yv[kEcsCtaRows];
  {
    int rw[kEcsCtaRows];
#pragma unroll
    for (int k = 0; k < kEcsCtaRows; ++k) {
      const int i = tid + k * nthr;
      rw[k] = (i < m) ? idx[i] : 0;
    }
#pragma unroll
    for (int k = 0; k < kEcsCtaRows; ++k) yv[k] = (tid + k * nthr < m) ? (float)M.y8[rw[k]] : 0.f;
  }
  float prior = 0.f, s1 = 0.f;
  for (int d = tid; d < D; d += nthr) {
    const float t = z[d];
    prior += 0.5f * t * t + B200MC_LOG_SQRT_2PI;
  }
  float dif[kEcsCtaRows], cf[kEcsCtaRows];
#pragma unroll
  for (int k = 0; k < kEcsCtaRows; ++k) {
    const int i = tid + k * nthr;
    dif[k] = 0.f;
    cf[k] = 0.f;
    if (i < m) {
      const float* x = rowp(i);
      float l = 0.f, ls = 0.f, dl = 0.f;
#pragma unroll 11
      for (int d = 0; d < D; ++d) {
        const float xv = x[d], zv = z[d];
        l = fmaf(xv, zv, l);
        if (deg) {
          const float rv = ref[d];
          ls = fmaf(xv, rv, ls);
          dl = fmaf(xv, zv - rv, dl);
        }
      }
      const float e = expf(-fabsf(l));
      const float bce = (fmaxf(l, 0.f) + log1pf(e)) - l * yv[k];  // distributions/util.py:295-298
      const float inv = 1.0f / (1.0f + e);
      const float sig = (l >= 0.f) ? inv : e * inv;
      if (!deg) {
        s1 += bce;
        cf[k] = scale * (yv[k] - sig);
      } else {
        const float es = expf(-fabsf(ls));
        const float bces = (fmaxf(ls, 0.f) + log1pf(es)) - ls * yv[k];
        const float invs = 1.0f / (1.0f + es);
        const float sigs = (ls >= 0.f) ? invs : es * invs;
        float proxy = -bces + (yv[k] - sigs) * dl;
        float dproxy = yv[k] - sigs;
        if (deg == 2) {
          const float w = sigs * (1.0f - sigs);
          proxy -= 0.5f * w * dl * dl;
          dproxy -= w * dl;
        }
        dif[k] = -bce - proxy;
        cf[k] = (yv[k] - sig) - dproxy;  // d diff_i / d logit_i
        s1 += dif[k];
      }
    }
  }
  g.sum2(prior, s1);
  float U;
  if (!deg) {
    U = prior + scale * s1;
  } else {
    const float mean = s1 / mf;
    float s2 = 0.f, ps = 0.f;
#pragma unroll
    for (int k = 0; k < kEcsCtaRows; ++k) {
      if (tid + k * nthr < m) {
        const float dv = dif[k] - mean;
        s2 += dv * dv;
        cf[k] *= scale - (scale * scale) * dv;
      }
    }
    // proxy_sum = A + b.dt + 1/2 dt^T H dt; its gradient b + H dt goes to grad[] (folded in below).
    // H dt: warp w takes rows w, w + nwarp, ... of H, lanes over the columns (coalesced)
    const float* bsum = M.ecs_sums + 1;
    const float* H = bsum + D;
    const float dz0 = (lane < D) ? z[lane] - ref[lane] : 0.f;
    const float dz1 = (lane + 32 < D) ? z[lane + 32] - ref[lane + 32] : 0.f;
    for (int d = warp; d < D; d += nwarp) {
      float hd = 0.f;
      if (deg == 2) {
        const float h0 = (lane < D) ? H[(long)d * D + lane] : 0.f;
        const float h1 = (lane + 32 < D) ? H[(long)d * D + lane + 32] : 0.f;
        hd = fmaf(h0, dz0, h1 * dz1);
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) hd += __shfl_xor_sync(0xffffffffu, hd, o);
      }
      if (lane == 0) {
        const float dzd = z[d] - ref[d];
        grad[d] = bsum[d] + hd;
        ps += bsum[d] * dzd + 0.5f * hd * dzd;
      }
    }
    g.sum2(s2, ps);
    const float var = s2 / mf;  // jnp.var: population variance
    const float loglik = (M.ecs_sums[0] + ps) + nf * mean - 0.5f * (nf * nf / mf) * var;
    U = prior - loglik;
  }
  // gradient: sum_i cf_i x_i, this thread's rows into registers, then across the CTA
  float acc[kEcsCtaCols];
#pragma unroll
  for (int d = 0; d < kEcsCtaCols; ++d) acc[d] = 0.f;
#pragma unroll
  for (int k = 0; k < kEcsCtaRows; ++k) {
    const int i = tid + k * nthr;
    if (i < m) {
      const float* x = rowp(i);
      const float c = cf[k];
#pragma unroll
      for (int d = 0; d < kEcsCtaCols; ++d)
        if (d < D) acc[d] = fmaf(c, x[d], acc[d]);
    }
  }
#pragma unroll
  for (int half = 0; half < 2; ++half) {
    float a[32];
#pragma unroll
    for (int i = 0; i < 32; ++i) a[i] = acc[32 * half + i];
    const float v = ecs_transpose_reduce32(a);
    part[warp * kEcsCtaCols + 32 * half + lane] = v;
  }
  __syncthreads();
  for (int d = tid; d < D; d += nthr) {
    float a = 0.f;
    for (int w = 0; w < nwarp; ++w) a += part[w * kEcsCtaCols + d];  // fixed order
    const float gp = deg ? grad[d] : 0.f;
    grad[d] = z[d] - gp - a;
  }
  __syncthreads();
  return U;
}

// Generalised linear models on a dense design matrix: coefs ~ Normal(0, 1)[ncols],
//   likelihood 0: obs ~ Poisson(rate = exp(X coefs))      (discrete.py:743: log(rate) y - lgamma(y + 1) - rate)
//   likelihood 1: prec_obs ~ Gamma(3, 1), obs ~ Normal(X coefs, prec_obs^-1/2)   (the horseshoe example's
//                 observation model, horseshoe_regression.py:60-66, without the shrinkage prior)
// z = [coefs (, log prec_obs)].  a = X[N][ncols] row-major, b = y[N] followed by ONE float: the sum
// over rows of lgamma(y + 1) (Poisson; constant in z, computed once by the binding).  Rows are
// swept like the horseshoe's: residuals dU/dmean_i parked in the chain's scratch row, then the
// transposed sweep.
template <class G>
__device__ float glm_potential(const ModelData& M, const G& g, const float* z, int zs, float* grad,
                               int gs, int chain) {
  const int Dc = M.ncols, N = M.N;
  const bool normal = M.likelihood == 1;
  g.sync();
  float pu = 0.f, prec = 1.f;
  if (normal) {
    pu = z[(long)Dc * zs];
    prec = expf(pu);
  }
  float lp = 0.f;
  for (int d = g.lane(); d < Dc; d += g.size()) {
    const float t = z[(long)d * zs];
    lp += -0.5f * t * t - B200MC_LOG_SQRT_2PI;
  }
  float* res = M.scratch + (long)chain * N;
  float lt = 0.f;  // Poisson: sum of rate - mean y; Normal: sum of squared residuals
  auto row_term = [&](int i, float m) {
    const float yv = M.b[i];
    if (normal) {
      const float r = yv - m;
      lt += r * r;
      res[i] = -(prec * r);
    } else {
      const float e = expf(m);
      lt += e - m * yv;
      res[i] = e - yv;
    }
  };
  if constexpr (G::kBlockWide) {
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5, nw = blockDim.x >> 5;
    for (int i = wid; i < N; i += nw) {  // warp per row: coalesced lines of X
      const float* x = M.a + (long)i * Dc;
      float m = 0.f;
      for (int d = lane; d < Dc; d += 32) m = fmaf(__ldg(x + d), z[(long)d * zs], m);
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) m += __shfl_xor_sync(0xffffffffu, m, o);
      if (lane == 0) row_term(i, m);
    }
  } else {
    for (int i = g.lane(); i < N; i += g.size()) {
      const float* x = M.a + (long)i * Dc;
      float m = 0.f;
      for (int d = 0; d < Dc; ++d) m = fmaf(x[d], z[(long)d * zs], m);
      row_term(i, m);
    }
  }
  g.sync();
  for (int d = g.lane(); d < Dc; d += g.size()) {  // dU/dcoefs_d = coefs_d + sum_i X[i][d] res_i
    const float* col = M.a + d;
    float a0 = 0.f, a1 = 0.f;
    int i = 0;
    for (; i + 1 < N; i += 2) {
      a0 = fmaf(__ldg(col + (long)i * Dc), res[i], a0);
      a1 = fmaf(__ldg(col + (long)(i + 1) * Dc), res[i + 1], a1);
    }
    if (i < N) a0 = fmaf(__ldg(col + (long)i * Dc), res[i], a0);
    grad[(long)d * gs] = z[(long)d * zs] + (a0 + a1);
  }
  g.sum2(lp, lt);
  const float fN = (float)N;
  if (normal) {
    lp += (2.0f * pu - prec - 0.6931471805599453f) + pu;  // Gamma(3, 1) + log|J|
    lp += -0.5f * prec * lt - fN * B200MC_LOG_SQRT_2PI + 0.5f * fN * pu;
    if (g.lane() == 0) grad[(long)Dc * gs] = -(3.0f - prec) + 0.5f * prec * lt - 0.5f * fN;
  } else {
    lp -= lt + M.b[N];
  }
  g.sync();
  return -lp;
}

template <class G>
__device__ __forceinline__ float family_potential(const ModelData& M, const G& g, const float* z,
                                                  int zs, float* grad, int gs, int chain,
                                                  const float* ecs_rows = nullptr) {
  if (M.family == FAM_EIGHT_SCHOOLS) return eight_schools_potential(M, g, z, zs, grad, gs);
  if (M.family == FAM_HORSESHOE) return horseshoe_potential(M, g, z, zs, grad, gs, chain);
  if (M.family == FAM_DIRICHLET_CATEGORICAL) return dirichlet_categorical_potential(M, g, z, zs, grad, gs);
  if (M.family == FAM_GLM) return glm_potential(M, g, z, zs, grad, gs, chain);
  if (M.family == FAM_BETA_BERNOULLI) return beta_bernoulli_potential(M, g, z, zs, grad, gs);
  if (M.family == FAM_EIGHT_SCHOOLS_LOCSCALE) return eight_schools_locscale_potential(M, g, z, zs, grad, gs);
  if constexpr (G::kBlockWide) {  // HMCECS handles run CTA per chain only (api.cu)
    if (M.family == FAM_LOGISTIC_ECS)
      return logistic_ecs_potential(M, g, z, zs, grad, gs, chain, M.ecs_idx + (long)chain * M.ecs_m,
                                    ecs_rows);
  }
  return sv_potential(M, g, z, zs, grad, gs);
}

template <class G>
__device__ __forceinline__ float model_potential(const ModelData& M, const G& g, const float* z,
                                                 int zs, float* grad, int gs, int chain,
                                                 const float* ecs_rows = nullptr) {
  if (M.cond_map == nullptr) return family_potential(M, g, z, zs, grad, gs, chain, ecs_rows);
  // conditioned sites: scatter the free coordinates into the chain's full vector, evaluate the
  // family there, gather the free coordinates of the gradient
  float* zf = M.cond_full + (long)chain * 2 * M.cond_dfull;
  float* gf = zf + M.cond_dfull;
  for (int d = g.lane(); d < M.cond_nfree; d += g.size()) zf[M.cond_map[d]] = z[(long)d * zs];
  g.sync();
  const float U = family_potential(M, g, zf, 1, gf, 1, chain, ecs_rows);
  g.sync();
  for (int d = g.lane(); d < M.cond_nfree; d += g.size()) grad[(long)d * gs] = gf[M.cond_map[d]];
  g.sync();
  return U + M.cond_offset[chain];
}

}  // namespace b200mc
