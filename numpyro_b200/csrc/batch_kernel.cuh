// Many-chain lock-step engine: once >= 64 chains share one design matrix, the per-chain
// matrix-vector sweeps of the potential become two dense contractions over ALL chains,
//     L = X . B^T   (N x C)      and      G = X^T . R   (Dx x C),
// run on the tensor cores (gemm3x_kernel.cuh) once per leapfrog for the whole batch.  Every
// chain still runs the same step machine as everywhere else (sampler.cuh): one global step =
//   [GEMM 1 + residual epilogue] -> [GEMM 2, split-K] -> [batch_step_kernel: one CTA per chain
//    finishes its potential (chain rule, priors), advances its sampler by one gradient and
//    writes the operand row of its next point].
// Chains are NOT synchronised at transition boundaries: a chain that ends a transition starts
// the next one in the same step, so no lane of the batch idles until its own last transition.
// The sampler object lives in global memory between launches (a plain-data snapshot).
#pragma once
#include "chain_kernel.cuh"
#include "umma.cuh"

namespace b200mc {

constexpr int kBatchThreads = 512;
constexpr int kBatchCols = 256;  // chains per operand image = columns of one GEMM tile

struct BatchLaunch {
  SamplerParams P;
  ModelData M;
  CollectSpec col;
  ChainScalars* cs;
  ChainVectors cv;
  float* work;  // [C][num_slots][D]
  int C;
  const uint32_t* init_keys;
  int user_init;
  int begin;            // 1: first launch of a b200mc_init / b200mc_sample call
  unsigned char* snap;  // [C][snap_stride] sampler snapshots
  int snap_stride;
  int* status;          // [C] 0 = waiting for a gradient, 1 = done
  unsigned* active;     // chains still waiting for a gradient after this launch
  // potential plumbing (one column group of kBatchCols chains)
  unsigned char *b_hi, *b_lo;  // images of B [256][K1]: the linear-predictor coefficients
  int K1;
  const float* gpart;   // [splits][256][RA2] split-K partials of X^T . R
  int splits, RA2;
  const float* ll_part; // [n_rowgroups][256] partial sums of the likelihood terms
  int n_rowgroups;
  float* col_scale;     // [256] Normal likelihood: precision per chain
  // evaluation hook (b200mc_potential_and_grad): points instead of sampler states
  const float* eval_z;  // [C][D]
  float* eval_U;
  float* eval_grad;
};

// ---- horseshoe regression (examples/horseshoe_regression.py:37-78), split around the GEMMs.
// Same arithmetic as horseshoe_potential (models.cuh); z, grad have unit stride.
template <class G>
__device__ void hs_prepare(const BatchLaunch& L, const G& g, const float* z, int c) {
  const ModelData& M = L.M;
  const int Dh = M.ncols;
  const bool normal = M.likelihood == 1;
  const int o_l = 0, o_p = Dh, o_t = Dh + (normal ? 1 : 0), o_u = o_t + 1;
  const float tau = expf(z[o_t]);
  float* bh = reinterpret_cast<float*>(L.b_hi);
  float* bl = reinterpret_cast<float*>(L.b_lo);
  for (int d = g.lane(); d < Dh; d += g.size()) {
    const float lam = expf(z[o_l + d]);
    const float beta = (tau * lam) * z[o_u + d];
    float hi, lo;
    umma::split_tf32(beta, hi, lo);
    const long off = umma::image_offset(kBatchCols, c, d) >> 2;
    bh[off] = hi;
    bl[off] = lo;
  }
  if (g.lane() == 0) L.col_scale[c] = normal ? expf(z[o_p]) : 1.0f;
}

template <class G>
__device__ float hs_finish(const BatchLaunch& L, const G& g, const float* z, float* grad, int c) {
  const ModelData& M = L.M;
  const int Dh = M.ncols, N = M.N;
  const bool normal = M.likelihood == 1;
  const int o_l = 0, o_p = Dh, o_t = Dh + (normal ? 1 : 0), o_u = o_t + 1;
  const float c_hc = 0.6931471805599453f - 1.1447298858494002f;  // log 2 - log pi
  const float tu = z[o_t];
  const float tau = expf(tu);
  float pu = 0.f, prec = 1.f;
  if (normal) {
    pu = z[o_p];
    prec = expf(pu);
  }
  float lp = 0.f, sbb = 0.f;
  for (int d = g.lane(); d < Dh; d += g.size()) {
    const float lu = z[o_l + d], ub = z[o_u + d];
    const float lam = expf(lu);
    lp += (c_hc - log1pf(lam * lam)) + lu;
    lp += -0.5f * ub * ub - B200MC_LOG_SQRT_2PI;
    const float beta = (tau * lam) * ub;
    float db = 0.f;  // dU/dbeta_d = sum_i X[i][d] res_i: split-K partials in a fixed order
    for (int s = 0; s < L.splits; ++s) db += L.gpart[((long)s * kBatchCols + c) * L.RA2 + d];
    const float l2 = lam * lam;
    grad[o_l + d] = (2.0f * l2 / (1.0f + l2) - 1.0f) + db * beta;
    grad[o_u + d] = ub + db * (tau * lam);
    sbb += db * beta;
  }
  float lt = 0.f;  // sum of BCE terms, or of squared residuals
  for (int r = g.lane(); r < L.n_rowgroups; r += g.size()) lt += L.ll_part[(long)r * kBatchCols + c];
  g.sum2(lp, sbb);
  float dummy = 0.f;
  g.sum2(lt, dummy);
  lp += (c_hc - log1pf(tau * tau)) + tu;
  const float fN = (float)N;
  if (normal) {
    lp += (2.0f * pu - prec - 0.6931471805599453f) + pu;  // Gamma(3,1): lgamma(3) = log 2
    lp += -0.5f * prec * lt - fN * B200MC_LOG_SQRT_2PI + 0.5f * fN * pu;
  } else {
    lp -= lt;
  }
  if (g.lane() == 0) {
    const float t2 = tau * tau;
    grad[o_t] = (2.0f * t2 / (1.0f + t2) - 1.0f) + sbb;
    if (normal) grad[o_p] = -(3.0f - prec) + 0.5f * prec * lt - 0.5f * fN;
  }
  g.sync();
  return -lp;
}

// One CTA per chain.  begin = 1: set the sampler up and publish its first point.  Otherwise:
// finish the potential at the published point, advance, publish the next point.
__global__ void __launch_bounds__(kBatchThreads) batch_step_kernel(const __grid_constant__ BatchLaunch L) {
  __shared__ float red_scratch[64];
  const int c = blockIdx.x;
  if (c >= L.C) return;
  const BlockGroup g{red_scratch};
  const int D = L.P.D;
  using S_t = Sampler<BlockGroup>;
  S_t* snap = reinterpret_cast<S_t*>(L.snap + (long)c * L.snap_stride);
  float* base = L.work + (long)c * num_slots(L.P.max_depth, L.P.dense_mass) * D;
  S_t S;
  StepResult r;
  if (L.begin) {
    S.setup(&L.P, g, base, D, 1, c, L.C, &L.col);
    S.load(L.cs, L.cv);
    if (L.init_keys)
      r = S.begin_init(Key{L.init_keys[2 * c], L.init_keys[2 * c + 1]}, L.user_init != 0);
    else
      r = S.begin_sampling();
  } else {
    if (L.status[c] != 0) return;
    S = *snap;
    S.P = &L.P; S.g = g; S.w = base; S.col = &L.col;
    if (S.dense) S.dn = L.cv.dense + (long)c * 4 * D * D;
    const float U = hs_finish(L, g, S.v(LZ), S.v(LG), c);
    r = S.advance(U);
  }
  if (r == NEED_GRAD) {
    g.sync();
    hs_prepare(L, g, S.v(LZ), c);
    if (threadIdx.x == 0) {
      L.status[c] = 0;
      atomicAdd(L.active, 1u);
    }
  } else {
    S.store(L.cs, L.cv);
    if (threadIdx.x == 0) L.status[c] = 1;
  }
  if (threadIdx.x == 0) *snap = S;
}

// Evaluation hook: phase 0 publishes the operand rows of the points eval_z[c], phase 1 turns
// the GEMM outputs into U and the gradient.
__global__ void __launch_bounds__(kBatchThreads) batch_potential_kernel(const __grid_constant__ BatchLaunch L,
                                                                        int phase) {
  __shared__ float red_scratch[64];
  const int c = blockIdx.x;
  if (c >= L.C) return;
  const BlockGroup g{red_scratch};
  const float* z = L.eval_z + (long)c * L.P.D;
  if (phase == 0) {
    hs_prepare(L, g, z, c);
  } else {
    const float U = hs_finish(L, g, z, L.eval_grad + (long)c * L.P.D, c);
    if (threadIdx.x == 0) L.eval_U[c] = U;
  }
}

}  // namespace b200mc
