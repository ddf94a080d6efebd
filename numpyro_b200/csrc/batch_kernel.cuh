// Many-chain lock-step engine: once >= 64 chains share one design matrix, the per-chain
// matrix-vector sweeps of the potential become two dense contractions over ALL chains,
//     L = X . B^T   (N x C)      and      G = X^T . R   (Dx x C),
// run on the tensor cores (gemm3x_kernel.cuh) once per leapfrog for the whole batch.  Every
// chain still runs the same step machine as everywhere else (sampler.cuh): one global step =
//   [GEMM 1 + residual epilogue] -> [GEMM 2, split-K] -> [batch_step_kernel: one CTA per chain
//    finishes its potential (chain rule, priors), advances its sampler by one gradient and
//    writes the operand row of its next point].
// With a dense mass matrix the sampler does not sweep its D x D matrices itself (one CTA
// cannot keep enough loads in flight): it yields a MatvecRequest, batch_matvec_kernel serves
// the requests of all chains at once with the whole machine (HBM-bound: D*D*4 bytes per
// chain per leapfrog, both right-hand sides of a leapfrog in one sweep), and a resume launch
// of batch_step_kernel continues the step machine.
// Chains are NOT synchronised at transition boundaries: a chain that ends a transition starts
// the next one in the same step, so no lane of the batch idles until its own last transition.
// The sampler object lives in global memory between launches (a plain-data snapshot).
#pragma once
#include "chain_kernel.cuh"
#include "umma.cuh"

namespace b200mc {

constexpr int kBatchThreads = 512;
constexpr int kBatchCols = 256;  // chains per operand image = columns of one GEMM tile
enum BatchKind { BATCH_BEGIN = 0, BATCH_GRAD = 1, BATCH_RESUME = 2 };
constexpr int kMatvecThreads = 256;
constexpr int kMatvecRows = 64;  // rows of one chain's matrix per CTA

struct BatchLaunch {
  SamplerParams P;
  ModelData M;
  CollectSpec col;
  ChainScalars* cs;
  ChainVectors cv;
  float* work;  // [C][num_slots][D]
  int C;                // chains of THIS column group (<= kBatchCols): the launch's grid
  int c0, C_total;      // first chain of the group / all chains of the handle: chain c0 + i is
                        // column i of the operand images
  const uint32_t* init_keys;
  int user_init;
  int kind;             // BATCH_BEGIN / BATCH_GRAD / BATCH_RESUME
  MatvecRequest* reqs;  // [C] pending dense-mass sweeps (status 2)
  unsigned char* snap;  // [C][snap_stride] sampler snapshots
  int snap_stride;
  int* status;          // [C] 0 = waiting for a gradient, 1 = done, 2 = waiting for a sweep,
                        // 3 = waiting for the window-end factorisation (host side)
  unsigned* active;     // chains not done after this launch
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

// ---- Bayesian logistic regression (examples/covtype.py:66-71): coefs ~ N(0, 1),
// y ~ Bernoulli(logits = X coefs).  Same arithmetic as the streaming kernel (glm_kernel.cuh).
template <class G>
__device__ void lg_prepare(const BatchLaunch& L, const G& g, const float* z, int c) {
  float* bh = reinterpret_cast<float*>(L.b_hi);
  float* bl = reinterpret_cast<float*>(L.b_lo);
  for (int d = g.lane(); d < L.P.D; d += g.size()) {
    float hi, lo;
    umma::split_tf32(z[d], hi, lo);
    const long off = umma::image_offset(kBatchCols, c, d) >> 2;
    bh[off] = hi;
    bl[off] = lo;
  }
  if (g.lane() == 0) L.col_scale[c] = 1.0f;
}

template <class G>
__device__ float lg_finish(const BatchLaunch& L, const G& g, const float* z, float* grad, int c) {
  float prior = 0.f, llik = 0.f;
  for (int d = g.lane(); d < L.P.D; d += g.size()) {
    const float t = z[d];
    float db = 0.f;
    for (int s = 0; s < L.splits; ++s) db += L.gpart[((long)s * kBatchCols + c) * L.RA2 + d];
    grad[d] = t + db;                              // grad U = theta + X^T (sigmoid - y)
    prior += 0.5f * t * t + B200MC_LOG_SQRT_2PI;   // -log N(theta; 0, 1)
  }
  for (int r = g.lane(); r < L.n_rowgroups; r += g.size()) llik += L.ll_part[(long)r * kBatchCols + c];
  g.sum2(prior, llik);
  g.sync();
  return prior + llik;
}

template <class G>
__device__ __forceinline__ void batch_prepare(const BatchLaunch& L, const G& g, const float* z, int c) {
  if (L.M.family == FAM_HORSESHOE) hs_prepare(L, g, z, c);
  else lg_prepare(L, g, z, c);
}
template <class G>
__device__ __forceinline__ float batch_finish(const BatchLaunch& L, const G& g, const float* z,
                                              float* grad, int c) {
  if (L.M.family == FAM_HORSESHOE) return hs_finish(L, g, z, grad, c);
  return lg_finish(L, g, z, grad, c);
}

#ifdef B200MC_TU_BATCH  // compiled in this family's translation unit only (launchers.h)
// One CTA per chain.  begin = 1: set the sampler up and publish its first point.  Otherwise:
// finish the potential at the published point, advance, publish the next point.
__global__ void __launch_bounds__(kBatchThreads, 2) batch_step_kernel(const __grid_constant__ BatchLaunch L) {
  __shared__ float red_scratch[64];
  const int cl = blockIdx.x;  // column of the operand images
  if (cl >= L.C) return;
  const int c = L.c0 + cl;    // chain
  const BlockGroup g{red_scratch};
  const int D = L.P.D;
  using S_t = Sampler<BlockGroup>;
  S_t* snap = reinterpret_cast<S_t*>(L.snap + (long)c * L.snap_stride);
  float* base = L.work + (long)c * num_slots(L.P.max_depth, L.P.dense_mass) * D;
  S_t S;
  StepResult r;
  if (L.kind == BATCH_BEGIN) {
    S.setup(&L.P, g, base, D, 1, c, L.C_total, &L.col);
    S.mv_yield = L.P.dense_mass != 0;
    S.load(L.cs, L.cv);
    if (L.init_keys)
      r = S.begin_init(Key{L.init_keys[2 * c], L.init_keys[2 * c + 1]}, L.user_init != 0);
    else
      r = S.begin_sampling();
  } else {
    const int st = L.status[c];
    if (st != ((L.kind == BATCH_GRAD) ? 0 : 2)) {  // not this launch's business
      if (threadIdx.x == 0 && st != 1) atomicAdd(L.active, 1u);
      return;
    }
    S = *snap;
    S.P = &L.P; S.g = g; S.w = base; S.col = &L.col;
    if (S.dense) S.dn = L.cv.dense + (long)c * 4 * D * D;
    if (L.kind == BATCH_GRAD) {
      const float U = batch_finish(L, g, S.v(LZ), S.v(LG), cl);
      r = S.advance(U);
    } else {
      r = S.resume();
    }
  }
  if (r == NEED_GRAD) {
    g.sync();
    batch_prepare(L, g, S.v(LZ), cl);
    if (threadIdx.x == 0) L.status[c] = 0;
  } else if (r == NEED_MATVEC || r == NEED_FACTOR) {
    if (threadIdx.x == 0) {
      L.reqs[c] = S.rq;
      L.status[c] = (r == NEED_MATVEC) ? 2 : 3;  // 3: parked until the host has factorised
    }
  } else {
    S.store(L.cs, L.cv);
    if (threadIdx.x == 0) L.status[c] = 1;
  }
  if (threadIdx.x == 0) {
    if (r != FINISHED) atomicAdd(L.active, 1u);
    *snap = S;
  }
}

// Serves the pending dense-mass sweeps of all chains: CTA (x, y) computes kMatvecRows rows
// of chain y's product(s).  Warp per row, lanes stride over the columns (coalesced 128-byte
// lines, 8 in flight per warp), the right-hand side(s) staged in shared memory; both products
// of a request come from ONE pass over the matrix.  Fixed summation order => deterministic.
__global__ void __launch_bounds__(kMatvecThreads, 4) batch_matvec_kernel(const __grid_constant__ BatchLaunch L) {
  extern __shared__ float mv_src[];  // [2][D]
  const int c = L.c0 + blockIdx.y;
  if (L.status[c] != 2) return;
  const MatvecRequest rq = L.reqs[c];
  if (rq.mat < 0) return;  // nothing to sweep (a chain resuming after its factorisation)
  const int D = L.P.D;
  const int i0 = blockIdx.x * kMatvecRows;
  float* base = L.work + (long)c * num_slots(L.P.max_depth, L.P.dense_mass) * D;
  const float* M = L.cv.dense + ((long)c * 4 + rq.mat) * D * D;
  const float* src1 = base + (long)rq.src1 * D;
  const float* src2 = rq.src2 >= 0 ? base + (long)rq.src2 * D : nullptr;
  float* dst1 = rq.dst1 >= 0 ? base + (long)rq.dst1 * D : nullptr;
  float* dst2 = rq.dst2 >= 0 ? base + (long)rq.dst2 * D : nullptr;
  float* s1 = mv_src;
  float* s2 = mv_src + D;
  for (int j = threadIdx.x; j < D; j += blockDim.x) {
    s1[j] = src1[j];
    s2[j] = src2 ? src2[j] : 0.f;
  }
  __syncthreads();
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5, nw = blockDim.x >> 5;
  const int i1 = (i0 + kMatvecRows < D) ? i0 + kMatvecRows : D;
  if (rq.lower == 2) {  // Welford: M += outer(src1, src2), coalesced read-modify-write
    float* Mw = L.cv.dense + ((long)c * 4 + rq.mat) * D * D;
    for (int i = i0 + wid; i < i1; i += nw) {
      const float dp = s1[i];
      float* row = Mw + (long)i * D;
      for (int j = lane; j < D; j += 32) {
        const float pr = dp * s2[j];
        row[j] = row[j] + pr;
      }
    }
    return;
  }
  for (int i = i0 + wid; i < i1; i += nw) {
    const float* row = M + (long)i * D;
    const int n = rq.lower ? i + 1 : D;
    float a[8], b[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) { a[k] = 0.f; b[k] = 0.f; }
    int j = lane;
    for (; j + 224 < n; j += 256) {
      float m[8];
#pragma unroll
      for (int k = 0; k < 8; ++k) m[k] = __ldcs(row + j + 32 * k);
#pragma unroll
      for (int k = 0; k < 8; ++k) {
        a[k] = fmaf(m[k], s1[j + 32 * k], a[k]);
        b[k] = fmaf(m[k], s2[j + 32 * k], b[k]);
      }
    }
    for (; j < n; j += 32) {
      const float m = __ldcs(row + j);
      a[0] = fmaf(m, s1[j], a[0]);
      b[0] = fmaf(m, s2[j], b[0]);
    }
    float ta = ((a[0] + a[1]) + (a[2] + a[3])) + ((a[4] + a[5]) + (a[6] + a[7]));
    float tb = ((b[0] + b[1]) + (b[2] + b[3])) + ((b[4] + b[5]) + (b[6] + b[7]));
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      ta += __shfl_xor_sync(0xffffffffu, ta, o);
      tb += __shfl_xor_sync(0xffffffffu, tb, o);
    }
    if (lane == 0) {
      dst1[i] = ta;
      if (dst2) dst2[i] = tb;
    }
  }
}

// ---- window-end factorisation of a dense mass matrix (hmc_util.py:198-237), element-wise
// pieces around the library Cholesky / triangular inverse.  mats = this chain's [4][D][D]:
// 0 inverse_mass_matrix (covariance), 1 mass_matrix_sqrt, 2 mass_matrix_sqrt_inv, 3 Welford m2.
__global__ void factor_cov_kernel(float* mats, int D, float n, int regularize) {
  // cov = m2 / (n - 1), shrunk towards the identity (:209-216); W = cov reversed in both axes
  const float scale = n / (n + 5.0f), shrink = 1e-3f * (5.0f / (n + 5.0f));
  float *cov = mats, *W = mats + 2L * D * D, *M2 = mats + 3L * D * D;
  const long total = (long)D * D;
  for (long e = blockIdx.x * (long)blockDim.x + threadIdx.x; e < total; e += (long)gridDim.x * blockDim.x) {
    const int i = (int)(e / D), j = (int)(e - (long)i * D);
    float cv = M2[e] / (n - 1.0f);
    if (regularize) {
      cv = scale * cv;
      if (i == j) cv = cv + shrink;
    }
    cov[e] = cv;
    W[(long)(D - 1 - i) * D + (D - 1 - j)] = cv;  // scratch: the slot of mass_matrix_sqrt_inv
  }
}
// After the column-major lower Cholesky of W (in place, slot 2): Lc[a][b] sits at
// chol[b * D + a].  tril_inv[i][j] = Lc[D-1-j][D-1-i] for j <= i (:229-230), written to the
// mass_matrix_sqrt slot (the triangular inverse then happens there in place) and to the m2
// slot, from where the host copies it over slot 2 before clearing m2.
__global__ void factor_tinv_kernel(float* mats, int D) {
  float *msq = mats + 1L * D * D, *tmp = mats + 3L * D * D;
  const float* chol = mats + 2L * D * D;
  const long total = (long)D * D;
  for (long e = blockIdx.x * (long)blockDim.x + threadIdx.x; e < total; e += (long)gridDim.x * blockDim.x) {
    const int i = (int)(e / D), j = (int)(e - (long)i * D);
    const float t = (j <= i) ? chol[(long)(D - 1 - i) * D + (D - 1 - j)] : 0.f;
    tmp[e] = t;
    msq[e] = t;
  }
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
    batch_prepare(L, g, z, c);
  } else {
    const float U = batch_finish(L, g, z, L.eval_grad + (long)c * L.P.D, c);
    if (threadIdx.x == 0) L.eval_U[c] = U;
  }
}

#endif  // B200MC_TU_BATCH

}  // namespace b200mc
