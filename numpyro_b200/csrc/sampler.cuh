// The HMC/NUTS transition as a device-side step machine.
//
// One call of advance(U) consumes exactly one potential+gradient evaluation (made
// at the point the previous call asked for) and carries the chain to the next point
// that needs one.  Everything between two gradient evaluations of the reference's
// nested loops -- leapfrog halves, energy test, multinomial / biased-progressive
// selection, checkpointed U-turn test, subtree and tree combination, warm-up
// adaptation, momentum refresh, init search -- happens inside advance().
//
// Reference functions restated here (numpyro/infer/):
//   hmc_util.py:289-309   velocity_verlet.update_fn      -> half_step / finish_leapfrog
//   hmc_util.py:1183-1200 euclidean_kinetic_energy       -> kinetic()
//   hmc_util.py:851-894   _build_basetree                -> advance_tree step 1
//   hmc_util.py:749-848   transition kernels/_combine_tree -> step 2 / step 5
//   hmc_util.py:941-981   ckpt indices / iterative turning -> step 3
//   hmc_util.py:984-1085  _iterative_build_subtree       -> steps 2-4
//   hmc_util.py:907-938   _double_tree                   -> next_doubling / step 5
//   hmc_util.py:1088-1180 build_tree                     -> start_transition / step 6
//   hmc_util.py:60-130    dual_averaging                 -> da_update
//   hmc_util.py:133-239   welford_covariance (diagonal)  -> adapt()
//   hmc_util.py:518-707   warmup_adapter                 -> adapt()
//   hmc.py:92-110         momentum_generator             -> draw_momentum
//   hmc.py:193-362        init_kernel                    -> init_kernel
//   hmc.py:364-530        _hmc_next/_nuts_next/sample_kernel -> advance_hmc / finish_transition
//   infer/util.py:386-450 find_valid_initial_params      -> init_candidate / advance_init
//   util.py:334-351       fori_collect body              -> collect
//
// Elementwise float arithmetic keeps the reference's operation order and is compiled
// with -fmad=false, so it is bit-identical to the NumPy oracle; only reductions and
// libm calls can differ in the last ulp.
#pragma once
#include <cfloat>
#include <cmath>

#include "group.cuh"
#include "rng.cuh"
#include "types.h"

namespace b200mc {

enum Slot {
  Z_LEFT = 0, R_LEFT, G_LEFT, Z_RIGHT, R_RIGHT, G_RIGHT,  // tree end points
  Z_PROP, G_PROP,                                          // tree proposal == HMCState.z, z_grad
  R_SUM,
  LZ, LR, LG,                                              // leaf under construction; LZ = point to evaluate
  SZ_PROP, SG_PROP, S_RSUM,                                // subtree proposal / momentum sum
  IMM, MSQRT, WMEAN, WM2,                                  // adaptation vectors
  EPS_NEXT,                                                // prefetched N(0, I) draw of the next transition
  CKPT0                                                    // r_ckpts[i], r_sum_ckpts[i] (, M^-1 r_ckpts[i]) = CKPT0 + stride*i + {0,1(,2)}
};
// Diagonal mass: 2 vectors per checkpoint.  Dense mass: 3 (the cached M^-1 r_ckpts[i]) plus
// six more slots behind the checkpoints (LV, V_LEFT, V_RIGHT: cached M^-1 r of LR, R_LEFT,
// R_RIGHT; LQ, Q_LEFT, Q_RIGHT: cached M^-1 grad of LG, G_LEFT, G_RIGHT, CTA-per-chain mode
// only); keeping them out of the diagonal layout buys shared memory for the glm ring.
__host__ __device__ inline int num_slots(int max_depth, int dense) {
  return CKPT0 + (dense ? 3 * max_depth + 6 : 2 * max_depth);
}

enum Phase { PH_INIT = 0, PH_TREE, PH_HMC, PH_IDLE, PH_FAIL, PH_FRSS };
// NEED_MATVEC: only with mv_yield (lock-step engine): the dense-mass product described by
// `rq` must be computed by the caller, then resume() continues where advance() stopped.
enum StepResult { NEED_GRAD = 0, FINISHED = 1, NEED_MATVEC = 2, NEED_FACTOR = 3,
                  STAGE_DONE = 4 /* internal: an adapt stage fell through */,
                  // only with end_yield (thread-per-chain kernel): a subtree is complete; the rare,
                  // long tail (combine, U-turn of the whole tree, next doubling's or next
                  // transition's Threefry draws, adaptation) runs in finish_subtree_yielded(),
                  // which the caller invokes when enough lanes of the warp wait for it
                  AT_SUBTREE_END = 5,
                  // only with transition_hook (HMCECS): a transition is complete (or sampling is
                  // about to begin) and another follows; the caller may change the potential --
                  // s.potential_energy and the gradient at v(G_PROP) -- and then calls
                  // resume_transition()
                  AT_TRANSITION_END = 6 };

struct MatvecRequest {
  int mat;         // which of the chain's four D x D matrices
  int lower;       // 1: triangular factor, columns j <= i only; 2: not a product but the
                   // rank-1 update M += outer(src1, src2) (Welford, hmc_util.py:195)
  int src1, dst1;  // slots: dst1 = M src1
  int src2, dst2;  // second right-hand side of the same sweep, or -1
};

__device__ __forceinline__ float logaddexp_f32(float a, float b) {
  // jnp.logaddexp: amax + log1p(exp(-|a-b|)), with the (inf - inf) case = a + b
  const float delta = a - b;
  if (isnan(delta)) return a + b;
  return fmaxf(a, b) + log1pf(expf(-fabsf(delta)));
}

template <class G>
struct Sampler {
  const SamplerParams* P;
  G g;
  float* w;            // this chain's slot 0, element 0
  long slot_stride;    // floats between slots
  int es_rt;           // floats between elements of one vector (1 for warp / CTA groups)
#define es (G::kUnitStride ? 1 : es_rt)
  int D;
  bool part;
  bool dense;
  float* dn;           // dense mass matrices [4][D][D] (null unless dense)

  bool writer;         // false: a replica that advances the same chain but writes nothing
#ifdef B200MC_SAMPLER_PROF
  unsigned long long pc[8];
#define B200MC_STAMP(var) const long long var = clock64()
#define B200MC_ACC(i, a, b) pc[i] += (unsigned long long)((b) - (a))
#else
#define B200MC_STAMP(var)
#define B200MC_ACC(i, a, b)
#endif
  ChainScalars s;      // persistent scalars (registers)

  // run control
  int iters_left, local_iter, chain, C;
  const CollectSpec* col;
  int phase;

  // transition state
  float e0, eps_s;
  int depth, nprop, turning, diverging, going_right, max_depth_cur;
  float weight, sum_acc, prop_pe, prop_energy;
  Key tree_key, sub_key, top_key;
  int s_n, s_div;
  float s_w, s_sum_acc, s_prop_pe, s_prop_energy;
  // precomputed leaf-level randomness (does not depend on the gradient)
  int pre_valid;
  int ck_stride, ck_end;  // checkpoint slot stride (2 diag / 3 dense), first slot behind them
  Key pre_sub_key;
  float pre_u;
  // doubling-level (hmc_util.py:1161-1162, :920) and transition-level (hmc.py:472, :94-102)
  // randomness, likewise computable ahead of the gradients that gate its use
  struct DoublingRng {
    Key tree_key, sub_key, top_key;
    int going_right;
    float u_top;
  };
  int pre_d_valid, pre_t_valid, pre_a_valid;
  DoublingRng pre_d, pre_t_d;
  Key pre_t_rng_key, pre_t_k_tr, pre_a_nk, pre_a_kss;
  // leaf state parked between part A and part B of the fused leaf step
  float fz[4], fg[4], f_U, f_e_new, f_dE;
  int deferred;  // 0 none, 1 leaf fold, 2 subtree end + doubling state, 3 transition end
  bool end_yield;      // return AT_SUBTREE_END instead of running the subtree tail in place
  bool y_sub_turning;  // the subtree's turning flag while it waits
  bool d_sub_turning, d_tree_turning;
  float u_top;
  // HMC
  int hmc_steps, hmc_done;
  float energy_old;
  Key mh_key;
  // init search
  Key init_key;
  int user_init;
  // find_reasonable_step_size (hmc_util.py:314-384)
  float fr_step, fr_e_cur;
  int fr_last_dir, fr_dir, fr_cont;
  Key fr_key;
  Key init_mom_key;
  // lock-step engine: dense-mass sweeps are handed to a batched kernel instead of run in place
  bool mv_yield, leaf_cached;
  int mv_resume;  // 0 none, 1 leaf finish, 2 / 3 transition start
  float pending_U;
  MatvecRequest rq;
  Key ad_kss;             // adapt stages: key of the window-end step-size search
  bool ad_middle, ad_reinit;

  // CTA-per-chain kernels may keep a few hot slots (the leaf under construction) in shared
  // memory instead of the chain's global work array: hot_mask has their bits, `hot` their copies
  float* hot;
  unsigned hot_mask;
  bool transition_hook;  // see AT_TRANSITION_END
  __device__ __forceinline__ float* v(int slot) const {
    if constexpr (G::kBlockWide) {
      if (slot < 32 && ((hot_mask >> slot) & 1u))
        return hot + (long)__popc(hot_mask & ((1u << slot) - 1u)) * D;
    }
    return w + (long)slot * slot_stride;
  }
  __device__ __forceinline__ float* ckpt_r(int i) const { return v(CKPT0 + ck_stride * i); }
  __device__ __forceinline__ float* ckpt_rs(int i) const { return v(CKPT0 + ck_stride * i + 1); }
  __device__ __forceinline__ float* ckpt_v(int i) const { return v(CKPT0 + 3 * i + 2); }  // dense only
#define LV (ck_end)  // dense-only slots, see num_slots()
#define V_LEFT (ck_end + 1)
#define V_RIGHT (ck_end + 2)
#define LQ (ck_end + 3)
#define Q_LEFT (ck_end + 4)
#define Q_RIGHT (ck_end + 5)
  // dense mass: this chain's (or this replica's) four D x D matrices in global memory
  __device__ __forceinline__ float* mat(int which) const { return dn + (long)which * D * D; }

#define B200MC_FOR_D for (int d = g.lane(); d < D; d += g.size())

  __device__ __forceinline__ void copy(int dst, int src) {
    float* a = v(dst);
    const float* b = v(src);
    B200MC_FOR_D a[d * es] = b[d * es];
  }

  // ------------------------------------------------------------------ setup / persistence
  __device__ void setup(const SamplerParams* p, G grp, float* work, long sstride, int estride,
                        int chain_, int C_, const CollectSpec* c) {
    P = p; g = grp; w = work; slot_stride = sstride; es_rt = estride; D = p->D;
    part = p->partitionable != 0;
    dense = p->dense_mass != 0;
    ck_stride = dense ? 3 : 2;
    ck_end = CKPT0 + ck_stride * p->max_depth;
    dn = nullptr;
    chain = chain_; C = C_; col = c;
    writer = true;
    iters_left = c ? c->num_iters : 0;
    local_iter = 0;
    phase = PH_IDLE;
    pre_valid = 0;
    pre_d_valid = 0; pre_t_valid = 0; pre_a_valid = 0;
    deferred = 0;
    end_yield = false; y_sub_turning = false;
    mv_yield = false; leaf_cached = false; mv_resume = 0;
    hot = nullptr; hot_mask = 0u;
    transition_hook = false;
  }

  __device__ void load(const ChainScalars* cs, const ChainVectors& cv) {
    s = cs[chain];
    if (dense && !dn) dn = cv.dense + (long)chain * 4 * D * D;
    const long o = (long)chain * D;
    float *zp = v(Z_PROP), *gp = v(G_PROP), *im = v(IMM), *ms = v(MSQRT), *wm = v(WMEAN),
          *w2 = v(WM2);
    B200MC_FOR_D {
      zp[d * es] = cv.z[o + d];
      gp[d * es] = cv.z_grad[o + d];
      im[d * es] = cv.inverse_mass_matrix[o + d];
      ms[d * es] = cv.mass_matrix_sqrt[o + d];
      wm[d * es] = cv.welford_mean[o + d];
      w2[d * es] = cv.welford_m2[o + d];
    }
  }

  __device__ void store(ChainScalars* cs, const ChainVectors& cv) {
    if (!writer) return;
    const long o = (long)chain * D;
    const float *zp = v(Z_PROP), *gp = v(G_PROP), *im = v(IMM), *ms = v(MSQRT), *wm = v(WMEAN),
                *w2 = v(WM2);
    B200MC_FOR_D {
      cv.z[o + d] = zp[d * es];
      cv.z_grad[o + d] = gp[d * es];
      cv.inverse_mass_matrix[o + d] = im[d * es];
      cv.mass_matrix_sqrt[o + d] = ms[d * es];
      cv.welford_mean[o + d] = wm[d * es];
      cv.welford_m2[o + d] = w2[d * es];
    }
    if (g.lane() == 0) cs[chain] = s;
  }

  // ------------------------------------------------------------------ small pieces
  // dst = M src for a dense row-major D x D matrix (rows over lanes); `lower` restricts
  // the sum to j <= i (triangular factor: the skipped terms are exact zeros).
  __device__ __noinline__ void matvec(const float* M, const float* src, float* dst, bool lower = false) {
    g.sync();  // src was written by its owner lanes
    if constexpr (G::kBlockWide) {
      // CTA per chain (large D): warp per row, lanes stride over the columns => every load is
      // a coalesced 128-byte line, 8 independent lines in flight per warp; fixed summation
      // order (per-lane partials, then an xor tree) => deterministic.
      const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5, nw = blockDim.x >> 5;
      for (int i = wid; i < D; i += nw) {
        const float* row = M + (long)i * D;
        const int n = lower ? i + 1 : D;
        float a[8];
#pragma unroll
        for (int k = 0; k < 8; ++k) a[k] = 0.f;
        int j = lane;
        for (; j + 224 < n; j += 256) {
#pragma unroll
          for (int k = 0; k < 8; ++k) a[k] = fmaf(__ldcg(row + j + 32 * k), src[j + 32 * k], a[k]);
        }
        for (; j < n; j += 32) a[0] = fmaf(__ldcg(row + j), src[j], a[0]);
        float t = ((a[0] + a[1]) + (a[2] + a[3])) + ((a[4] + a[5]) + (a[6] + a[7]));
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) t += __shfl_xor_sync(0xffffffffu, t, o);
        if (lane == 0) dst[i] = t;
      }
      g.sync();
      return;
    }
    B200MC_FOR_D {
      const float* row = M + (long)d * D;
      const int n = lower ? d + 1 : D;
      float a0 = 0.f, a1 = 0.f;
      int j = 0;
      for (; j + 1 < n; j += 2) {
        a0 = fmaf(row[j], src[j * es], a0);
        a1 = fmaf(row[j + 1], src[(j + 1) * es], a1);
      }
      if (j < n) a0 = fmaf(row[j], src[j * es], a0);
      dst[d * es] = a0 + a1;
    }
    g.sync();
  }

  // Two right-hand sides in ONE sweep over the matrix (CTA per chain only): dst1 = M src1,
  // dst2 = M src2.  The matrix (16 MB at D = 2001) is the HBM traffic of a dense-mass
  // leapfrog; reading it once instead of twice is the point.  Same per-row summation order
  // as matvec().
  __device__ __noinline__ void matvec2(const float* M, const float* src1, const float* src2,
                                       float* dst1, float* dst2) {
    g.sync();
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5, nw = blockDim.x >> 5;
    for (int i = wid; i < D; i += nw) {
      const float* row = M + (long)i * D;
      float a[8], b[8];
#pragma unroll
      for (int k = 0; k < 8; ++k) { a[k] = 0.f; b[k] = 0.f; }
      int j = lane;
      for (; j + 224 < D; j += 256) {
        float m[8];
#pragma unroll
        for (int k = 0; k < 8; ++k) m[k] = __ldcg(row + j + 32 * k);
#pragma unroll
        for (int k = 0; k < 8; ++k) {
          a[k] = fmaf(m[k], src1[j + 32 * k], a[k]);
          b[k] = fmaf(m[k], src2[j + 32 * k], b[k]);
        }
      }
      for (; j < D; j += 32) {
        const float m = __ldcg(row + j);
        a[0] = fmaf(m, src1[j], a[0]);
        b[0] = fmaf(m, src2[j], b[0]);
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
        dst2[i] = tb;
      }
    }
    g.sync();
  }

  // Dense mass, leaf (LZ, LR, LG) complete: cache LV = M^-1 LR for the energy and the U-turn
  // tests and, CTA per chain, LQ = M^-1 LG from the same sweep: the next half step then needs
  // no sweep of its own, M^-1 (r - h g) = LV - h LQ (one rounding apart from the reference's
  // matmul of the updated momentum, hmc_util.py:297-302, 1203-1223).
  __device__ __forceinline__ void dense_leaf_cache() {
    if constexpr (G::kBlockWide) matvec2(mat(0), v(LR), v(LG), v(LV), v(LQ));
    else matvec(mat(0), v(LR), v(LV));
  }

  // 0.5 * dot(M^-1 r, r): hmc_util.py:1183-1200.  Dense: vslot holds the cached M^-1 r.
  __device__ __forceinline__ float kinetic(int rslot, int vslot = -1) {
    const float *r = v(rslot), *im = v(IMM);
    const float* vv_ = dense ? v(vslot) : nullptr;
    float acc = 0.f;
    B200MC_FOR_D {
      const float rv = r[d * es];
      const float vv = dense ? vv_[d * es] : im[d * es] * rv;
      acc += vv * rv;
    }
    return 0.5f * g.sum(acc);
  }

  // hmc_util.py:710-746 with r_sum already the (sub)tree sum; vl_/vr_: cached M^-1 r (dense)
  __device__ __forceinline__ bool is_turning(const float* rl, const float* rr, const float* rsum,
                                             const float* sub_a, const float* add_b,
                                             const float* vl_ = nullptr,
                                             const float* vr_ = nullptr) {
    // s = rsum (- sub_a + add_b); s' = s - (rl + rr)/2; (M^-1 rl).s' <= 0 | (M^-1 rr).s' <= 0
    const float* im = v(IMM);
    float la = 0.f, ra = 0.f;
    B200MC_FOR_D {
      float sv = rsum[d * es];
      if (sub_a) {
        sv = sv - sub_a[d * es];
        sv = sv + add_b[d * es];
      }
      const float l = rl[d * es], r = rr[d * es];
      const float h = (l + r) / 2.0f;
      const float sp = sv - h;
      const float vl = dense ? vl_[d * es] : im[d * es] * l;
      const float vr = dense ? vr_[d * es] : im[d * es] * r;
      la += vl * sp;
      ra += vr * sp;
    }
    g.sum2(la, ra);
    return (la <= 0.f) | (ra <= 0.f);
  }

  // r = M^{1/2} eps, eps ~ N(0, I): hmc.py:92-110 (one dict block => split(key, 1)).
  // use_imm: the reference's find_reasonable_step_size hands the INVERSE mass matrix to
  // momentum_generator (hmc_util.py:362), so r = M^-1 eps there; restated as is.
  // Structured mass matrix (hmc.py:92-102): keys = split(k_mom, num_blocks) in the dict order
  // of the blocks, block b draws normal(keys[b], (size_b,)); coordinate d takes element
  // blk_pos[d] of its block's draw.
  __device__ __forceinline__ float block_eps(Key k_mom, int d) const {
    const int b = P->blk_of[d];
    const Key kb = split_at(k_mom, (uint32_t)P->num_blocks, (uint32_t)b, part);
    return normal_at(kb, (uint32_t)P->blk_size[b], (uint32_t)P->blk_pos[d], part);
  }

  // eps ~ N(0, I) of momentum_generator into slot `dst`
  __device__ __forceinline__ void draw_eps(Key k_mom, int dst) {
    float* eps = v(dst);
    if (P->num_blocks > 1) {
      B200MC_FOR_D eps[d * es] = block_eps(k_mom, d);
      return;
    }
    const Key kb = split_at(k_mom, 1u, 0u, part);
    B200MC_FOR_D eps[d * es] = normal_at(kb, (uint32_t)D, (uint32_t)d, part);
  }

  // M^{1/2} is lower triangular only for ONE block in flat order; with several blocks the
  // factor is block-structured under a permutation and the sweep covers whole rows (the
  // entries outside the blocks are exact zeros).
  __device__ __forceinline__ bool sqrt_is_lower() const { return P->num_blocks <= 1; }

  __device__ __forceinline__ void draw_momentum(Key k_mom, int dst, bool use_imm = false) {
    float* r = v(dst);
    if (dense) {
      draw_eps(k_mom, LV);
      matvec(use_imm ? mat(0) : mat(1), v(LV), r, !use_imm && sqrt_is_lower());
      return;
    }
    const float* ms = use_imm ? v(IMM) : v(MSQRT);
    if (P->num_blocks > 1) {
      B200MC_FOR_D r[d * es] = ms[d * es] * block_eps(k_mom, d);
      return;
    }
    const Key kb = split_at(k_mom, 1u, 0u, part);
    B200MC_FOR_D {
      const float eps = normal_at(kb, (uint32_t)D, (uint32_t)d, part);
      r[d * es] = ms[d * es] * eps;
    }
  }

  // first half of velocity Verlet from the leaf in (LZ, LR, LG): hmc_util.py:297-302
  __device__ __forceinline__ void half_step() {
    float *z = v(LZ), *r = v(LR);
    const float *gr = v(LG), *im = v(IMM);
    const float hs = 0.5f * eps_s;
    if (dense) {
      if constexpr (G::kBlockWide) {
        float* lv = v(LV);
        const float* lq = v(LQ);
        B200MC_FOR_D {
          const float t = hs * gr[d];
          r[d] = r[d] - t;
          const float tq = hs * lq[d];
          const float vh = lv[d] - tq;  // M^-1 of the half-stepped momentum, by linearity
          lv[d] = vh;
          const float dz = eps_s * vh;
          z[d] = z[d] + dz;
        }
        return;
      }
      B200MC_FOR_D {
        const float t = hs * gr[d * es];
        r[d * es] = r[d * es] - t;
      }
      matvec(mat(0), r, v(LV));
      const float* lv = v(LV);
      B200MC_FOR_D {
        const float dz = eps_s * lv[d * es];
        z[d * es] = z[d * es] + dz;
      }
      return;
    }
    B200MC_FOR_D {
      const float t = hs * gr[d * es];
      const float rn = r[d * es] - t;
      r[d * es] = rn;
      const float vv = im[d * es] * rn;
      const float dz = eps_s * vv;
      z[d * es] = z[d * es] + dz;
    }
  }

  // second half: hmc_util.py:306-308
  __device__ __forceinline__ void finish_leapfrog() {
    if (leaf_cached) return;  // resume(): momentum and its M^-1 products are already in place
    float* r = v(LR);
    const float* gr = v(LG);
    const float hs = 0.5f * eps_s;
    B200MC_FOR_D {
      const float t = hs * gr[d * es];
      r[d * es] = r[d * es] - t;
    }
    if (dense) dense_leaf_cache();  // cached for the energy and the U-turn tests
  }

  // ------------------------------------------------------------------ warm-up adaptation
  __device__ __forceinline__ void da_init(float prox_center) {  // hmc_util.py:97-107
    s.ss_x_t = 0.f; s.ss_x_avg = 0.f; s.ss_g_avg = 0.f; s.ss_t = 0;
    s.ss_prox_center = prox_center;
  }

  __device__ __forceinline__ void da_update(float gstat) {  // hmc_util.py:109-128
    const float t0 = 10.f, kappa = 0.75f, gamma = 0.05f;
    s.ss_t += 1;
    const float t = (float)s.ss_t;
    const float inv = 1.0f / (t + t0);
    const float a = (1.0f - inv) * s.ss_g_avg;
    const float b = gstat / (t + t0);
    s.ss_g_avg = a + b;
    const float c = sqrtf(t) / gamma;
    s.ss_x_t = s.ss_prox_center - c * s.ss_g_avg;
    const float wt = powf(t, -kappa);
    const float e = (1.0f - wt) * s.ss_x_avg;
    const float f = wt * s.ss_x_t;
    s.ss_x_avg = e + f;
  }

  // (inverse_mass_matrix, mass_matrix_sqrt, mass_matrix_sqrt_inv) from a dense covariance in
  // mat(0): hmc_util.py:224-233 (welford final) == :497-507 (_initialize_mass_matrix):
  //   tril_inv     = swapaxes(cholesky(cov[::-1, ::-1])[::-1, ::-1])
  //   cov_inv_sqrt = solve_triangular(tril_inv, identity, lower=True)
  // mat(3) (the Welford m2 buffer) is the workspace and is left zeroed.
  // Structured mass matrix: the same recipe runs in BLOCK ORDER (position a <-> flat index
  // blk_perm[a]); there the masked covariance is block diagonal, so the reversed Cholesky of
  // the whole matrix is the reversed Cholesky of every block (a diagonal block gives
  // sqrt(cov) and 1/sqrt(cov), hmc_util.py:234-236), and the factors are scattered back to
  // flat order.
  __device__ __noinline__ void factor_dense() {
    float *cov = mat(0), *msq = mat(1), *tinv = mat(2), *W = mat(3);
    const int n = D;
    const int* perm = (P->num_blocks > 1) ? P->blk_perm : nullptr;
#define B200MC_PI(a) (perm ? perm[a] : (a))
    g.sync();
    B200MC_FOR_D  // W = cov (block order) reversed in both axes
      for (int j = 0; j < n; ++j)
        W[(long)d * n + j] = cov[(long)B200MC_PI(n - 1 - d) * n + B200MC_PI(n - 1 - j)];
    g.sync();
    for (int k = 0; k < n; ++k) {  // left-looking Cholesky, lower triangle, in place
      float dk = W[(long)k * n + k];
      for (int p = 0; p < k; ++p) {
        const float l = W[(long)k * n + p];
        dk -= l * l;
      }
      dk = sqrtf(dk);
      g.sync();  // every lane has read row k before its diagonal changes
      for (int i = k + g.lane(); i < n; i += g.size()) {
        if (i == k) {
          W[(long)k * n + k] = dk;
        } else {
          float a = W[(long)i * n + k];
          for (int p = 0; p < k; ++p) a -= W[(long)i * n + p] * W[(long)k * n + p];
          W[(long)i * n + k] = a / dk;
        }
      }
      g.sync();
    }
    B200MC_FOR_D {  // tril_inv[i][j] = Lc[n-1-j][n-1-i] for j <= i, else 0
      const long row = (long)B200MC_PI(d) * n;
      for (int j = 0; j < n; ++j)
        tinv[row + B200MC_PI(j)] = (j <= d) ? W[(long)(n - 1 - j) * n + (n - 1 - d)] : 0.f;
    }
    g.sync();
    B200MC_FOR_D {  // column d of the inverse of the lower-triangular tril_inv
      const int pd = B200MC_PI(d);
      for (int i = 0; i < n; ++i) {
        const long ri = (long)B200MC_PI(i) * n;
        float a = (i == d) ? 1.f : 0.f;
        if (i < d) {
          msq[ri + pd] = 0.f;
          continue;
        }
        for (int p = d; p < i; ++p) a -= tinv[ri + B200MC_PI(p)] * msq[(long)B200MC_PI(p) * n + pd];
        msq[ri + pd] = a / tinv[ri + B200MC_PI(i)];
      }
    }
    g.sync();
#undef B200MC_PI
    B200MC_FOR_D
      for (int j = 0; j < n; ++j) W[(long)d * n + j] = 0.f;
    g.sync();
  }

  // warmup_adapter.update_fn: hmc_util.py:637-705.  Returns true when a step-size search
  // (find_heuristic_step_size) was started and the transition tail must wait for it.
  // warmup_adapter.update_fn (hmc_util.py:637-705) in three resumable stages, so that the
  // lock-step engine (mv_yield) can hand the two O(D^2) pieces of a dense mass matrix -- the
  // Welford rank-1 update and the window-end factorisation -- to whole-machine kernels.
  // Everywhere else the stages simply run back to back.
  __device__ __noinline__ StepResult adapt_a() {
    const int t = s.i;
    const float accept_prob = s.accept_prob;
    if (!pre_a_valid) split2(Key{s.adapt_key[0], s.adapt_key[1]}, part, pre_a_nk, pre_a_kss);  // :656
    pre_a_valid = 0;
    const Key nk = pre_a_nk;
    ad_kss = pre_a_kss;
    s.adapt_key[0] = nk.k0; s.adapt_key[1] = nk.k1;
    if (P->adapt_step_size) {
      da_update(P->target_accept_prob - accept_prob);
      float ss = (t == P->num_warmup - 1) ? expf(s.ss_x_avg) : expf(s.ss_x_t);
      ss = fminf(fmaxf(ss, FLT_MIN), FLT_MAX);  // :671-672
      s.step_size = ss;
    }
    const bool is_middle = (0 < s.window_idx) && (s.window_idx < P->num_windows - 1);
    ad_middle = is_middle;
    if (P->adapt_mass_matrix && is_middle) {  // welford update, :172-196
      s.mm_n += 1;
      const float n = (float)s.mm_n;
      const float* z = v(Z_PROP);
      float *mean = v(WMEAN), *m2 = v(WM2);
      float *dpre_v = v(LV), *dpost_v = v(LR);  // free between transitions
      B200MC_FOR_D {
        const float x = z[d * es];
        const float dpre = x - mean[d * es];
        const float q = dpre / n;
        const float mnew = mean[d * es] + q;
        mean[d * es] = mnew;
        const float dpost = x - mnew;
        if (dense) {
          dpre_v[d * es] = dpre;
          dpost_v[d * es] = dpost;
        } else {
          const float pr = dpre * dpost;
          m2[d * es] = m2[d * es] + pr;
        }
      }
      if (dense) {  // m2 += outer(delta_post, delta_pre), :195
        if constexpr (G::kBlockWide) {
          if (mv_yield) {
            rq = MatvecRequest{3, 2 /* rank-1 update */, LR, -1, LV, -1};
            mv_resume = 5;
            return NEED_MATVEC;
          }
        }
        g.sync();
        float* M2 = mat(3);
        if constexpr (G::kBlockWide) {  // warp per row: coalesced read-modify-write
          const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5, nw = blockDim.x >> 5;
          for (int i = wid; i < D; i += nw) {
            const float dp = dpost_v[i];
            float* row = M2 + (long)i * D;
            for (int j = lane; j < D; j += 32) {
              const float pr = dp * dpre_v[j];
              row[j] = row[j] + pr;
            }
          }
        } else {
          B200MC_FOR_D {
            const float dp = dpost_v[d * es];
            float* row = M2 + (long)d * D;
            for (int j = 0; j < D; ++j) {
              const float pr = dp * dpre_v[j * es];
              row[j] = row[j] + pr;
            }
          }
        }
        g.sync();
      }
    }
    return STAGE_DONE;
  }

  __device__ __noinline__ StepResult adapt_b() {
    const int t = s.i;
    const bool is_middle = ad_middle;
    const int wend = P->window_end[min(s.window_idx, P->num_windows - 1)];
    const bool at_end = (t == wend);
    if (at_end) s.window_idx += 1;
    ad_reinit = at_end && is_middle;
    if (at_end && is_middle && P->adapt_mass_matrix) {  // welford final, :198-237
      const float n = (float)s.mm_n;
      float *mean = v(WMEAN), *m2 = v(WM2), *im = v(IMM), *ms = v(MSQRT);
      const float scale = n / (n + 5.0f);
      const float shrink = 1e-3f * (5.0f / (n + 5.0f));
      if (dense) {
        if constexpr (G::kBlockWide) {
          if (mv_yield) {  // covariance + factorisation by the caller (rq.src1 carries n)
            rq = MatvecRequest{-1, 0, s.mm_n, -1, -1, -1};
            mv_resume = 6;
            return NEED_FACTOR;
          }
        }
        float *cov = mat(0), *M2 = mat(3);
        const bool blocks = P->num_blocks > 1;
        B200MC_FOR_D {
          const int bd = blocks ? P->blk_of[d] : 0;
          const bool full = blocks ? (P->blk_dense[bd] != 0) : true;
          for (int j = 0; j < D; ++j) {
            float c = M2[(long)d * D + j] / (n - 1.0f);
            if (P->regularize_mass_matrix) {
              c = scale * c;
              if (j == d) c = c + shrink;
            }
            // structured mass matrix: covariances across blocks (and the off-diagonal of a
            // diagonal block) are not part of the model of M (hmc_util.py:176-182)
            if (blocks && j != d && !(full && P->blk_of[j] == bd)) c = 0.f;
            cov[(long)d * D + j] = c;
          }
        }
        factor_dense();
      } else {
        B200MC_FOR_D {
          float cov = m2[d * es] / (n - 1.0f);
          if (P->regularize_mass_matrix) {
            const float sc = scale * cov;
            cov = sc + shrink;
          }
          im[d * es] = cov;
          ms[d * es] = 1.0f / sqrtf(cov);
          m2[d * es] = 0.f;
        }
      }
    }
    return STAGE_DONE;
  }

  __device__ __noinline__ StepResult adapt_c() {
    if (ad_reinit) {  // _update_at_window_end, :596-635
      if (P->adapt_mass_matrix) {
        float* mean = v(WMEAN);
        B200MC_FOR_D mean[d * es] = 0.f;
        s.mm_n = 0;
      }
      if (P->adapt_step_size) {
        if (P->find_heuristic) {
          start_frss(ad_kss, 1);
          if (frss_try()) return NEED_GRAD;               // step-size search needs gradients
        }
        da_init(logf(10.0f) + logf(s.step_size));  // :624 (or: the search ended at once)
      }
    }
    return STAGE_DONE;
  }

  // The adapt stages from `stage` on, then the transition tail.  Stages return to this frame
  // (no nesting: the device stack is small).
  __device__ StepResult run_adapt(int stage) {
    StepResult r;
    if (stage <= 0) {
      r = adapt_a();
      if (r != STAGE_DONE) return r;
    }
    if (stage <= 1) {
      r = adapt_b();
      if (r != STAGE_DONE) return r;
    }
    r = adapt_c();
    if (r != STAGE_DONE) return r;
    return finish_transition_tail();
  }

  // ------------------------------------------------------------------ find_reasonable_step_size
  // hmc_util.py:314-384 as a phase of the step machine: each try draws a fresh momentum,
  // takes ONE leapfrog from the current position (Z_PROP, G_PROP) and doubles or halves the
  // step size until the direction of change flips.  cont: 0 = called from init_kernel,
  // 1 = called at a window end (the transition tail resumes afterwards).
  __device__ void start_frss(Key key, int cont) {
    fr_step = s.step_size;
    fr_last_dir = 0;
    fr_dir = 0;
    fr_key = key;
    fr_cont = cont;
  }

  // true: a leapfrog was started (NEED_GRAD); false: the search is over (s.step_size set)
  __device__ __noinline__ bool frss_try() {
    const bool not_small = (fr_step > FLT_MIN) || (fr_dir >= 0);   // :373
    const bool not_large = (fr_step < FLT_MAX) || (fr_dir <= 0);   // :375
    const bool go = not_small && not_large && ((fr_last_dir == 0) || (fr_dir == fr_last_dir));
    if (!go) {
      s.step_size = fr_step;
      return false;
    }
    Key nk, k_mom;
    split2(fr_key, part, nk, k_mom);  // :355
    fr_key = nk;
    const float scale = (fr_dir > 0) ? 2.0f : ((fr_dir < 0) ? 0.5f : 1.0f);
    fr_step = scale * fr_step;
    draw_momentum(k_mom, LR, true);
    copy(LZ, Z_PROP); copy(LG, G_PROP);
    if (dense) dense_leaf_cache();
    fr_e_cur = kinetic(LR, LV) + s.potential_energy;
    eps_s = fr_step;
    phase = PH_FRSS;
    half_step();
    return true;
  }

  __device__ __noinline__ StepResult advance_frss(float U) {
    finish_leapfrog();
    const float e_new = kinetic(LR, LV) + U;
    const float dE = e_new - fr_e_cur;
    const int dir_new = (-0.2231435513142097558f < -dE) ? 1 : -1;  // log(0.8); NaN -> -1
    fr_last_dir = fr_dir;
    fr_dir = dir_new;
    if (frss_try()) return NEED_GRAD;
    if (fr_cont == 0) {
      da_init(logf(10.0f * s.step_size));  // :577
      return after_init();
    }
    da_init(logf(10.0f) + logf(s.step_size));  // :624
    return finish_transition_tail();
  }

  // ------------------------------------------------------------------ collection (util.py:334-351)
  __device__ void collect() {
    const int k = local_iter;  // index of the transition just finished, within this call
    if (!writer || !col || col->n_keep <= 0 || k < col->collect_start) return;
    const int rel = k - col->collect_start;
    if (rel % col->thinning != col->thinning - 1) return;  // the last write to a slot wins
    const int idx = rel / col->thinning;
    if (idx >= col->n_keep) return;
    const long row = (long)idx * C + chain;
    if (col->z) {
      const float* z = v(Z_PROP);
      float* out = col->z + row * D;
      B200MC_FOR_D out[d] = z[d * es];
    }
    if (g.lane() == 0) {
      if (col->potential_energy) col->potential_energy[row] = s.potential_energy;
      if (col->energy) col->energy[row] = s.energy;
      if (col->num_steps) col->num_steps[row] = s.num_steps;
      if (col->accept_prob) col->accept_prob[row] = s.accept_prob;
      if (col->mean_accept_prob) col->mean_accept_prob[row] = s.mean_accept_prob;
      if (col->diverging) col->diverging[row] = s.diverging;
      if (col->step_size) col->step_size[row] = s.step_size;
    }
  }

  // tail of sample_kernel: hmc.py:500-530
  __device__ StepResult finish_transition(float accept_prob, int num_steps, int div) {
    s.accept_prob = accept_prob;
    s.num_steps = num_steps;
    s.diverging = div;
    if (s.i < P->num_warmup) return run_adapt(0);
    return finish_transition_tail();
  }

  __device__ StepResult finish_transition_tail() {
    const int t = s.i;
    const float accept_prob = s.accept_prob;
    const int itr = t + 1;
    const int n = (t < P->num_warmup) ? itr : itr - P->num_warmup;
    const float dm = accept_prob - s.mean_accept_prob;
    s.mean_accept_prob = s.mean_accept_prob + dm / (float)n;
    s.i = itr;
    collect();
    local_iter += 1;
    iters_left -= 1;
    if (iters_left <= 0) {
      phase = PH_IDLE;
      return FINISHED;
    }
    if (transition_hook) return AT_TRANSITION_END;
    return start_transition();
  }

  __device__ StepResult resume_transition() { return start_transition(); }

  // ------------------------------------------------------------------ transition start
  // sample_kernel head (hmc.py:472-490) + build_tree head (hmc_util.py:1129-1153)
  __device__ StepResult start_transition() {
    if (!pre_t_valid) compute_transition_rng();
    pre_t_valid = 0;
    s.rng_key[0] = pre_t_rng_key.k0; s.rng_key[1] = pre_t_rng_key.k1;
    if (P->algo == ALGO_HMC) return start_hmc(pre_t_k_tr);
    if (mv_yield && dense) {  // r = M^{1/2} eps as a batched sweep
      rq = MatvecRequest{1, 1, EPS_NEXT, R_LEFT, -1, -1};
      mv_resume = 2;
      return NEED_MATVEC;
    }
    momentum_from_eps(R_LEFT);
    return start_transition_b();
  }

  __device__ StepResult start_transition_b() {
    pre_d = pre_t_d;  // the first doubling of this transition
    pre_d_valid = 1;
    {
      float *rl = v(R_LEFT), *rr = v(R_RIGHT), *rs = v(R_SUM);
      float *zl = v(Z_LEFT), *zr = v(Z_RIGHT), *gl = v(G_LEFT), *gr = v(G_RIGHT);
      const float *zp = v(Z_PROP), *gp = v(G_PROP);
      B200MC_FOR_D {
        const float r = rl[d * es];
        rr[d * es] = r;
        rs[d * es] = r;
        const float z = zp[d * es], gg = gp[d * es];
        zl[d * es] = z; zr[d * es] = z;
        gl[d * es] = gg; gr[d * es] = gg;
      }
    }
    if (dense) {
      if constexpr (G::kBlockWide) {
        if (mv_yield) {
          rq = MatvecRequest{0, 0, R_LEFT, V_LEFT, G_PROP, Q_LEFT};
          mv_resume = 3;
          return NEED_MATVEC;
        }
        matvec2(mat(0), v(R_LEFT), v(G_PROP), v(V_LEFT), v(Q_LEFT));
      } else {
        matvec(mat(0), v(R_LEFT), v(V_LEFT));
      }
    }
    return start_transition_c();
  }

  __device__ StepResult start_transition_c() {
    if (dense) {
      if constexpr (G::kBlockWide) copy(Q_RIGHT, Q_LEFT);
      copy(V_RIGHT, V_LEFT);
    }
    e0 = s.potential_energy + kinetic(R_LEFT, V_LEFT);
    depth = 0; weight = 0.f; turning = 0; diverging = 0; sum_acc = 0.f; nprop = 0;
    prop_pe = s.potential_energy;
    prop_energy = e0;
    tree_key = pre_t_k_tr;
    max_depth_cur = (s.i < P->num_warmup) ? P->max_depth_warmup : P->max_depth_sample;
    phase = PH_TREE;
    if (!(depth < max_depth_cur)) return finish_transition(nanf(""), 0, 0);
    next_doubling();
    return NEED_GRAD;
  }

  // build_tree._body_fn (hmc_util.py:1159-1176) + _double_tree head (:920)
  // the state part of next_doubling (no vector work)
  __device__ void next_doubling_state() {
    if (!pre_d_valid) compute_doubling_rng(tree_key, pre_d);
    pre_d_valid = 0;
    tree_key = pre_d.tree_key;
    going_right = pre_d.going_right;
    sub_key = pre_d.sub_key;
    top_key = pre_d.top_key;
    u_top = pre_d.u_top;
    s_n = 0;
    pre_valid = 0;
    eps_s = going_right ? s.step_size : -s.step_size;
  }

  __device__ void next_doubling() {
    next_doubling_state();
    if (going_right) {
      copy(LZ, Z_RIGHT); copy(LR, R_RIGHT); copy(LG, G_RIGHT);
    } else {
      copy(LZ, Z_LEFT); copy(LR, R_LEFT); copy(LG, G_LEFT);
    }
    if constexpr (G::kBlockWide) {
      if (dense) {
        copy(LV, going_right ? V_RIGHT : V_LEFT);
        copy(LQ, going_right ? Q_RIGHT : Q_LEFT);
      }
    }
    half_step();
  }

  // Leaf-level randomness of _iterative_build_subtree (hmc_util.py:1005, :804): does
  // not depend on the gradient, so a kernel may call this while the gradient is still
  // being computed.
  __device__ __forceinline__ void prefetch_leaf_rng() {
    if (phase == PH_TREE && !pre_valid) {
      Key tkey;
      split2(sub_key, part, pre_sub_key, tkey);
      pre_u = uniform01(tkey, part);
      pre_valid = 1;
    }
  }

  __device__ void prefetch_rng() {
    prefetch_leaf_rng();
    if (phase != PH_TREE && phase != PH_HMC) return;
    if (phase == PH_TREE && !pre_d_valid) {
      compute_doubling_rng(tree_key, pre_d);
      pre_d_valid = 1;
    }
    if (!pre_t_valid) {
      compute_transition_rng();
      pre_t_valid = 1;
    }
    if (!pre_a_valid && s.i < P->num_warmup) {
      split2(Key{s.adapt_key[0], s.adapt_key[1]}, part, pre_a_nk, pre_a_kss);
      pre_a_valid = 1;
    }
  }

  // build_tree._body_fn randomness (hmc_util.py:1161-1162) + _double_tree's split (:920) +
  // the top-level combine's bernoulli draw (:804), all functions of the tree key alone
  __device__ __forceinline__ void compute_doubling_rng(Key tk, DoublingRng& out) {
    Key dir_key, dbl_key;
    split3(tk, part, out.tree_key, dir_key, dbl_key);
    out.going_right = uniform01(dir_key, part) < 0.5f;
    split2(dbl_key, part, out.sub_key, out.top_key);
    out.u_top = uniform01(out.top_key, part);
  }

  // sample_kernel's split (hmc.py:472) + the N(0, I) draw of momentum_generator (hmc.py:94-102,
  // one dict block => split(key, 1)) + the first doubling's randomness
  __device__ __forceinline__ void compute_transition_rng() {
    Key k_mom;
    split3(Key{s.rng_key[0], s.rng_key[1]}, part, pre_t_rng_key, k_mom, pre_t_k_tr);
    draw_eps(k_mom, EPS_NEXT);
    if (P->algo == ALGO_NUTS) compute_doubling_rng(pre_t_k_tr, pre_t_d);
  }

  // r = M^{1/2} eps with the prefetched eps
  __device__ __forceinline__ void momentum_from_eps(int dst) {
    float* r = v(dst);
    const float* eps = v(EPS_NEXT);
    if (dense) {
      matvec(mat(1), eps, r, sqrt_is_lower());
      return;
    }
    const float* ms = v(MSQRT);
    B200MC_FOR_D r[d * es] = ms[d * es] * eps[d * es];
  }

  // Part B of the fused leaf step: fold the leaf (position fz, gradient fg, energy f_e_new)
  // into the subtree -- uniform / multinomial kernel, hmc_util.py:749-753, :1020-1032.
  __device__ void leaf_deferred() {
    deferred = 0;
    const float dE = f_dE;
    const float wleaf = -dE;
    const int leaf_div = dE > P->max_delta_energy;
    const float acc = fminf(expf(-dE), 1.0f);
    prefetch_leaf_rng();
    sub_key = pre_sub_key;
    pre_valid = 0;
    bool take = true;
    if (s_n == 0) {
      s_w = wleaf; s_div = leaf_div; s_sum_acc = acc;
      s_prop_pe = f_U; s_prop_energy = f_e_new;
      s_n = 1;
    } else {
      const float p = 1.0f / (1.0f + expf(-(wleaf - s_w)));  // expit
      take = pre_u < p;
      if (take) {
        s_prop_pe = f_U; s_prop_energy = f_e_new;
      }
      s_w = logaddexp_f32(s_w, wleaf);
      s_div = leaf_div;
      s_sum_acc = s_sum_acc + acc;
      s_n += 1;
    }
    if (take) {
      const int lane = g.lane();
      float *sz = v(SZ_PROP), *sg = v(SG_PROP);
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const int d = lane + 32 * k;
        if (d < D) {
          sz[d] = fz[k];
          sg[d] = fg[k];
        }
      }
    }
  }

  // _combine_tree(biased=True) minus the end / r_sum merge: transition kernel, proposal
  // choice and the scalar tree state (hmc_util.py:756-764, :796-848)
  __device__ void combine_bookkeeping(bool sub_turning, bool tree_turning) {
    float p = expf(s_w - weight);
    p = (sub_turning || s_div) ? 0.0f : fminf(p, 1.0f);
    if (u_top < p) {
      copy(Z_PROP, SZ_PROP); copy(G_PROP, SG_PROP);
      prop_pe = s_prop_pe; prop_energy = s_prop_energy;
    }
    depth += 1;
    weight = logaddexp_f32(weight, s_w);
    diverging = s_div;
    turning = tree_turning;
    sum_acc = sum_acc + s_sum_acc;
    nprop += s_n;
  }

  // What the critical path of advance_early left undone (see `deferred`).
  __device__ void run_deferred() {
    const int what = deferred;
    deferred = 0;
    if (what == 1) {
      leaf_deferred();
    } else if (what == 2) {  // a subtree ended and the next doubling's first point is out
      leaf_deferred();
      combine_bookkeeping(d_sub_turning, d_tree_turning);
      next_doubling_state();
    } else if (what == 3) {  // a transition ended and the next one's first point is out
      finish_transition(sum_acc / (float)nprop, nprop, diverging);
    }
  }

  // ------------------------------------------------------------------ NUTS: one leaf
  __device__ StepResult advance_tree(float U) {
    bool sub_turning = false;
    if (G::kRegVec && !dense && D <= 128) {
      // ---- steps 1-4 fused for one-chain-per-warp groups: the leaf's vectors stay in
      // registers (<= 4 elements per lane).  Part A is the critical path to the next point:
      // finish the leapfrog, energy and divergence test, checkpoint U-turn tests, next half
      // step.  Part B (multinomial fold, proposal copy: leaf_deferred) does not influence the
      // next point when the subtree continues, so a caller may run it AFTER handing the
      // point to whoever computes the gradient.  Same arithmetic, same order as below.
      B200MC_STAMP(t0_);
      const int lane = g.lane();
      float *pLZ = v(LZ), *pLR = v(LR), *pRS = v(S_RSUM);
      const float *pLG = v(LG), *pIM = v(IMM);
      const float hs = 0.5f * eps_s;
      float r[4], im[4], rs[4];
      float kacc = 0.f;
      const int leaf_idx = s_n;
      const bool first = (s_n == 0);
      const int idx_max = __popc(leaf_idx >> 1);
      const int num_sub = __popc(((~leaf_idx) & (leaf_idx + 1)) - 1);
      const int idx_min = idx_max - num_sub + 1;
      const bool even = (leaf_idx & 1) == 0;
      {
        float *cr = ckpt_r(idx_max), *crs = ckpt_rs(idx_max);
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          const int d = lane + 32 * k;
          r[k] = 0.f; im[k] = 0.f; rs[k] = 0.f; fz[k] = 0.f; fg[k] = 0.f;
          if (d < D) {
            fg[k] = pLG[d];
            im[k] = pIM[d];
            fz[k] = pLZ[d];
            const float t = hs * fg[k];
            r[k] = pLR[d] - t;                 // finish_leapfrog, hmc_util.py:306-308
            const float vv = im[k] * r[k];
            kacc += vv * r[k];
            rs[k] = first ? r[k] : pRS[d] + r[k];
            pRS[d] = rs[k];
            if (even) {
              cr[d] = r[k];
              crs[d] = rs[k];
            }
          }
        }
      }
      f_U = U;
      f_e_new = U + 0.5f * g.sum(kacc);
      float dE = f_e_new - e0;
      if (isnan(dE)) dE = INFINITY;
      f_dE = dE;
      const int leaf_div = dE > P->max_delta_energy;
      B200MC_STAMP(t1_);
      B200MC_ACC(0, t0_, t1_);
      for (int i = idx_max; i >= idx_min && !sub_turning; --i) {
        const float *cr = ckpt_r(i), *crs = ckpt_rs(i);
        float la = 0.f, ra = 0.f;
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          const int d = lane + 32 * k;
          if (d < D) {
            const float l = cr[d];
            float sv = rs[k] - crs[d];
            sv = sv + l;
            const float h = (l + r[k]) / 2.0f;
            const float sp = sv - h;
            const float vl = im[k] * l, vr = im[k] * r[k];
            la += vl * sp;
            ra += vr * sp;
          }
        }
        g.sum2(la, ra);
        sub_turning = (la <= 0.f) | (ra <= 0.f);
      }
      B200MC_STAMP(t3_);
      B200MC_ACC(2, t1_, t3_);
      if (s_n + 1 < (1 << depth) && !sub_turning && !leaf_div) {
#pragma unroll
        for (int k = 0; k < 4; ++k) {  // half_step of the next leaf, hmc_util.py:297-302
          const int d = lane + 32 * k;
          if (d < D) {
            const float t = hs * fg[k];
            const float rn = r[k] - t;
            pLR[d] = rn;
            const float vv = im[k] * rn;
            const float dz = eps_s * vv;
            pLZ[d] = fz[k] + dz;
          }
        }
        g.sync();
        B200MC_STAMP(t4_);
        B200MC_ACC(3, t3_, t4_);
        B200MC_ACC(4, 0, 1);
        deferred = 1;  // leaf_deferred() still has to fold this leaf into the subtree
        return NEED_GRAD;
      }
      // ---- the subtree is finished.  Critical path first: merge the tree ends, the
      // tree-level U-turn test (:710-746 on the merged tree) and, if the tree doubles again,
      // the first half step of the next doubling; the bookkeeping follows in run_deferred().
      {
        float *ze = v(going_right ? Z_RIGHT : Z_LEFT), *re = v(going_right ? R_RIGHT : R_LEFT),
              *ge = v(going_right ? G_RIGHT : G_LEFT), *pSUM = v(R_SUM);
        const float* ro = v(going_right ? R_LEFT : R_RIGHT);
        float la = 0.f, ra = 0.f;
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          const int d = lane + 32 * k;
          if (d < D) {
            pLR[d] = r[k];
            ze[d] = fz[k]; re[d] = r[k]; ge[d] = fg[k];
            const float rsum = pSUM[d] + rs[k];
            pSUM[d] = rsum;
            const float o = ro[d];
            const float l = going_right ? o : r[k], rr_ = going_right ? r[k] : o;
            const float h = (l + rr_) / 2.0f;
            const float sp = rsum - h;
            const float vl = im[k] * l, vr = im[k] * rr_;
            la += vl * sp;
            ra += vr * sp;
          }
        }
        g.sum2(la, ra);
        const bool tree_turning = sub_turning | ((la <= 0.f) | (ra <= 0.f));
        d_sub_turning = sub_turning;
        d_tree_turning = tree_turning;
        if (depth + 1 < max_depth_cur && !tree_turning && !leaf_div) {
          if (!pre_d_valid) {
            compute_doubling_rng(tree_key, pre_d);
            pre_d_valid = 1;
          }
          const bool gr = pre_d.going_right != 0;
          const float e2 = gr ? s.step_size : -s.step_size;
          const float hs2 = 0.5f * e2;
          const float *zs_ = v(gr ? Z_RIGHT : Z_LEFT), *rs_ = v(gr ? R_RIGHT : R_LEFT),
                      *gs_ = v(gr ? G_RIGHT : G_LEFT);
          float* wLG = v(LG);
          g.sync();
#pragma unroll
          for (int k = 0; k < 4; ++k) {  // next_doubling's copies + half_step
            const int d = lane + 32 * k;
            if (d < D) {
              const float gg = gs_[d];
              wLG[d] = gg;
              const float t = hs2 * gg;
              const float rn = rs_[d] - t;
              pLR[d] = rn;
              const float vv = im[k] * rn;
              const float dz = e2 * vv;
              pLZ[d] = zs_[d] + dz;
            }
          }
          g.sync();
          deferred = 2;
          return NEED_GRAD;
        }
        // the transition ends here
        leaf_deferred();
        combine_bookkeeping(sub_turning, tree_turning);
        s.potential_energy = prop_pe;
        s.energy = prop_energy;
        if (s.i >= P->num_warmup && iters_left > 1 && !transition_hook) {
          // sampling phase: nothing adapts between transitions, so the first point of the
          // next transition is known now (prefetched momentum draw and direction);
          // finish_transition / start_transition run in run_deferred() and recompute it
          // bit-identically.
          if (!pre_t_valid) {
            compute_transition_rng();
            pre_t_valid = 1;
          }
          const bool gr = pre_t_d.going_right != 0;
          const float e2 = gr ? s.step_size : -s.step_size;
          const float hs2 = 0.5f * e2;
          const float *zp = v(Z_PROP), *gp = v(G_PROP), *ms = v(MSQRT), *eps = v(EPS_NEXT);
          g.sync();
#pragma unroll
          for (int k = 0; k < 4; ++k) {
            const int d = lane + 32 * k;
            if (d < D) {
              const float r0 = ms[d] * eps[d];
              const float t = hs2 * gp[d];
              const float rn = r0 - t;
              const float vv = im[k] * rn;
              const float dz = e2 * vv;
              pLZ[d] = zp[d] + dz;
            }
          }
          g.sync();
          deferred = 3;
          return NEED_GRAD;
        }
        return finish_transition(sum_acc / (float)nprop, nprop, diverging);
      }
    } else {
      // 1. finish the leapfrog, energy test (_build_basetree, hmc_util.py:863-875)
      B200MC_STAMP(t0_);
      finish_leapfrog();
      const float e_new = U + kinetic(LR, LV);
      float dE = e_new - e0;
      if (isnan(dE)) dE = INFINITY;
      const float wleaf = -dE;
      const int leaf_div = dE > P->max_delta_energy;
      const float acc = fminf(expf(-dE), 1.0f);

      // 2. fold the leaf into the subtree (uniform / multinomial kernel, :749-753, :1020-1032)
      B200MC_STAMP(t1_);
      B200MC_ACC(0, t0_, t1_);
      prefetch_leaf_rng();
      sub_key = pre_sub_key;
      pre_valid = 0;
      const int leaf_idx = s_n;
      if (s_n == 0) {
        s_w = wleaf; s_div = leaf_div; s_sum_acc = acc;
        s_prop_pe = U; s_prop_energy = e_new;
        copy(SZ_PROP, LZ); copy(SG_PROP, LG); copy(S_RSUM, LR);
        s_n = 1;
      } else {
        const float p = 1.0f / (1.0f + expf(-(wleaf - s_w)));  // expit
        const bool take = pre_u < p;
        if (take) {
          s_prop_pe = U; s_prop_energy = e_new;
          copy(SZ_PROP, LZ); copy(SG_PROP, LG);
        }
        s_w = logaddexp_f32(s_w, wleaf);
        s_div = leaf_div;
        s_sum_acc = s_sum_acc + acc;
        s_n += 1;
        float* rs = v(S_RSUM);
        const float* r = v(LR);
        B200MC_FOR_D rs[d * es] = rs[d * es] + r[d * es];
      }

      // 3. checkpoints and iterative U-turn test (:941-981, :1034-1058)
      B200MC_STAMP(t2_);
      B200MC_ACC(1, t1_, t2_);
      const int idx_max = __popc(leaf_idx >> 1);
      const int num_sub = __popc(((~leaf_idx) & (leaf_idx + 1)) - 1);
      const int idx_min = idx_max - num_sub + 1;
      if ((leaf_idx & 1) == 0) {
        float *cr = ckpt_r(idx_max), *crs = ckpt_rs(idx_max);
        const float *r = v(LR), *rs = v(S_RSUM);
        B200MC_FOR_D {
          cr[d * es] = r[d * es];
          crs[d * es] = rs[d * es];
        }
        if (dense) {
          float* cv_ = ckpt_v(idx_max);
          const float* lv = v(LV);
          B200MC_FOR_D cv_[d * es] = lv[d * es];
        }
      }
      for (int i = idx_max; i >= idx_min && !sub_turning; --i)
        sub_turning = is_turning(ckpt_r(i), v(LR), v(S_RSUM), ckpt_rs(i), ckpt_r(i), ckpt_v(i), v(LV));

      // 4. keep extending the subtree? (_cond_fn, :999-1001)
      B200MC_STAMP(t3_);
      B200MC_ACC(2, t2_, t3_);
      if (s_n < (1 << depth) && !sub_turning && !s_div) {
        half_step();
        B200MC_STAMP(t4_);
        B200MC_ACC(3, t3_, t4_);
        B200MC_ACC(4, 0, 1);
        return NEED_GRAD;
      }

    }

    if (end_yield) {
      y_sub_turning = sub_turning;
      return AT_SUBTREE_END;
    }
    return finish_subtree(sub_turning);
  }

  __device__ StepResult finish_subtree_yielded() { return finish_subtree(y_sub_turning); }

  __device__ StepResult finish_subtree(bool sub_turning) {
    // 5. subtree finished: biased-progressive combine with the tree (:767-848, biased=True)
    B200MC_STAMP(t3_);
    if (going_right) {
      copy(Z_RIGHT, LZ); copy(R_RIGHT, LR); copy(G_RIGHT, LG);
      if (dense) copy(V_RIGHT, LV);
      if (dense && G::kBlockWide) copy(Q_RIGHT, LQ);
    } else {
      copy(Z_LEFT, LZ); copy(R_LEFT, LR); copy(G_LEFT, LG);
      if (dense) copy(V_LEFT, LV);
      if (dense && G::kBlockWide) copy(Q_LEFT, LQ);
    }
    {
      float* rs = v(R_SUM);
      const float* srs = v(S_RSUM);
      B200MC_FOR_D rs[d * es] = rs[d * es] + srs[d * es];
    }
    const bool tree_turning =
        sub_turning | is_turning(v(R_LEFT), v(R_RIGHT), v(R_SUM), nullptr, nullptr, v(V_LEFT),
                                 v(V_RIGHT));
    combine_bookkeeping(sub_turning, tree_turning);

    // 6. double again? (build_tree._cond_fn, :1155-1157)
    B200MC_STAMP(t5_);
    B200MC_ACC(5, t3_, t5_);
    if (depth < max_depth_cur && !turning && !diverging) {
      next_doubling();
      B200MC_STAMP(t6_);
      B200MC_ACC(6, t5_, t6_);
      B200MC_ACC(7, 0, 1);
      return NEED_GRAD;
    }

    // 7. _nuts_next tail (hmc.py:441-455)
    s.potential_energy = prop_pe;
    s.energy = prop_energy;
    return finish_transition(sum_acc / (float)nprop, nprop, diverging);
  }

  // ------------------------------------------------------------------ HMC (hmc.py:364-414)
  __device__ __noinline__ StepResult start_hmc(Key k_tr) {
    momentum_from_eps(LR);
    copy(LZ, Z_PROP); copy(LG, G_PROP);
    mh_key = k_tr;
    float eps = s.step_size;
    int n;
    if (P->fixed_num_steps > 0) n = P->fixed_num_steps;
    else n = (int)ceilf(P->trajectory_length / eps);  // hmc.py:85-89
    if (P->has_trajectory_length) eps = P->trajectory_length / (float)n;
    hmc_steps = n; hmc_done = 0;
    eps_s = eps;
    if (dense) dense_leaf_cache();
    energy_old = s.potential_energy + kinetic(LR, LV);
    phase = PH_HMC;
    half_step();
    return NEED_GRAD;
  }

  __device__ __noinline__ StepResult advance_hmc(float U) {
    finish_leapfrog();
    hmc_done += 1;
    if (hmc_done < hmc_steps) {
      half_step();
      return NEED_GRAD;
    }
    const float e_new = U + kinetic(LR, LV);
    float dE = e_new - energy_old;
    if (isnan(dE)) dE = INFINITY;
    const float acc = fminf(expf(-dE), 1.0f);
    const int div = dE > P->max_delta_energy;
    if (uniform01(mh_key, part) < acc) {
      copy(Z_PROP, LZ); copy(G_PROP, LG);
      s.potential_energy = U;
      s.energy = e_new;
    } else {
      s.energy = energy_old;
    }
    return finish_transition(acc, hmc_steps, div);
  }

  // ------------------------------------------------------------------ initialisation
  // HMC.init (hmc.py:744-750): rng_key, rng_key_init_model = split(rng_key)
  __device__ __noinline__ StepResult begin_init(Key chain_key, bool has_user_z) {
    Key k_run, k_init;
    split2(chain_key, part, k_run, k_init);
    s.rng_key[0] = k_run.k0; s.rng_key[1] = k_run.k1;
    s.status = 0; s.init_tries = 0; s.grad_evals = 0;
    init_key = k_init;
    user_init = has_user_z ? 1 : 0;
    phase = PH_INIT;
    if (user_init) copy(LZ, Z_PROP);
    else init_candidate();
    return NEED_GRAD;
  }

  // find_valid_initial_params body_fn, fast branch (infer/util.py:390-432)
  __device__ __noinline__ void init_candidate() {
    Key key, subkey;
    split2(init_key, part, key, subkey);
    float* z = v(LZ);
    for (int k = 0; k < P->nsites; ++k) {
      const int off = P->site_off[k], n = P->site_size[k];
      for (int j = g.lane(); j < n; j += g.size())
        z[(off + j) * es] =
            uniform_at(subkey, (uint32_t)n, (uint32_t)j, -P->init_radius, P->init_radius, part);
      Key nk, ns;
      split2(key, part, nk, ns);
      key = nk; subkey = ns;
    }
    init_key = key;
    s.init_tries += 1;
  }

  __device__ __noinline__ StepResult advance_init(float U) {
    if (!user_init) {
      const float* gr = v(LG);
      float bad = 0.f;
      B200MC_FOR_D bad += isfinite(gr[d * es]) ? 0.f : 1.f;
      bad = g.sum(bad);
      const bool valid = isfinite(U) && bad == 0.f;
      if (!valid) {
        if (s.init_tries >= 100) {  // RuntimeError("Cannot find valid initial parameters")
          s.status = 1;
          phase = PH_FAIL;
          return FINISHED;
        }
        init_candidate();
        return NEED_GRAD;
      }
    }
    copy(Z_PROP, LZ); copy(G_PROP, LG);
    s.potential_energy = U;
    return init_kernel();
  }

  // init_kernel (hmc.py:335-362) + warmup_adapter.init_fn (hmc_util.py:552-594); the
  // host has put the initial mass matrix into IMM / MSQRT (diagonal) or mat(0..2) (dense).
  __device__ __noinline__ StepResult init_kernel() {
    Key k_hmc, k_wa, k_adapt, k_ss;
    split3(Key{s.rng_key[0], s.rng_key[1]}, part, k_hmc, k_wa, init_mom_key);
    split2(k_wa, part, k_adapt, k_ss);
    s.adapt_key[0] = k_adapt.k0; s.adapt_key[1] = k_adapt.k1;
    s.rng_key[0] = k_hmc.k0; s.rng_key[1] = k_hmc.k1;
    s.step_size = P->init_step_size;
    {
      float *mean = v(WMEAN), *m2 = v(WM2);
      B200MC_FOR_D { mean[d * es] = 0.f; m2[d * es] = 0.f; }
    }
    s.mm_n = 0;
    s.window_idx = 0;
    if (dense && P->factor_at_init) factor_dense();  // user-supplied dense M^-1, :497-507
    if (P->adapt_step_size && P->find_heuristic) {   // :573-576
      start_frss(k_ss, 0);
      if (frss_try()) return NEED_GRAD;
    }
    da_init(logf(10.0f * s.step_size));
    return after_init();
  }

  __device__ __noinline__ StepResult after_init() {
    draw_momentum(init_mom_key, LR);  // value discarded by the reference except for `energy`
    if (dense) matvec(mat(0), v(LR), v(LV));
    s.energy = s.potential_energy + kinetic(LR, LV);
    s.i = 0; s.num_steps = 0; s.accept_prob = 0.f; s.mean_accept_prob = 0.f; s.diverging = 0;
    if (iters_left <= 0) {
      phase = PH_IDLE;
      return FINISHED;
    }
    if (transition_hook) return AT_TRANSITION_END;
    return start_transition();
  }

  // ------------------------------------------------------------------ driver interface
  __device__ StepResult begin_sampling() {
    if (iters_left <= 0 || s.status != 0) return FINISHED;
    if (transition_hook) return AT_TRANSITION_END;
    return start_transition();
  }

  // The gradient of the potential at v(LZ) has been written to v(LG); U is its value.
  __device__ StepResult advance(float U) {
    if constexpr (G::kBlockWide) {
      if (mv_yield && dense && (phase == PH_TREE || phase == PH_HMC || phase == PH_FRSS)) {
        // second half of the leapfrog, then LV = M^-1 LR and LQ = M^-1 LG from the caller
        float* r = v(LR);
        const float* gr = v(LG);
        const float hs = 0.5f * eps_s;
        B200MC_FOR_D {
          const float t = hs * gr[d];
          r[d] = r[d] - t;
        }
        rq = MatvecRequest{0, 0, LR, LV, LG, LQ};
        pending_U = U;
        mv_resume = 1;
        return NEED_MATVEC;
      }
    }
    const StepResult r = advance_early(U);
    if (deferred) run_deferred();
    return r;
  }

  // The product requested through `rq` is in place: continue.
  __device__ StepResult resume() {
    const int where = mv_resume;
    mv_resume = 0;
    if (where == 1) {
      leaf_cached = true;
      const StepResult r = advance_early(pending_U);
      leaf_cached = false;
      if (deferred) run_deferred();
      return r;
    }
    if (where == 2) return start_transition_b();
    if (where == 3) return start_transition_c();
    if (where == 5) return run_adapt(1);
    if (where == 6) return run_adapt(2);
    return FINISHED;
  }

  // advance() in two parts: when `deferred` is set on return, the next point is ready in
  // v(LZ) but leaf_deferred() must still run before the next call (it may overlap with the
  // gradient evaluation of that point).
  __device__ StepResult advance_early(float U) {
    s.grad_evals += 1;
    if (phase == PH_TREE) return advance_tree(U);
    if (phase == PH_HMC) return advance_hmc(U);
    if (phase == PH_INIT) return advance_init(U);
    if (phase == PH_FRSS) return advance_frss(U);
    return FINISHED;
  }
#undef B200MC_FOR_D
#undef es
#undef LV
#undef V_LEFT
#undef V_RIGHT
#undef LQ
#undef Q_LEFT
#undef Q_RIGHT
};

}  // namespace b200mc
