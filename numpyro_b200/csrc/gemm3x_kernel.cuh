// Split-precision ("3xTF32") GEMM on the 5th-generation tensor cores, for the many-chain
// potentials: once >= 64 chains share one design matrix, X.Theta and X^T.R are dense
// contractions, not matrix-vector sweeps.
//
//   D[m, c] = sum_k A[m, k] * B[c, k]          (A: [RA][K], B: [256][K], both K-major)
//
// with every f32 operand split as x = hi + lo (TF32 each) and three tcgen05.mma per k-step,
// hi*hi + hi*lo + lo*hi, accumulated in f32 in TMEM: the result holds f32-level accuracy
// (~2^-22 per product) at 3 tensor-core passes.  Operands arrive as pre-swizzled "operand
// images" (umma.cuh) through 1-D TMA bulk copies; one CTA = one 128 x 256 output tile over
// a k-block range (split-K); warp 0 = TMA producer, warp 1 = MMA issuer (one lane) and TMEM
// owner, warps 2.. = epilogue (tcgen05.ld; TMEM lane quadrant = warp % 4, kGemmEpiGroups warps
// per quadrant share its 256 columns: the residual epilogue is transcendental-bound).
//
// Epilogues:
//   EPI_STORE  partial tile -> out[split][c][m] (chain-major; the consumer sums the splits in a
//              fixed order: deterministic, no float atomics)
//   EPI_RESID  rows are data points: turns the linear predictor into the likelihood residual
//              dU/dl (Bernoulli logits, distributions/util.py:295-298, or Normal with the
//              chain's precision, continuous.py:2203-2207) written straight into the hi/lo
//              operand images of the next GEMM (X^T.R), plus per-32-row partial sums of the
//              log-likelihood terms.
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

#include "glm_kernel.cuh"  // mbarrier / bulk-copy helpers
#include "umma.cuh"

namespace b200mc {

constexpr int kGemmM = 128;        // rows of A per CTA = TMEM lanes
constexpr int kGemmN = 256;        // chains per tile = TMEM columns
constexpr int kGemmStages = 2;
constexpr int kGemmEpiGroups = 4;  // epilogue warps per TMEM lane quadrant (columns split among them)
constexpr int kGemmThreads = 64 + 128 * kGemmEpiGroups;  // producer, MMA issuer, epilogue warps
constexpr int kGemmABytes = kGemmM * 128;  // one operand tile of A: 128 rows x 32 f32
constexpr int kGemmBBytes = kGemmN * 128;
constexpr int kGemmStageBytes = 2 * kGemmABytes + 2 * kGemmBBytes;  // hi + lo of both = 96 KB
constexpr int kGemmSmem = kGemmStages * kGemmStageBytes + 1024 /*align*/ + 256 /*barriers*/;

enum GemmEpilogue { EPI_STORE = 0, EPI_RESID = 1 };

// binary_cross_entropy_with_logits (distributions/util.py:295-298) and its derivative for ONE
// (row, chain) element of the residual epilogue, on the special-function unit only: three MUFU
// ops (ex2, rcp, lg2) and no libdevice slow paths -- the epilogue runs once per element of the
// N x C linear predictor and is the bound of the whole contraction.  Every intermediate lies
// in [0, 2], so the approximations' errors are ABSOLUTE errors of ~1e-7 on quantities of
// order 1 (e = exp(-|l|) <= 1, 1 + e in [1, 2], log1p(e) <= ln 2): the same size as the f32
// round-off of the exact formula and far inside tier 2 (rtol 1e-5 on U and grad U).
__device__ __forceinline__ void bce_logits_fast(float l, float y, float& term, float& res) {
  float e, inv, lg;
  const float t = fabsf(l) * -1.4426950408889634f;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(e) : "f"(t));
  const float d = 1.0f + e;
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(inv) : "f"(d));
  asm("lg2.approx.ftz.f32 %0, %1;" : "=f"(lg) : "f"(d));
  // (rcp.approx is good to 1 ulp on [1, 2]: no Newton step)
  term = fmaf(-l, y, fmaf(lg, 0.6931471805599453f, fmaxf(l, 0.f)));
  res = ((l >= 0.f) ? inv : e * inv) - y;
}

struct Gemm3xLaunch {
  const unsigned char *a_hi, *a_lo;  // images of A [RA][32*KB]
  const unsigned char *b_hi, *b_lo;  // images of B [256][32*KB]
  int RA;      // rows of A, multiple of 128
  int KB;      // k-blocks (K / 32)
  int splits;  // split-K: grid = (RA / 128) * splits, every split non-empty (splits <= KB)
  // EPI_STORE
  float* out;  // [splits][256][RA]
  // EPI_RESID (splits == 1)
  const float* y;          // [n_valid] observations
  int n_valid;             // rows >= n_valid are padding: residual 0, no log-lik
  int likelihood;          // 0 Bernoulli(logits), 1 Normal
  const float* col_scale;  // Normal: precision per chain [256]
  unsigned char *r_hi, *r_lo;  // images of R [256][RA] (K-major in the data index)
  float* ll_part;          // [RA / 32][256] partial sums: BCE terms, or squared residuals
};

template <int EPI>
__global__ void __launch_bounds__(kGemmThreads, 1) gemm3x_kernel(const __grid_constant__ Gemm3xLaunch L) {
  extern __shared__ unsigned char gsm_raw[];
  unsigned char* gsm = reinterpret_cast<unsigned char*>(((uintptr_t)gsm_raw + 1023) & ~(uintptr_t)1023);
  uint64_t* full = reinterpret_cast<uint64_t*>(gsm + kGemmStages * kGemmStageBytes);
  uint64_t* empty = full + kGemmStages;
  uint64_t* acc_full = empty + kGemmStages;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(acc_full + 1);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int mt = blockIdx.x / L.splits, sp = blockIdx.x - mt * L.splits;
  const int kb0 = (int)((long)sp * L.KB / L.splits), kb1 = (int)((long)(sp + 1) * L.KB / L.splits);

  if (threadIdx.x == 0) {
    for (int s = 0; s < kGemmStages; ++s) {
      mbar_init(&full[s], 1);
      mbar_init(&empty[s], 1);
    }
    mbar_init(acc_full, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 1) umma::tmem_alloc(tmem_slot, kGemmN);
  umma::fence_before_sync();
  __syncthreads();
  umma::fence_after_sync();
  const uint32_t tmem = *tmem_slot;

  if (warp == 0) {
    // ---------------- producer: one lane streams the k-blocks of this split
    if (lane == 0) {
      const long a_rowgroups = L.RA >> 3;
      for (int kb = kb0, it = 0; kb < kb1; ++kb, ++it) {
        const int s = it % kGemmStages;
        const uint32_t n = (uint32_t)(it / kGemmStages);
        mbar_wait(&empty[s], (n & 1u) ^ 1u);
        unsigned char* st = gsm + (long)s * kGemmStageBytes;
        mbar_arrive_expect_tx(&full[s], (uint32_t)kGemmStageBytes);
        const long ao = ((long)kb * a_rowgroups + (long)mt * (kGemmM >> 3)) * umma::kAtomBytes;
        const long bo = (long)kb * (kGemmN >> 3) * umma::kAtomBytes;
        tma_load_1d(st, L.a_hi + ao, kGemmABytes, &full[s]);
        tma_load_1d(st + kGemmABytes, L.a_lo + ao, kGemmABytes, &full[s]);
        tma_load_1d(st + 2 * kGemmABytes, L.b_hi + bo, kGemmBBytes, &full[s]);
        tma_load_1d(st + 2 * kGemmABytes + kGemmBBytes, L.b_lo + bo, kGemmBBytes, &full[s]);
      }
    }
  } else if (warp == 1) {
    // ---------------- MMA issuer
    if (lane == 0) {
      constexpr uint32_t idesc = umma::make_idesc_tf32(kGemmM, kGemmN);
      for (int kb = kb0, it = 0; kb < kb1; ++kb, ++it) {
        const int s = it % kGemmStages;
        const uint32_t n = (uint32_t)(it / kGemmStages);
        mbar_wait(&full[s], n & 1u);
        umma::fence_after_sync();
        const uint32_t sa = umma::smem_addr(gsm + (long)s * kGemmStageBytes);
        const uint64_t a_hi = umma::make_desc_k_sw128(sa);
        const uint64_t a_lo = umma::make_desc_k_sw128(sa + kGemmABytes);
        const uint64_t b_hi = umma::make_desc_k_sw128(sa + 2 * kGemmABytes);
        const uint64_t b_lo = umma::make_desc_k_sw128(sa + 2 * kGemmABytes + kGemmBBytes);
#pragma unroll
        for (int k = 0; k < umma::kBlockK / 8; ++k) {
          const uint64_t adv = (uint64_t)(k * 2);  // 8 tf32 = 32 bytes = 2 x 16-byte units
          // small terms first, then the leading one
          umma::mma_tf32(tmem, a_lo + adv, b_hi + adv, idesc, (it > 0 || k > 0) ? 1u : 0u);
          umma::mma_tf32(tmem, a_hi + adv, b_lo + adv, idesc, 1u);
          umma::mma_tf32(tmem, a_hi + adv, b_hi + adv, idesc, 1u);
        }
        umma::mma_commit(&empty[s]);  // the stage is free once these MMAs have read it
      }
      umma::mma_commit(acc_full);
    }
  } else {
    // ---------------- epilogue: warp w may touch TMEM lanes 32 * (w % 4) ..
    const int q = warp & 3;
    const int c_begin = ((warp - 2) >> 2) * (kGemmN / kGemmEpiGroups), c_end = c_begin + kGemmN / kGemmEpiGroups;
    mbar_wait(acc_full, 0u);
    umma::fence_after_sync();
    const int m = mt * kGemmM + q * 32 + lane;  // this thread's row of A
    const uint32_t trow = tmem + ((uint32_t)(q * 32) << 16);
    if (EPI == EPI_STORE) {
      float* out = L.out + (long)sp * kGemmN * L.RA;
#pragma unroll 1
      for (int c0 = c_begin; c0 < c_end; c0 += 32) {
        uint32_t v[32];
        umma::tmem_ld32(trow + c0, v);
        umma::tmem_ld_wait();
#pragma unroll
        for (int j = 0; j < 32; ++j) out[(long)(c0 + j) * L.RA + m] = __uint_as_float(v[j]);
      }
    } else {
      const bool valid = m < L.n_valid;
      const float yv = valid ? L.y[m] : 0.f;
      const long kbR = m >> 5;  // this warp's 32 rows are one k-block of the R images
      unsigned char* rh = L.r_hi + kbR * (kGemmN >> 3) * umma::kAtomBytes;
      unsigned char* rl = L.r_lo + kbR * (kGemmN >> 3) * umma::kAtomBytes;
      float* llp = L.ll_part + kbR * kGemmN;
#pragma unroll 1
      for (int c0 = c_begin; c0 < c_end; c0 += 32) {
        uint32_t v[32];
        umma::tmem_ld32(trow + c0, v);
        umma::tmem_ld_wait();
        float part[32];
#pragma unroll
        for (int j = 0; j < 32; ++j) {
          const int c = c0 + j;
          const float l = __uint_as_float(v[j]);
          float res = 0.f, term = 0.f;
          if (valid) {
            if (L.likelihood == 1) {
              const float r = yv - l;
              term = r * r;
              res = -(L.col_scale[c] * r);
            } else {
              bce_logits_fast(l, yv, term, res);
            }
          }
          float hi, lo;
          umma::split_tf32(res, hi, lo);
          // image element (row c, k = m): same 128-byte line for the whole warp
          const long off = (long)(c >> 3) * umma::kAtomBytes + (c & 7) * 128 +
                           (((lane >> 2) ^ (c & 7)) << 4) + ((lane & 3) << 2);
          *reinterpret_cast<float*>(rh + off) = hi;
          *reinterpret_cast<float*>(rl + off) = lo;
          part[j] = term;
        }
        const float tot = warp_transpose_reduce32(part);  // lane j: sum over the 32 rows of column c0 + j
        llp[c0 + lane] = tot;
      }
    }
    umma::fence_before_sync();
  }
  __syncthreads();
  if (warp == 1) {
    umma::fence_after_sync();
    umma::tmem_dealloc(tmem, kGemmN);
  }
}

// ---------------------------------------------------------------- persistent residual GEMM
// Same contraction + residual epilogue as gemm3x_kernel<EPI_RESID>, for a SHORT inner dimension
// (KB <= kPersistMaxKB k-blocks: the logistic design matrix has K = 64): one CTA per SM walks
// the row tiles mt = blockIdx.x, blockIdx.x + gridDim.x, ...
//   * B (the 256 coefficient rows, hi + lo) is loaded ONCE and stays in shared memory (128 KB);
//   * A streams through a 3-stage ring of k-blocks (32 KB each, hi + lo);
//   * the accumulator is double-buffered in TMEM (2 x 256 columns = the whole 512): the MMAs of
//     tile i + 1 run while the 16 epilogue warps drain tile i, so loads, tensor-core work and
//     the transcendental-bound epilogue overlap instead of running back to back per CTA.
constexpr int kPersistMaxKB = 2;
constexpr int kPersistStages = 3;
constexpr int kPersistABytes = 2 * kGemmABytes;                      // one k-block of A, hi + lo
constexpr int kPersistBBytes = kPersistMaxKB * 2 * kGemmBBytes;      // all of B, hi + lo
constexpr int kPersistSmem = kPersistBBytes + kPersistStages * kPersistABytes + 1024 + 256;
constexpr int kPersistEpiWarps = 4 * kGemmEpiGroups;

// One 32-column chunk of the residual epilogue for this thread's row: v = the linear predictor
// of 32 chains.  LIK: 0 Bernoulli(logits), 1 Normal; CHECK: the tile holds padding rows.  The
// eight line pointers bh/bl[k] (k = chain & 7) already carry the swizzle term of this lane, so
// a store is [pointer + (j >> 3) * 1024] with no address arithmetic left in the loop.
template <int LIK, bool CHECK>
__device__ __forceinline__ void resid_chunk(const uint32_t (&v)[32], float yv, bool valid,
                                            const float* __restrict__ col_scale_c0,
                                            unsigned char* const (&bh)[8], unsigned char* const (&bl)[8],
                                            float (&part)[32]) {
#pragma unroll
  for (int j = 0; j < 32; ++j) {
    const float l = __uint_as_float(v[j]);
    float res, term;
    if (LIK == 1) {
      const float r = yv - l;
      term = r * r;
      res = -(col_scale_c0[j] * r);
    } else {
      bce_logits_fast(l, yv, term, res);
    }
    if (CHECK) {
      res = valid ? res : 0.f;
      term = valid ? term : 0.f;
    }
    // x = hi + lo in integer arithmetic (res is finite and small: no inf/NaN path needed):
    // hi = res rounded to TF32, ties away (== cvt.rna.tf32.f32); lo = res - hi is exact and
    // its low 13 bits are dropped exactly as the tensor core would drop them
    const uint32_t hb = (__float_as_uint(res) + 0x1000u) & 0xFFFFE000u;
    const float hi = __uint_as_float(hb);
    const uint32_t lb = __float_as_uint(res - hi) & 0xFFFFE000u;
    *reinterpret_cast<uint32_t*>(bh[j & 7] + (j >> 3) * umma::kAtomBytes) = hb;
    *reinterpret_cast<uint32_t*>(bl[j & 7] + (j >> 3) * umma::kAtomBytes) = lb;
    part[j] = term;
  }
}

template <int LIK>
__global__ void __launch_bounds__(kGemmThreads, 1)
gemm3x_resid_persistent_kernel(const __grid_constant__ Gemm3xLaunch L) {
  extern __shared__ unsigned char gsm_raw[];
  unsigned char* gsm = reinterpret_cast<unsigned char*>(((uintptr_t)gsm_raw + 1023) & ~(uintptr_t)1023);
  unsigned char* smB = gsm;
  unsigned char* smA = gsm + kPersistBBytes;
  uint64_t* a_full = reinterpret_cast<uint64_t*>(smA + kPersistStages * kPersistABytes);
  uint64_t* a_empty = a_full + kPersistStages;
  uint64_t* acc_full = a_empty + kPersistStages;  // [2]
  uint64_t* acc_empty = acc_full + 2;             // [2]
  uint64_t* b_full = acc_empty + 2;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(b_full + 1);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int n_tiles = L.RA / kGemmM;
  const int KB = L.KB;

  if (threadIdx.x == 0) {
    for (int s = 0; s < kPersistStages; ++s) {
      mbar_init(&a_full[s], 1);
      mbar_init(&a_empty[s], 1);
    }
    for (int b = 0; b < 2; ++b) {
      mbar_init(&acc_full[b], 1);
      mbar_init(&acc_empty[b], kPersistEpiWarps);
    }
    mbar_init(b_full, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 1) umma::tmem_alloc(tmem_slot, 2 * kGemmN);
  umma::fence_before_sync();
  __syncthreads();
  umma::fence_after_sync();
  const uint32_t tmem = *tmem_slot;

  if (warp == 0) {
    // ---------------- producer
    if (lane == 0) {
      mbar_arrive_expect_tx(b_full, (uint32_t)(KB * 2 * kGemmBBytes));
      for (int kb = 0; kb < KB; ++kb) {
        const long bo = (long)kb * (kGemmN >> 3) * umma::kAtomBytes;
        tma_load_1d(smB + (long)kb * 2 * kGemmBBytes, L.b_hi + bo, kGemmBBytes, b_full);
        tma_load_1d(smB + (long)kb * 2 * kGemmBBytes + kGemmBBytes, L.b_lo + bo, kGemmBBytes, b_full);
      }
      const long a_rowgroups = L.RA >> 3;
      int it = 0;
      for (int mt = blockIdx.x; mt < n_tiles; mt += gridDim.x) {
        for (int kb = 0; kb < KB; ++kb, ++it) {
          const int s = it % kPersistStages;
          const uint32_t n = (uint32_t)(it / kPersistStages);
          mbar_wait(&a_empty[s], (n & 1u) ^ 1u);
          unsigned char* st = smA + (long)s * kPersistABytes;
          mbar_arrive_expect_tx(&a_full[s], (uint32_t)kPersistABytes);
          const long ao = ((long)kb * a_rowgroups + (long)mt * (kGemmM >> 3)) * umma::kAtomBytes;
          tma_load_1d(st, L.a_hi + ao, kGemmABytes, &a_full[s]);
          tma_load_1d(st + kGemmABytes, L.a_lo + ao, kGemmABytes, &a_full[s]);
        }
      }
    }
  } else if (warp == 1) {
    // ---------------- MMA issuer
    if (lane == 0) {
      constexpr uint32_t idesc = umma::make_idesc_tf32(kGemmM, kGemmN);
      mbar_wait(b_full, 0u);
      umma::fence_after_sync();
      const uint32_t sb = umma::smem_addr(smB);
      int it = 0, i = 0;
      for (int mt = blockIdx.x; mt < n_tiles; mt += gridDim.x, ++i) {
        const int buf = i & 1;
        const uint32_t u = (uint32_t)(i >> 1);
        mbar_wait(&acc_empty[buf], (u & 1u) ^ 1u);  // the epilogue has drained this accumulator
        umma::fence_after_sync();
        const uint32_t d_tmem = tmem + (uint32_t)(buf * kGemmN);
        for (int kb = 0; kb < KB; ++kb, ++it) {
          const int s = it % kPersistStages;
          const uint32_t n = (uint32_t)(it / kPersistStages);
          mbar_wait(&a_full[s], n & 1u);
          umma::fence_after_sync();
          const uint32_t sa = umma::smem_addr(smA + (long)s * kPersistABytes);
          const uint64_t a_hi = umma::make_desc_k_sw128(sa);
          const uint64_t a_lo = umma::make_desc_k_sw128(sa + kGemmABytes);
          const uint64_t b_hi = umma::make_desc_k_sw128(sb + (uint32_t)(kb * 2 * kGemmBBytes));
          const uint64_t b_lo = umma::make_desc_k_sw128(sb + (uint32_t)(kb * 2 * kGemmBBytes + kGemmBBytes));
#pragma unroll
          for (int k = 0; k < umma::kBlockK / 8; ++k) {
            const uint64_t adv = (uint64_t)(k * 2);
            umma::mma_tf32(d_tmem, a_lo + adv, b_hi + adv, idesc, (kb > 0 || k > 0) ? 1u : 0u);
            umma::mma_tf32(d_tmem, a_hi + adv, b_lo + adv, idesc, 1u);
            umma::mma_tf32(d_tmem, a_hi + adv, b_hi + adv, idesc, 1u);
          }
          umma::mma_commit(&a_empty[s]);
        }
        umma::mma_commit(&acc_full[buf]);
      }
    }
  } else {
    // ---------------- epilogue: warp w may touch TMEM lanes 32 * (w % 4) ..
    const int q = warp & 3;
    const int c_begin = ((warp - 2) >> 2) * (kGemmN / kGemmEpiGroups), c_end = c_begin + kGemmN / kGemmEpiGroups;
    int i = 0;
    for (int mt = blockIdx.x; mt < n_tiles; mt += gridDim.x, ++i) {
      const int buf = i & 1;
      const uint32_t u = (uint32_t)(i >> 1);
      const int m = mt * kGemmM + q * 32 + lane;  // this thread's row of A
      const bool valid = m < L.n_valid;
      const float yv = valid ? L.y[m] : 0.f;
      const long kbR = m >> 5;
      unsigned char* rh = L.r_hi + kbR * (kGemmN >> 3) * umma::kAtomBytes;
      unsigned char* rl = L.r_lo + kbR * (kGemmN >> 3) * umma::kAtomBytes;
      float* llp = L.ll_part + kbR * kGemmN;
      mbar_wait(&acc_full[buf], u & 1u);
      umma::fence_after_sync();
      const uint32_t trow = tmem + (uint32_t)(buf * kGemmN) + ((uint32_t)(q * 32) << 16);
      const bool check = (mt + 1) * kGemmM > L.n_valid;  // CTA-uniform: padding rows in this tile
      const int swz = (lane >> 2) << 4, sub = (lane & 3) << 2;
#pragma unroll 1
      for (int c0 = c_begin; c0 < c_end; c0 += 32) {
        uint32_t v[32];
        umma::tmem_ld32(trow + c0, v);
        unsigned char *bh[8], *bl[8];
#pragma unroll
        for (int k = 0; k < 8; ++k) {
          const long o = (long)(c0 >> 3) * umma::kAtomBytes + k * 128 + (swz ^ (k << 4)) + sub;
          bh[k] = rh + o;
          bl[k] = rl + o;
        }
        umma::tmem_ld_wait();
        float part[32];
        if (check) resid_chunk<LIK, true>(v, yv, valid, L.col_scale + c0, bh, bl, part);
        else resid_chunk<LIK, false>(v, yv, valid, L.col_scale + c0, bh, bl, part);
        const float tot = warp_transpose_reduce32(part);
        llp[c0 + lane] = tot;
      }
      // this warp is done reading the accumulator: hand it back to the MMA issuer
      umma::fence_before_sync();
      __syncwarp();
      if (lane == 0) mbar_arrive(&acc_empty[buf]);
    }
  }
  umma::fence_before_sync();
  __syncthreads();
  if (warp == 1) {
    umma::fence_after_sync();
    umma::tmem_dealloc(tmem, 2 * kGemmN);
  }
}

#ifdef B200MC_TU_GEMM  // compiled in this family's translation unit only (launchers.h)
// ---------------------------------------------------------------- operand image builders
// Row-major src[R_src][K_src] (optionally read transposed) -> hi/lo images of Mtx[R][K]
// (R, K padded with zeros).  One thread per element; writes are 4 bytes into swizzled lines.
__global__ void image_build_kernel(const float* __restrict__ src, int R_src, int K_src, int transpose,
                                   int R, int K, unsigned char* __restrict__ hi,
                                   unsigned char* __restrict__ lo) {
  const long total = (long)R * K;
  for (long i = blockIdx.x * (long)blockDim.x + threadIdx.x; i < total; i += (long)gridDim.x * blockDim.x) {
    // enumerate in image order so writes are contiguous; derive (r, k) from the offset
    const long atom = i >> 8;  // 256 floats per atom
    const int within = (int)(i & 255);
    const int rr = within >> 5, chunk_sw = (within >> 2) & 7, e = within & 3;
    const int kk = ((chunk_sw ^ rr) << 2) + e;
    const long rg = atom % (R >> 3), kb = atom / (R >> 3);
    const int r = (int)(rg * 8 + rr);
    const int k = (int)(kb * 32 + kk);
    float x = 0.f;
    if (!transpose) {
      if (r < R_src && k < K_src) x = src[(long)r * K_src + k];
    } else {
      if (k < R_src && r < K_src) x = src[(long)k * K_src + r];  // Mtx[r][k] = src[k][r]
    }
    float h, l;
    umma::split_tf32(x, h, l);
    reinterpret_cast<float*>(hi)[i] = h;
    reinterpret_cast<float*>(lo)[i] = l;
  }
}

#endif  // B200MC_TU_GEMM

}  // namespace b200mc
