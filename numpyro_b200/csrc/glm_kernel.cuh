// Persistent streaming kernel for tall-data Bayesian logistic regression
// (examples/covtype.py:66-71): one launch runs whole NUTS/HMC trajectories.
//
// Every CTA owns a contiguous slice of the rows of X.  Per leapfrog:
//   * a producer warp keeps a ring of shared-memory stages full with TMA bulk copies
//     (cp.async.bulk + mbarrier complete_tx) of 32-row tiles of X (+ their labels);
//     it runs AHEAD across leapfrog boundaries (tile addresses do not depend on
//     theta), so HBM keeps streaming while the sampler decides the next point;
//   * consumer warps compute, in ONE pass over the tile (the reference's XLA program
//     reads X twice: forward dot + transposed dot), logits = X theta, the Bernoulli
//     log-likelihood and the partial gradient X^T (sigmoid(l) - y);
//   * CTA partials go to global memory; a grid-wide arrive counter hands them to the
//     leader CTA, which reduces them in a fixed order, runs the sampler step machine
//     (sampler.cuh) on one warp with its state in shared memory, and publishes the
//     next theta behind a generation flag.
//
// Replaces: potential_energy/log_density for this family (infer/util.py:302-327),
// value_and_grad (hmc_util.py:242-252) and the whole sample_kernel loop (hmc.py:459-530).
#pragma once
#include <cuda_runtime.h>

#include "sampler.cuh"

namespace b200mc {

constexpr int kGlmConsumerWarps = 14;
constexpr int kGlmThreads = (kGlmConsumerWarps + 1) * 32;
constexpr int kGlmTileRows = 32;
constexpr int kGlmMaxD = 128;

struct GlmLaunch {
  SamplerParams P;
  CollectSpec col;
  ChainScalars* cs;
  ChainVectors cv;
  const float* X;      // [N][D] row-major, 16-byte aligned
  const uint8_t* y;    // [N] 0/1, 16-byte aligned
  int N, D;
  int rows_per_cta;    // multiple of 32
  int n_stages;
  // grid-level exchange (global memory)
  float* partials;       // [nCTA][DP1]
  unsigned* arrive;      // monotonic arrival counter
  unsigned* pub_gen;     // generation of theta_pub / ctrl
  float* theta_pub;      // [D]
  int* ctrl;             // 1 continue, 0 stop
  int* error;            // set to 1 on a spin time-out
  const uint32_t* init_keys;  // [C][2] or null
  int user_init;
  int chain;             // which chain of cs/cv this launch advances
  int C;
  // eval mode (tier-2 hook): evaluate U, grad at eval_z and stop
  const float* eval_z;
  float* eval_U;
  float* eval_grad;
};

// ---------------------------------------------------------------- PTX helpers
__device__ __forceinline__ uint32_t smem_u32(const void* p) {
  return (uint32_t)__cvta_generic_to_shared(p);
}
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)),
               "r"(bytes)
               : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint64_t* bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t}"
      : "=r"(ok)
      : "r"(smem_u32(bar)), "r"(parity)
      : "memory");
  return ok != 0;
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  // bounded: a copy that never lands is a bug; fail loudly (launch error) instead of hanging
  for (long it = 0; !mbar_try_wait(bar, parity); ++it)
    if (it > (1L << 28)) __trap();
}
// TMA bulk copy global -> shared, completion counted in bytes on an mbarrier
__device__ __forceinline__ void tma_load_1d(void* dst, const void* src, uint32_t bytes,
                                            uint64_t* bar) {
  asm volatile(
      "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::
          "r"(smem_u32(dst)),
      "l"(src), "r"(bytes), "r"(smem_u32(bar))
      : "memory");
}
__device__ __forceinline__ unsigned ld_acquire_u32(const unsigned* p) {
  unsigned v;
  asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_release_u32(unsigned* p, unsigned v) {
  asm volatile("st.release.gpu.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ void red_release_add(unsigned* p, unsigned v) {
  asm volatile("red.release.gpu.global.add.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ void consumer_bar() {  // named barrier 1: consumer warps only
  asm volatile("bar.sync 1, %0;" ::"n"(kGlmConsumerWarps * 32) : "memory");
}

// Spin until *p >= target (monotonic counters); false on time-out (~ seconds).
__device__ __forceinline__ bool spin_until_ge(const unsigned* p, unsigned target) {
  for (long it = 0; it < (1L << 27); ++it) {
    if ((int)(ld_acquire_u32(p) - target) >= 0) return true;
    if (it > 64) __nanosleep(20);
  }
  return false;
}

// One 32-row tile: logits, log-likelihood terms, gradient contribution.
// xs: [rows][D] floats (shared), ys: labels, th: theta (shared, 16-byte aligned).
// Lane r owns row r for the logits; lane c owns columns c, c+32, ... for the gradient.
template <int K>
__device__ __forceinline__ void glm_tile(const float* __restrict__ xs,
                                         const uint8_t* __restrict__ ys, int rows,
                                         const float* __restrict__ th, int D, float (&acc)[K],
                                         float& ll) {
  const int lane = threadIdx.x & 31;
  const float* xr = xs + lane * D;
  float l0 = 0.f, l1 = 0.f, l2 = 0.f, l3 = 0.f;
  int d = 0;
  for (; d + 3 < D; d += 4) {
    const float4 t4 = *reinterpret_cast<const float4*>(th + d);
    l0 = fmaf(xr[d], t4.x, l0);
    l1 = fmaf(xr[d + 1], t4.y, l1);
    l2 = fmaf(xr[d + 2], t4.z, l2);
    l3 = fmaf(xr[d + 3], t4.w, l3);
  }
  for (; d < D; ++d) l0 = fmaf(xr[d], th[d], l0);
  const float l = (l0 + l1) + (l2 + l3);
  float gres = 0.f;
  if (lane < rows) {
    const float yv = (float)ys[lane];
    const float e = expf(-fabsf(l));
    const float inv = 1.0f / (1.0f + e);
    const float sig = (l >= 0.f) ? inv : e * inv;
    // binary_cross_entropy_with_logits (distributions/util.py:295-298)
    ll += (fmaxf(l, 0.f) + log1pf(e)) - l * yv;
    gres = sig - yv;
  }
  for (int r = 0; r < rows; ++r) {
    const float gr = __shfl_sync(0xffffffffu, gres, r);
    const float* row = xs + r * D;
#pragma unroll
    for (int k = 0; k < K; ++k) {
      const int c = lane + 32 * k;
      if (c < D) acc[k] = fmaf(gr, row[c], acc[k]);
    }
  }
}

// ---------------------------------------------------------------- shared-memory layout
struct GlmSmem {
  int stage_bytes, off_tail, off_full, off_empty, off_theta, off_red, off_red2, off_work,
      off_flags, total;
};
constexpr int kGlmParts = kGlmConsumerWarps * 32 / 64;  // leader reduction: rows split in parts

__host__ __device__ inline GlmSmem glm_smem_layout(int D, int n_stages, int max_depth) {
  GlmSmem s;
  const int dp1 = D + 1;
  s.stage_bytes = kGlmTileRows * D * 4 + kGlmTileRows;  // X tile then labels; multiple of 16
  int o = n_stages * s.stage_bytes;
  s.off_tail = o;  o += s.stage_bytes;
  s.off_full = o;  o += n_stages * 8;
  s.off_empty = o; o += n_stages * 8;
  o = (o + 15) & ~15;
  s.off_theta = o; o += ((D + 3) & ~3) * 4;
  s.off_red = o;   o += kGlmConsumerWarps * dp1 * 4;
  s.off_red2 = o;  o += kGlmParts * dp1 * 4;
  o = (o + 15) & ~15;
  s.off_work = o;  o += num_slots(max_depth) * D * 4;
  s.off_flags = o; o += 16;
  s.total = o;
  return s;
}

// One leapfrog's worth of tiles for one consumer warp; leaves its partial gradient and
// log-likelihood in red[warp][0..D].
template <int K>
__device__ __forceinline__ void glm_consume_pass(const GlmLaunch& L, unsigned char* smem,
                                                 const GlmSmem& lay, int warp, long& q_next,
                                                 long q_end, int row0, int ntiles, int tail_rows) {
  const int lane = threadIdx.x & 31;
  const int D = L.D, NS = L.n_stages;
  uint64_t* full = reinterpret_cast<uint64_t*>(smem + lay.off_full);
  uint64_t* empty = reinterpret_cast<uint64_t*>(smem + lay.off_empty);
  const float* th = reinterpret_cast<const float*>(smem + lay.off_theta);
  float acc[K];
#pragma unroll
  for (int k = 0; k < K; ++k) acc[k] = 0.f;
  float ll = 0.f;
  for (; q_next < q_end; q_next += kGlmConsumerWarps) {
    const int stage = (int)(q_next % NS);
    const uint32_t parity = (uint32_t)((q_next / NS) & 1);
    mbar_wait(&full[stage], parity);
    const unsigned char* sp = smem + (long)stage * lay.stage_bytes;
    glm_tile<K>(reinterpret_cast<const float*>(sp), sp + kGlmTileRows * D * 4, kGlmTileRows, th, D,
                acc, ll);
    __syncwarp();
    if (lane == 0) mbar_arrive(&empty[stage]);
  }
  if (tail_rows > 0 && warp == kGlmConsumerWarps - 1) {
    // the last (< 32)-row tile of the data set: plain loads into the tail buffer
    unsigned char* tp = smem + lay.off_tail;
    float* tx = reinterpret_cast<float*>(tp);
    unsigned char* ty = tp + kGlmTileRows * D * 4;
    const long r0 = (long)row0 + (long)ntiles * kGlmTileRows;
    for (int i = lane; i < tail_rows * D; i += 32) tx[i] = L.X[r0 * D + i];
    if (lane < tail_rows) ty[lane] = L.y[r0 + lane];
    __syncwarp();
    glm_tile<K>(tx, ty, tail_rows, th, D, acc, ll);
    __syncwarp();
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) ll += __shfl_xor_sync(0xffffffffu, ll, o);
  float* red = reinterpret_cast<float*>(smem + lay.off_red) + warp * (D + 1);
#pragma unroll
  for (int k = 0; k < K; ++k) {
    const int c = lane + 32 * k;
    if (c < D) red[c] = acc[k];
  }
  if (lane == 0) red[D] = ll;
}

__global__ void __launch_bounds__(kGlmThreads, 1) glm_kernel(const __grid_constant__ GlmLaunch L) {
  extern __shared__ __align__(128) unsigned char smem[];
  const int D = L.D, NS = L.n_stages, DP1 = D + 1;
  const GlmSmem lay = glm_smem_layout(D, NS, L.P.max_depth);
  uint64_t* full = reinterpret_cast<uint64_t*>(smem + lay.off_full);
  uint64_t* empty = reinterpret_cast<uint64_t*>(smem + lay.off_empty);
  float* th = reinterpret_cast<float*>(smem + lay.off_theta);
  float* red = reinterpret_cast<float*>(smem + lay.off_red);
  float* red2 = reinterpret_cast<float*>(smem + lay.off_red2);
  volatile int* flags = reinterpret_cast<volatile int*>(smem + lay.off_flags);  // [0] done, [1] passes

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int cta = blockIdx.x, nCTA = gridDim.x;
  const long row0l = (long)cta * L.rows_per_cta;
  const int row0 = (int)(row0l < L.N ? row0l : L.N);
  const int row1 = (int)((row0l + L.rows_per_cta) < L.N ? (row0l + L.rows_per_cta) : L.N);
  const int ntiles = (row1 - row0) / kGlmTileRows;           // full tiles: TMA
  const int tail_rows = (row1 - row0) - ntiles * kGlmTileRows;  // < 32, last CTA only

  if (threadIdx.x == 0) {
    for (int i = 0; i < NS; ++i) {
      mbar_init(&full[i], 1);
      mbar_init(&empty[i], 1);
    }
    flags[0] = 0;
    flags[1] = 0;
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  const unsigned gen0 = 0u, arr0 = 0u;  // the host zeroes pub_gen / arrive before every launch

  if (warp == kGlmConsumerWarps) {
    // ================= producer warp: TMA ring, runs ahead across leapfrogs =================
    if (lane == 0 && ntiles > 0) {
      const uint32_t xbytes = kGlmTileRows * D * 4, ybytes = kGlmTileRows;
      long q = 0;
      bool stop = false;
      while (!stop) {
        for (int j = 0; j < ntiles && !stop; ++j, ++q) {
          const int stage = (int)(q % NS);
          const uint32_t parity = (uint32_t)(((q / NS) & 1) ^ 1);
          while (!mbar_try_wait(&empty[stage], parity)) {
            if (flags[0]) { stop = true; break; }
          }
          if (stop || flags[0]) { stop = true; break; }
          unsigned char* sp = smem + (long)stage * lay.stage_bytes;
          mbar_arrive_expect_tx(&full[stage], xbytes + ybytes);
          tma_load_1d(sp, L.X + ((long)row0 + (long)j * kGlmTileRows) * D, xbytes, &full[stage]);
          tma_load_1d(sp + xbytes, L.y + row0 + (long)j * kGlmTileRows, ybytes, &full[stage]);
        }
      }
      // drain: wait for every copy that was issued but never consumed
      const long consumed = (long)flags[1] * ntiles;
      for (long t = consumed; t < q; ++t)
        mbar_wait(&full[(int)(t % NS)], (uint32_t)((t / NS) & 1));
    }
    return;
  }

  // ================= consumer warps =================
  const int tid = threadIdx.x;  // 0 .. kGlmConsumerWarps*32-1
  const bool leader = (cta == 0);
  Sampler<WarpGroup> S;
  StepResult sres = FINISHED;
  const bool eval_mode = (L.eval_z != nullptr);
  if (leader && warp == 0) {
    float* work = reinterpret_cast<float*>(smem + lay.off_work);
    S.setup(&L.P, WarpGroup{}, work, D, 1, L.chain, L.C, &L.col);
    if (eval_mode) {
      for (int d = lane; d < D; d += 32) S.v(LZ)[d] = L.eval_z[d];
      sres = NEED_GRAD;
    } else {
      S.load(L.cs, L.cv);
      if (L.init_keys)
        sres = S.begin_init(Key{L.init_keys[2 * L.chain], L.init_keys[2 * L.chain + 1]},
                            L.user_init != 0);
      else
        sres = S.begin_sampling();
    }
    for (int d = lane; d < D; d += 32) __stcg(&L.theta_pub[d], S.v(LZ)[d]);
    if (lane == 0) *L.ctrl = (sres == NEED_GRAD) ? 1 : 0;
    __syncwarp();
    __threadfence();
    if (lane == 0) st_release_u32(L.pub_gen, gen0 + 1u);
  }

  const int K = (D + 31) / 32;
  long q_next = warp;
  int pass = 0;
  bool ok = true;
  while (true) {
    // ---- wait for theta of this pass
    if (tid == 0) {
      ok = spin_until_ge(L.pub_gen, gen0 + 1u + (unsigned)pass);
      if (!ok) atomicExch(L.error, 1);
      flags[2] = ok ? __ldcg(L.ctrl) : 0;
    }
    consumer_bar();
    const int go = flags[2];
    if (!go) break;
    if (tid < D) th[tid] = __ldcg(&L.theta_pub[tid]);
    consumer_bar();

    // ---- stream this CTA's rows
    const long q_end = (long)(pass + 1) * ntiles;
    switch (K) {
      case 1: glm_consume_pass<1>(L, smem, lay, warp, q_next, q_end, row0, ntiles, tail_rows); break;
      case 2: glm_consume_pass<2>(L, smem, lay, warp, q_next, q_end, row0, ntiles, tail_rows); break;
      case 3: glm_consume_pass<3>(L, smem, lay, warp, q_next, q_end, row0, ntiles, tail_rows); break;
      default: glm_consume_pass<4>(L, smem, lay, warp, q_next, q_end, row0, ntiles, tail_rows); break;
    }
    consumer_bar();
    // ---- CTA partial: fixed-order sum over warps
    if (tid < DP1) {
      float a = 0.f;
#pragma unroll
      for (int w2 = 0; w2 < kGlmConsumerWarps; ++w2) a += red[w2 * DP1 + tid];
      __stcg(&L.partials[(long)cta * DP1 + tid], a);
    }
    consumer_bar();
    if (tid == 0) {
      flags[1] = pass + 1;
      __threadfence();
      red_release_add(L.arrive, 1u);
    }

    // ---- leader: grid reduction in a fixed order + one sampler step
    if (leader) {
      if (tid == 0) {
        ok = spin_until_ge(L.arrive, arr0 + (unsigned)(pass + 1) * (unsigned)nCTA);
        if (!ok) atomicExch(L.error, 1);
      }
      consumer_bar();
      for (int item = tid; item < kGlmParts * DP1; item += kGlmConsumerWarps * 32) {
        const int c = item % DP1, part = item / DP1;
        float a = 0.f;
        for (int b = part; b < nCTA; b += kGlmParts) a += __ldcg(&L.partials[(long)b * DP1 + c]);
        red2[part * DP1 + c] = a;
      }
      consumer_bar();
      if (warp == 0) {
        float* lz = S.v(LZ);
        float* lg = S.v(LG);
        float prior = 0.f;
        for (int d = lane; d < D; d += 32) {
          float a = 0.f;
#pragma unroll
          for (int p2 = 0; p2 < kGlmParts; ++p2) a += red2[p2 * DP1 + d];
          const float t = lz[d];
          lg[d] = t + a;                           // grad U = theta + X^T (sigmoid - y)
          prior += 0.5f * t * t + 0.9189385332046727f;  // -log N(theta; 0, 1)
        }
        prior = WarpGroup{}.sum(prior);
        float llik = 0.f;
#pragma unroll
        for (int p2 = 0; p2 < kGlmParts; ++p2) llik += red2[p2 * DP1 + D];
        const float U = prior + llik;
        if (eval_mode) {
          for (int d = lane; d < D; d += 32) L.eval_grad[d] = lg[d];
          if (lane == 0) *L.eval_U = U;
          sres = FINISHED;
        } else {
          sres = ok ? S.advance(U) : FINISHED;
        }
        for (int d = lane; d < D; d += 32) __stcg(&L.theta_pub[d], S.v(LZ)[d]);
        if (lane == 0) *L.ctrl = (sres == NEED_GRAD) ? 1 : 0;
        __syncwarp();
        __threadfence();
        if (lane == 0) st_release_u32(L.pub_gen, gen0 + 2u + (unsigned)pass);
        if (sres == NEED_GRAD) S.prefetch_rng();
      }
    }
    ++pass;
  }
  // ---- shutdown: tell the producer, persist the chain
  if (tid == 0) {
    flags[1] = pass;
    __threadfence_block();
    flags[0] = 1;
  }
  if (leader && warp == 0 && !eval_mode) S.store(L.cs, L.cv);
}

}  // namespace b200mc
