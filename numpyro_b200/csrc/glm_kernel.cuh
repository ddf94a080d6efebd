// Persistent streaming kernel for tall-data Bayesian logistic regression
// (examples/covtype.py:66-71): one cooperative launch runs whole NUTS/HMC trajectories.
//
// Data layout in HBM (built once by glm_retile_kernel when the handle is created): X and
// the labels are re-tiled into 32-row tiles, each one contiguous block of
//     [G column groups][32 rows][4 floats]  +  32 label bytes        (G = ceil(D / 4))
// so that ONE TMA bulk copy brings a whole tile and lane r reads its row as conflict-free
// 128-bit shared-memory loads.  Rows past N and columns past D are zero.
//
// Every CTA (one per SM) owns a contiguous range of tiles.  Per leapfrog ("pass"):
//   * the producer warp keeps a ring of shared-memory stages full: lane s owns stage s,
//     polls its "empty" mbarrier and issues cp.async.bulk (mbarrier complete_tx) for the
//     next tile of that stage.  It runs AHEAD across pass boundaries (tile addresses do
//     not depend on theta), so memory keeps streaming while the sampler decides the next
//     point.  Odd passes sweep the range backwards, so most of X (128 MB) is still in the
//     126 MB L2 when the next pass reads it.
//   * consumer warps own one row per lane: the row goes shared memory -> registers once,
//     and logit = x . theta (packed FFMA2), the Bernoulli log-likelihood and the gradient
//     contribution (sigmoid(l) - y) x are all computed from those registers (the
//     reference's XLA program reads X twice: forward dot + transposed dot).  Per-lane
//     gradient accumulators are folded across lanes once per pass (shuffle transpose-reduce).
//   * CTA partials go to global memory behind a grid-wide arrival counter.  EVERY CTA then
//     reduces all partials in the same fixed order and advances its own replica of the
//     sampler step machine (sampler.cuh) in warp 0 -- bit-identical in all CTAs -- so the
//     next theta is available locally without a publish/poll round trip.  Only CTA 0
//     writes chain state and collected draws.
//
// Replaces: potential_energy/log_density for this family (infer/util.py:302-327),
// value_and_grad (hmc_util.py:242-252) and the whole sample_kernel loop (hmc.py:459-530).
#pragma once
#include <cuda_runtime.h>

#include "sampler.cuh"

namespace b200mc {

// warps 0..9 consume tiles, warp 10 is the control (sampler) warp, warp 11 the TMA producer:
// 12 warps = 3 per SM sub-partition (168 registers per thread)
constexpr int kGlmConsumerWarps = 10;
constexpr int kGlmConsumerThreads = kGlmConsumerWarps * 32;
constexpr int kGlmControlWarp = kGlmConsumerWarps;
constexpr int kGlmProducerWarp = kGlmConsumerWarps + 1;
constexpr int kGlmThreads = kGlmConsumerThreads + 64;
constexpr int kGlmTileRows = 32;
constexpr int kGlmMaxD = 128;
constexpr int kGlmMaxStages = 32;      // one producer lane per stage
constexpr int kGlmProfSlots = 16;
constexpr int kGlmMaxRanks = 8;

__host__ __device__ inline int glm_groups(int D) { return (D + 3) / 4; }
__host__ __device__ inline int glm_tile_bytes(int D) { return glm_groups(D) * 512 + 32; }
__host__ __device__ inline int glm_partial_stride(int D) { return (D + 1 + 3) & ~3; }

struct GlmLaunch {
  SamplerParams P;
  CollectSpec col;
  ChainScalars* cs;
  ChainVectors cv;
  const unsigned char* Xt;  // re-tiled data, glm_tile_bytes(D) per tile, 16-byte aligned
  int N, D;
  int nt_total;        // ceil(N / 32)
  int n_stages;        // multiple of kGlmConsumerWarps, <= kGlmMaxStages
  int n_pinned;        // tiles per CTA that live in shared memory for the whole launch
  int reverse_odd;     // sweep odd passes backwards (L2 reuse)
  // grid-level exchange (global memory)
  float* partials;       // [2][nCTA][glm_partial_stride(D)], double-buffered by pass parity
  unsigned* arrive;      // monotonic arrival counter, zeroed by the host before the launch
  int* error;            // set to 1 on a spin time-out
  unsigned long long* prof;  // [kGlmProfSlots] cycle counters of CTA 0 (or null)
  float* dense_replicas;     // dense mass: private [nCTA-1][4][D][D] copies for CTAs > 0
  // row sharding over GPUs (one rank per GPU, rows split, sampler replicated on every rank):
  // per pass every rank writes its D+1 partial sums into every rank's exchange buffer over
  // NVLink (P2P stores of (value, sequence) pairs, no fence or flag); all ranks then add the same
  // nranks vectors in rank order, so the reduced gradient is bit-identical everywhere.
  int nranks, rank;
  float* xchg[kGlmMaxRanks];        // rank r's buffer [2][nranks][PS] of (value, seq) pairs
                                    // (peer-mapped), own included
  unsigned* xarrive[kGlmMaxRanks];  // rank r's counters: [1] exchanges completed before this
                                    // launch (only the local one is used)
  const uint32_t* init_keys;  // [C][2] or null
  int user_init;
  int chain;             // which chain of cs/cv this launch advances
  int C;
  // eval mode (tier-2 hook): evaluate U, grad at eval_z and stop.  eval_passes > 1 evaluates
  // the SAME point on that many consecutive passes (forward sweep, reversed sweep, ... with the
  // TMA ring wrapping in between) and writes every pass's result: eval_U[p], eval_grad[p][D]
  const float* eval_z;
  float* eval_U;
  float* eval_grad;
  int eval_passes;
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
    if (it > (1L << 24)) __trap();
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
__device__ __forceinline__ void red_release_add(unsigned* p, unsigned v) {
  asm volatile("red.release.gpu.global.add.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ unsigned ld_acquire_sys_u32(const unsigned* p) {
  unsigned v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void red_release_sys_add(unsigned* p, unsigned v) {
  asm volatile("red.release.sys.global.add.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ float ld_relaxed_sys_f32(const float* p) {
  float v;
  asm volatile("ld.relaxed.sys.global.f32 %0, [%1];" : "=f"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_relaxed_sys_f32(float* p, float v) {
  asm volatile("st.relaxed.sys.global.f32 [%0], %1;" ::"l"(p), "f"(v) : "memory");
}
__device__ __forceinline__ void st_relaxed_sys_v2(uint2* p, unsigned x, unsigned y) {
  asm volatile("st.relaxed.sys.global.v2.u32 [%0], {%1, %2};" ::"l"(p), "r"(x), "r"(y) : "memory");
}
__device__ __forceinline__ uint2 ld_relaxed_sys_v2(const uint2* p) {
  uint2 v;
  asm volatile("ld.relaxed.sys.global.v2.u32 {%0, %1}, [%2];" : "=r"(v.x), "=r"(v.y) : "l"(p) : "memory");
  return v;
}
// Clock read that the hardware cannot issue before `dep` is available: bar.sync is deferred-
// blocking, so a bare clock read placed after it may issue before the barrier completes.
__device__ __forceinline__ long long clk_after(int dep) {
  long long t;
  asm volatile("{\n\t.reg .b32 z;\n\tand.b32 z, %1, 0;\n\tmov.u64 %0, %%clock64;\n\t}"
               : "=l"(t) : "r"(dep) : "memory");
  return t;
}
__device__ __forceinline__ long long clk_ordered() { return clk_after(0); }
__device__ __forceinline__ void consumer_bar() {  // named barrier 1: consumer warps only
  asm volatile("bar.sync 1, %0;" ::"n"(kGlmConsumerThreads) : "memory");
}
__device__ __forceinline__ void handoff_bar() {  // named barrier 2: consumer warps + control warp
  asm volatile("bar.sync 2, %0;" ::"n"(kGlmConsumerThreads + 32) : "memory");
}

// Spin until *p >= target (monotonic counter); false on time-out (seconds).
__device__ __forceinline__ bool spin_until_ge(const unsigned* p, unsigned target) {
  for (long it = 0; it < (1L << 24); ++it) {
    if ((int)(ld_acquire_u32(p) - target) >= 0) return true;
    if (it > 4096) __nanosleep(100);
  }
  return false;
}

// 1/d for d in [1, 2]: MUFU.RCP + one Newton step (<= 1 ulp), no slow-path call -- a call in
// the tile loop would force the register-resident row and accumulators through the stack.
__device__ __forceinline__ float rcp_1to2(float d) {
  float r;
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(d));
  return fmaf(r, fmaf(-d, r, 1.0f), r);
}

#ifdef B200MC_TU_GLM  // compiled in this family's translation unit only (launchers.h)
// Re-tile X [N][D] row-major + labels [N] into the layout described at the top of the file.
__global__ void glm_retile_kernel(const float* __restrict__ X, const uint8_t* __restrict__ y,
                                  int N, int D, int nt_total, unsigned char* __restrict__ Xt) {
  const int G = glm_groups(D);
  const long tb = glm_tile_bytes(D);
  const long per_tile = (long)G * 32 + 2;  // float4 slots + two 16-byte label slots
  const long total = per_tile * nt_total;
  for (long i = blockIdx.x * (long)blockDim.x + threadIdx.x; i < total;
       i += (long)gridDim.x * blockDim.x) {
    const long t = i / per_tile;
    const int s = (int)(i - t * per_tile);
    unsigned char* tile = Xt + t * tb;
    if (s < G * 32) {
      const int g = s >> 5, r = s & 31;
      const long row = t * 32 + r;
      float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
      if (row < N) {
        const float* xr = X + row * D + 4 * g;
        v.x = xr[0];
        if (4 * g + 1 < D) v.y = xr[1];
        if (4 * g + 2 < D) v.z = xr[2];
        if (4 * g + 3 < D) v.w = xr[3];
      }
      reinterpret_cast<float4*>(tile)[s] = v;
    } else {
      const int h = s - G * 32;  // 0 or 1: rows 16h .. 16h+15
      unsigned char* dst = tile + (long)G * 512 + 16 * h;
      for (int k = 0; k < 16; ++k) {
        const long row = t * 32 + 16 * h + k;
        dst[k] = (row < N) ? y[row] : (unsigned char)0;
      }
    }
  }
}

#endif  // B200MC_TU_GLM

// ---------------------------------------------------------------- per-tile arithmetic
// One 32-row tile, lane r <-> row r.  xt: [G][32] float4 (shared), ys: 32 labels, th4: theta
// zero-padded to 4G floats (shared).  acc[k] accumulates, per lane, (sigmoid(l) - y) * x[k]
// over this lane's rows.  GT = G for the register-resident variant (G <= 16); GT = 0 selects
// the wide variant (G <= 32), which re-reads the row for the gradient.
template <int GT, typename Release>
__device__ __forceinline__ void glm_tile(const float4* __restrict__ xt,
                                         const uint8_t* __restrict__ ys, int rows,
                                         const float4* __restrict__ th4, int G,
                                         float (&acc)[GT ? 4 * GT : 128], float& ll,
                                         Release release) {
  constexpr int GM = GT ? GT : 32;
  const int lane = threadIdx.x & 31;
  const float4* xr = xt + lane;
  float4 x[GT ? GT : 1];
  float2 l01 = make_float2(0.f, 0.f), l23 = make_float2(0.f, 0.f);
  const float yv = (lane < rows) ? (float)ys[lane] : 0.f;
#pragma unroll
  for (int g = 0; g < GM; ++g) {
    if (GT || g < G) {
#if B200MC_GLM_DBG  // timing experiments only (profiles/r02_glm_stream_experiments.md): wrong numerics
      const float4 v = xr[((B200MC_GLM_DBG & 2) ? 0 : g) * 32];
      const float4 t = (B200MC_GLM_DBG & 1) ? make_float4(0.01f, 0.02f, 0.03f, 0.04f) : th4[g];
#else
      const float4 v = xr[g * 32];
      const float4 t = th4[g];
#endif
      if (GT) x[g] = v;
      l01 = __ffma2_rn(make_float2(v.x, v.y), make_float2(t.x, t.y), l01);
      l23 = __ffma2_rn(make_float2(v.z, v.w), make_float2(t.z, t.w), l23);
    }
  }
  const float l = (l01.x + l01.y) + (l23.x + l23.y);
  // Register-resident variant: the row now lives in x[] (l depends on every load, so they have
  // all returned), the stage can go back to the producer before the sigmoid and the gradient
  // FMAs -- its time out of flight is what bounds the ring's bytes in flight (Little's law)
  if (GT) release(l + yv);
  float gres = 0.f;
  if (lane < rows) {
    const float e = expf(-fabsf(l));
    const float inv = rcp_1to2(1.0f + e);
    const float sig = (l >= 0.f) ? inv : e * inv;
    // binary_cross_entropy_with_logits (distributions/util.py:295-298)
    ll += (fmaxf(l, 0.f) + log1pf(e)) - l * yv;
    gres = sig - yv;
  }
  const float2 g2 = make_float2(gres, gres);
#pragma unroll
  for (int g = 0; g < GM; ++g) {
    if (GT || g < G) {
      const float4 v = GT ? x[GT ? g : 0] : xr[g * 32];
      const float2 a01 = __ffma2_rn(g2, make_float2(v.x, v.y), make_float2(acc[4 * g], acc[4 * g + 1]));
      const float2 a23 =
          __ffma2_rn(g2, make_float2(v.z, v.w), make_float2(acc[4 * g + 2], acc[4 * g + 3]));
      acc[4 * g] = a01.x; acc[4 * g + 1] = a01.y;
      acc[4 * g + 2] = a23.x; acc[4 * g + 3] = a23.y;
    }
  }
  if (!GT) release(0.f);
}

// Recursive-halving transpose-reduce: on return a[0] of lane L holds the sum over all
// lanes of the original a[L] (31 shuffles for 32 columns; fixed order => deterministic).
__device__ __forceinline__ float warp_transpose_reduce32(float (&a)[32]) {
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

// ---------------------------------------------------------------- shared-memory layout
struct GlmSmem {
  int stage_bytes, off_full, off_empty, off_pin, off_theta, off_red, off_red2, off_work, off_flags, total;
};
constexpr int kGlmRedGroups = 16;  // grid reduction: CTAs are summed in strided groups

__host__ __device__ inline GlmSmem glm_smem_layout(int D, int n_stages, int n_pinned, int max_depth, int dense) {
  GlmSmem s;
  const int ps = glm_partial_stride(D);
  s.stage_bytes = glm_tile_bytes(D);  // multiple of 16
  int o = (n_stages + n_pinned) * s.stage_bytes;  // ring stages, then the pinned tiles
  s.off_full = o;  o += kGlmMaxStages * 8;
  s.off_empty = o; o += kGlmMaxStages * 8;
  s.off_pin = o;   o += kGlmMaxStages * 8;
  s.off_theta = o; o += kGlmMaxD * 4;  // zero-padded
  s.off_red = o;   o += kGlmConsumerWarps * ps * 4;
  s.off_red2 = o;  o += kGlmRedGroups * ps * 4;
  s.off_work = o;  o += num_slots(max_depth, dense) * D * 4;
  o = (o + 15) & ~15;
  s.off_flags = o; o += 16;
  s.total = o;
  return s;
}

__device__ __forceinline__ bool mbar_test_wait(uint64_t* bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "mbarrier.test_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t}"
      : "=r"(ok)
      : "r"(smem_u32(bar)), "r"(parity)
      : "memory");
  return ok != 0;
}

// One pass worth of tiles for one consumer warp; leaves its partial gradient and
// log-likelihood in red[warp][0..D].  The CTA's tile stream is numbered p = 0, 1, ... over
// all passes; position p is consumed by warp p % W from stage p % NS.  NS = m * W, so warp w
// only ever touches stages w, w + W, ..., w + (m-1) W, strictly in phase order (si, ph).
// `nt` counts the CTA's STREAMED tiles; its last `n_pin` tiles (local index nt + k) were
// loaded once into the slots behind the ring and are consumed from there in every pass
// (slot k by warp k % W), after the warp has put its ring stages back in flight.
template <int GT>
__device__ __forceinline__ void glm_consume_pass(const GlmLaunch& L, unsigned char* smem,
                                                 const GlmSmem& lay, int warp, long& p_next,
                                                 long p_begin, long p_end, int& si, uint32_t& ph,
                                                 int m, int tile0, int nt, bool rev, int n_pin) {
  constexpr int NA = GT ? 4 * GT : 128;
  const int lane = threadIdx.x & 31;
  const int G = glm_groups(L.D), PS = glm_partial_stride(L.D);
  uint64_t* full = reinterpret_cast<uint64_t*>(smem + lay.off_full);
  uint64_t* empty = reinterpret_cast<uint64_t*>(smem + lay.off_empty);
  uint64_t* pin = reinterpret_cast<uint64_t*>(smem + lay.off_pin);
  const float4* th4 = reinterpret_cast<const float4*>(smem + lay.off_theta);
  float acc[NA];
#pragma unroll
  for (int k = 0; k < NA; ++k) acc[k] = 0.f;
  float ll = 0.f;
  int pin_k = warp, done = 0;
  for (;;) {
    const bool ring_left = p_next < p_end;
    const bool pin_now = pin_k < n_pin && (done >= m || !ring_left);
    if (!pin_now && !ring_left) break;
    const unsigned char* sp;
    int tile, stage = -1;
    if (pin_now) {
      mbar_wait(&pin[pin_k], 0u);  // completes once per launch
      sp = smem + (long)(L.n_stages + pin_k) * lay.stage_bytes;
      tile = tile0 + nt + pin_k;
      pin_k += kGlmConsumerWarps;
    } else {
      stage = warp + si * kGlmConsumerWarps;
      mbar_wait(&full[stage], ph);
      sp = smem + (long)stage * lay.stage_bytes;
      const int j = (int)(p_next - p_begin);
      tile = tile0 + (rev ? (nt - 1 - j) : j);
    }
    const int rows = (tile == L.nt_total - 1) ? (L.N - 32 * tile) : 32;
    auto release = [&](float dep) {
      // `dep` carries the data dependence on this lane's loads of the tile
      asm volatile("" ::"f"(dep) : "memory");
      __syncwarp();
      if (stage >= 0 && lane == 0) mbar_arrive(&empty[stage]);
    };
    glm_tile<GT>(reinterpret_cast<const float4*>(sp), sp + G * 512, rows, th4, G, acc, ll, release);
    if (stage >= 0) {
      p_next += kGlmConsumerWarps;
      if (++si == m) {
        si = 0;
        ph ^= 1u;
      }
    }
    ++done;
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) ll += __shfl_xor_sync(0xffffffffu, ll, o);
  float* red = reinterpret_cast<float*>(smem + lay.off_red) + warp * PS;
#pragma unroll
  for (int c = 0; c < (NA + 31) / 32; ++c) {
    if (GT || 32 * c < L.D) {
      float a[32];
#pragma unroll
      for (int i = 0; i < 32; ++i) a[i] = (32 * c + i < NA) ? acc[(32 * c + i < NA) ? 32 * c + i : 0] : 0.f;
      const float v = warp_transpose_reduce32(a);
      if (32 * c + lane < L.D) red[32 * c + lane] = v;
    }
  }
  if (lane == 0) red[L.D] = ll;
}

#ifdef B200MC_TU_GLM  // compiled in this family's translation unit only (launchers.h)
__global__ void __launch_bounds__(kGlmThreads, 1) glm_kernel(const __grid_constant__ GlmLaunch L) {
  extern __shared__ __align__(128) unsigned char smem[];
  const int D = L.D, NS = L.n_stages, PS = glm_partial_stride(D);
  const GlmSmem lay = glm_smem_layout(D, NS, L.n_pinned, L.P.max_depth, L.P.dense_mass);
  uint64_t* full = reinterpret_cast<uint64_t*>(smem + lay.off_full);
  uint64_t* empty = reinterpret_cast<uint64_t*>(smem + lay.off_empty);
  float* th = reinterpret_cast<float*>(smem + lay.off_theta);
  float* red = reinterpret_cast<float*>(smem + lay.off_red);
  float* red2 = reinterpret_cast<float*>(smem + lay.off_red2);
  // [0] consumers done, [1] passes consumed, [2] go (1) / stop (0), [3] spin ok
  volatile int* flags = reinterpret_cast<volatile int*>(smem + lay.off_flags);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int cta = blockIdx.x, nCTA = gridDim.x;
  const int tile0 = (int)((long)cta * L.nt_total / nCTA);
  const int nt_all = (int)((long)(cta + 1) * L.nt_total / nCTA) - tile0;
  const int n_pin = L.n_pinned < nt_all ? L.n_pinned : nt_all;  // pinned: the last n_pin tiles
  const int nt = nt_all - n_pin;                                 // streamed through the ring
  uint64_t* pin = reinterpret_cast<uint64_t*>(smem + lay.off_pin);
  const long tb = lay.stage_bytes;

  if (threadIdx.x == 0) {
    for (int i = 0; i < NS; ++i) {
      mbar_init(&full[i], 1);
      mbar_init(&empty[i], 1);
    }
    for (int i = 0; i < L.n_pinned; ++i) mbar_init(&pin[i], 1);
    flags[0] = 0; flags[1] = 0; flags[2] = 0; flags[3] = 1;
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  for (int i = threadIdx.x; i < (kGlmConsumerWarps + kGlmRedGroups) * PS; i += blockDim.x)
    red[i] = 0.f;  // red and red2 are adjacent; their pad columns stay zero
  __syncthreads();

  if (warp == kGlmProducerWarp) {
    // ================= producer warp: lane s feeds stage s, runs ahead across passes ========
    const int s = lane;
    if (s < n_pin) {  // the pinned tiles: loaded once, stay for the whole launch
      mbar_arrive_expect_tx(&pin[s], (uint32_t)tb);
      tma_load_1d(smem + (long)(NS + s) * tb, L.Xt + (long)(tile0 + nt + s) * tb, (uint32_t)tb,
                  &pin[s]);
    }
    if (s < NS && nt > 0) {
      int pass_l = 0, j = s;
      while (j >= nt) { j -= nt; ++pass_l; }
      uint32_t n_issued = 0;
      unsigned char* sp = smem + (long)s * tb;
      // (polling, not a blocking mbarrier.try_wait: sleeping lanes woke late enough to cost 5 %
      // of the pass, profiles/r02_glm_stream_experiments.md)
      while (!flags[0]) {
        if (mbar_test_wait(&empty[s], (n_issued & 1u) ^ 1u)) {
          const int jj = (L.reverse_odd && (pass_l & 1)) ? (nt - 1 - j) : j;
          mbar_arrive_expect_tx(&full[s], (uint32_t)tb);
          tma_load_1d(sp, L.Xt + (long)(tile0 + jj) * tb, (uint32_t)tb, &full[s]);
          ++n_issued;
          j += NS;
          while (j >= nt) { j -= nt; ++pass_l; }
        }
      }
      // drain: only the last copy of this stage can still be unconsumed
      if (n_issued > 0) {
        const long p_last = (long)s + (long)(n_issued - 1) * NS;
        if (p_last >= (long)flags[1] * nt) mbar_wait(&full[s], (n_issued - 1) & 1u);
      }
    }
    if (s < n_pin) mbar_wait(&pin[s], 0u);
    return;
  }

  const bool eval_mode = (L.eval_z != nullptr);
  const int G = glm_groups(D);
  const unsigned xbase = (L.nranks > 1) ? __ldcg(&L.xarrive[L.rank][1]) : 0u;

  if (warp == kGlmControlWarp) {
    // ================= control warp: the sampler replica of this CTA =================
    // It owns the chain state (registers + shared memory), finishes the reduction, runs the
    // cross-rank exchange and the sampler step, and precomputes all randomness of the coming
    // leaves / doublings / transitions while the consumer warps stream the next pass.
    Sampler<WarpGroup> S;  // a replica in every CTA; only CTA 0 writes
    StepResult sres = FINISHED;
    float* work = reinterpret_cast<float*>(smem + lay.off_work);
    S.setup(&L.P, WarpGroup{}, work, D, 1, L.chain, L.C, &L.col);
    S.writer = (cta == 0);
#ifdef B200MC_SAMPLER_PROF
    for (int i = 0; i < 8; ++i) S.pc[i] = 0;
#endif
    if (eval_mode) {
      for (int d = lane; d < D; d += 32) S.v(LZ)[d] = L.eval_z[d];
      sres = NEED_GRAD;
    } else {
      if (L.P.dense_mass) {
        // replicas must not share read-modify-write state: CTA 0 owns the chain's matrices,
        // every other CTA works on a private copy taken at launch
        float* canon = L.cv.dense + (long)L.chain * 4 * D * D;
        if (cta == 0) {
          S.dn = canon;
        } else {
          S.dn = L.dense_replicas + (long)(cta - 1) * 4 * D * D;
          for (int i = lane; i < 4 * D * D; i += 32) S.dn[i] = canon[i];
          __syncwarp();
        }
      }
      S.load(L.cs, L.cv);
      if (L.init_keys)
        sres = S.begin_init(Key{L.init_keys[2 * L.chain], L.init_keys[2 * L.chain + 1]},
                            L.user_init != 0);
      else
        sres = S.begin_sampling();
    }
    __syncwarp();
    for (int d = lane; d < kGlmMaxD; d += 32) th[d] = (d < D) ? S.v(LZ)[d] : 0.f;
    if (lane == 0) flags[2] = (sres == NEED_GRAD) ? 1 : 0;
    handoff_bar();  // theta of pass 0 is ready
    if (sres == NEED_GRAD) S.prefetch_rng();
    int pass = 0;
#ifdef B200MC_SAMPLER_PROF
    unsigned long long c_pro = 0, c_fast = 0, c_other = 0;
#endif
    while (sres == NEED_GRAD) {
      handoff_bar();  // red[0..PS) holds this rank's grid-reduced partial sums
#ifdef B200MC_SAMPLER_PROF
      const long long tp0 = (red2[0] == 12345.678f) ? 0 : clock64();  // after the barrier released
#endif
      float* lz = S.v(LZ);
      float* lg = S.v(LG);
      // this rank's sums over its CTAs, columns lane, lane+32, ... (< PS <= 132)
      float sl[5];
#pragma unroll
      for (int k = 0; k < 5; ++k) {
        const int d = lane + 32 * k;
        sl[k] = (d < PS) ? red[d] : 0.f;
      }
      if (L.nranks > 1) {
        // Low-latency exchange: every 8-byte slot carries (value, sequence number), so one
        // one-way NVLink store per element is the whole message -- no fence, no flag.  Slots
        // are double-buffered by exchange parity; a rank can only be one exchange ahead.
        const unsigned e = xbase + (unsigned)pass;  // exchange index, monotonic over launches
        const unsigned seq = e + 1u;
        const long slot = (long)(e & 1u) * L.nranks * PS;
        if (cta == 0) {
          for (int r = 0; r < L.nranks; ++r) {
            uint2* dst = reinterpret_cast<uint2*>(L.xchg[r]) + slot + (long)L.rank * PS;
#pragma unroll
            for (int k = 0; k < 5; ++k)
              if (lane + 32 * k < PS) st_relaxed_sys_v2(dst + lane + 32 * k, __float_as_uint(sl[k]), seq);
          }
        }
        const uint2* mine = reinterpret_cast<const uint2*>(L.xchg[L.rank]) + slot;
        // poll ALL ranks' slots together (independent loads in flight), not rank after rank:
        // the wait is one round trip after the slowest rank, whatever the rank count
        bool ok = true;
#pragma unroll
        for (int k = 0; k < 5; ++k) {
          const int d = lane + 32 * k;
          float a = 0.f;
          if (d < PS) {
            uint2 v[kGlmMaxRanks];
            bool all = false;
            for (long it = 0; !all && it < (1L << 22); ++it) {
              if (it > 65536) __nanosleep(200);
#pragma unroll
              for (int r = 0; r < kGlmMaxRanks; ++r)
                if (r < L.nranks) v[r] = ld_relaxed_sys_v2(mine + (long)r * PS + d);
              all = true;
#pragma unroll
              for (int r = 0; r < kGlmMaxRanks; ++r)
                if (r < L.nranks) all = all && (v[r].y == seq);
            }
            ok = ok && all;
#pragma unroll
            for (int r = 0; r < kGlmMaxRanks; ++r)  // rank order: identical sums on every rank
              if (r < L.nranks) a += __uint_as_float(v[r].x);
          }
          sl[k] = a;
        }
        if (!__all_sync(0xffffffffu, ok)) {
          if (lane == 0) {
            atomicExch(L.error, 2);
            flags[3] = 0;
          }
          __syncwarp();
        }
      }
      float prior = 0.f, llik = 0.f;
#pragma unroll
      for (int k = 0; k < 5; ++k) {
        const int d = lane + 32 * k;
        if (d < D) {
          const float t = lz[d];
          lg[d] = t + sl[k];                            // grad U = theta + X^T (sigmoid - y)
          prior += 0.5f * t * t + 0.9189385332046727f;  // -log N(theta; 0, 1)
        }
        llik += (d == D) ? sl[k] : 0.f;
      }
      WarpGroup{}.sum2(prior, llik);
      const float U = prior + llik;
      if (eval_mode) {
        if (cta == 0) {
          for (int d = lane; d < D; d += 32) L.eval_grad[(long)pass * D + d] = lg[d];
          if (lane == 0) L.eval_U[pass] = U;
        }
        sres = (pass + 1 < L.eval_passes) ? NEED_GRAD : FINISHED;
      } else {
        // the critical part only: bookkeeping of a continuing leaf is deferred below
#ifdef B200MC_SAMPLER_PROF
        const long long tp1 = clock64();
        c_pro += (unsigned long long)(tp1 - tp0);
#endif
        sres = flags[3] ? S.advance_early(U) : FINISHED;
#ifdef B200MC_SAMPLER_PROF
        const long long tp2 = clock64();
        if (S.deferred == 1) c_fast += (unsigned long long)(tp2 - tp1);
        else if (S.deferred == 2) { S.pc[5] += (unsigned long long)(tp2 - tp1); S.pc[1] += 1; }
        else if (S.deferred == 3) { S.pc[6] += (unsigned long long)(tp2 - tp1); S.pc[7] += 1; }
        else c_other += (unsigned long long)(tp2 - tp1);
#endif
      }
      __syncwarp();
      for (int d = lane; d < D; d += 32) th[d] = S.v(LZ)[d];
      if (lane == 0) flags[2] = (sres == NEED_GRAD) ? 1 : 0;
      ++pass;
      handoff_bar();  // theta of the next pass is ready (or stop)
      // off the critical path, overlapped with the consumers' streaming of the next pass
      if (S.deferred) S.run_deferred();
      if (sres == NEED_GRAD) S.prefetch_rng();
    }
    if (lane == 0 && cta == 0) {
      if (L.nranks > 1) L.xarrive[L.rank][1] = xbase + (unsigned)pass;
#ifdef B200MC_SAMPLER_PROF
      if (L.prof) {
        for (int i = 0; i < 8; ++i) atomicAdd(&L.prof[8 + i], S.pc[i]);
        atomicAdd(&L.prof[5], c_pro);
        atomicAdd(&L.prof[6], c_fast);
        atomicAdd(&L.prof[7], c_other);
      }
#endif
    }
    if (!eval_mode) S.store(L.cs, L.cv);
    return;
  }

  // ================= consumer warps =================
  const int tid = threadIdx.x;  // 0 .. kGlmConsumerThreads-1
  const int m = NS / kGlmConsumerWarps;
  long p_next = warp;
  int si = 0;
  uint32_t ph = 0;
  int pass = 0;
  handoff_bar();  // theta of pass 0
  long long t_mark = clk_ordered();
  unsigned long long c_stream = 0, c_wait = 0, c_reduce = 0, c_sampler = 0;
  while (flags[2]) {
    // ---- stream this CTA's tiles
    const long p_begin = (long)pass * nt, p_end = p_begin + nt;
    const bool rev = L.reverse_odd && (pass & 1);
#define B200MC_GLM_CASE(GT) \
  case GT: glm_consume_pass<GT>(L, smem, lay, warp, p_next, p_begin, p_end, si, ph, m, tile0, nt, rev, n_pin); break;
    switch (G <= 16 ? G : 0) {
      B200MC_GLM_CASE(1) B200MC_GLM_CASE(2) B200MC_GLM_CASE(3) B200MC_GLM_CASE(4)
      B200MC_GLM_CASE(5) B200MC_GLM_CASE(6) B200MC_GLM_CASE(7) B200MC_GLM_CASE(8)
      B200MC_GLM_CASE(9) B200MC_GLM_CASE(10) B200MC_GLM_CASE(11) B200MC_GLM_CASE(12)
      B200MC_GLM_CASE(13) B200MC_GLM_CASE(14) B200MC_GLM_CASE(15) B200MC_GLM_CASE(16)
      default:
        glm_consume_pass<0>(L, smem, lay, warp, p_next, p_begin, p_end, si, ph, m, tile0, nt, rev, n_pin);
        break;
    }
#undef B200MC_GLM_CASE
    consumer_bar();
    // ---- CTA partial: fixed-order sum over warps
    float* my_partials = L.partials + ((long)(pass & 1) * nCTA + cta) * PS;
    if (tid < PS) {
      float a = 0.f;
#pragma unroll
      for (int w2 = 0; w2 < kGlmConsumerWarps; ++w2) a += red[w2 * PS + tid];
      __stcg(&my_partials[tid], a);
    }
    consumer_bar();
    if (tid == 0) {
      flags[1] = pass + 1;
      __threadfence();
      red_release_add(L.arrive, 1u);
      const long long t1 = clk_ordered();
      c_stream += (unsigned long long)(t1 - t_mark);
      const bool ok = spin_until_ge(L.arrive, (unsigned)(pass + 1) * (unsigned)nCTA);
      if (!ok) {
        atomicExch(L.error, 1);
        flags[3] = 0;
      }
      t_mark = clk_ordered();
      c_wait += (unsigned long long)(t_mark - t1);
    }
    consumer_bar();
    // ---- every CTA: grid reduction in a fixed order (128-bit L2 loads, all independent)
    const float4* all4 = reinterpret_cast<const float4*>(L.partials + (long)(pass & 1) * nCTA * PS);
    const int PS4 = PS >> 2;
    for (int item = tid; item < kGlmRedGroups * PS4; item += kGlmConsumerThreads) {
      const int c4 = item % PS4, grp = item / PS4;
      float4 a = make_float4(0.f, 0.f, 0.f, 0.f);
      for (int b0 = grp; b0 < nCTA; b0 += kGlmRedGroups * 10) {
        float4 v[10];
#pragma unroll
        for (int i = 0; i < 10; ++i) {
          const int b = b0 + kGlmRedGroups * i;
          v[i] = (b < nCTA) ? __ldcg(&all4[(long)b * PS4 + c4]) : make_float4(0.f, 0.f, 0.f, 0.f);
        }
#pragma unroll
        for (int i = 0; i < 10; ++i) {
          a.x += v[i].x; a.y += v[i].y; a.z += v[i].z; a.w += v[i].w;
        }
      }
      reinterpret_cast<float4*>(red2 + grp * PS)[c4] = a;
    }
    consumer_bar();
    if (tid < PS) {  // second stage, fixed order; red[0..PS) is free again and feeds the control warp
      float a = 0.f;
#pragma unroll
      for (int p2 = 0; p2 < kGlmRedGroups; ++p2) a += red2[p2 * PS + tid];
      red[tid] = a;
    }
    long long t2 = 0;
    if (tid == 0) {
      t2 = clk_ordered();
      c_reduce += (unsigned long long)(t2 - t_mark);
    }
    handoff_bar();  // red2 -> control warp
    handoff_bar();  // control warp -> next theta
    if (tid == 0) {
      t_mark = clk_after(flags[2]);  // a value the control warp wrote before the barrier
      c_sampler += (unsigned long long)(t_mark - t2);
    }
    ++pass;
  }
  // ---- shutdown: tell the producer
  if (tid == 0) {
    flags[1] = pass;
    __threadfence_block();
    flags[0] = 1;
    if (cta == 0 && L.prof) {
      atomicAdd(&L.prof[0], c_stream);
      atomicAdd(&L.prof[1], c_wait);
      atomicAdd(&L.prof[2], c_reduce);
      atomicAdd(&L.prof[3], c_sampler);
      atomicAdd(&L.prof[4], (unsigned long long)pass);
    }
  }
}

#endif  // B200MC_TU_GLM

}  // namespace b200mc
