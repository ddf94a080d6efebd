// Memory-roofline microbenchmark with the access pattern of glm_kernel (glm_kernel.cuh): one
// persistent CTA per SM streams its contiguous range of fixed-size tiles through a shared-memory
// ring with 1-D TMA bulk copies (cp.async.bulk + mbarrier complete_tx), producer lane per stage,
// stage ownership by consumer warp, odd passes swept backwards ("ping-pong").  The consumers only
// touch each tile (one 128-bit shared-memory load per lane) and hand the stage back, so the
// figure is the rate at which this pattern can be FED: from L2 when the buffer is L2-resident
// (the regime of the covtype-shape design matrix, 125 MB streamed per pass), from HBM when it
// is several times the 126 MB L2.  bench.py measures both in the run and reports glm_kernel
// against them (roofline.peak); no reference counterpart (diagnostic).
#pragma once
#include <cuda_runtime.h>

#include "glm_kernel.cuh"

namespace b200mc {

struct MemBenchLaunch {
  const unsigned char* buf;
  int nt_total, tile_bytes, n_stages, passes, reverse_odd;
  unsigned* sink;
};

#ifdef B200MC_TU_GLM  // compiled in this family's translation unit only (launchers.h)
__global__ void __launch_bounds__(kGlmThreads, 1) membench_kernel(const __grid_constant__ MemBenchLaunch L) {
  extern __shared__ __align__(128) unsigned char smem[];
  const int NS = L.n_stages;
  const long tb = L.tile_bytes;
  uint64_t* full = reinterpret_cast<uint64_t*>(smem + (long)NS * tb);
  uint64_t* empty = full + kGlmMaxStages;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int cta = blockIdx.x, nCTA = gridDim.x;
  const int tile0 = (int)((long)cta * L.nt_total / nCTA);
  const int nt = (int)((long)(cta + 1) * L.nt_total / nCTA) - tile0;
  if (threadIdx.x == 0) {
    for (int i = 0; i < NS; ++i) {
      mbar_init(&full[i], 1);
      mbar_init(&empty[i], 1);
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  const long total = (long)L.passes * nt;
  if (warp == kGlmProducerWarp) {
    const int s = lane;
    if (s < NS && nt > 0) {
      unsigned char* sp = smem + (long)s * tb;
      uint32_t k = 0;
      int pass = 0, j = s;
      while (j >= nt) { j -= nt; ++pass; }
      for (long p = s; p < total; p += NS, ++k) {
        mbar_wait(&empty[s], (k & 1u) ^ 1u);
        const int jj = (L.reverse_odd && (pass & 1)) ? (nt - 1 - j) : j;
        mbar_arrive_expect_tx(&full[s], (uint32_t)tb);
        tma_load_1d(sp, L.buf + (long)(tile0 + jj) * tb, (uint32_t)tb, &full[s]);
        j += NS;
        while (j >= nt) { j -= nt; ++pass; }
      }
      if (k > 0) mbar_wait(&full[s], (k - 1) & 1u);  // the last copy has landed before smem is released
    }
    return;
  }
  if (warp < kGlmConsumerWarps) {
    unsigned acc = 0;
    // position p is consumed by warp p % W from stage p % NS (NS is a multiple of W)
    for (long p = warp; p < total; p += kGlmConsumerWarps) {
      const int stage = (int)(p % NS);
      mbar_wait(&full[stage], (uint32_t)((p / NS) & 1));
      const uint4 v = reinterpret_cast<const uint4*>(smem + (long)stage * tb)[lane];
      acc ^= v.x ^ v.y ^ v.z ^ v.w;
      __syncwarp();
      if (lane == 0) mbar_arrive(&empty[stage]);
    }
    if (acc == 0x9e3779b9u) atomicAdd(L.sink, 1u);  // keeps the loads alive
  }
}

#endif  // B200MC_TU_GLM

}  // namespace b200mc
