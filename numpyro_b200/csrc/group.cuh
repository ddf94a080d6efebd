// Execution groups that own one Markov chain: a single thread, a warp, or a CTA.
// A D-vector of chain state is distributed over the group's lanes
// (element d belongs to lane d % size) so every elementwise update is lane-local;
// only dot products cross lanes, through deterministic reductions whose result is
// bit-identical in every lane.
#pragma once
#include <cuda_runtime.h>

namespace b200mc {

struct ThreadGroup {  // one chain per thread (tiny D: eight schools)
  static constexpr bool kUnitStride = false;  // vectors are strided [D][chain]
  static constexpr bool kRegVec = false;
  static constexpr bool kBlockWide = false;
  __device__ __forceinline__ int lane() const { return 0; }
  __device__ __forceinline__ int size() const { return 1; }
  __device__ __forceinline__ float sum(float x) const { return x; }
  __device__ __forceinline__ void sum2(float& a, float& b) const {}
  __device__ __forceinline__ void sync() const {}
};

struct WarpGroup {  // one chain per warp (D up to a few hundred)
  static constexpr bool kUnitStride = true;
  static constexpr bool kRegVec = true;  // <= 4 elements per lane fit registers (D <= 128)
  static constexpr bool kBlockWide = false;
  __device__ __forceinline__ int lane() const { return threadIdx.x & 31; }
  __device__ __forceinline__ int size() const { return 32; }
  __device__ __forceinline__ float sum(float x) const {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) x += __shfl_xor_sync(0xffffffffu, x, o);
    return x;
  }
  __device__ __forceinline__ void sum2(float& a, float& b) const {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      a += __shfl_xor_sync(0xffffffffu, a, o);
      b += __shfl_xor_sync(0xffffffffu, b, o);
    }
  }
  __device__ __forceinline__ void sync() const { __syncwarp(); }
};

// One chain per CTA (large D: stochastic volatility).  `scratch` is 2*32 floats of
// shared memory private to the CTA.
struct BlockGroup {
  static constexpr bool kUnitStride = true;
  static constexpr bool kRegVec = false;
  static constexpr bool kBlockWide = true;  // whole warps available: warp-per-row sweeps
  float* scratch;
  __device__ __forceinline__ int lane() const { return threadIdx.x; }
  __device__ __forceinline__ int size() const { return blockDim.x; }
  __device__ __forceinline__ void sum2(float& a, float& b) const {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      a += __shfl_xor_sync(0xffffffffu, a, o);
      b += __shfl_xor_sync(0xffffffffu, b, o);
    }
    const int w = threadIdx.x >> 5, nw = (blockDim.x + 31) >> 5;
    __syncthreads();  // previous readers of scratch are done
    if ((threadIdx.x & 31) == 0) {
      scratch[w] = a;
      scratch[32 + w] = b;
    }
    __syncthreads();
    float ta = 0.f, tb = 0.f;
    for (int i = 0; i < nw; ++i) {  // fixed order: identical in every thread
      ta += scratch[i];
      tb += scratch[32 + i];
    }
    a = ta;
    b = tb;
  }
  __device__ __forceinline__ float sum(float x) const {
    float y = 0.f;
    sum2(x, y);
    return x;
  }
  __device__ __forceinline__ void sync() const { __syncthreads(); }
};

}  // namespace b200mc
