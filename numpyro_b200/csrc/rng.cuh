// Threefry-2x32 and the jax.random draws built on it, as device functions.
//
// Replaces (third-party, not under /root/reference) jax/_src/prng.py
// threefry2x32 / threefry_split / threefry_random_bits and jax/_src/random.py
// uniform / normal / bernoulli, which the reference calls at
// numpyro/infer/hmc.py:102,335,472 and numpyro/infer/hmc_util.py:355,565,656,
// 804,920,1005,1161-1162.  Both counter layouts are implemented:
//   legacy        (jax_threefry_partitionable=False, JAX < 0.5 default)
//   partitionable (JAX >= 0.5 default)
// Everything here is integer-exact; the float transforms are written so that
// every operation is a single IEEE fp32 op (compile with -fmad=false).
#pragma once
#include <cstdint>

namespace b200mc {

struct Key {
  uint32_t k0, k1;
};

__host__ __device__ __forceinline__ uint32_t rotl32(uint32_t x, int r) {
  return (x << r) | (x >> (32 - r));
}

// Threefry-2x32, 20 rounds (Random123 threefry2x32_20).
__host__ __device__ __forceinline__ void threefry2x32(uint32_t k0, uint32_t k1, uint32_t& x0,
                                                      uint32_t& x1) {
  const uint32_t k2 = k0 ^ k1 ^ 0x1BD11BDAu;
  x0 += k0;
  x1 += k1;
#define B200MC_TF_ROUND(r) \
  x0 += x1;                \
  x1 = rotl32(x1, r);      \
  x1 ^= x0;
  B200MC_TF_ROUND(13) B200MC_TF_ROUND(15) B200MC_TF_ROUND(26) B200MC_TF_ROUND(6)
  x0 += k1; x1 += k2 + 1u;
  B200MC_TF_ROUND(17) B200MC_TF_ROUND(29) B200MC_TF_ROUND(16) B200MC_TF_ROUND(24)
  x0 += k2; x1 += k0 + 2u;
  B200MC_TF_ROUND(13) B200MC_TF_ROUND(15) B200MC_TF_ROUND(26) B200MC_TF_ROUND(6)
  x0 += k0; x1 += k1 + 3u;
  B200MC_TF_ROUND(17) B200MC_TF_ROUND(29) B200MC_TF_ROUND(16) B200MC_TF_ROUND(24)
  x0 += k1; x1 += k2 + 4u;
  B200MC_TF_ROUND(13) B200MC_TF_ROUND(15) B200MC_TF_ROUND(26) B200MC_TF_ROUND(6)
  x0 += k2; x1 += k0 + 5u;
#undef B200MC_TF_ROUND
}

// Word `w` (0..2n-1) of the flat legacy output of threefry_2x32(key, iota(2n)):
// counts are split in halves (x0 = first half, x1 = second half) and the outputs
// concatenated [o0..., o1...].
__host__ __device__ __forceinline__ uint32_t legacy_word(Key key, uint32_t w, uint32_t half) {
  const uint32_t j = (w < half) ? w : w - half;
  uint32_t a = j, b = j + half;
  threefry2x32(key.k0, key.k1, a, b);
  return (w < half) ? a : b;
}

// jax.random.split(key, num)[i]
__host__ __device__ __forceinline__ Key split_at(Key key, uint32_t num, uint32_t i,
                                                 bool partitionable) {
  Key out;
  if (partitionable) {
    uint32_t a = 0u, b = i;
    threefry2x32(key.k0, key.k1, a, b);
    out.k0 = a;
    out.k1 = b;
  } else {
    out.k0 = legacy_word(key, 2u * i, num);
    out.k1 = legacy_word(key, 2u * i + 1u, num);
  }
  return out;
}

// split(key) -> (a, b): the common 2-way split done with two blocks.
__host__ __device__ __forceinline__ void split2(Key key, bool partitionable, Key& a, Key& b) {
  if (partitionable) {
    uint32_t x0 = 0u, x1 = 0u, y0 = 0u, y1 = 1u;
    threefry2x32(key.k0, key.k1, x0, x1);
    threefry2x32(key.k0, key.k1, y0, y1);
    a.k0 = x0; a.k1 = x1; b.k0 = y0; b.k1 = y1;
  } else {
    // blocks (0,2) and (1,3); flat = [o0(0), o0(1), o1(0), o1(1)]
    uint32_t x0 = 0u, x1 = 2u, y0 = 1u, y1 = 3u;
    threefry2x32(key.k0, key.k1, x0, x1);
    threefry2x32(key.k0, key.k1, y0, y1);
    a.k0 = x0; a.k1 = y0; b.k0 = x1; b.k1 = y1;
  }
}

// split(key, 3) -> (a, b, c): three blocks.
__host__ __device__ __forceinline__ void split3(Key key, bool partitionable, Key& a, Key& b,
                                                Key& c) {
  if (partitionable) {
    a = split_at(key, 3, 0, true);
    b = split_at(key, 3, 1, true);
    c = split_at(key, 3, 2, true);
  } else {
    // blocks (0,3),(1,4),(2,5); flat = [a0,b0,c0,a1,b1,c1]
    uint32_t a0 = 0u, a1 = 3u, b0 = 1u, b1 = 4u, c0 = 2u, c1 = 5u;
    threefry2x32(key.k0, key.k1, a0, a1);
    threefry2x32(key.k0, key.k1, b0, b1);
    threefry2x32(key.k0, key.k1, c0, c1);
    a.k0 = a0; a.k1 = b0; b.k0 = c0; b.k1 = a1; c.k0 = b1; c.k1 = c1;
  }
}

// jax.random.bits(key, (n,), uint32)[i]
__host__ __device__ __forceinline__ uint32_t random_bits_at(Key key, uint32_t n, uint32_t i,
                                                            bool partitionable) {
  if (partitionable) {
    uint32_t a = 0u, b = i;
    threefry2x32(key.k0, key.k1, a, b);
    return a ^ b;
  }
  const uint32_t half = (n + 1u) >> 1;  // odd sizes are zero-padded to even
  const uint32_t j = (i < half) ? i : i - half;
  uint32_t a = j, b = j + half;
  if (b >= n) b = 0u;  // the pad element
  threefry2x32(key.k0, key.k1, a, b);
  return (i < half) ? a : b;
}

__host__ __device__ __forceinline__ float bits_to_unit(uint32_t bits) {
#ifdef __CUDA_ARCH__
  return __uint_as_float((bits >> 9) | 0x3F800000u) - 1.0f;
#else
  union { uint32_t u; float f; } v;
  v.u = (bits >> 9) | 0x3F800000u;
  return v.f - 1.0f;
#endif
}

// jax.random.uniform(key, (n,), minval, maxval)[i]
__host__ __device__ __forceinline__ float uniform_at(Key key, uint32_t n, uint32_t i, float lo,
                                                     float hi, bool partitionable) {
  const float u = bits_to_unit(random_bits_at(key, n, i, partitionable));
  const float scale = hi - lo;
  const float v = u * scale;
  const float w = v + lo;
  return fmaxf(lo, w);
}

__host__ __device__ __forceinline__ float uniform01(Key key, bool partitionable) {
  return fmaxf(0.0f, bits_to_unit(random_bits_at(key, 1u, 0u, partitionable)));
}

// XLA's float32 erf_inv (Giles' single-precision polynomial).
__host__ __device__ __forceinline__ float erf_inv_f32(float x) {
  const float xx = x * x;
  float w = -log1pf(-xx);
  float p;
  if (w < 5.0f) {
    w = w - 2.5f;
    p = 2.81022636e-08f;
    p = 3.43273939e-07f + p * w;
    p = -3.5233877e-06f + p * w;
    p = -4.39150654e-06f + p * w;
    p = 0.00021858087f + p * w;
    p = -0.00125372503f + p * w;
    p = -0.00417768164f + p * w;
    p = 0.246640727f + p * w;
    p = 1.50140941f + p * w;
  } else {
    w = sqrtf(w) - 3.0f;
    p = -0.000200214257f;
    p = 0.000100950558f + p * w;
    p = 0.00134934322f + p * w;
    p = -0.00367342844f + p * w;
    p = 0.00573950773f + p * w;
    p = -0.0076224613f + p * w;
    p = 0.00943887047f + p * w;
    p = 1.00167406f + p * w;
    p = 2.83297682f + p * w;
  }
  const float r = p * x;
  return (fabsf(x) == 1.0f) ? copysignf(INFINITY, x) : r;
}

// jax.random.normal(key, (n,))[i]
__host__ __device__ __forceinline__ float normal_at(Key key, uint32_t n, uint32_t i,
                                                    bool partitionable) {
  const float lo = -0.99999994f;  // nextafter(-1, 0)
  const float u = uniform_at(key, n, i, lo, 1.0f, partitionable);
  return 1.41421356f * erf_inv_f32(u);
}

}  // namespace b200mc
