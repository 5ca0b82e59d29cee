// Device-side jax.random primitives shared by prng.cu (one update's key chain) and predictive.cu
// (MixtureGuidePredictive draws): threefry2x32 in the "partitionable" stream layout, Giles' erfinv,
// jax.random.normal's bits -> float map.  Known-answer tests: tests/test_gpu_optim_prng.py.
#pragma once
#include <math.h>
#include <stdint.h>

namespace smi_prng {

__device__ __forceinline__ uint32_t rotl32(uint32_t x, int r) { return __funnelshift_l(x, x, r); }

__device__ __forceinline__ void threefry2x32(uint32_t k0, uint32_t k1, uint32_t x0, uint32_t x1, uint32_t& o0,
                                             uint32_t& o1) {
  const uint32_t ks0 = k0, ks1 = k1, ks2 = k0 ^ k1 ^ 0x1BD11BDAu;
  x0 += ks0;
  x1 += ks1;
#define TF_ROUND(r)   \
  x0 += x1;           \
  x1 = rotl32(x1, r); \
  x1 ^= x0;
  TF_ROUND(13) TF_ROUND(15) TF_ROUND(26) TF_ROUND(6)
  x0 += ks1;
  x1 += ks2 + 1u;
  TF_ROUND(17) TF_ROUND(29) TF_ROUND(16) TF_ROUND(24)
  x0 += ks2;
  x1 += ks0 + 2u;
  TF_ROUND(13) TF_ROUND(15) TF_ROUND(26) TF_ROUND(6)
  x0 += ks0;
  x1 += ks1 + 3u;
  TF_ROUND(17) TF_ROUND(29) TF_ROUND(16) TF_ROUND(24)
  x0 += ks1;
  x1 += ks2 + 4u;
  TF_ROUND(13) TF_ROUND(15) TF_ROUND(26) TF_ROUND(6)
  x0 += ks2;
  x1 += ks0 + 5u;
#undef TF_ROUND
  o0 = x0;
  o1 = x1;
}

struct Key {
  uint32_t a, b;
};

// Two stream layouts (jax_threefry_partitionable flipped its default in JAX 0.5.0 and the reference leaves
// JAX unpinned):
//   partitionable: word j of any flat shape = out0 ^ out1 of threefry(key, (0, j)); split(key, n)[j] = threefry(key, (0, j))
//   legacy       : counters iota(n) (zero-padded to even) cut in two halves (x0, x1); the outputs are
//                  concat(out0, out1)[:n]; split(key, n) = the 2n words reshaped (n, 2)
// `legacy` is warp-uniform everywhere it is used.
__device__ __forceinline__ uint32_t bits_at(Key k, uint32_t j, uint32_t n, int legacy) {
  uint32_t o0, o1;
  if (!legacy) {
    threefry2x32(k.a, k.b, 0u, j, o0, o1);
    return o0 ^ o1;
  }
  const uint32_t half = (n + 1u) >> 1;
  if (j < half) {
    threefry2x32(k.a, k.b, j, j + half < n ? j + half : 0u, o0, o1);
    return o0;
  }
  threefry2x32(k.a, k.b, j - half, j, o0, o1);
  return o1;
}
// jax.random.split(key, num)[j]
__device__ __forceinline__ Key split_at(Key k, uint32_t j, uint32_t num, int legacy) {
  Key r;
  if (!legacy) {
    threefry2x32(k.a, k.b, 0u, j, r.a, r.b);
  } else {
    r.a = bits_at(k, 2u * j, 2u * num, 1);
    r.b = bits_at(k, 2u * j + 1u, 2u * num, 1);
  }
  return r;
}

// Giles' single-precision erfinv (the polynomial XLA lowers erf_inv(f32) to).
__device__ __forceinline__ float erfinv_giles(float x) {
  float w = -log1pf(-(x * x));
  float p;
  if (w < 5.0f) {
    w -= 2.5f;
    p = 2.81022636e-08f;
    p = fmaf(p, w, 3.43273939e-07f);
    p = fmaf(p, w, -3.5233877e-06f);
    p = fmaf(p, w, -4.39150654e-06f);
    p = fmaf(p, w, 0.00021858087f);
    p = fmaf(p, w, -0.00125372503f);
    p = fmaf(p, w, -0.00417768164f);
    p = fmaf(p, w, 0.246640727f);
    p = fmaf(p, w, 1.50140941f);
  } else {
    w = sqrtf(w) - 3.0f;
    p = -0.000200214257f;
    p = fmaf(p, w, 0.000100950558f);
    p = fmaf(p, w, 0.00134934322f);
    p = fmaf(p, w, -0.00367342844f);
    p = fmaf(p, w, 0.00573950773f);
    p = fmaf(p, w, -0.0076224613f);
    p = fmaf(p, w, 0.00943887047f);
    p = fmaf(p, w, 1.00167406f);
    p = fmaf(p, w, 2.83297682f);
  }
  return p * x;
}

// jax.random.normal: uniform on [nextafter(-1, 0), 1) from 23 mantissa bits, sqrt(2) erfinv(u).
__device__ __forceinline__ float normal_from_bits(uint32_t bits) {
  const float f = __uint_as_float((bits >> 9) | 0x3F800000u) - 1.0f;
  const float lo = -0.99999994f;  // nextafter(-1, 0)
  const float u = fmaxf(lo, __fadd_rn(__fmul_rn(f, 1.0f - lo), lo));
  return 1.41421356237f * erfinv_giles(u);
}

}  // namespace smi_prng
