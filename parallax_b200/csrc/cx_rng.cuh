// cx_rng.cuh — device RNG: jax.random's threefry2x32 key chain (bit-exact mode) and
// Philox4x32-10 (throughput mode).
//
// The reference draws every random number through jax.random (call sites:
// cotix/_colliders.py:142,175,222-224,254,264,284,295; cotix/_lunar_lander.py:109-123;
// examples/test_viz.py:39).  jax is an un-vendored dependency; what is implemented here is
// its published algorithm for the non-partitionable threefry PRNG (the default of every jax
// release the reference can run on):
//   threefry2x32          Random123 Threefry-2x32, 20 rounds
//   random bits, n words  counters iota(n) zero-padded to even, first half = x0 lane,
//                         second half = x1 lane, outputs concatenated
//   split(key, m)         2m words of random bits, reshaped (m, 2)
//   uniform               bitcast((bits >> 9) | 0x3f800000) - 1, scaled, max(lo, .)
//   bernoulli(p)          uniform < p
// Only the words that are actually consumed are generated (lazy evaluation): a split key
// needs two threefry calls, whatever the fan-out.
#pragma once
#include "cx_math.cuh"

struct Key2 {
  uint32_t a, b;
};

__host__ __device__ __forceinline__ uint32_t rotl32(uint32_t x, int r) {
#ifdef __CUDA_ARCH__
  return __funnelshift_l(x, x, r);
#else
  return (x << r) | (x >> (32 - r));
#endif
}

__host__ __device__ __forceinline__ void threefry2x32(uint32_t k0, uint32_t k1, uint32_t x0, uint32_t x1, uint32_t& o0,
                         uint32_t& o1) {
  const uint32_t ks2 = k0 ^ k1 ^ 0x1BD11BDAu;
  x0 += k0;
  x1 += k1;
#define CX_TF_ROUND(r) \
  x0 += x1;            \
  x1 = rotl32(x1, r);  \
  x1 ^= x0;
  CX_TF_ROUND(13) CX_TF_ROUND(15) CX_TF_ROUND(26) CX_TF_ROUND(6)
  x0 += k1;
  x1 += ks2 + 1u;
  CX_TF_ROUND(17) CX_TF_ROUND(29) CX_TF_ROUND(16) CX_TF_ROUND(24)
  x0 += ks2;
  x1 += k0 + 2u;
  CX_TF_ROUND(13) CX_TF_ROUND(15) CX_TF_ROUND(26) CX_TF_ROUND(6)
  x0 += k0;
  x1 += k1 + 3u;
  CX_TF_ROUND(17) CX_TF_ROUND(29) CX_TF_ROUND(16) CX_TF_ROUND(24)
  x0 += k1;
  x1 += ks2 + 4u;
  CX_TF_ROUND(13) CX_TF_ROUND(15) CX_TF_ROUND(26) CX_TF_ROUND(6)
  x0 += ks2;
  x1 += k0 + 5u;
#undef CX_TF_ROUND
  o0 = x0;
  o1 = x1;
}

// One out-of-line copy for the key-chain code: the 20 rounds are ~110 SASS instructions and
// the step kernel has ~50 call sites; inlining them all made the kernel 197 KB of code and
// I-cache-miss bound (profiles/r1a).  Returns (o0, o1) packed as a uint2.
static __device__ __noinline__ uint2 tf_call(uint32_t k0, uint32_t k1, uint32_t x0, uint32_t x1) {
  uint32_t o0, o1;
  threefry2x32(k0, k1, x0, x1, o0, o1);
  return make_uint2(o0, o1);
}
// word w of random_bits(key, n) (n words, n even or odd)
CX_DEV uint32_t tf_bits_word(Key2 key, uint32_t n, uint32_t w) {
  const uint32_t half = (n + 1u) >> 1;
  const bool lo = w < half;
  uint32_t c0 = lo ? w : w - half;
  uint32_t c1 = lo ? half + w : w;
  if (c1 >= n) c1 = 0u;  // the zero padding of an odd count
  uint2 o = tf_call(key.a, key.b, c0, c1);
  return lo ? o.x : o.y;
}
// two independent blocks under one key in one out-of-line call: the scheduler interleaves the two
// dependency chains, so a split costs the latency of one block instead of two
static __device__ __noinline__ uint4 tf_call2(uint32_t k0, uint32_t k1, uint32_t a0, uint32_t a1, uint32_t b0,
                                              uint32_t b1) {
  uint32_t oa0, oa1, ob0, ob1;
  threefry2x32(k0, k1, a0, a1, oa0, oa1);
  threefry2x32(k0, k1, b0, b1, ob0, ob1);
  return make_uint4(oa0, oa1, ob0, ob1);
}
// key `idx` of jax.random.split(key, num): words 2 idx and 2 idx + 1 of random_bits(key, 2 num)
CX_DEV Key2 tf_split(Key2 key, uint32_t num, uint32_t idx) {
  const uint32_t half = num, wa = 2u * idx, wb = wa + 1u;
  const bool la = wa < half, lb = wb < half;
  const uint4 o = tf_call2(key.a, key.b, la ? wa : wa - half, la ? half + wa : wa, lb ? wb : wb - half,
                           lb ? half + wb : wb);
  Key2 r;
  r.a = la ? o.x : o.y;
  r.b = lb ? o.z : o.w;
  return r;
}
CX_DEV float bits_to_uniform(uint32_t bits, float lo, float hi) {
  float f = __uint_as_float((bits >> 9) | 0x3F800000u) - 1.0f;
  float v = f * (hi - lo) + lo;
  return jmax(lo, v);
}
// jax.random.uniform(key, ()) / bernoulli(key, p, ())
CX_DEV float tf_uniform(Key2 key, float lo, float hi) {
  return bits_to_uniform(tf_bits_word(key, 1u, 0u), lo, hi);
}
CX_DEV bool tf_bernoulli(Key2 key, float p) { return tf_uniform(key, 0.0f, 1.0f) < p; }

// Philox4x32-10 (Random123)
CX_DEV void philox4x32_10(uint32_t k0, uint32_t k1, uint32_t c0, uint32_t c1, uint32_t c2,
                          uint32_t c3, uint32_t out[4]) {
#pragma unroll 1
  for (int r = 0; r < 10; ++r) {
    uint32_t hi0 = __umulhi(0xD2511F53u, c0), lo0 = 0xD2511F53u * c0;
    uint32_t hi1 = __umulhi(0xCD9E8D57u, c2), lo1 = 0xCD9E8D57u * c2;
    uint32_t n0 = hi1 ^ c1 ^ k0;
    uint32_t n2 = hi0 ^ c3 ^ k1;
    c0 = n0;
    c1 = lo1;
    c2 = n2;
    c3 = lo0;
    k0 += 0x9E3779B9u;
    k1 += 0xBB67AE85u;
  }
  out[0] = c0;
  out[1] = c1;
  out[2] = c2;
  out[3] = c3;
}

#define CX_TAG_COIN 0x434F494Eu
#define CX_TAG_CHOICE 0x43484F49u
#define CX_TAG_RESET 0x52535400u

// word `w` of the engine's Philox stream (steps, episode, (g << 16) | (w >> 2), tag)
CX_DEV uint32_t philox_word(Key2 key, uint32_t steps, uint32_t episode, uint32_t g, uint32_t w,
                            uint32_t tag) {
  uint32_t o[4];
  philox4x32_10(key.a, key.b, steps, episode, (g << 16) | (w >> 2), tag, o);
  const uint32_t s = w & 3u;
  return s == 0 ? o[0] : (s == 1 ? o[1] : (s == 2 ? o[2] : o[3]));
}
