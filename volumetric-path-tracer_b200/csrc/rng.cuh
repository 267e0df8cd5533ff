// rng.cuh — PCG32 streams, one per pixel, carried across samples.
// Follows yocto_sampling.h:184-229: 64-bit LCG state, xorshift-rotate output,
// rand1f by the mantissa trick. rand2f draws x then y.
#pragma once

#include "vmath.cuh"

namespace vpt {

struct rng_t {
  uint64_t state, inc;
};

constexpr uint64_t pcg_mult = 6364136223846793005ULL;

VPT_FN uint32_t rng_next(rng_t& rng) {
  uint64_t old = rng.state;
  rng.state    = old * pcg_mult + rng.inc;
  auto xorshifted = (uint32_t)(((old >> 18u) ^ old) >> 27u);
  auto rot        = (uint32_t)(old >> 59u);
  return (xorshifted >> rot) | (xorshifted << ((~rot + 1u) & 31));
}

VPT_FN float rand1f(rng_t& rng) {
  uint32_t u = (rng_next(rng) >> 9) | 0x3f800000u;
  return i2f((int)u) - 1.0f;
}
VPT_FN f2 rand2f(rng_t& rng) {
  auto x = rand1f(rng);
  auto y = rand1f(rng);
  return {x, y};
}

// Jump the generator ahead by `delta` draws in O(log delta) (Brown, "Random
// number generation with arbitrary strides"): used to hand each GPU rank a
// disjoint stretch of every pixel's stream.
VPT_FN void rng_advance(rng_t& rng, uint64_t delta) {
  uint64_t cur_mult = pcg_mult, cur_plus = rng.inc;
  uint64_t acc_mult = 1u, acc_plus = 0u;
  while (delta > 0) {
    if (delta & 1) {
      acc_mult *= cur_mult;
      acc_plus = acc_plus * cur_mult + cur_plus;
    }
    cur_plus = (cur_mult + 1) * cur_plus;
    cur_mult *= cur_mult;
    delta >>= 1;
  }
  rng.state = acc_mult * rng.state + acc_plus;
}

}  // namespace vpt
