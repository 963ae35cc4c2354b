// rng.cuh — counter-keyed random numbers for the device path tracer.
//
// The reference draws from a per-thread rand::SmallRng seeded from OS entropy (src/core/rng.rs:8-16), so only the
// distribution of uniform_1d() is contractual: 24-bit uniform in [0,1).  Here every camera sample owns a stream
// keyed by (seed, pixel, sample): the image does not depend on scheduling, pool size, tile ownership or GPU count.
// The stream is a 64-bit LCG with PCG-XSH-RR output (O'Neill 2014); 8 bytes of state per path.
#pragma once
#include "vecmath.cuh"

SPT_HD uint64_t rng_mix64(uint64_t z) { // SplitMix64 finaliser
  z += 0x9E3779B97F4A7C15ull;
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
  return z ^ (z >> 31);
}
SPT_HD uint64_t rng_init(uint64_t seed, uint64_t pixel, uint64_t sample) {
  return rng_mix64(rng_mix64(seed ^ (pixel * 0xD1B54A32D192ED03ull)) + sample);
}
SPT_HD uint32_t rng_next_u32(uint64_t& s) {
  uint64_t old = s;
  s = old * 6364136223846793005ull + 1442695040888963407ull;
  uint32_t xs = (uint32_t)(((old >> 18u) ^ old) >> 27u);
  uint32_t rot = (uint32_t)(old >> 59u);
  return (xs >> rot) | (xs << ((32u - rot) & 31u));
}
// Rng::uniform_1d (rng.rs:14-16): rand's Standard f32 = top 24 bits * 2^-24
SPT_HD float rng_uniform(uint64_t& s) { return (float)(rng_next_u32(s) >> 8) * (1.0f / 16777216.0f); }

// A tiny draw source so the same BxDF code can run from a path stream or from caller-supplied numbers (test hook).
struct DrawSrc {
  uint64_t state;
  const float* list; // non-null: take numbers from here instead
  int idx, n;
  SPT_HD float next() {
    if (list) { float v = idx < n ? list[idx] : 0.5f; idx++; return v; }
    return rng_uniform(state);
  }
};

// The draw source of the render kernels: just the 8-byte stream state.
struct PathRng {
  uint64_t state;
  SPT_HD float next() { return rng_uniform(state); }
  SPT_HD void skip2() { state = state * 0x685f98a2018fade9ull + 0x1a08ee1184ba6d32ull; } // two LCG steps at once (A^2, (A+1)C)
};

// Sub-pixel offsets of the additive-recurrence (R2) sampler, pixel_sampler/recurrence.rs:41-56, in closed form:
// x_n = frac(0.5 + n*a1), y_n = frac(0.5 + n*a2) with 32-bit fixed-point a1 = 1/phi2, a2 = 1/phi2^2, and n the
// global ordinal of the sample (the reference never resets the recurrence per pixel, Q17).
SPT_HD void r2_offset(uint32_t n, float& ox, float& oy) {
  const uint32_t A1 = 3242174889u, A2 = 2447445413u;
  uint32_t X = 0x80000000u + n * A1, Y = 0x80000000u + n * A2;
  ox = (float)(X >> 8) * (1.0f / 16777216.0f);
  oy = (float)(Y >> 8) * (1.0f / 16777216.0f);
}
